// Hand-written sm_100a kernels of the physics tick.  One tick =
//   k_integrate_prep   Euler step (euler_physics_relay.h:21-97) fused with per-collider
//                      composite-transform / bound / grid-bucket derivation
//   k_scan_*           exclusive scan of the bucket histogram (counting sort)
//   k_scatter          packets -> bucket-sorted bound array
//   k_pairs            uniform-grid candidate lists (the broadphase the reference lacks)
//   k_resolve_round    narrowphase + response + interception, replayed in the reference's
//                      sequential pair order by a dependency-round schedule
// All kernels are HBM / latency bound FP32 CUDA-core work; there is no dense contraction,
// hence no tensor cores.
#include "icx_world.cuh"
#include "icx_math.cuh"
#define ICX_UNIT_RADIUS 0.5f

namespace icx {

__device__ __forceinline__ uint32_t hash_cell(int cx, int cy, int cz, uint32_t wl, uint32_t mask) {
    uint32_t h = (uint32_t)cx * 73856093u ^ (uint32_t)cy * 19349663u ^ (uint32_t)cz * 83492791u ^ wl * 2654435761u;
    h ^= h >> 15; h *= 2246822519u; h ^= h >> 13;
    return h & mask;
}

__device__ __forceinline__ bool can_intersect(int a, int b, uint32_t cfg_flags) {
    // non-null entries of intersectProcedures, collision_procedures.h:411-417
    if (a == ICX_COLLIDER_BOX) return b != ICX_COLLIDER_NONE;
    if (b == ICX_COLLIDER_BOX) return a != ICX_COLLIDER_NONE;
    return (cfg_flags & ICX_F_SPHERE_SPHERE) && a == ICX_COLLIDER_SPHERE && b == ICX_COLLIDER_SPHERE;
}

// ---------------------------------------------------------------------------------------
// K1: Euler integration + per-packet derivation.  One thread per entity.
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_integrate_prep(DeviceWorld W, double deltaTime, int do_euler)
{
    const uint32_t e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= W.n_entities) return;
    const float4 p1 = W.prm1[e];
    const uint32_t ef = __float_as_uint(p1.w);
    float4 pos = W.in.pos[e], rot = W.in.rot[e], vel = W.in.vel[e], acc = W.in.acc[e];
    float4 av = W.in.angvel[e], aa = W.in.angacc[e];

    if (do_euler && (ef & EF_HAS_PHYSICS) && !(ef & EF_ASLEEP)) {
        // EulerPhysicsRelay::engage, euler_physics_relay.h:25-95
        V3 org = v3(pos), v = v3(vel), a = v3(acc), w = v3(av), al = v3(aa);
        Q4 q; q.w = rot.x; q.x = rot.y; q.y = rot.z; q.z = rot.w;
        a = a + v3(W.gx, W.gy, W.gz);                                               // :32
        if (ef & EF_LOCK_ORIGIN) { v = v3(0.0f, 0.0f, 0.0f); a = v3(0.0f, 0.0f, 0.0f); org = v3(W.locked_pos[e]); }   // :36-46
        if (ef & EF_LOCK_ORIENTATION) {                                             // :47-57
            w = v3(0.0f, 0.0f, 0.0f); al = v3(0.0f, 0.0f, 0.0f);
            float4 lq = W.locked_rot[e]; q.w = lq.x; q.x = lq.y; q.y = lq.z; q.z = lq.w;
        }
        const float halfDt = (float)(deltaTime / 2.0f);                             // :62 double divide, then narrowed
        const float dt = (float)deltaTime;
#pragma unroll
        for (int half = 0; half < 2; half++) {
            org = org + v * halfDt;                                                 // :62 / :80
            float angularSpeed = length(w);
            if (angularSpeed != 0.0f) q = qmul(angle_axis(angularSpeed * halfDt, w / angularSpeed), q);   // :65-70
            if (half == 0) { v = v + a * dt; w = w + al * dt; }                     // :74-75
        }
        a = v3(0.0f, 0.0f, 0.0f); al = v3(0.0f, 0.0f, 0.0f);                         // :90-91
        v = v * (float)(1.0f - p1.x * deltaTime);                                   // :94 evaluated in double, narrowed by glm's *=
        w = w * (float)(1.0f - p1.y * deltaTime);                                   // :95
        pos = make_float4(org.x, org.y, org.z, 0.0f);
        rot = make_float4(q.w, q.x, q.y, q.z);
        vel = make_float4(v.x, v.y, v.z, 0.0f); acc = make_float4(a.x, a.y, a.z, 0.0f);
        av = make_float4(w.x, w.y, w.z, 0.0f); aa = make_float4(al.x, al.y, al.z, 0.0f);
    }
    W.out.pos[e] = pos; W.out.rot[e] = rot; W.out.vel[e] = vel; W.out.acc[e] = acc;
    W.out.angvel[e] = av; W.out.angacc[e] = aa;
    W.out.ts[e] = W.in.ts[e];
    W.pos0[e] = pos;
    W.head[0][e] = 0u;
    W.ent_active[e] = 0u;

    // ---- per-collider derivation (ColliderSynchro::update + what every later pair test needs) ----
    const uint32_t fp = W.first_packet[e];
    const uint32_t npk = fp >> 30, p0 = fp & 0x3FFFFFFFu;
    if (npk == 0) return;
    const float4 p0c = W.prm0[e];
    const V3 escale = v3(p0c);
    const uint32_t wl = W.worldline_stride ? e / W.worldline_stride : 0u;
    for (uint32_t k = 0; k < npk; k++) {
        const uint32_t p = p0 + k;
        const float4 lpos = W.pk_lpos[p], lrot = W.pk_lrot[p], lscale = W.pk_lscale[p];
        const uint32_t bits = __float_as_uint(lpos.w);
        const int type = bits & PF_TYPE_MASK;
        Q4 pq; pq.w = rot.x; pq.x = rot.y; pq.y = rot.z; pq.z = rot.w;
        if (!(bits & PF_INHERIT_ORIENT)) { pq.w = 1.0f; pq.x = 0.0f; pq.y = 0.0f; pq.z = 0.0f; }   // collider.h:150-153
        Q4 lq; lq.w = lrot.x; lq.x = lrot.y; lq.y = lrot.z; lq.z = lrot.w;
        const F4 P = model_transform(v3(pos), pq, escale);
        const F4 Lm = model_transform(v3(lpos), lq, v3(lscale));
        const F4 f = mul(P, Lm);                                                    // collider.h:155
        // translation column = base + (origin + 0): only `origin` changes during the collision pass
        const V4 b4 = P.c[0] * Lm.c[3].x + P.c[1] * Lm.c[3].y + P.c[2] * Lm.c[3].z;
        M3 L; L.c[0] = v3(f.c[0].x, f.c[0].y, f.c[0].z); L.c[1] = v3(f.c[1].x, f.c[1].y, f.c[1].z); L.c[2] = v3(f.c[2].x, f.c[2].y, f.c[2].z);
        const M3 inv = inverse(L);
        // Collider::get_inertia_tensor, collider.h:97-142
        const V3 s = escale * v3(lscale);
        V3 I;
        if (type == ICX_COLLIDER_BOX) I = v3((s.y * s.y + s.z * s.z) / 4.0f, (s.x * s.x + s.z * s.z) / 4.0f, (s.x * s.x + s.y * s.y) / 4.0f);
        else if (type == ICX_COLLIDER_RAY) I = v3((s.z * s.z) / 3.0f, (s.z * s.z) / 3.0f, 0.0001f);
        else { float r = smin(smin(s.x, s.y), s.z); float d = r * r * (2.0f / 5.0f); I = v3(d, d, d); }
        float4* d = W.pk_d + (size_t)p * 6;
        d[0] = make_float4(L.c[0].x, L.c[0].y, L.c[0].z, b4.x);
        d[1] = make_float4(L.c[1].x, L.c[1].y, L.c[1].z, b4.y);
        d[2] = make_float4(L.c[2].x, L.c[2].y, L.c[2].z, b4.z);
        d[3] = make_float4(inv.c[0].x, inv.c[0].y, inv.c[0].z, I.x);
        d[4] = make_float4(inv.c[1].x, inv.c[1].y, inv.c[1].z, I.y);
        d[5] = make_float4(inv.c[2].x, inv.c[2].y, inv.c[2].z, I.z);
        W.pk_minparam[p] = 1.0f;                                                    // Collider::reset, collider.h:75-78

        // conservative bound (DESIGN.md "Broadphase bound"); must match oracle packet_aabb bit for bit
        M4 M; M.c[0] = L.c[0]; M.c[1] = L.c[1]; M.c[2] = L.c[2]; M.t = v3(f.c[3].x, f.c[3].y, f.c[3].z);
        V3 ctr, h;
        if (type == ICX_COLLIDER_SPHERE) {
            ctr = M.t;
            float r = length(xform(M, ICX_UNIT_RADIUS, 0.0f, 0.0f) - xform(M, 0.0f, 0.0f, 0.0f)) * 1.7320509f;
            h = v3(r, r, r);
        } else if (type == ICX_COLLIDER_RAY) {
            V3 o = xform(M, 0.0f, 0.0f, 0.0f), en = xform(M, 0.0f, 0.0f, 1.0f);
            ctr = (o + en) * 0.5f;
            float r = length(en - o) * 0.5f * 1.000001f;
            h = v3(r, r, r);
        } else {
            ctr = M.t;
            h = v3((fabsf(L.c[0].x) + fabsf(L.c[1].x) + fabsf(L.c[2].x)) * 0.5f,
                   (fabsf(L.c[0].y) + fabsf(L.c[1].y) + fabsf(L.c[2].y)) * 0.5f,
                   (fabsf(L.c[0].z) + fabsf(L.c[1].z) + fabsf(L.c[2].z)) * 0.5f);
        }
        const float m = W.margin;
        const float4 lo = make_float4(ctr.x - h.x - m, ctr.y - h.y - m, ctr.z - h.z - m, 0.0f);
        const float4 hi = make_float4(ctr.x + h.x + m, ctr.y + h.y + m, ctr.z + h.z + m, 0.0f);
        W.pk_lo[p] = lo; W.pk_hi[p] = hi;

        uint32_t cell = 0xFFFFFFFFu;
        if (type != ICX_COLLIDER_NONE) {
            const float ext = fmaxf(fmaxf(hi.x - lo.x, hi.y - lo.y), hi.z - lo.z);
            if (ext > W.cell_size || !(ext == ext)) {
                if (ef & EF_IMMUTABLE) {
                    uint32_t idx = atomicAdd(&W.counters[CTR_LARGE], 1u);
                    if (idx < W.max_large) W.large_list[idx] = p; else atomicOr(W.flags, RF_LARGE_OVERFLOW);
                    cell = 0xFFFFFFFEu;     // large
                } else {
                    atomicOr(W.flags, RF_CELL_TOO_SMALL);
                    atomicMax((int*)&W.counters[CTR_ROUND_BASE - 1], __float_as_int(ext));   // needed cell size (positive floats order as ints)
                }
            } else {
                const float cxf = (lo.x + hi.x) * 0.5f, cyf = (lo.y + hi.y) * 0.5f, czf = (lo.z + hi.z) * 0.5f;
                cell = hash_cell((int)floorf(cxf * W.inv_cell), (int)floorf(cyf * W.inv_cell), (int)floorf(czf * W.inv_cell), wl, W.n_buckets - 1);
                atomicAdd(&W.cell_count[cell], 1u);
            }
        }
        W.pk_cell[p] = cell;
    }
}


// ---------------------------------------------------------------------------------------
// K2: exclusive scan of the bucket histogram (counting sort of packets by grid bucket).
// 1024 threads x 4 elements per block; partials scanned by one block; offsets added in place.
// ---------------------------------------------------------------------------------------
#define SCAN_TPB 1024
#define SCAN_EPB 4096

__device__ __forceinline__ uint32_t block_exclusive_scan(uint32_t v, uint32_t* total)
{
    __shared__ uint32_t warp_sums[32];
    const uint32_t lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    uint32_t x = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { uint32_t y = __shfl_up_sync(0xFFFFFFFFu, x, o); if (lane >= (uint32_t)o) x += y; }
    if (lane == 31) warp_sums[wid] = x;
    __syncthreads();
    if (wid == 0) {
        uint32_t s = warp_sums[lane];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { uint32_t y = __shfl_up_sync(0xFFFFFFFFu, s, o); if (lane >= (uint32_t)o) s += y; }
        warp_sums[lane] = s;
    }
    __syncthreads();
    const uint32_t warp_off = wid ? warp_sums[wid - 1] : 0u;
    if (total) *total = warp_sums[31];
    __syncthreads();
    return warp_off + x - v;
}

__global__ void __launch_bounds__(SCAN_TPB) k_scan_blocks(uint32_t* __restrict__ data, uint32_t* __restrict__ partial, uint32_t n)
{
    const uint32_t base = blockIdx.x * SCAN_EPB + threadIdx.x * 4;
    uint4 v = make_uint4(0, 0, 0, 0);
    if (base + 3 < n) v = *reinterpret_cast<const uint4*>(data + base);
    else { if (base < n) v.x = data[base]; if (base + 1 < n) v.y = data[base + 1]; if (base + 2 < n) v.z = data[base + 2]; }
    const uint32_t sum = v.x + v.y + v.z + v.w;
    uint32_t total;
    const uint32_t off = block_exclusive_scan(sum, &total);
    uint4 o; o.x = off; o.y = off + v.x; o.z = o.y + v.y; o.w = o.z + v.z;
    if (base + 3 < n) *reinterpret_cast<uint4*>(data + base) = o;
    else { if (base < n) data[base] = o.x; if (base + 1 < n) data[base + 1] = o.y; if (base + 2 < n) data[base + 2] = o.z; }
    if (threadIdx.x == 0) partial[blockIdx.x] = total;
}

// one block: exclusive scan of up to 4096 partials; writes the grand total to partial[n_partial]
__global__ void __launch_bounds__(SCAN_TPB) k_scan_partials(uint32_t* __restrict__ partial, uint32_t n_partial)
{
    const uint32_t base = threadIdx.x * 4;
    uint32_t v[4];
#pragma unroll
    for (int i = 0; i < 4; i++) v[i] = (base + i < n_partial) ? partial[base + i] : 0u;
    uint32_t total;
    const uint32_t off = block_exclusive_scan(v[0] + v[1] + v[2] + v[3], &total);
    uint32_t run = off;
#pragma unroll
    for (int i = 0; i < 4; i++) { if (base + i < n_partial) partial[base + i] = run; run += v[i]; }
    if (threadIdx.x == 0) partial[n_partial] = total;
}

__global__ void __launch_bounds__(256) k_scan_add(uint32_t* __restrict__ data, const uint32_t* __restrict__ partial, uint32_t n, uint32_t n_partial)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) data[i] += partial[i / SCAN_EPB];
    else if (i == n) data[n] = partial[n_partial];      // cell_start[n_buckets] = number of grid packets
}

// ---------------------------------------------------------------------------------------
// K3: scatter packets into bucket order.  Order inside a bucket is arbitrary (atomics); the
// candidate lists are sorted afterwards, so results do not depend on it.
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_scatter(DeviceWorld W)
{
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= W.n_packets) return;
    const uint32_t cell = W.pk_cell[p];
    if (cell >= 0xFFFFFFFEu) { if (cell == 0xFFFFFFFFu) W.pos_of[p] = 0xFFFFFFFFu; return; }   // large columns are assigned by k_pairs_large
    const uint32_t pos = W.cell_count[cell] + atomicAdd(&W.cell_fill[cell], 1u);
    const uint32_t e = W.pk_entity[p];
    const uint32_t bits = __float_as_uint(W.pk_lpos[p].w);
    const uint32_t ef = __float_as_uint(W.prm1[e].w);
    const float4 lo = W.pk_lo[p], hi = W.pk_hi[p];
    const uint32_t tag = p | ((bits & PF_TYPE_MASK) << 28) | ((ef & EF_IMMUTABLE) ? 0x40000000u : 0u);
    W.s_lo[pos] = make_float4(lo.x, lo.y, lo.z, __uint_as_float(tag));
    W.s_hi[pos] = make_float4(hi.x, hi.y, hi.z, __uint_as_float(e));
    W.pos_of[p] = pos;
}

// ---------------------------------------------------------------------------------------
// K4: candidate lists.  One thread per bucket-sorted packet: visit the 27 neighbouring buckets
// (deduplicated after hashing) plus the large-collider list, keep partners whose inflated
// bounds overlap, sorted ascending by packet index (= the order the reference's pair loop
// meets them, collision_physics_relay.h:29-37).
// ---------------------------------------------------------------------------------------
#define ICX_KMAX 192

__device__ __forceinline__ bool overlap(float4 alo, float4 ahi, float4 blo, float4 bhi) {
    return alo.x <= bhi.x && blo.x <= ahi.x && alo.y <= bhi.y && blo.y <= ahi.y && alo.z <= bhi.z && blo.z <= ahi.z;
}

__device__ __forceinline__ void list_insert(uint32_t* list, uint32_t& cnt, uint32_t cap, uint32_t entry, bool& overflow)
{
    if (cnt >= cap) { overflow = true; return; }
    uint32_t i = cnt;
    const uint32_t key = entry & ICX_NB_INDEX_MASK;
    while (i > 0 && (list[i - 1] & ICX_NB_INDEX_MASK) > key) { list[i] = list[i - 1]; i--; }
    list[i] = entry; cnt++;
}

__global__ void __launch_bounds__(128) k_pairs(DeviceWorld W)
{
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t n_grid = W.cell_count[W.n_buckets];
    const uint32_t n_large = min(W.counters[CTR_LARGE], W.max_large);
    if (t >= n_grid + n_large) return;
    uint32_t list[ICX_KMAX];
    uint32_t cnt = 0; bool overflow = false;
    const uint32_t cap = W.max_neighbors;
    uint32_t p, e; int type; bool immut; float4 lo, hi;
    if (t < n_grid) {
        lo = W.s_lo[t]; hi = W.s_hi[t];
        const uint32_t tag = __float_as_uint(lo.w);
        p = tag & ICX_NB_INDEX_MASK; type = (tag >> 28) & 3; immut = (tag & 0x40000000u) != 0;
        e = __float_as_uint(hi.w);
    } else {
        p = W.large_list[t - n_grid];
        lo = W.pk_lo[p]; hi = W.pk_hi[p];
        e = W.pk_entity[p]; type = __float_as_uint(W.pk_lpos[p].w) & PF_TYPE_MASK; immut = true;
        W.pos_of[p] = t;
    }
    const uint32_t wl = W.worldline_stride ? e / W.worldline_stride : 0u;

    if (t < n_grid) {
        const float cxf = (lo.x + hi.x) * 0.5f, cyf = (lo.y + hi.y) * 0.5f, czf = (lo.z + hi.z) * 0.5f;
        const int cx = (int)floorf(cxf * W.inv_cell), cy = (int)floorf(cyf * W.inv_cell), cz = (int)floorf(czf * W.inv_cell);
        uint32_t seen[27]; int n_seen = 0;
        for (int dz = -1; dz <= 1; dz++) for (int dy = -1; dy <= 1; dy++) for (int dx = -1; dx <= 1; dx++) {
            const uint32_t hsh = hash_cell(cx + dx, cy + dy, cz + dz, wl, W.n_buckets - 1);
            bool dup = false;
            for (int s = 0; s < n_seen; s++) dup |= (seen[s] == hsh);
            if (dup) continue;
            seen[n_seen++] = hsh;
            const uint32_t beg = W.cell_count[hsh], end = W.cell_count[hsh + 1];
            for (uint32_t j = beg; j < end; j++) {
                if (j == t) continue;
                const float4 blo = W.s_lo[j], bhi = W.s_hi[j];
                const uint32_t btag = __float_as_uint(blo.w);
                const uint32_t be = __float_as_uint(bhi.w);
                if (be == e) continue;                                               // same entity, relay :51
                const bool bimmut = (btag & 0x40000000u) != 0;
                if (immut && !bimmut) continue;         // an immutable body never drives a pair with a mutable one
                if (W.worldline_stride && be / W.worldline_stride != wl) continue;
                if (!can_intersect(type, (btag >> 28) & 3, W.cfg_flags)) continue;
                if (!overlap(lo, hi, blo, bhi)) continue;
                list_insert(list, cnt, cap, (btag & ICX_NB_INDEX_MASK) | (bimmut ? ICX_NB_IMMUTABLE : 0u), overflow);
            }
        }
    }
    // large colliders (all immutable): everyone tests them; large ones test only each other
    for (uint32_t k = 0; k < n_large; k++) {
        const uint32_t q = W.large_list[k];
        if (q == p) continue;
        const uint32_t qe = W.pk_entity[q];
        if (qe == e) continue;
        if (W.worldline_stride && qe / W.worldline_stride != wl) continue;
        const int qtype = __float_as_uint(W.pk_lpos[q].w) & PF_TYPE_MASK;
        if (!can_intersect(type, qtype, W.cfg_flags)) continue;
        if (!overlap(lo, hi, W.pk_lo[q], W.pk_hi[q])) continue;
        list_insert(list, cnt, cap, q | ICX_NB_IMMUTABLE | ICX_NB_LARGE, overflow);
    }
    if (overflow) atomicOr(W.flags, RF_NEIGHBOR_OVERFLOW);
    W.nbr_count[t] = cnt;
    uint32_t upper = 0;
    for (uint32_t i = 0; i < cnt; i++) {
        W.nbr[(size_t)i * W.n_packets + t] = list[i];
        // count each pair once: from its lower packet, or from the driving side of an immutable pair
        const uint32_t o = list[i] & ICX_NB_INDEX_MASK;
        const bool oimm = (list[i] & ICX_NB_IMMUTABLE) != 0;
        if (!immut && oimm) upper++;                    // only I list it
        else if (immut && oimm) { const bool olarge = (list[i] & ICX_NB_LARGE) != 0, ilarge = t >= n_grid;
                                  if (ilarge == olarge ? p < o : !ilarge) upper++; }
        else if (o > p) upper++;
    }
    if (upper) atomicAdd(&W.counters[CTR_CANDIDATES], upper);
    if (cnt > 0 && atomicExch(&W.ent_active[e], 1u) == 0u) {
        const uint32_t idx = atomicAdd(&W.counters[CTR_ACTIVE0], 1u);
        W.active[0][idx] = e;
    }
}

// ---- launchers ------------------------------------------------------------------------
void launch_integrate_prep(const DeviceWorld& W, double dt, int do_euler, cudaStream_t st)
{
    if (W.n_entities == 0) return;
    k_integrate_prep<<<(W.n_entities + 255) / 256, 256, 0, st>>>(W, dt, do_euler);
}

void launch_grid(const DeviceWorld& W, cudaStream_t st)
{
    if (W.n_packets == 0) return;
    const uint32_t nb = W.n_buckets;
    const uint32_t n_part = (nb + SCAN_EPB - 1) / SCAN_EPB;
    k_scan_blocks<<<n_part, SCAN_TPB, 0, st>>>(W.cell_count, W.scan_partial, nb);
    k_scan_partials<<<1, SCAN_TPB, 0, st>>>(W.scan_partial, n_part);
    k_scan_add<<<(nb + 1 + 255) / 256, 256, 0, st>>>(W.cell_count, W.scan_partial, nb, n_part);
    k_scatter<<<(W.n_packets + 255) / 256, 256, 0, st>>>(W);
}

void launch_pairs(const DeviceWorld& W, cudaStream_t st)
{
    if (W.n_packets == 0) return;
    k_pairs<<<(W.n_packets + 127) / 128, 128, 0, st>>>(W);
}

} // namespace icx
