// Hand-written sm_100a kernels of the physics tick, part 1 (part 2: icx_resolve.cu).  One tick =
//   k_integrate_prep   Euler step (euler_physics_relay.h:21-97) fused with per-collider
//                      composite-transform / bound / grid-bucket derivation
//   k_scan_*           exclusive scan of the bucket histogram (counting sort)
//   k_scatter          packets -> bucket-sorted bound array
//   k_pairs            uniform-grid pair generator (the broadphase the reference lacks): flat,
//                      block-staged, compacted list of candidate pairs
//   k_narrow           speculative narrowphase of every candidate on the post-integration state;
//                      marks the bodies that are hit ("dirty")
//   k_resolve          narrowphase + response + interception for the dirty bodies only, replayed
//                      in the reference's sequential pair order (DESIGN.md "Ordered resolve")
// All kernels are HBM / latency bound FP32 CUDA-core work; there is no dense contraction,
// hence no tensor cores.
#include "icx_world.cuh"
#include "icx_math.cuh"
#include "icx_device.cuh"
#define ICX_UNIT_RADIUS 0.5f

namespace icx {

// ---------------------------------------------------------------------------------------
// K1: Euler integration + per-packet derivation.  One thread per entity.
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_integrate_prep(DeviceWorld W, double deltaTime, int do_euler)
{
    const uint32_t e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= W.n_entities) return;
    const float4 p1 = W.prm1[e];
    const uint32_t ef = __float_as_uint(p1.w);
    const EntityRegs r = euler_entity(W, e, ef, p1, deltaTime, do_euler);
    const float4 pos = r.pos, rot = r.rot;
    W.out.pos[e] = r.pos; W.out.rot[e] = r.rot; W.out.vel[e] = r.vel; W.out.acc[e] = r.acc;
    W.out.angvel[e] = r.av; W.out.angacc[e] = r.aa;
    W.out.ts[e] = W.in.ts[e];
    W.pos0[e] = r.pos;
    W.dirty_slot[e] = ICX_NONE;
    W.push[e] = 0.0f;

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
                    atomicMax((int*)&W.counters[CTR_NEED_CELL], __float_as_int(ext));   // needed cell size (positive floats order as ints)
                }
            } else {
                cell = bucket_of(cell_coord(lo.x, hi.x, W.inv_cell), cell_coord(lo.y, hi.y, W.inv_cell),
                                 cell_coord(lo.z, hi.z, W.inv_cell), wl, W.n_buckets);
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
    if (cell >= 0xFFFFFFFEu) return;                    // inert, or in the large-collider list
    const uint32_t pos = W.cell_count[cell] + atomicAdd(&W.cell_fill[cell], 1u);
    const uint32_t e = W.pk_entity[p];
    const uint32_t bits = __float_as_uint(W.pk_lpos[p].w);
    const uint32_t ef = __float_as_uint(W.prm1[e].w);
    const float4 lo = W.pk_lo[p], hi = W.pk_hi[p];
    const uint32_t tag = p | ((bits & PF_TYPE_MASK) << 28) | ((ef & EF_IMMUTABLE) ? 0x40000000u : 0u);
    W.s_lo[pos] = make_float4(lo.x, lo.y, lo.z, __uint_as_float(tag));
    W.s_hi[pos] = make_float4(hi.x, hi.y, hi.z, __uint_as_float(e));
}

// ---------------------------------------------------------------------------------------
// K4: uniform-grid pair generator.  One thread per bucket-sorted packet (then one per large
// collider): walk the 27 surrounding cells as 9-18 contiguous ranges, keep partners whose inflated
// bounds overlap, emit each unordered pair once (from its lower sorted position).  Pairs are
// appended warp-aggregated to a shared-memory stage and flushed with one global atomic per block,
// so the list is written in coalesced runs.  Entry = (lower packet, higher packet | flags).
// ---------------------------------------------------------------------------------------
#define PAIRS_TPB 256
#define STAGE_TIGHT 1024
#define STAGE_LOOSE 2048

struct PairStage { uint2 tight[STAGE_TIGHT]; uint2 loose[STAGE_LOOSE]; uint32_t n_tight, n_loose, base_all, base_tight; };

__device__ __forceinline__ void put_global(const DeviceWorld& W, uint2 entry, bool tight)
{
    const uint32_t g = atomicAdd(&W.counters[CTR_CANDIDATES], 1u);
    if (g < W.max_pairs) W.pairs[g] = entry; else atomicOr(W.flags, RF_PAIR_OVERFLOW);
    if (tight) {
        const uint32_t q = atomicAdd(&W.counters[CTR_TIGHT], 1u);
        if (q < W.max_pairs) W.tight[q] = entry; else atomicOr(W.flags, RF_PAIR_OVERFLOW);
    }
}

// warp-aggregated append to the block's shared-memory stage (tight and loose pairs separately)
__device__ __forceinline__ void emit_pair(const DeviceWorld& W, PairStage& S, uint32_t a, uint32_t b, bool tight, uint32_t flags)
{
    const uint2 entry = make_uint2(min(a, b), max(a, b) | flags | (tight ? ICX_PAIR_TIGHT : 0u));
    const uint32_t m = __activemask(), lane = threadIdx.x & 31u;
    const uint32_t mt = __ballot_sync(m, tight);
    const uint32_t grp = tight ? mt : (m & ~mt);
    const int leader = __ffs(grp) - 1;
    uint32_t base = 0;
    if ((int)lane == leader) base = atomicAdd(tight ? &S.n_tight : &S.n_loose, (uint32_t)__popc(grp));
    base = __shfl_sync(grp, base, leader);
    const uint32_t idx = base + __popc(grp & ((1u << lane) - 1u));
    if (tight) { if (idx < STAGE_TIGHT) S.tight[idx] = entry; else put_global(W, entry, true); }
    else { if (idx < STAGE_LOOSE) S.loose[idx] = entry; else put_global(W, entry, false); }
}

// un-inflated bounds overlap (with 10 % of the margin kept as slack against rounding): only then can
// the narrowphase report a hit on the post-integration state (DESIGN.md "Broadphase bound")
__device__ __forceinline__ bool tight_le(float a, float b, float m2) {
    const float s = m2 - 2.0e-6f * (fabsf(a) + fabsf(b));
    return a + fmaxf(s, 0.0f) <= b;
}
__device__ __forceinline__ bool tight_overlap(float4 alo, float4 ahi, float4 blo, float4 bhi, float margin) {
    const float m2 = 1.8f * margin;
    return tight_le(alo.x, bhi.x, m2) && tight_le(blo.x, ahi.x, m2) && tight_le(alo.y, bhi.y, m2) && tight_le(blo.y, ahi.y, m2) &&
           tight_le(alo.z, bhi.z, m2) && tight_le(blo.z, ahi.z, m2);
}

__global__ void __launch_bounds__(PAIRS_TPB) k_pairs(DeviceWorld W)
{
    __shared__ PairStage S;
    if (threadIdx.x == 0) { S.n_tight = 0; S.n_loose = 0; }
    __syncthreads();
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t n_grid = W.cell_count[W.n_buckets];
    const uint32_t n_large = min(W.counters[CTR_LARGE], W.max_large);
    if (t < n_grid + n_large) {
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
        }
        const uint32_t wl = W.worldline_stride ? e / W.worldline_stride : 0u;
        if (t < n_grid) {
            // forward half of the 27-cell neighbourhood: rows (dz, dy) >= (0, 0); a pair is emitted by
            // the packet whose cell comes first in (z, y, x) order, ties by sorted position
            const int cx = cell_coord(lo.x, hi.x, W.inv_cell), cy = cell_coord(lo.y, hi.y, W.inv_cell), cz = cell_coord(lo.z, hi.z, W.inv_cell);
            const int xb0 = (cx - 1) >> ICX_XB, xb1 = (cx + 1) >> ICX_XB;
            uint32_t seen[10]; int n_seen = 0;
#pragma unroll 1
            for (int r = 4; r < 9; r++) {
                const int dy = r % 3 - 1, dz = r / 3 - 1;
                for (int xb = xb0; xb <= xb1; xb++) {
                    const uint32_t block = row_block(xb, cy + dy, cz + dz, wl, W.n_buckets);
                    const uint32_t key = block * 2u + (uint32_t)(xb - xb0);
                    bool dup = false;
                    for (int q = 0; q < n_seen; q++) dup |= (seen[q] == key);
                    if (dup) continue;
                    seen[n_seen++] = key;
                    const int x_lo = max(cx - 1, xb << ICX_XB), x_hi = min(cx + 1, (xb << ICX_XB) + ICX_XMASK);
                    const uint32_t beg = W.cell_count[(block << ICX_XB) | (uint32_t)(x_lo & ICX_XMASK)];
                    const uint32_t end = W.cell_count[((block << ICX_XB) | (uint32_t)(x_hi & ICX_XMASK)) + 1u];
                    for (uint32_t j = beg; j < end; j++) {
                        if (j == t) continue;
                        const float4 blo = W.s_lo[j], bhi = W.s_hi[j];
                        if (!overlap(lo, hi, blo, bhi)) continue;
                        const uint32_t btag = __float_as_uint(blo.w), be = __float_as_uint(bhi.w);
                        if (be == e) continue;                                      // same entity, relay :51
                        if (W.worldline_stride && be / W.worldline_stride != wl) continue;
                        if (!can_intersect(type, (btag >> 28) & 3, W.cfg_flags)) continue;
                        const int bx = cell_coord(blo.x, bhi.x, W.inv_cell), by = cell_coord(blo.y, bhi.y, W.inv_cell), bz = cell_coord(blo.z, bhi.z, W.inv_cell);
                        const bool mine = bz != cz ? bz > cz : (by != cy ? by > cy : (bx != cx ? bx > cx : j > t));
                        if (!mine) continue;
                        const bool bimmut = (btag & 0x40000000u) != 0;
                        emit_pair(W, S, p, btag & ICX_NB_INDEX_MASK, tight_overlap(lo, hi, blo, bhi, W.margin), (immut && bimmut) ? ICX_PAIR_STATIC : 0u);
                    }
                }
            }
        }
        // large colliders (all immutable): every grid packet tests them; large ones test the later large ones
        for (uint32_t k = (t < n_grid ? 0u : t - n_grid + 1u); k < n_large; k++) {
            const uint32_t q = W.large_list[k];
            const uint32_t qe = W.pk_entity[q];
            if (qe == e) continue;
            if (W.worldline_stride && qe / W.worldline_stride != wl) continue;
            const int qtype = __float_as_uint(W.pk_lpos[q].w) & PF_TYPE_MASK;
            if (!can_intersect(type, qtype, W.cfg_flags)) continue;
            const float4 qlo = W.pk_lo[q], qhi = W.pk_hi[q];
            if (!overlap(lo, hi, qlo, qhi)) continue;
            emit_pair(W, S, p, q, tight_overlap(lo, hi, qlo, qhi, W.margin), immut ? ICX_PAIR_STATIC : 0u);
        }
    }
    __syncthreads();
    const uint32_t nt = min(S.n_tight, (uint32_t)STAGE_TIGHT), nl = min(S.n_loose, (uint32_t)STAGE_LOOSE);
    if (threadIdx.x == 0) {
        if (nt + nl) S.base_all = atomicAdd(&W.counters[CTR_CANDIDATES], nt + nl);
        if (nt) S.base_tight = atomicAdd(&W.counters[CTR_TIGHT], nt);
    }
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < nt + nl; i += blockDim.x) {
        const uint2 entry = i < nt ? S.tight[i] : S.loose[i - nt];
        const uint32_t g = S.base_all + i;
        if (g < W.max_pairs) W.pairs[g] = entry; else atomicOr(W.flags, RF_PAIR_OVERFLOW);
        if (i < nt) { const uint32_t q = S.base_tight + i; if (q < W.max_pairs) W.tight[q] = entry; else atomicOr(W.flags, RF_PAIR_OVERFLOW); }
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
    k_pairs<<<(W.n_packets + PAIRS_TPB - 1) / PAIRS_TPB, PAIRS_TPB, 0, st>>>(W);
}

} // namespace icx
