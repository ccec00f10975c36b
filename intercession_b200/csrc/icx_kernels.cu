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
    W.push[e] = 0.0f; W.pushv[e] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    W.bm[e] = W.margin;                                 // displacement budget: the uniform margin unless the ordered pass decides otherwise (budget_entity)
    { const float k = W.bm_keep[e]; if (k > 0.0f) { const float d = k * 0.85f; W.bm_keep[e] = d > W.margin ? d : 0.0f; } }   // the carried budget floor fades

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
            if (!(ext < 3.0e38f)) {
                // non-finite bound (a NaN / Inf transform): the reference carries such a body along, every test against it
                // fails; here it takes no part in the grid
            } else if (ext > W.cell_size) {
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
        uint32_t s = lane < (blockDim.x >> 5) ? warp_sums[lane] : 0u;
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
// K4: uniform-grid pair generator.  Phase A (one thread per bucket-sorted packet) resolves the
// forward half of the 27-cell neighbourhood into <= 10 contiguous ranges of the sorted array (all
// table look-ups issued back to back) and stages them, with the packet's bound, in shared memory.
// Phase B: every warp walks the candidates of its own 32 packets with 8 lanes per packet, four
// packets at a time, so a packet's ~40 candidates cost ~5 warp iterations instead of 40 divergent
// ones; partners whose inflated bounds overlap are kept and each unordered pair is emitted once
// (from the packet whose cell comes first in (z, y, x) order, ties by sorted position).  Pairs are
// staged per warp (ballot / popc, no atomics) and copied out in coalesced runs with one global
// atomic per flush.  Entry = (lower packet, higher packet | flags).
// ---------------------------------------------------------------------------------------
#define PAIRS_TPB 256
#define PAIRS_MAXR 10
#define CAND_CAP 96                                 // 31 left over + 2 per lane per iteration
#define WSTAGE 64                                   // per-warp stage capacity; flushed when more than WSTAGE - 32 entries are staged

struct PairStage {
    float4 lo[PAIRS_TPB], hi[PAIRS_TPB];            // own bounds (+ tag / entity in .w)
    uint32_t rbeg[PAIRS_TPB][PAIRS_MAXR];           // range start in the sorted array
    uint32_t rcum[PAIRS_TPB][PAIRS_MAXR];           // candidates of this packet up to the end of the range (inclusive scan)
    uint2 tight[PAIRS_TPB / 32][WSTAGE], loose[PAIRS_TPB / 32][WSTAGE];
    uint2 cand[PAIRS_TPB / 32][CAND_CAP];           // per warp: (packet slot in the block, sorted position of the partner) whose bounds overlap
};

// Per-warp pair stage: every lane of a converged warp calls with its verdict; entries are appended
// by ballot / popc (no atomics) and the stage is copied out in one coalesced run with ONE global
// atomic per list when it could overflow on the next call.
struct WarpStage { uint32_t n_tight, n_loose; };     // warp-uniform registers

__device__ __forceinline__ void warp_flush(const DeviceWorld& W, uint2* stage, uint32_t& n, uint32_t ctr, uint2* dst, uint2* dst2, uint32_t ctr2)
{
    const uint32_t FULL = 0xFFFFFFFFu, lane = threadIdx.x & 31u;
    if (n == 0) return;
    uint32_t base = 0, base2 = 0;
    if (lane == 0) { base = atomicAdd(&W.counters[ctr], n); if (dst2) base2 = atomicAdd(&W.counters[ctr2], n); }
    base = __shfl_sync(FULL, base, 0); base2 = __shfl_sync(FULL, base2, 0);
    __syncwarp();
    for (uint32_t i = lane; i < n; i += 32u) {
        const uint2 entry = stage[i];
        if (base + i < W.max_pairs) dst[base + i] = entry; else atomicOr(W.flags, RF_PAIR_OVERFLOW);
        if (dst2) { if (base2 + i < W.max_pairs) dst2[base2 + i] = entry; else atomicOr(W.flags, RF_PAIR_OVERFLOW); }
    }
    __syncwarp();
    n = 0;
}

__device__ __forceinline__ void warp_emit(const DeviceWorld& W, PairStage& S, WarpStage& ws, bool emit, bool tight, uint2 entry)
{
    const uint32_t FULL = 0xFFFFFFFFu, lane = threadIdx.x & 31u, wid = threadIdx.x >> 5;
    const uint32_t mt = __ballot_sync(FULL, emit && tight), ml = __ballot_sync(FULL, emit && !tight);
    const uint32_t below = (1u << lane) - 1u;
    if (emit && tight) S.tight[wid][ws.n_tight + __popc(mt & below)] = entry;
    if (emit && !tight) S.loose[wid][ws.n_loose + __popc(ml & below)] = entry;
    ws.n_tight += __popc(mt); ws.n_loose += __popc(ml);
    // tight pairs go to both lists (all candidates; narrowphase input)
    if (ws.n_tight > WSTAGE - 32) warp_flush(W, S.tight[wid], ws.n_tight, CTR_CANDIDATES, W.pairs, W.tight, CTR_TIGHT);
    if (ws.n_loose > WSTAGE - 32) warp_flush(W, S.loose[wid], ws.n_loose, CTR_CANDIDATES, W.pairs, nullptr, 0);
}

#ifndef PAIRS_MINB
#define PAIRS_MINB 4
#endif
__global__ void __launch_bounds__(PAIRS_TPB, PAIRS_MINB) k_pairs(DeviceWorld W)
{
    __shared__ PairStage S;
    const uint32_t tid = threadIdx.x;
    const uint32_t t = blockIdx.x * blockDim.x + tid;
    const uint32_t n_grid = W.cell_count[W.n_buckets];
    const uint32_t n_large = min(W.counters[CTR_LARGE], W.max_large);
    WarpStage ws; ws.n_tight = 0; ws.n_loose = 0;

    // ---- phase A: own bound and the ranges of the forward half neighbourhood ----
    uint32_t p = 0, e = 0; int type = 0; bool immut = false;
    float4 lo = make_float4(0, 0, 0, 0), hi = lo;
    uint32_t count = 0;
    const bool have = t < n_grid + n_large;
    if (have) {
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
    }
    const uint32_t wl = W.worldline_stride ? e / W.worldline_stride : 0u;
    if (have && t < n_grid) {
        // rows (dz, dy) >= (0, 0) of the 27-cell neighbourhood; the 3 x-neighbours of a row are one or two
        // contiguous ranges of the sorted array (icx_world.cuh, ICX_XB).  All table look-ups are issued
        // before any is used (independent loads), duplicates (hash collisions between rows) are dropped.
        const int cx = cell_coord(lo.x, hi.x, W.inv_cell), cy = cell_coord(lo.y, hi.y, W.inv_cell), cz = cell_coord(lo.z, hi.z, W.inv_cell);
        // A partner's bound overlaps mine only if its centre lies within my bound widened by its half extent
        // (<= cell / 2): cells outside that box are skipped (slack: 0.1 % of a cell + rounding of the coordinates).
        const float half = 0.5005f * W.cell_size;
        const float sx = half + 4.0e-6f * fmaxf(fabsf(lo.x), fabsf(hi.x)), sy = half + 4.0e-6f * fmaxf(fabsf(lo.y), fabsf(hi.y)),
                    sz = half + 4.0e-6f * fmaxf(fabsf(lo.z), fabsf(hi.z));
        const int nx0 = max(cx - 1, (int)floorf((lo.x - sx) * W.inv_cell)), nx1 = min(cx + 1, (int)floorf((hi.x + sx) * W.inv_cell));
        const int ny0 = max(cy - 1, (int)floorf((lo.y - sy) * W.inv_cell)), ny1 = min(cy + 1, (int)floorf((hi.y + sy) * W.inv_cell));
        const int nz1 = min(cz + 1, (int)floorf((hi.z + sz) * W.inv_cell));
        const int xb0 = nx0 >> ICX_XB, xb1 = nx1 >> ICX_XB;
        uint32_t key[PAIRS_MAXR], beg[PAIRS_MAXR], end[PAIRS_MAXR];
#pragma unroll
        for (int r = 0; r < 5; r++) {
            const int dy = (r + 4) % 3 - 1, dz = (r + 4) / 3 - 1;
            const bool row_ok = cy + dy >= ny0 && cy + dy <= ny1 && cz + dz <= nz1;
#pragma unroll
            for (int part = 0; part < 2; part++) {
                const int xb = xb0 + part, i = r * 2 + part;
                key[i] = ICX_NONE; beg[i] = 0; end[i] = 0;
                if (row_ok && xb <= xb1) {
                    const uint32_t block = row_block(xb, cy + dy, cz + dz, wl, W.n_buckets);
                    key[i] = block * 2u + (uint32_t)part;
                    const int x_lo = max(nx0, xb << ICX_XB), x_hi = min(nx1, (xb << ICX_XB) + ICX_XMASK);
                    beg[i] = W.cell_count[(block << ICX_XB) | (uint32_t)(x_lo & ICX_XMASK)];
                    end[i] = W.cell_count[((block << ICX_XB) | (uint32_t)(x_hi & ICX_XMASK)) + 1u];
                }
            }
        }
        int n_r = 0;
#pragma unroll
        for (int i = 0; i < PAIRS_MAXR; i++) {
            bool keep = key[i] != ICX_NONE && end[i] > beg[i];
#pragma unroll
            for (int q = 0; q < i; q++) keep = keep && key[q] != key[i];
            if (keep) { count += end[i] - beg[i]; S.rbeg[tid][n_r] = beg[i]; S.rcum[tid][n_r] = count; n_r++; }
        }
    }
    S.lo[tid] = lo; S.hi[tid] = make_float4(hi.x, hi.y, hi.z, __uint_as_float(e));
    __syncwarp();

    // ---- phase B: 8 lanes per packet, 4 packets of this warp at a time, 2 candidates per lane and iteration ----
    // Only ~8 % of the candidates overlap; what follows the overlap test (emission rule, tight test,
    // staging) would run with 1-3 active lanes per iteration, so overlapping candidates are queued per
    // warp and finished 32 at a time with all lanes busy.
    const uint32_t FULL = 0xFFFFFFFFu, lane = tid & 31u, sub = lane & 7u, grp = lane >> 3, wid = tid >> 5;
    const uint32_t wbase = tid & ~31u, tb0 = blockIdx.x * blockDim.x;
    uint32_t n_cand = 0;                                                                // warp-uniform
    auto finish = [&](uint32_t first, uint32_t n) {                                     // queue entries [first, first + n), n <= 32
        bool emit = false, tight = false; uint2 entry = make_uint2(0, 0);
        if (lane < n) {
            const uint2 cq = S.cand[wid][first + lane];
            const uint32_t q = cq.x, j = cq.y, tq = tb0 + q;
            const float4 alo = S.lo[q], ahi = S.hi[q];
            const float4 blo = W.s_lo[j], bhi = W.s_hi[j];
            const uint32_t atag = __float_as_uint(alo.w), ae = __float_as_uint(ahi.w);
            const uint32_t btag = __float_as_uint(blo.w), be = __float_as_uint(bhi.w);
            bool ok = be != ae;                                                         // same entity, relay :51
            if (W.worldline_stride) ok = ok && (be / W.worldline_stride == ae / W.worldline_stride);
            ok = ok && can_intersect((atag >> 28) & 3, (btag >> 28) & 3, W.cfg_flags);
            if (ok) {
                const int ax = cell_coord(alo.x, ahi.x, W.inv_cell), ay = cell_coord(alo.y, ahi.y, W.inv_cell), az = cell_coord(alo.z, ahi.z, W.inv_cell);
                const int bx = cell_coord(blo.x, bhi.x, W.inv_cell), by = cell_coord(blo.y, bhi.y, W.inv_cell), bz = cell_coord(blo.z, bhi.z, W.inv_cell);
                emit = bz != az ? bz > az : (by != ay ? by > ay : (bx != ax ? bx > ax : j > tq));
                const uint32_t a = atag & ICX_NB_INDEX_MASK, b = btag & ICX_NB_INDEX_MASK;
                tight = tight_overlap(alo, ahi, blo, bhi, W.margin);
                const bool ia = (atag & 0x40000000u) != 0, ib = (btag & 0x40000000u) != 0;
                const bool lo_immut = a < b ? ia : ib, hi_immut = a < b ? ib : ia;
                entry = make_uint2(min(a, b) | (lo_immut ? ICX_PAIR_A_IMMUT : 0u),
                                   max(a, b) | ((ia && ib) ? ICX_PAIR_STATIC : 0u) | (tight ? ICX_PAIR_TIGHT : 0u) | (hi_immut ? ICX_PAIR_B_IMMUT : 0u));
            }
        }
        warp_emit(W, S, ws, emit, tight, entry);
        __syncwarp();
    };
#pragma unroll 1
    for (uint32_t round = 0; round < 8u; round++) {
        const uint32_t ql = round * 4u + grp, q = wbase + ql;                          // packet of this lane group
        const uint32_t cnt = __shfl_sync(FULL, count, ql);
        uint32_t mx = max(cnt, __shfl_xor_sync(FULL, cnt, 8)); mx = max(mx, __shfl_xor_sync(FULL, mx, 16));
        if (mx == 0u) continue;
        const float4 alo = S.lo[q], ahi = S.hi[q];
        const uint32_t tq = tb0 + q;
        int k = 0; uint32_t kbase = 0, kend = cnt ? S.rcum[q][0] : 0u;                  // current range: candidates [kbase, kend)
        for (uint32_t rb = 0; rb < mx; rb += 16u) {                                     // warp-uniform trip count
            uint32_t j[2]; bool pass[2];
#pragma unroll
            for (int u = 0; u < 2; u++) {
                const uint32_t r = rb + u * 8u + sub;
                j[u] = ICX_NONE;
                if (r < cnt) {
                    while (r >= kend) { k++; kbase = kend; kend = S.rcum[q][k]; }
                    j[u] = S.rbeg[q][k] + (r - kbase);
                    if (j[u] == tq) j[u] = ICX_NONE;
                }
            }
            float4 blo[2], bhi[2];
#pragma unroll
            for (int u = 0; u < 2; u++) if (j[u] != ICX_NONE) { blo[u] = W.s_lo[j[u]]; bhi[u] = W.s_hi[j[u]]; }
#pragma unroll
            for (int u = 0; u < 2; u++) {
                pass[u] = j[u] != ICX_NONE && overlap(alo, ahi, blo[u], bhi[u]);
                const uint32_t m = __ballot_sync(FULL, pass[u]);
                if (pass[u]) S.cand[wid][n_cand + __popc(m & ((1u << lane) - 1u))] = make_uint2(q, j[u]);
                n_cand += __popc(m);
            }
            __syncwarp();
            while (n_cand >= 32u) { n_cand -= 32u; finish(n_cand, 32u); }
        }
    }
    __syncwarp();
    if (n_cand) finish(0u, n_cand);

    // ---- large colliders (all immutable): every grid packet tests them; large ones test the later large ones ----
    for (uint32_t k = 0; k < n_large; k++) {
        bool emit = false, tight = false; uint2 entry = make_uint2(0, 0);
        if (have && k >= (t < n_grid ? 0u : t - n_grid + 1u)) {
            const uint32_t q = W.large_list[k];
            const uint32_t qe = W.pk_entity[q];
            const int qtype = __float_as_uint(W.pk_lpos[q].w) & PF_TYPE_MASK;
            if (qe != e && (!W.worldline_stride || qe / W.worldline_stride == wl) && can_intersect(type, qtype, W.cfg_flags)) {
                const float4 qlo = W.pk_lo[q], qhi = W.pk_hi[q];
                if (overlap(lo, hi, qlo, qhi)) {
                    emit = true; tight = tight_overlap(lo, hi, qlo, qhi, W.margin);
                    // q (large list) is always immutable
                    entry = make_uint2(min(p, q) | ((p < q ? immut : true) ? ICX_PAIR_A_IMMUT : 0u),
                                       max(p, q) | (immut ? ICX_PAIR_STATIC : 0u) | (tight ? ICX_PAIR_TIGHT : 0u) | ((p < q ? true : immut) ? ICX_PAIR_B_IMMUT : 0u));
                }
            }
        }
        warp_emit(W, S, ws, emit, tight, entry);
    }
    warp_flush(W, S.tight[tid >> 5], ws.n_tight, CTR_CANDIDATES, W.pairs, W.tight, CTR_TIGHT);
    warp_flush(W, S.loose[tid >> 5], ws.n_loose, CTR_CANDIDATES, W.pairs, nullptr, 0);
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

// exclusive scan of data[0..n) in place; data[n] receives the total (n <= 4096 * 4096)
void launch_exclusive_scan(uint32_t* data, uint32_t* partial, uint32_t n, cudaStream_t st)
{
    const uint32_t n_part = (n + SCAN_EPB - 1) / SCAN_EPB;
    if (n) {
        k_scan_blocks<<<n_part, SCAN_TPB, 0, st>>>(data, partial, n);
        k_scan_partials<<<1, SCAN_TPB, 0, st>>>(partial, n_part);
    } else cudaMemsetAsync(partial, 0, 4, st);
    k_scan_add<<<(n + 1 + 255) / 256, 256, 0, st>>>(data, partial, n, n_part);
}

void launch_pairs(const DeviceWorld& W, cudaStream_t st)
{
    if (W.n_packets == 0) return;
    k_pairs<<<(W.n_packets + PAIRS_TPB - 1) / PAIRS_TPB, PAIRS_TPB, 0, st>>>(W);
}

} // namespace icx
