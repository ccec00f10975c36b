// World-line record / replay / divergence scan on a device-resident
// snapshot ring.
//
// The reference records by serialising every entity into a heap message on a per-entity mutexed
// list each tick (server_network_dynamo.cpp:642-679, parallel_network_dynamo.cpp:369-385) and
// replays by popping them again (server_network_dynamo.cpp:377-421).  Here the dense ENTITY_UPDATE
// stream is a ring of SoA snapshots in HBM: slot = coherency % depth, stamped with the coherency
// it holds.  Rare, variable-length events (create / remove / jump) stay on the host queues.
//   k_record          7 SoA arrays -> ring slot, 128-bit loads/stores, one launch
//   k_replay          ring slot -> world, predicated on the timestream state (non-divergent only)
//   k_divergence_scan TMA-staged (cp.async.bulk + mbarrier, 3-stage ring in shared memory) diff of
//                     live vs recorded origins, congruence predicate of timejump_conditions.h:21-25,
//                     reduced per world-line to {incongruent bodies, fixed-point score}
//   k_fork_age        forking -> forked promotion and the divergence notification
//                     (server_network_dynamo.cpp:681-705)
// All four are pure HBM streaming kernels.
#include <algorithm>
#include <cstring>
#include <new>
#include "icx_host.h"

using namespace icx;

struct IcxTimeline {
    int device = 0;
    uint32_t depth = 0, n_max = 0;
    float4* state = nullptr;            // [depth][6][n_max]: pos, rot, vel, acc, angvel, angacc
    uint32_t* ts = nullptr;             // [depth][n_max]
    std::vector<uint16_t> stamp;        // coherency held by each slot
    std::vector<uint8_t> valid;
    std::vector<uint8_t> behind;        // slot lies behind the breakpoint (old history, TsBreakpointQueue)
    bool breakpoint_active = false;
    std::vector<uint32_t> count;        // entities recorded in each slot
    // scan outputs (device) + pinned host mirror
    unsigned long long* d_score = nullptr; uint32_t* d_count = nullptr; uint32_t cap_wl = 0;
    uint32_t* d_age = nullptr;          // [0] promoted, [1] resolution needed
    // optional kernel timing (icx_timeline_profile): events around the last record / replay / scan launch
    bool profile = false;
    cudaEvent_t pev[3][2] = {};
    bool pev_used[3] = {};
    std::string error;
};

namespace {

#define TCK(call) do { cudaError_t _e = (call); if (_e != cudaSuccess) { if (tl) tl->error = std::string(#call) + ": " + cudaGetErrorString(_e); return ICX_E_CUDA; } } while (0)

__global__ void __launch_bounds__(256) k_record(EntityState S, float4* __restrict__ dst, uint32_t* __restrict__ dst_ts, uint32_t n, uint32_t n_max)
{
    // thread i moves one float4 of each array: a warp moves 512 contiguous bytes per array
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const float4 a = S.pos[i], b = S.rot[i], c = S.vel[i], d = S.acc[i], e = S.angvel[i], f = S.angacc[i];
        const uint32_t t = S.ts[i];
        dst[i] = a; dst[(size_t)n_max + i] = b; dst[2 * (size_t)n_max + i] = c; dst[3 * (size_t)n_max + i] = d;
        dst[4 * (size_t)n_max + i] = e; dst[5 * (size_t)n_max + i] = f;
        dst_ts[i] = t;
    }
}

// ENTITY_UPDATE replay, server_network_dynamo.cpp:396-420: a non-divergent entity takes every
// component of the record; a divergent one only the `upstream` category, which holds no physics
// component (hard_config_cosmos.h:47-58 registers only SpacialInputComponent upstream).
__global__ void __launch_bounds__(256) k_replay(EntityState S, const float4* __restrict__ src, uint32_t n, uint32_t n_src, uint32_t n_max)
{
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const uint32_t state = S.ts[i] & 0xFFu;
        if (state == ICX_TS_FORKING || state == ICX_TS_FORKED) continue;            // is_divergent, timestream_state.h:41-44
        const uint32_t j = n_src == n ? i : i % n_src;
        S.pos[i] = src[j]; S.rot[i] = src[(size_t)n_max + j]; S.vel[i] = src[2 * (size_t)n_max + j];
        S.acc[i] = src[3 * (size_t)n_max + j]; S.angvel[i] = src[4 * (size_t)n_max + j]; S.angacc[i] = src[5 * (size_t)n_max + j];
    }
}

// ---- TMA helpers (1-D bulk copies, mbarrier completion) ----
__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_addr(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smem_addr(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_load(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 :: "r"(smem_addr(dst)), "l"(src), "r"(bytes), "r"(smem_addr(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile("{\n .reg .pred p;\n WAIT_%=:\n mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n @p bra DONE_%=;\n bra WAIT_%=;\n DONE_%=:\n}"
                 :: "r"(smem_addr(bar)), "r"(parity) : "memory");
}

#ifndef SCAN_TILE
#define SCAN_TILE 1024          // bodies per stage: 2 x 16 KB
#define SCAN_STAGES 3
#define SCAN_CTAS 2             // persistent CTAs per SM
#endif
#define SCAN_TPB 256

// TimejumpConditions::operator== (timejump_conditions.h:21-25): congruent <=> length2(a - b) <= 2.0f.
// Per world-line: number of incongruent bodies and sum of min(length2, 2^20) in 48.16 fixed point
// (integer accumulation: the result does not depend on the reduction order).
__global__ void __launch_bounds__(SCAN_TPB) k_divergence_scan(const float4* __restrict__ live, const float4* __restrict__ snap,
                                                             uint32_t n, uint32_t n_snap, uint32_t stride,
                                                             uint32_t* __restrict__ wl_count, unsigned long long* __restrict__ wl_score)
{
    extern __shared__ __align__(128) unsigned char smem[];
    float4* s_live = reinterpret_cast<float4*>(smem);                                   // [STAGES][TILE]
    float4* s_snap = s_live + SCAN_STAGES * SCAN_TILE;
    uint64_t* bar = reinterpret_cast<uint64_t*>(s_snap + SCAN_STAGES * SCAN_TILE);      // [STAGES + 1]
    const uint32_t n_tiles = (n + SCAN_TILE - 1) / SCAN_TILE;
    // A base history shared by all world-lines (config 4: n_snap = one world-line) that fits the snapshot stages is
    // loaded ONCE per CTA and stays resident: re-fetching the same 16 KB for every tile made all CTAs hammer the same
    // few L2 slices, and the batch scan ran at half the speed of the single-world case.
    const bool resident = n_snap != n && n_snap <= SCAN_STAGES * SCAN_TILE;
    if (threadIdx.x == 0) {
        for (int s = 0; s <= SCAN_STAGES; s++) mbar_init(&bar[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (resident && threadIdx.x == 0) {
        mbar_expect_tx(&bar[SCAN_STAGES], n_snap * 16u);
        bulk_load(s_snap, snap, n_snap * 16u, &bar[SCAN_STAGES]);
    }
    // tile -> source range of the snapshot: the whole snapshot is addressed modulo n_snap (a base
    // history shared by all world-lines); SCAN_TILE divides n_snap or n_snap == n is required by the host
    auto issue = [&](uint32_t tile, int stage) {
        const uint32_t first = tile * SCAN_TILE, cnt = min((uint32_t)SCAN_TILE, n - first);
        if (resident) {
            mbar_expect_tx(&bar[stage], cnt * 16u);
            bulk_load(s_live + stage * SCAN_TILE, live + first, cnt * 16u, &bar[stage]);
            return;
        }
        const uint32_t sfirst = n_snap == n ? first : first % n_snap;
        mbar_expect_tx(&bar[stage], cnt * 32u);
        bulk_load(s_live + stage * SCAN_TILE, live + first, cnt * 16u, &bar[stage]);
        bulk_load(s_snap + stage * SCAN_TILE, snap + sfirst, cnt * 16u, &bar[stage]);
    };
    uint32_t my = 0;                                        // tiles this CTA has consumed
    if (threadIdx.x == 0)
        for (int s = 0; s < SCAN_STAGES - 1; s++) { const uint32_t t = blockIdx.x + s * gridDim.x; if (t < n_tiles) issue(t, s); }
    // When a tile never straddles two world-lines (one world, or world-lines of whole tiles: the config-4 batch), every
    // thread accumulates in registers and a warp reduces + publishes ONCE per world-line it has seen -- once per kernel for
    // a single world.  Integer accumulation: the sums do not depend on the order, so both paths give the same bits.
    const bool tile_uniform = stride == 0u || stride % SCAN_TILE == 0u;
    const uint32_t FULL = 0xFFFFFFFFu;
    uint32_t acc_bad = 0, acc_wl = 0xFFFFFFFFu; unsigned long long acc_q = 0;
    auto publish = [&]() {                                  // warp-uniform call
        if (acc_wl == 0xFFFFFFFFu) return;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) { acc_bad += __shfl_xor_sync(FULL, acc_bad, o); acc_q += __shfl_xor_sync(FULL, acc_q, o); }
        if ((threadIdx.x & 31u) == 0) { if (acc_bad) atomicAdd(&wl_count[acc_wl], acc_bad); if (acc_q) atomicAdd(&wl_score[acc_wl], acc_q); }
        acc_bad = 0; acc_q = 0;
    };
    for (uint32_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, my++) {
        const int stage = my % SCAN_STAGES;
        if (threadIdx.x == 0) {                             // keep STAGES-1 tiles in flight
            const uint32_t t = tile + (SCAN_STAGES - 1) * gridDim.x;
            if (t < n_tiles) issue(t, (my + SCAN_STAGES - 1) % SCAN_STAGES);
        }
        const uint32_t first = tile * SCAN_TILE, cnt = min((uint32_t)SCAN_TILE, n - first);
        if (tile_uniform) {
            const uint32_t wl_t = stride ? first / stride : 0u;
            if (wl_t != acc_wl) { publish(); acc_wl = wl_t; }
        }
        if (resident && my == 0) mbar_wait(&bar[SCAN_STAGES], 0u);
        mbar_wait(&bar[stage], (my / SCAN_STAGES) & 1u);
        const float4* a = s_live + stage * SCAN_TILE;
        const float4* b = resident ? s_snap + first % n_snap : s_snap + stage * SCAN_TILE;
#pragma unroll
        for (uint32_t k = 0; k < SCAN_TILE / SCAN_TPB; k++) {
            const uint32_t i = k * SCAN_TPB + threadIdx.x;
            uint32_t bad = 0; unsigned long long q = 0; uint32_t wl = 0xFFFFFFFFu;
            if (i < cnt) {
                const float4 x = a[i], y = b[i];
                const float dx = x.x - y.x, dy = x.y - y.y, dz = x.z - y.z;
                const float len2 = dx * dx + dy * dy + dz * dz;                     // glm::length2 = dot(v, v)
                bad = (len2 <= 2.0f) ? 0u : 1u;
                const float c = (len2 <= 1048576.0f) ? len2 : 1048576.0f;
                q = __float2ull_rz(c * 65536.0f);
                if (!tile_uniform) wl = (first + i) / stride;
            }
            if (tile_uniform) { acc_bad += bad; acc_q += q; continue; }
            // ragged world-lines: one atomic per warp when the whole warp is in one world-line
            const uint32_t wl0 = __shfl_sync(FULL, wl, 0);
            if (__all_sync(FULL, wl == wl0)) {
                if (wl0 != 0xFFFFFFFFu) {
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) { bad += __shfl_xor_sync(FULL, bad, o); q += __shfl_xor_sync(FULL, q, o); }
                    if ((threadIdx.x & 31u) == 0) { if (bad) atomicAdd(&wl_count[wl0], bad); if (q) atomicAdd(&wl_score[wl0], q); }
                }
            } else if (wl != 0xFFFFFFFFu) {
                if (bad) atomicAdd(&wl_count[wl], bad);
                if (q) atomicAdd(&wl_score[wl], q);
            }
        }
        __syncthreads();                                    // the stage may be refilled
    }
    if (tile_uniform) publish();
}

// Fork ageing, server_network_dynamo.cpp:681-705 (thresholds :13-15).  The comparison is the
// reference's plain integer one (`state.second + 13 <= coherency`, not wrap-safe).
__global__ void __launch_bounds__(256) k_fork_age(uint32_t* __restrict__ ts, uint32_t n, uint32_t coherency, uint32_t* __restrict__ out)
{
    uint32_t promoted = 0, resolve = 0;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const uint32_t v = ts[i];
        const uint32_t state = v & 0xFFu, since = v >> 16;
        if (state == ICX_TS_FORKING && since + 13u <= coherency) {                  // FORKING_THRESHOLD
            ts[i] = ICX_TS_FORKED | (coherency << 16);                              // set_timestream_state stamps `now` (cosmos.cpp:155-173)
            promoted++;
        }
        // the reference tests the state it read BEFORE the promotion (`state` is a copy)
        if (state == ICX_TS_FORKED && since + 13u <= coherency) resolve = 1;        // FORKED_THRESHOLD
    }
    promoted = __reduce_add_sync(0xFFFFFFFFu, promoted);
    resolve = __reduce_or_sync(0xFFFFFFFFu, resolve);
    if ((threadIdx.x & 31u) == 0) { if (promoted) atomicAdd(&out[0], promoted); if (resolve) atomicOr(&out[1], 1u); }
}

int slot_of(const IcxTimeline* tl, uint16_t coherency) { return (int)(coherency % tl->depth); }

} // namespace

extern "C" {

int icx_timeline_create(int device, uint32_t depth_ticks, uint32_t n_max, IcxTimeline** out)
{
    if (!out || depth_ticks == 0 || n_max == 0) return ICX_E_INVALID;
    *out = nullptr;
    if (icx_device_count() <= device) return ICX_E_CUDA;        // no CPU fallback
    IcxTimeline* tl = new (std::nothrow) IcxTimeline();
    if (!tl) return ICX_E_NOMEM;
    tl->device = device; tl->depth = depth_ticks; tl->n_max = n_max;
    tl->stamp.assign(depth_ticks, 0); tl->valid.assign(depth_ticks, 0); tl->behind.assign(depth_ticks, 0); tl->count.assign(depth_ticks, 0);
    cudaSetDevice(device);
    cudaError_t e = cudaMalloc((void**)&tl->state, (size_t)depth_ticks * 6 * n_max * sizeof(float4));
    if (e == cudaSuccess) e = cudaMalloc((void**)&tl->ts, (size_t)depth_ticks * n_max * sizeof(uint32_t));
    if (e == cudaSuccess) e = cudaMalloc((void**)&tl->d_age, 2 * sizeof(uint32_t));
    if (e != cudaSuccess) { icx_timeline_destroy(tl); return e == cudaErrorMemoryAllocation ? ICX_E_NOMEM : ICX_E_CUDA; }
    *out = tl;
    return ICX_OK;
}

void icx_timeline_destroy(IcxTimeline* tl)
{
    if (!tl) return;
    cudaSetDevice(tl->device);
    if (tl->state) cudaFree(tl->state);
    if (tl->ts) cudaFree(tl->ts);
    if (tl->d_score) cudaFree(tl->d_score);
    if (tl->d_count) cudaFree(tl->d_count);
    if (tl->d_age) cudaFree(tl->d_age);
    for (auto& pair : tl->pev) for (cudaEvent_t e : pair) if (e) cudaEventDestroy(e);
    delete tl;
}

const char* icx_timeline_last_error(const IcxTimeline* tl) { return tl ? tl->error.c_str() : "null timeline"; }

int icx_timeline_profile(IcxTimeline* tl, int enable)
{
    if (!tl) return ICX_E_INVALID;
    cudaSetDevice(tl->device);
    if (enable && !tl->pev[0][0])
        for (auto& pair : tl->pev) for (cudaEvent_t& e : pair) TCK(cudaEventCreate(&e));
    tl->profile = enable != 0;
    return ICX_OK;
}

int icx_timeline_kernel_ms(IcxTimeline* tl, int which, float* ms)
{
    if (!tl || !ms || which < 0 || which > 2) return ICX_E_INVALID;
    if (!tl->pev_used[which]) { tl->error = "kernel_ms: that kernel has not run with profiling on"; return ICX_E_STATE; }
    cudaSetDevice(tl->device);
    TCK(cudaEventSynchronize(tl->pev[which][1]));
    TCK(cudaEventElapsedTime(ms, tl->pev[which][0], tl->pev[which][1]));
    return ICX_OK;
}

static int record_impl(IcxWorld* w, IcxTimeline* tl, uint16_t coherency, bool at_breakpoint);
int icx_record(IcxWorld* w, IcxTimeline* tl, uint16_t coherency) { return record_impl(w, tl, coherency, false); }
int icx_record_at_breakpoint(IcxWorld* w, IcxTimeline* tl, uint16_t coherency)
{
    if (tl && !tl->breakpoint_active) { tl->error = "record_at_breakpoint: no breakpoint set"; return ICX_E_STATE; }   // push_at_breakpoint returns false
    return record_impl(w, tl, coherency, true);
}
// TsBreakpointQueue::set_breakpoint_at_begin on every timestream (EntityTimestreamMap, link_timestreams):
// everything recorded so far becomes old history, reachable only through the breakpoint.
int icx_breakpoint_set(IcxTimeline* tl)
{
    if (!tl) return ICX_E_INVALID;
    for (uint32_t s = 0; s < tl->depth; s++) tl->behind[s] = tl->valid[s];
    tl->breakpoint_active = true;
    return ICX_OK;
}
int icx_breakpoint_clear(IcxTimeline* tl)
{
    if (!tl) return ICX_E_INVALID;
    std::fill(tl->behind.begin(), tl->behind.end(), 0);
    tl->breakpoint_active = false;
    return ICX_OK;
}

static int record_impl(IcxWorld* w, IcxTimeline* tl, uint16_t coherency, bool at_breakpoint)
{
    if (!w || !tl) return ICX_E_INVALID;
    if (w->n_entities > tl->n_max || w->device != tl->device) { tl->error = "record: world does not fit the timeline"; return ICX_E_INVALID; }
    cudaSetDevice(w->device);
    const int s = slot_of(tl, coherency);
    const uint32_t n = w->n_entities;
    if (n) {
        const int blocks = (int)std::min<uint32_t>((n + 255) / 256, 148 * 16);
        if (tl->profile) cudaEventRecord(tl->pev[0][0], w->stream);
        k_record<<<blocks, 256, 0, w->stream>>>(w->W.in, tl->state + (size_t)s * 6 * tl->n_max, tl->ts + (size_t)s * tl->n_max, n, tl->n_max);
        if (tl->profile) { cudaEventRecord(tl->pev[0][1], w->stream); tl->pev_used[0] = true; }
        TCK(cudaGetLastError());
        w->launches += 1;
    }
    tl->stamp[s] = coherency; tl->valid[s] = 1; tl->count[s] = n;
    // push_back lands after the breakpoint (old side), push_at_breakpoint before it (ts_breakpoint_queue.h:44-52, :140-152)
    tl->behind[s] = (tl->breakpoint_active && !at_breakpoint) ? 1 : 0;
    return ICX_OK;
}

int icx_replay(IcxWorld* w, IcxTimeline* tl, uint16_t coherency, int mode, int* applied)
{
    if (!w || !tl) return ICX_E_INVALID;
    if (applied) *applied = 0;
    if (mode != ICX_REPLAY_SERVER && mode != ICX_REPLAY_PARALLEL) return ICX_E_INVALID;
    // pop_future_timestream hands out every record with coherency <= now (cyclic, message.h:179-183) and
    // the relay drops those older than now (server_network_dynamo.cpp:391): only the record stamped `now`
    // is applied, older ones are consumed unseen.  pop_front stops at the breakpoint; pop_at_breakpoint
    // only takes what lies behind it.
    const int s = slot_of(tl, coherency);
    bool hit = false;
    for (uint32_t k = 0; k < tl->depth; k++) {
        if (!tl->valid[k]) continue;
        const uint16_t a = coherency, b = tl->stamp[k];
        const bool le_now = (a >= b && a - b <= 32768) || (a <= b && b - a > 32768);
        if (!le_now) continue;
        const bool behind = tl->breakpoint_active && tl->behind[k];
        if (mode == ICX_REPLAY_SERVER ? behind : !behind) continue;
        tl->valid[k] = 0; tl->behind[k] = 0;
        if ((int)k == s && b == coherency) hit = true;
    }
    if (!hit) return ICX_OK;
    if (mode == ICX_REPLAY_PARALLEL) return ICX_OK;             // parallel_network_dynamo.cpp:78-86: upstream components only
    const uint32_t n = w->n_entities, n_src = tl->count[s];
    if (!n || !n_src) return ICX_OK;
    if (n_src != n && !(w->W.worldline_stride && n_src == w->W.worldline_stride)) { tl->error = "replay: record size matches neither the world nor one world-line"; return ICX_E_INVALID; }
    cudaSetDevice(w->device);
    const int blocks = (int)std::min<uint32_t>((n + 255) / 256, 148 * 16);
    if (tl->profile) cudaEventRecord(tl->pev[1][0], w->stream);
    k_replay<<<blocks, 256, 0, w->stream>>>(w->W.in, tl->state + (size_t)s * 6 * tl->n_max, n, n_src, tl->n_max);
    if (tl->profile) { cudaEventRecord(tl->pev[1][1], w->stream); tl->pev_used[1] = true; }
    TCK(cudaGetLastError());
    w->launches += 1;
    if (applied) *applied = 1;
    return ICX_OK;
}

int icx_divergence_scan(IcxWorld* w, IcxTimeline* tl, uint16_t coherency, uint32_t n_worldlines,
                        uint8_t* flags, uint32_t* incongruent, double* score)
{
    if (!w || !tl || n_worldlines == 0) return ICX_E_INVALID;
    const int s = slot_of(tl, coherency);
    if (!tl->valid[s] || tl->stamp[s] != coherency) { tl->error = "scan: no record stamped with this coherency"; return ICX_E_STATE; }
    const uint32_t n = w->n_entities, n_snap = tl->count[s], stride = w->W.worldline_stride;
    if ((stride ? (n + stride - 1) / stride : 1u) != n_worldlines) { tl->error = "scan: n_worldlines does not match the world"; return ICX_E_INVALID; }
    if (n_snap != n && !(stride && n_snap == stride && stride % SCAN_TILE == 0)) { tl->error = "scan: record must cover the world, or one world-line whose size is a multiple of 1024"; return ICX_E_INVALID; }
    cudaSetDevice(w->device);
    if (tl->cap_wl < n_worldlines) {
        if (tl->d_score) cudaFree(tl->d_score);
        if (tl->d_count) cudaFree(tl->d_count);
        tl->d_score = nullptr; tl->d_count = nullptr; tl->cap_wl = 0;
        TCK(cudaMalloc((void**)&tl->d_score, (size_t)n_worldlines * 8)); TCK(cudaMalloc((void**)&tl->d_count, (size_t)n_worldlines * 4));
        tl->cap_wl = n_worldlines;
    }
    cudaStream_t st = w->stream;
    TCK(cudaMemsetAsync(tl->d_score, 0, (size_t)n_worldlines * 8, st)); TCK(cudaMemsetAsync(tl->d_count, 0, (size_t)n_worldlines * 4, st));
    if (n) {
        const size_t smem = (size_t)2 * SCAN_STAGES * SCAN_TILE * sizeof(float4) + (SCAN_STAGES + 1) * sizeof(uint64_t);
        static bool attr_set[64] = {};
        if (!attr_set[w->device & 63]) { TCK(cudaFuncSetAttribute(k_divergence_scan, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); attr_set[w->device & 63] = true; }
        const uint32_t n_tiles = (n + SCAN_TILE - 1) / SCAN_TILE;
        const int blocks = (int)std::min<uint32_t>(n_tiles, 148 * SCAN_CTAS);
        if (tl->profile) cudaEventRecord(tl->pev[2][0], st);
        k_divergence_scan<<<blocks, SCAN_TPB, smem, st>>>(w->W.in.pos, tl->state + (size_t)s * 6 * tl->n_max, n, n_snap, stride, tl->d_count, tl->d_score);
        if (tl->profile) { cudaEventRecord(tl->pev[2][1], st); tl->pev_used[2] = true; }
        TCK(cudaGetLastError());
        w->launches += 1;
    }
    std::vector<unsigned long long> hs(n_worldlines); std::vector<uint32_t> hc(n_worldlines);
    TCK(cudaMemcpyAsync(hs.data(), tl->d_score, (size_t)n_worldlines * 8, cudaMemcpyDeviceToHost, st));
    TCK(cudaMemcpyAsync(hc.data(), tl->d_count, (size_t)n_worldlines * 4, cudaMemcpyDeviceToHost, st));
    TCK(cudaStreamSynchronize(st));
    for (uint32_t k = 0; k < n_worldlines; k++) {
        if (flags) flags[k] = hc[k] ? 1 : 0;
        if (incongruent) incongruent[k] = hc[k];
        if (score) score[k] = (double)hs[k] / 65536.0;
    }
    return ICX_OK;
}

int icx_fork_age(IcxWorld* w, uint32_t* n_promoted, int* resolution_needed)
{
    if (!w) return ICX_E_INVALID;
    IcxTimeline* tl = nullptr;
    cudaSetDevice(w->device);
    cudaStream_t st = w->stream;
    uint32_t* d = w->W.counters + CTR_PROF;         // two spare words of the world's counter block
    TCK(cudaMemsetAsync(d, 0, 8, st));
    const uint32_t n = w->n_entities;
    if (n) {
        const int blocks = (int)std::min<uint32_t>((n + 255) / 256, 148 * 8);
        k_fork_age<<<blocks, 256, 0, st>>>(w->W.in.ts, n, (uint32_t)w->W.coherency, d);
        TCK(cudaGetLastError());
        w->launches += 1;
    }
    uint32_t h[2] = {0, 0};
    TCK(cudaMemcpyAsync(h, d, 8, cudaMemcpyDeviceToHost, st));
    TCK(cudaStreamSynchronize(st));
    if (n_promoted) *n_promoted = h[0];
    if (resolution_needed) *resolution_needed = h[1] ? 1 : 0;
    return ICX_OK;
}

// The snapshot stream between adjacent timeslices (future -> past, server_network_dynamo.cpp:642-679
// push_past_timestream) when the two timeslices live on different GPUs / processes: the slot stamped
// `coherency` packed as [6][n] float4 + [n] uint32 = 100 B per entity in a caller-owned DEVICE buffer
// (what a ncclSend / cudaMemcpyPeer moves), and its inverse on the receiving side, which lands in
// the ring exactly like icx_record (push_back).  Copies run on `stream` (cudaStream_t, NULL = the
// default stream) and are complete on return.
int icx_timeline_slot_export(IcxTimeline* tl, uint16_t coherency, void* device_buffer, size_t capacity_bytes, uint32_t* n_out, void* stream)
{
    if (!tl || !device_buffer) return ICX_E_INVALID;
    const int s = slot_of(tl, coherency);
    if (!(tl->valid[s] && tl->stamp[s] == coherency)) { tl->error = "slot_export: no record with this coherency"; return ICX_E_STATE; }
    const uint32_t n = tl->count[s];
    if (n_out) *n_out = n;
    if ((size_t)n * ICX_SNAPSHOT_BYTES > capacity_bytes) { tl->error = "slot_export: buffer too small"; return ICX_E_CAPACITY; }
    if (!n) return ICX_OK;
    cudaSetDevice(tl->device);
    cudaStream_t st = (cudaStream_t)stream;
    TCK(cudaMemcpy2DAsync(device_buffer, (size_t)n * 16, tl->state + (size_t)s * 6 * tl->n_max, (size_t)tl->n_max * 16, (size_t)n * 16, 6,
                          cudaMemcpyDeviceToDevice, st));
    TCK(cudaMemcpyAsync((char*)device_buffer + (size_t)n * 96, tl->ts + (size_t)s * tl->n_max, (size_t)n * 4, cudaMemcpyDeviceToDevice, st));
    TCK(cudaStreamSynchronize(st));
    return ICX_OK;
}

int icx_timeline_slot_import(IcxTimeline* tl, uint16_t coherency, uint32_t n, const void* device_buffer, void* stream)
{
    if (!tl || (!device_buffer && n)) return ICX_E_INVALID;
    if (n > tl->n_max) { tl->error = "slot_import: more entities than the timeline holds"; return ICX_E_CAPACITY; }
    const int s = slot_of(tl, coherency);
    cudaSetDevice(tl->device);
    cudaStream_t st = (cudaStream_t)stream;
    if (n) {
        TCK(cudaMemcpy2DAsync(tl->state + (size_t)s * 6 * tl->n_max, (size_t)tl->n_max * 16, device_buffer, (size_t)n * 16, (size_t)n * 16, 6,
                              cudaMemcpyDeviceToDevice, st));
        TCK(cudaMemcpyAsync(tl->ts + (size_t)s * tl->n_max, (const char*)device_buffer + (size_t)n * 96, (size_t)n * 4, cudaMemcpyDeviceToDevice, st));
        TCK(cudaStreamSynchronize(st));
    }
    tl->stamp[s] = coherency; tl->valid[s] = 1; tl->count[s] = n;
    tl->behind[s] = tl->breakpoint_active ? 1 : 0;
    return ICX_OK;
}

int icx_timeline_slot_info(const IcxTimeline* tl, uint16_t coherency, int* valid, uint32_t* n_recorded)
{
    if (!tl) return ICX_E_INVALID;
    const int s = slot_of(tl, coherency);
    const bool ok = tl->valid[s] && tl->stamp[s] == coherency;
    if (valid) *valid = ok ? 1 : 0;
    if (n_recorded) *n_recorded = ok ? tl->count[s] : 0;
    return ICX_OK;
}

} // extern "C"
