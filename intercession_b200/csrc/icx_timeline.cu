// World-line record / replay / divergence scan on a device-resident
// snapshot ring.
//
// The reference records by serialising every entity -- ALL of its components (ComponentCategory::all,
// server_network_dynamo.cpp:666-677, parallel_network_dynamo.cpp:369-385) -- into a heap message on a
// per-entity mutexed list each tick and replays by popping them again and overwriting all components
// of the non-divergent entities (server_network_dynamo.cpp:377-421).  Here the dense ENTITY_UPDATE
// stream is a ring of SoA snapshots in HBM: slot = coherency % depth, stamped with the coherency it
// holds.  A slot carries, per entity, the whole TransformComponent + PhysicsComponent (10 float4: the
// six mutable ones plus scale|mass, drags|flag bits, lockedOrigin|collisionLinearDrag, lockedOrientation)
// and the id the receiving cosmos knows the entity by (causal chain-link + 1, :673); the
// ColliderComponent, which changes rarely (BipedBehaviors writes restLength / damping of its own
// entities, biped_behaviors.h:96-100), travels as a SPARSE list of {entity, 168-byte row} per slot for
// the entities whose component changed since the previous record into this ring (DeviceWorld::c_ver
// against IcxTimeline::seen).  Rare, variable-length events (create / remove / jump) stay on the host queues.
//   k_record          11 SoA arrays -> ring slot, 128-bit loads/stores, one launch
//   k_replay          ring slot -> world, predicated on the timestream state (non-divergent only) and on the id
//   k_delta_*         sparse collider rows: collect on record, apply (or stash while divergent) on replay
//   k_divergence_scan TMA-staged (cp.async.bulk + mbarrier, 3-stage ring in shared memory) diff of
//                     live vs recorded origins, congruence predicate of timejump_conditions.h:21-25,
//                     reduced per world-line to {incongruent bodies, fixed-point score}
//   k_fork_age        forking -> forked promotion and the divergence notification
//                     (server_network_dynamo.cpp:681-705)
// All of them are pure HBM streaming kernels.
#include <algorithm>
#include <cstring>
#include <new>
#include "icx_host.h"
#include "icx_rows.cuh"

using namespace icx;

#define TL_ARRAYS 10                    // float4 arrays per slot: pos rot vel acc angvel angacc | prm0 prm1 locked_pos locked_rot

struct DeltaRow { uint32_t idx, id; IcxColliderComponent C; };         // 176 B, 16-byte aligned: sender's row (a hint), recorded id (the address)
static_assert(sizeof(DeltaRow) == ICX_SNAPSHOT_DELTA_BYTES, "delta row layout");
static_assert(sizeof(IcxTimelineEvent) == 32, "event record layout");
static_assert(16 + 32 * ICX_TIMELINE_EVENTS_PER_TICK == ICX_SNAPSHOT_HEADER_BYTES, "transport header = row count word + events block");

struct IcxTimeline {
    int device = 0, sms = 148;
    uint32_t depth = 0, n_max = 0;
    float4* state = nullptr;            // [depth][TL_ARRAYS][n_max]
    uint32_t* ids = nullptr;            // [depth][n_max]: id under which the RECEIVER knows the entity
    std::vector<uint16_t> stamp;        // coherency held by each slot
    std::vector<uint8_t> valid;
    std::vector<uint8_t> behind;        // slot lies behind the breakpoint (old history, TsBreakpointQueue)
    std::vector<uint64_t> seq;          // push order (FIFO order of the per-entity lists)
    uint64_t next_seq = 1;
    bool breakpoint_active = false;
    std::vector<uint32_t> count;        // entities recorded in each slot
    // sparse collider rows
    DeltaRow* delta = nullptr; uint32_t delta_cap = 0;      // [depth][delta_cap]
    uint32_t* d_delta_n = nullptr;      // [depth + 1] rows held by each slot (device), + one scratch counter
    std::vector<int64_t> delta_n;       // host view: rows in the slot, -1 = unknown (imported)
    uint32_t* seen = nullptr;           // [n_max]: DeviceWorld::c_ver of each entity at its last collected row
    const IcxWorld* last_recorder = nullptr; uint32_t last_epoch = 0;
    unsigned long long* d_mismatch = nullptr;   // entities a replay skipped because the recorded id is not theirs
    // event records (ENTITY_CREATED / REMOVED, JUMP_*): one block per slot on the device (it travels in the transport header),
    // mirrored in pinned host memory; the slot's snapshot and its events are consumed independently
    struct EvBlock { uint32_t rows_unused, n_events, pad[2]; IcxTimelineEvent ev[ICX_TIMELINE_EVENTS_PER_TICK]; };
    EvBlock* d_ev = nullptr; EvBlock* h_ev = nullptr;       // [depth]
    void* d_fetch = nullptr;            // staging of icx_timeline_entity_fetch (one recorded entity)
    std::vector<uint8_t> ev_valid, ev_behind, ev_pending;   // pending: the host mirror is being filled by the import's copy (slot_ev)
    std::vector<uint16_t> ev_stamp;
    std::vector<uint64_t> ev_seq;
    std::vector<cudaEvent_t> slot_ev;   // slot contents complete (record / import) -- waited on by replay / export on other streams
    // scan outputs (device) + pinned host mirror
    unsigned long long* d_score = nullptr; uint32_t* d_count = nullptr; uint32_t cap_wl = 0;
    uint8_t* d_ever = nullptr;          // per world-line: flagged by some scan since the last reset
    cudaStream_t scan_stream = nullptr; uint32_t scan_wl = 0;
    uint32_t* d_age = nullptr;          // [0] promoted, [1] resolution needed
    // optional kernel timing (icx_timeline_profile): events around the last record / replay / scan launch
    bool profile = false;
    cudaEvent_t pev[3][2] = {};
    bool pev_used[3] = {};
    std::string error;
};

namespace {

#define TCK(call) do { cudaError_t _e = (call); if (_e != cudaSuccess) { if (tl) tl->error = std::string(#call) + ": " + cudaGetErrorString(_e); return ICX_E_CUDA; } } while (0)

struct RecordSrc { EntityState S; const float4* prm0; const float4* prm1; const float4* locked_pos; const float4* locked_rot; const uint32_t* entity_id; };

// increment_causal_chain_link, ecs_types.h:118-136: refused for link >= 14 (14 + 1 would be the null link)
__device__ __forceinline__ uint32_t inc_link(uint32_t id) { return (id & 0xFu) + 1u >= 15u ? id : id + 1u; }

__global__ void __launch_bounds__(256) k_record(RecordSrc R, float4* __restrict__ dst, uint32_t* __restrict__ dst_ids, uint32_t n, uint32_t n_max, int link_inc)
{
    // thread i moves one float4 of each array: a warp moves 512 contiguous bytes per array
    const size_t m = n_max;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const float4 a = R.S.pos[i], b = R.S.rot[i], c = R.S.vel[i], d = R.S.acc[i], e = R.S.angvel[i], f = R.S.angacc[i];
        const float4 g = R.prm0[i], h = R.prm1[i], k = R.locked_pos[i], l = R.locked_rot[i];
        const uint32_t id = R.entity_id[i];
        dst[i] = a; dst[m + i] = b; dst[2 * m + i] = c; dst[3 * m + i] = d; dst[4 * m + i] = e; dst[5 * m + i] = f;
        dst[6 * m + i] = g; dst[7 * m + i] = h; dst[8 * m + i] = k; dst[9 * m + i] = l;
        dst_ids[i] = link_inc ? inc_link(id) : id;                                  // server_network_dynamo.cpp:673
    }
}

// ENTITY_UPDATE replay, server_network_dynamo.cpp:396-420: a non-divergent entity takes every
// component of the record; a divergent one only the `upstream` category, which holds no physics
// component (hard_config_cosmos.h:47-58 registers only SpacialInputComponent upstream).  The record is
// addressed to the id it carries: an entity whose id differs is left alone and counted.
// Row of `id` in an ascending id array (the order entities are submitted in: std::set<Entity>, ecs/i_synchro.h:48), hint first.
__device__ __forceinline__ uint32_t find_id(const uint32_t* __restrict__ ids, uint32_t n, uint32_t id, uint32_t hint)
{
    if (hint < n && ids[hint] == id) return hint;
    uint32_t lo = 0, hi = n;
    while (lo < hi) { const uint32_t mid = (lo + hi) >> 1; if (ids[mid] < id) lo = mid + 1; else hi = mid; }
    return (lo < n && ids[lo] == id) ? lo : ICX_NONE;
}

__global__ void __launch_bounds__(256) k_replay(DeviceWorld W, const float4* __restrict__ src, const uint32_t* __restrict__ src_ids,
                                                uint32_t n, uint32_t n_src, uint32_t n_max, int by_id, unsigned long long* __restrict__ mismatch)
{
    const EntityState S = W.in;
    const size_t m = n_max;
    uint32_t bad = 0;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const uint32_t state = S.ts[i] & 0xFFu;
        if (state == ICX_TS_FORKING || state == ICX_TS_FORKED) continue;            // is_divergent, timestream_state.h:41-44
        const uint32_t id = W.entity_id[i];
        uint32_t j;
        if (by_id) {                                                                // the record is addressed to ids: find mine (same row when the two
            j = find_id(src_ids, n_src, id, i);                                     // cosmoses hold the same entities; a stream that has no record for me
            if (j == ICX_NONE) { bad++; continue; }                                 // leaves me alone, as EntityTimestreamMap does)
        } else j = i % n_src;                                                       // one world-line's record shared by all candidates
        S.pos[i] = src[j]; S.rot[i] = src[m + j];                                   // TransformComponent
        const float4 r0 = src[6 * m + j], r1 = src[7 * m + j];
        const float4 o0 = W.prm0[i], o1 = W.prm1[i];
        const uint32_t recf = __float_as_uint(r1.w), oldf = __float_as_uint(o1.w);
        const bool both = (recf & EF_HAS_PHYSICS) && (oldf & EF_HAS_PHYSICS);       // deserialize_data_for does nothing for a missing component
        float4 p0 = make_float4(r0.x, r0.y, r0.z, both ? r0.w : o0.w);
        float4 p1 = o1; uint32_t lockbits = oldf;
        if (both) {                                                                 // PhysicsComponent
            S.vel[i] = src[2 * m + j]; S.acc[i] = src[3 * m + j]; S.angvel[i] = src[4 * m + j]; S.angacc[i] = src[5 * m + j];
            W.locked_pos[i] = src[8 * m + j]; W.locked_rot[i] = src[9 * m + j];
            p1 = r1; lockbits = recf;
        }
        const bool hasP = (oldf & EF_HAS_PHYSICS) != 0;
        p1.w = __uint_as_float(entity_flags(id, hasP, (lockbits & EF_LOCK_ORIGIN) != 0, (lockbits & EF_LOCK_ORIENTATION) != 0,
                                            (lockbits & EF_ASLEEP) != 0, p0.w, oldf));
        W.prm0[i] = p0; W.prm1[i] = p1;
    }
    bad = __reduce_add_sync(0xFFFFFFFFu, bad);
    if (bad && (threadIdx.x & 31u) == 0) atomicAdd(mismatch, (unsigned long long)bad);
}

// ---- sparse ColliderComponent rows ----
__global__ void __launch_bounds__(256) k_delta_count(const uint32_t* __restrict__ c_ver, const uint32_t* __restrict__ seen, uint32_t n, int all,
                                                     uint32_t* __restrict__ out)
{
    uint32_t c = 0;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) c += (all || c_ver[i] != seen[i]) ? 1u : 0u;
    c = __reduce_add_sync(0xFFFFFFFFu, c);
    if (c && (threadIdx.x & 31u) == 0) atomicAdd(out, c);
}
__global__ void __launch_bounds__(256) k_delta_collect(const uint32_t* __restrict__ c_ver, uint32_t* __restrict__ seen, const IcxColliderComponent* __restrict__ C,
                                                       const uint32_t* __restrict__ ids, int link_inc,
                                                       uint32_t n, int all, DeltaRow* __restrict__ dst, uint32_t cap, uint32_t* __restrict__ counter)
{
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const uint32_t v = c_ver[i];
        if (!all && v == seen[i]) continue;
        seen[i] = v;
        const uint32_t at = atomicAdd(counter, 1u);
        if (at < cap) { dst[at].idx = i; dst[at].id = link_inc ? inc_link(ids[i]) : ids[i]; dst[at].C = C[i]; }
    }
}

// one collider row into entity e: AoS mirror (the source of later downloads / records), per-packet constants, flag word
__device__ __forceinline__ void apply_collider_row(const DeviceWorld& W, IcxColliderComponent* __restrict__ MC, uint32_t e, IcxColliderComponent row,
                                                   uint32_t epoch, unsigned long long* __restrict__ mismatch)
{
    const uint32_t npk = W.first_packet[e] >> 30;
    uint32_t act = 0;
    for (int s = 0; s < ICX_COLLIDERS_PER_ENTITY; s++) {
        if (row.colliders[s].isActive) act++;
        row.colliders[s].minParametricValue = MC[e].colliders[s].minParametricValue;   // per-tick scratch (Collider::reset), not state
    }
    if (act != npk) { atomicAdd(mismatch, 1ull); return; }                           // packet structure differs: needs a re-upload
    MC[e] = row;
    const float4 p0 = W.prm0[e]; float4 p1 = W.prm1[e];
    const uint32_t oldf = __float_as_uint(p1.w);
    const uint32_t ef = entity_flags(W.entity_id[e], (oldf & EF_HAS_PHYSICS) != 0, (oldf & EF_LOCK_ORIGIN) != 0, (oldf & EF_LOCK_ORIENTATION) != 0,
                                     (oldf & EF_ASLEEP) != 0, p0.w, collider_flag_bits(&row));
    p1.w = __uint_as_float(ef); W.prm1[e] = p1;
    import_collider_row(W, e, row, p0.x, p0.y, p0.z, ef, false);
    W.c_ver[e] = epoch;
}

// Rows of one slot: applied to non-divergent entities (ComponentCategory::all), kept aside for divergent ones -- the
// reference's records carry every component every tick, so the first record such an entity takes again restores them.
__global__ void __launch_bounds__(128) k_replay_delta(DeviceWorld W, IcxColliderComponent* __restrict__ MC, const DeltaRow* __restrict__ rows,
                                                      const uint32_t* __restrict__ d_n, uint32_t cap, uint32_t epoch,
                                                      IcxColliderComponent* __restrict__ pend_rows, uint8_t* __restrict__ pend_flag,
                                                      unsigned long long* __restrict__ mismatch, int keep_all)
{
    const uint32_t n = min(*d_n, cap);
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const uint32_t e = find_id(W.entity_id, W.n_entities, rows[i].id, rows[i].idx);
        if (e == ICX_NONE) continue;                                                    // a stream nobody here pops (its entity does not exist in this cosmos)
        const uint32_t state = W.in.ts[e] & 0xFFu;
        // keep_all: the record itself is dropped as stale (server_network_dynamo.cpp:391), but the NEXT record the entity takes
        // carries this row unchanged (every record holds every component): kept until then
        if (keep_all || state == ICX_TS_FORKING || state == ICX_TS_FORKED) { pend_rows[e] = rows[i].C; pend_flag[e] = 1; continue; }
        pend_flag[e] = 0;
        apply_collider_row(W, MC, e, rows[i].C, epoch, mismatch);
    }
}
__global__ void __launch_bounds__(256) k_replay_pending(DeviceWorld W, IcxColliderComponent* __restrict__ MC, uint32_t epoch,
                                                        const IcxColliderComponent* __restrict__ pend_rows, uint8_t* __restrict__ pend_flag,
                                                        unsigned long long* __restrict__ mismatch)
{
    for (uint32_t e = blockIdx.x * blockDim.x + threadIdx.x; e < W.n_entities; e += gridDim.x * blockDim.x) {
        if (!pend_flag[e]) continue;
        const uint32_t state = W.in.ts[e] & 0xFFu;
        if (state == ICX_TS_FORKING || state == ICX_TS_FORKED) continue;
        pend_flag[e] = 0;
        apply_collider_row(W, MC, e, pend_rows[e], epoch, mismatch);
    }
}

// transport: the used part of a buffer's collider rows (count on the device), 16 bytes per thread
__global__ void __launch_bounds__(256) k_copy_rows(uint4* __restrict__ dst, const uint4* __restrict__ src, const uint32_t* __restrict__ d_n, uint32_t room)
{
    const uint32_t words = min(*d_n, room) * (uint32_t)(sizeof(DeltaRow) / 16);
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < words; i += gridDim.x * blockDim.x) dst[i] = src[i];
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
float4* slot_state(const IcxTimeline* tl, int s) { return tl->state + (size_t)s * TL_ARRAYS * tl->n_max; }
uint32_t* slot_ids(const IcxTimeline* tl, int s) { return tl->ids + (size_t)s * tl->n_max; }
int stream_blocks(const IcxTimeline* tl, uint32_t n, int per_sm) { return (int)std::min<uint32_t>((n + 255) / 256, (uint32_t)(tl->sms * per_sm)); }

// Per-slot capacity of the sparse collider rows (grown on demand; existing rows are kept).
int ensure_delta_cap(IcxTimeline* tl, uint32_t rows)
{
    if (rows <= tl->delta_cap) return ICX_OK;
    uint32_t cap = std::max<uint32_t>(64, tl->delta_cap);
    while (cap < rows) cap *= 2;
    DeltaRow* fresh = nullptr;
    TCK(cudaMalloc((void**)&fresh, (size_t)tl->depth * cap * sizeof(DeltaRow)));
    if (tl->delta) {
        TCK(cudaDeviceSynchronize());
        TCK(cudaMemcpy2D(fresh, (size_t)cap * sizeof(DeltaRow), tl->delta, (size_t)tl->delta_cap * sizeof(DeltaRow),
                         (size_t)tl->delta_cap * sizeof(DeltaRow), tl->depth, cudaMemcpyDeviceToDevice));
        cudaFree(tl->delta);
    }
    tl->delta = fresh; tl->delta_cap = cap;
    return ICX_OK;
}

// A world that replays collider rows needs somewhere to keep the rows of entities that are divergent at that moment.
int ensure_pending(IcxWorld* w, IcxTimeline* tl)
{
    if (w->pend_flag.ptr) return ICX_OK;
    TCK(w->pend_rows.ensure((size_t)w->cap_entities * sizeof(IcxColliderComponent)));
    TCK(w->pend_flag.ensure(w->cap_entities));
    TCK(cudaMemsetAsync(w->pend_flag.ptr, 0, w->cap_entities, w->stream));
    return ICX_OK;
}

} // namespace

// accessors for icx_nccl.cu (IcxTimeline is opaque outside this file)
int icx_tl_scan_buffers(IcxTimeline* tl, uint32_t n_worldlines, const uint32_t** d_count, const unsigned long long** d_score, const uint8_t** d_ever,
                        cudaStream_t* scan_stream, int* device)
{
    if (!tl->scan_wl || tl->scan_wl != n_worldlines) { tl->error = "no scan of that many world-lines was enqueued on this timeline"; return ICX_E_STATE; }
    *d_count = tl->d_count; *d_score = tl->d_score; *d_ever = tl->d_ever; *scan_stream = tl->scan_stream; *device = tl->device;
    return ICX_OK;
}
void icx_tl_set_error(IcxTimeline* tl, const char* msg) { if (tl) tl->error = msg; }
int icx_tl_device(const IcxTimeline* tl) { return tl->device; }

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
    tl->seq.assign(depth_ticks, 0); tl->delta_n.assign(depth_ticks, 0); tl->slot_ev.assign(depth_ticks, nullptr);
    tl->ev_valid.assign(depth_ticks, 0); tl->ev_behind.assign(depth_ticks, 0); tl->ev_pending.assign(depth_ticks, 0);
    tl->ev_stamp.assign(depth_ticks, 0); tl->ev_seq.assign(depth_ticks, 0);
    cudaSetDevice(device);
    cudaDeviceGetAttribute(&tl->sms, cudaDevAttrMultiProcessorCount, device);
    cudaError_t e = cudaMalloc((void**)&tl->state, (size_t)depth_ticks * TL_ARRAYS * n_max * sizeof(float4));
    if (e == cudaSuccess) e = cudaMalloc((void**)&tl->ids, (size_t)depth_ticks * n_max * sizeof(uint32_t));
    if (e == cudaSuccess) e = cudaMalloc((void**)&tl->seen, (size_t)n_max * sizeof(uint32_t));
    if (e == cudaSuccess) e = cudaMalloc((void**)&tl->d_delta_n, ((size_t)depth_ticks + 1) * sizeof(uint32_t));
    if (e == cudaSuccess) e = cudaMalloc((void**)&tl->d_mismatch, sizeof(unsigned long long));
    if (e == cudaSuccess) e = cudaMalloc((void**)&tl->d_age, 2 * sizeof(uint32_t));
    if (e == cudaSuccess) e = cudaMalloc((void**)&tl->d_ev, (size_t)depth_ticks * sizeof(IcxTimeline::EvBlock));
    if (e == cudaSuccess) e = cudaMallocHost((void**)&tl->h_ev, (size_t)depth_ticks * sizeof(IcxTimeline::EvBlock));
    if (e == cudaSuccess) e = cudaMemset(tl->d_ev, 0, (size_t)depth_ticks * sizeof(IcxTimeline::EvBlock));
    if (e == cudaSuccess) std::memset(tl->h_ev, 0, (size_t)depth_ticks * sizeof(IcxTimeline::EvBlock));
    if (e == cudaSuccess) e = cudaMemset(tl->seen, 0, (size_t)n_max * sizeof(uint32_t));
    if (e == cudaSuccess) e = cudaMemset(tl->d_delta_n, 0, ((size_t)depth_ticks + 1) * sizeof(uint32_t));
    if (e == cudaSuccess) e = cudaMemset(tl->d_mismatch, 0, sizeof(unsigned long long));
    for (uint32_t k = 0; k < depth_ticks && e == cudaSuccess; k++) e = cudaEventCreateWithFlags(&tl->slot_ev[k], cudaEventDisableTiming);
    if (e != cudaSuccess) { icx_timeline_destroy(tl); return e == cudaErrorMemoryAllocation ? ICX_E_NOMEM : ICX_E_CUDA; }
    *out = tl;
    return ICX_OK;
}

void icx_timeline_destroy(IcxTimeline* tl)
{
    if (!tl) return;
    cudaSetDevice(tl->device);
    cudaDeviceSynchronize();
    void* bufs[] = {tl->state, tl->ids, tl->seen, tl->d_delta_n, tl->d_mismatch, tl->delta, tl->d_score, tl->d_count, tl->d_ever, tl->d_age, tl->d_ev, tl->d_fetch};
    for (void* b : bufs) if (b) cudaFree(b);
    if (tl->h_ev) cudaFreeHost(tl->h_ev);
    for (cudaEvent_t e : tl->slot_ev) if (e) cudaEventDestroy(e);
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

int icx_timeline_id_mismatches(IcxTimeline* tl, uint64_t* total)
{
    if (!tl || !total) return ICX_E_INVALID;
    cudaSetDevice(tl->device);
    unsigned long long v = 0;
    TCK(cudaDeviceSynchronize());
    TCK(cudaMemcpy(&v, tl->d_mismatch, sizeof(v), cudaMemcpyDeviceToHost));
    *total = v;
    return ICX_OK;
}

int icx_record(IcxWorld* w, IcxTimeline* tl, uint16_t coherency) { return icx_record_ex(w, tl, coherency, 0); }
int icx_record_at_breakpoint(IcxWorld* w, IcxTimeline* tl, uint16_t coherency)
{
    return icx_record_ex(w, tl, coherency, ICX_RECORD_AT_BREAKPOINT | ICX_RECORD_KEEP_LINK);
}
// TsBreakpointQueue::set_breakpoint_at_begin on every timestream (EntityTimestreamMap, link_timestreams):
// everything recorded so far becomes old history, reachable only through the breakpoint.
int icx_breakpoint_set(IcxTimeline* tl)
{
    if (!tl) return ICX_E_INVALID;
    for (uint32_t s = 0; s < tl->depth; s++) { tl->behind[s] = tl->valid[s]; tl->ev_behind[s] = tl->ev_valid[s]; }
    tl->breakpoint_active = true;
    return ICX_OK;
}
int icx_breakpoint_clear(IcxTimeline* tl)
{
    if (!tl) return ICX_E_INVALID;
    std::fill(tl->behind.begin(), tl->behind.end(), 0); std::fill(tl->ev_behind.begin(), tl->ev_behind.end(), 0);
    tl->breakpoint_active = false;
    return ICX_OK;
}

int icx_record_ex(IcxWorld* w, IcxTimeline* tl, uint16_t coherency, uint32_t flags)
{
    if (!w || !tl) return ICX_E_INVALID;
    const bool at_breakpoint = (flags & ICX_RECORD_AT_BREAKPOINT) != 0;
    if (at_breakpoint && !tl->breakpoint_active) { tl->error = "record_at_breakpoint: no breakpoint set"; return ICX_E_STATE; }   // push_at_breakpoint returns false
    if (w->n_entities > tl->n_max || w->device != tl->device) { tl->error = "record: world does not fit the timeline"; return ICX_E_INVALID; }
    cudaSetDevice(w->device);
    cudaStream_t st = w->stream;
    const int s = slot_of(tl, coherency);
    const uint32_t n = w->n_entities;
    if (n) {
        RecordSrc R; R.S = w->W.in; R.prm0 = w->W.prm0; R.prm1 = w->W.prm1; R.locked_pos = w->W.locked_pos; R.locked_rot = w->W.locked_rot;
        R.entity_id = w->W.entity_id;
        if (tl->profile) cudaEventRecord(tl->pev[0][0], st);
        k_record<<<stream_blocks(tl, n, 16), 256, 0, st>>>(R, slot_state(tl, s), slot_ids(tl, s), n, tl->n_max, (flags & ICX_RECORD_KEEP_LINK) ? 0 : 1);
        if (tl->profile) { cudaEventRecord(tl->pev[0][1], st); tl->pev_used[0] = true; }
        TCK(cudaGetLastError());
        w->launches += 1;
    }
    // ColliderComponents that changed since the last record into this ring (everything when another cosmos takes over the ring)
    int64_t rows = 0;
    if (n && w->have_C && (tl->last_recorder != w || tl->last_epoch != w->c_epoch)) {
        const int all = tl->last_recorder != w ? 1 : 0;
        uint32_t* scratch = tl->d_delta_n + tl->depth;
        uint32_t cnt = n;
        if (!all) {
            TCK(cudaMemsetAsync(scratch, 0, 4, st));
            k_delta_count<<<stream_blocks(tl, n, 8), 256, 0, st>>>(w->W.c_ver, tl->seen, n, 0, scratch);
            TCK(cudaMemcpyAsync(&cnt, scratch, 4, cudaMemcpyDeviceToHost, st));
            TCK(cudaStreamSynchronize(st));                                         // rare path: some collider changed this tick
        }
        if (cnt) { int rc = ensure_delta_cap(tl, cnt); if (rc != ICX_OK) return rc; }
        TCK(cudaMemsetAsync(tl->d_delta_n + s, 0, 4, st));
        if (cnt) {
            k_delta_collect<<<stream_blocks(tl, n, 8), 256, 0, st>>>(w->W.c_ver, tl->seen, (const IcxColliderComponent*)w->mir_C.ptr, w->W.entity_id,
                                                                   (flags & ICX_RECORD_KEEP_LINK) ? 0 : 1, n, all,
                                                                   tl->delta + (size_t)s * tl->delta_cap, tl->delta_cap, tl->d_delta_n + s);
            TCK(cudaGetLastError());
            w->launches += 2;
        }
        rows = cnt;
        tl->last_recorder = w; tl->last_epoch = w->c_epoch;
    } else if (tl->delta_n[s] != 0) TCK(cudaMemsetAsync(tl->d_delta_n + s, 0, 4, st));
    tl->delta_n[s] = rows;
    TCK(cudaEventRecord(tl->slot_ev[s], st));
    tl->stamp[s] = coherency; tl->valid[s] = 1; tl->count[s] = n; tl->seq[s] = tl->next_seq++;
    // push_back lands after the breakpoint (old side), push_at_breakpoint before it (ts_breakpoint_queue.h:44-52, :140-152)
    tl->behind[s] = (tl->breakpoint_active && !at_breakpoint) ? 1 : 0;
    return ICX_OK;
}

int icx_replay(IcxWorld* w, IcxTimeline* tl, uint16_t coherency, int mode, int* applied)
{
    if (!w || !tl) return ICX_E_INVALID;
    if (applied) *applied = 0;
    if (mode != ICX_REPLAY_SERVER && mode != ICX_REPLAY_PARALLEL) return ICX_E_INVALID;
    // pop_future_timestream hands out, oldest first, every record with coherency <= now (cyclic compare, message.h:179-183);
    // the relay then drops those with `coherency < now` -- a PLAIN integer compare (server_network_dynamo.cpp:391), so across
    // the uint16 wrap an older record (65535 with now = 2) is still deserialised, before the current one.  pop_front stops at
    // the breakpoint; pop_at_breakpoint only takes what lies behind it.
    std::vector<int> popped;
    for (uint32_t k = 0; k < tl->depth; k++) {
        if (!tl->valid[k]) continue;
        const uint16_t a = coherency, b = tl->stamp[k];
        const bool le_now = (a >= b && a - b <= 32768) || (a <= b && b - a > 32768);
        if (!le_now) continue;
        const bool behind = tl->breakpoint_active && tl->behind[k];
        if (mode == ICX_REPLAY_SERVER ? behind : !behind) continue;
        tl->valid[k] = 0; tl->behind[k] = 0;
        popped.push_back((int)k);
    }
    if (popped.empty() || mode == ICX_REPLAY_PARALLEL) return ICX_OK;       // parallel_network_dynamo.cpp:78-86: upstream components only
    std::sort(popped.begin(), popped.end(), [&](int x, int y) { return tl->seq[x] < tl->seq[y]; });
    int last = -1;                                                          // position of the newest record that is not dropped as stale
    for (size_t i = 0; i < popped.size(); i++) if (!(tl->stamp[popped[i]] < coherency)) last = (int)i;
    const uint32_t n = w->n_entities;
    if (!n) return ICX_OK;
    cudaSetDevice(w->device);
    cudaStream_t st = w->stream;
    const uint32_t n_src = last >= 0 ? tl->count[popped[last]] : n;
    if (last >= 0 && !n_src) return ICX_OK;
    // a record of one world-line replayed into every candidate is addressed by position; everything else by id
    const bool by_stride = w->W.worldline_stride && n_src == w->W.worldline_stride && n_src != n;
    for (int k : popped) TCK(cudaStreamWaitEvent(st, tl->slot_ev[k], 0));
    // Collider rows first, oldest record first.  The reference's records repeat every component, so the record that is applied
    // in the end carries the rows of ALL earlier ones: rows of records dropped as stale are applied too (when an applied record
    // follows in this very replay) or kept for the next record the entity takes (when none does).
    bool rows_any = false;
    for (int k : popped) rows_any |= tl->delta_n[k] != 0;
    if (w->have_C && !by_stride && (rows_any || (w->pend_any && last >= 0))) {
        int rc = ensure_pending(w, tl); if (rc != ICX_OK) return rc;
        const uint32_t epoch = ++w->c_epoch;
        IcxColliderComponent* MC = (IcxColliderComponent*)w->mir_C.ptr;
        if (w->pend_any && last >= 0) k_replay_pending<<<stream_blocks(tl, n, 8), 256, 0, st>>>(w->W, MC, epoch, (const IcxColliderComponent*)w->pend_rows.ptr,
                                                                                              (uint8_t*)w->pend_flag.ptr, tl->d_mismatch);
        for (size_t i = 0; i < popped.size(); i++) {
            const int k = popped[i];
            if (tl->delta_n[k] == 0) continue;
            const uint32_t guess = tl->delta_n[k] > 0 ? (uint32_t)tl->delta_n[k] : tl->delta_cap;
            k_replay_delta<<<std::max(1, std::min<int>((int)((guess + 127) / 128), tl->sms * 4)), 128, 0, st>>>(
                w->W, MC, tl->delta + (size_t)k * tl->delta_cap, tl->d_delta_n + k, tl->delta_cap, epoch,
                (IcxColliderComponent*)w->pend_rows.ptr, (uint8_t*)w->pend_flag.ptr, tl->d_mismatch, (int)i > last ? 1 : 0);
            w->pend_any = true;         // a row may have been kept (conservative: costs one flag pass per replay)
        }
        TCK(cudaGetLastError());
    }
    if (last < 0) return ICX_OK;                                            // every popped record was stale
    const int last_slot = popped[last];
    if (tl->profile) cudaEventRecord(tl->pev[1][0], st);
    k_replay<<<stream_blocks(tl, n, 16), 256, 0, st>>>(w->W, slot_state(tl, last_slot), slot_ids(tl, last_slot), n, n_src, tl->n_max, by_stride ? 0 : 1, tl->d_mismatch);
    if (tl->profile) { cudaEventRecord(tl->pev[1][1], st); tl->pev_used[1] = true; }
    TCK(cudaGetLastError());
    w->launches += 1;
    if (applied) *applied = 1;
    return ICX_OK;
}

// per world-line: ever[k] |= (count[k] != 0) -- lets a caller scan every tick and read the verdicts back once
__global__ void k_scan_fold(const uint32_t* __restrict__ count, uint8_t* __restrict__ ever, uint32_t n)
{
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k < n && count[k]) ever[k] = 1;
}

int icx_divergence_scan_enqueue(IcxWorld* w, IcxTimeline* tl, uint16_t coherency, uint32_t n_worldlines)
{
    if (!w || !tl || n_worldlines == 0) return ICX_E_INVALID;
    const int s = slot_of(tl, coherency);
    if (!tl->valid[s] || tl->stamp[s] != coherency) { tl->error = "scan: no record stamped with this coherency"; return ICX_E_STATE; }
    const uint32_t n = w->n_entities, n_snap = tl->count[s], stride = w->W.worldline_stride;
    if ((stride ? (n + stride - 1) / stride : 1u) != n_worldlines) { tl->error = "scan: n_worldlines does not match the world"; return ICX_E_INVALID; }
    if (n_snap != n && !(stride && n_snap == stride && stride % SCAN_TILE == 0)) { tl->error = "scan: record must cover the world, or one world-line whose size is a multiple of 1024"; return ICX_E_INVALID; }
    cudaSetDevice(w->device);
    if (tl->cap_wl < n_worldlines) {
        TCK(cudaDeviceSynchronize());
        if (tl->d_score) cudaFree(tl->d_score);
        if (tl->d_count) cudaFree(tl->d_count);
        if (tl->d_ever) cudaFree(tl->d_ever);
        tl->d_score = nullptr; tl->d_count = nullptr; tl->d_ever = nullptr; tl->cap_wl = 0;
        TCK(cudaMalloc((void**)&tl->d_score, (size_t)n_worldlines * 8)); TCK(cudaMalloc((void**)&tl->d_count, (size_t)n_worldlines * 4));
        TCK(cudaMalloc((void**)&tl->d_ever, n_worldlines)); TCK(cudaMemset(tl->d_ever, 0, n_worldlines));
        tl->cap_wl = n_worldlines;
    }
    cudaStream_t st = w->stream;
    TCK(cudaStreamWaitEvent(st, tl->slot_ev[s], 0));
    TCK(cudaMemsetAsync(tl->d_score, 0, (size_t)n_worldlines * 8, st)); TCK(cudaMemsetAsync(tl->d_count, 0, (size_t)n_worldlines * 4, st));
    if (n) {
        const size_t smem = (size_t)2 * SCAN_STAGES * SCAN_TILE * sizeof(float4) + (SCAN_STAGES + 1) * sizeof(uint64_t);
        static bool attr_set[64] = {};
        if (!attr_set[w->device & 63]) { TCK(cudaFuncSetAttribute(k_divergence_scan, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); attr_set[w->device & 63] = true; }
        const uint32_t n_tiles = (n + SCAN_TILE - 1) / SCAN_TILE;
        const int blocks = (int)std::min<uint32_t>(n_tiles, (uint32_t)(tl->sms * SCAN_CTAS));
        if (tl->profile) cudaEventRecord(tl->pev[2][0], st);
        k_divergence_scan<<<blocks, SCAN_TPB, smem, st>>>(w->W.in.pos, slot_state(tl, s), n, n_snap, stride, tl->d_count, tl->d_score);
        if (tl->profile) { cudaEventRecord(tl->pev[2][1], st); tl->pev_used[2] = true; }
        k_scan_fold<<<(n_worldlines + 255) / 256, 256, 0, st>>>(tl->d_count, tl->d_ever, n_worldlines);
        TCK(cudaGetLastError());
        w->launches += 2;
    }
    tl->scan_stream = st; tl->scan_wl = n_worldlines;
    return ICX_OK;
}

int icx_divergence_scan_fetch(IcxTimeline* tl, uint32_t n_worldlines, uint8_t* flags, uint8_t* flags_ever, uint32_t* incongruent, double* score, int reset_ever)
{
    if (!tl || n_worldlines == 0) return ICX_E_INVALID;
    if (!tl->scan_wl || n_worldlines != tl->scan_wl) { tl->error = "scan_fetch: no scan of that many world-lines was enqueued"; return ICX_E_STATE; }
    cudaSetDevice(tl->device);
    cudaStream_t st = tl->scan_stream;
    std::vector<unsigned long long> hs(n_worldlines); std::vector<uint32_t> hc(n_worldlines);
    TCK(cudaMemcpyAsync(hs.data(), tl->d_score, (size_t)n_worldlines * 8, cudaMemcpyDeviceToHost, st));
    TCK(cudaMemcpyAsync(hc.data(), tl->d_count, (size_t)n_worldlines * 4, cudaMemcpyDeviceToHost, st));
    if (flags_ever) TCK(cudaMemcpyAsync(flags_ever, tl->d_ever, n_worldlines, cudaMemcpyDeviceToHost, st));
    if (reset_ever) TCK(cudaMemsetAsync(tl->d_ever, 0, n_worldlines, st));
    TCK(cudaStreamSynchronize(st));
    for (uint32_t k = 0; k < n_worldlines; k++) {
        if (flags) flags[k] = hc[k] ? 1 : 0;
        if (incongruent) incongruent[k] = hc[k];
        if (score) score[k] = (double)hs[k] / 65536.0;
    }
    return ICX_OK;
}

int icx_divergence_scan(IcxWorld* w, IcxTimeline* tl, uint16_t coherency, uint32_t n_worldlines,
                        uint8_t* flags, uint32_t* incongruent, double* score)
{
    int rc = icx_divergence_scan_enqueue(w, tl, coherency, n_worldlines);
    if (rc != ICX_OK) return rc;
    return icx_divergence_scan_fetch(tl, n_worldlines, flags, nullptr, incongruent, score, 0);
}

// ---- the scan without any allocation (no counterpart in the reference) ----
// For callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops): the per-world-line accumulators the
// scan kernel adds into live in a file-scope device array instead of the timeline's lazily grown buffers.  Every GPU that
// loads the library holds the array (96 KB) until exit; all calls share it, so calls that overlap on one GPU race.
static __device__ __align__(256) unsigned char g_scan_scratch[ICX_SCAN_STATIC_MAX_WORLDLINES * 12];   // [MAX] u64 scores, then [MAX] u32 counts
static void* g_scan_scratch_ptr[64] = {};       // device address of the array per device (the name above is only a host-side stand-in)

int icx_divergence_scan_static(IcxWorld* w, IcxTimeline* tl, uint16_t coherency, uint32_t n_worldlines,
                               uint8_t* flags, uint32_t* incongruent, double* score)
{
    if (!w || !tl || n_worldlines == 0) return ICX_E_INVALID;
    if (n_worldlines > ICX_SCAN_STATIC_MAX_WORLDLINES) { tl->error = "scan_static: more world-lines than ICX_SCAN_STATIC_MAX_WORLDLINES (the static scratch holds 8192)"; return ICX_E_CAPACITY; }
    if (w->device < 0 || w->device >= 64) { tl->error = "scan_static: device index beyond the address table"; return ICX_E_INVALID; }
    const int s = slot_of(tl, coherency);
    if (!tl->valid[s] || tl->stamp[s] != coherency) { tl->error = "scan: no record stamped with this coherency"; return ICX_E_STATE; }
    const uint32_t n = w->n_entities, n_snap = tl->count[s], stride = w->W.worldline_stride;
    if ((stride ? (n + stride - 1) / stride : 1u) != n_worldlines) { tl->error = "scan: n_worldlines does not match the world"; return ICX_E_INVALID; }
    if (n_snap != n && !(stride && n_snap == stride && stride % SCAN_TILE == 0)) { tl->error = "scan: record must cover the world, or one world-line whose size is a multiple of 1024"; return ICX_E_INVALID; }
    TCK(cudaSetDevice(w->device));
    const size_t smem = (size_t)2 * SCAN_STAGES * SCAN_TILE * sizeof(float4) + (SCAN_STAGES + 1) * sizeof(uint64_t);
    if (!g_scan_scratch_ptr[w->device]) {                                   // set-up, once per device, with that device current
        void* ptr = nullptr;
        TCK(cudaGetSymbolAddress(&ptr, g_scan_scratch));
        TCK(cudaFuncSetAttribute(k_divergence_scan, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        g_scan_scratch_ptr[w->device] = ptr;
    }
    unsigned long long* d_score = (unsigned long long*)g_scan_scratch_ptr[w->device];
    uint32_t* d_count = (uint32_t*)((unsigned char*)g_scan_scratch_ptr[w->device] + (size_t)ICX_SCAN_STATIC_MAX_WORLDLINES * 8);
    cudaStream_t st = w->stream;
    TCK(cudaStreamWaitEvent(st, tl->slot_ev[s], 0));
    // the array keeps the last call's sums: both accumulators start at zero
    TCK(cudaMemsetAsync(d_score, 0, (size_t)n_worldlines * 8, st)); TCK(cudaMemsetAsync(d_count, 0, (size_t)n_worldlines * 4, st));
    if (n) {
        const uint32_t n_tiles = (n + SCAN_TILE - 1) / SCAN_TILE;
        const int blocks = (int)std::min<uint32_t>(n_tiles, (uint32_t)(tl->sms * SCAN_CTAS));
        k_divergence_scan<<<blocks, SCAN_TPB, smem, st>>>(w->W.in.pos, slot_state(tl, s), n, n_snap, stride, d_count, d_score);
        TCK(cudaGetLastError());
        w->launches += 1;
    }
    // results straight into the caller's arrays (no host allocation either): the fixed-point sums land in `score` and are
    // converted in place, the counts in `incongruent` -- or, without it, pass through a bounded buffer on the stack
    uint32_t local_count[256];
    if (score) TCK(cudaMemcpyAsync(score, d_score, (size_t)n_worldlines * 8, cudaMemcpyDeviceToHost, st));
    if (incongruent) TCK(cudaMemcpyAsync(incongruent, d_count, (size_t)n_worldlines * 4, cudaMemcpyDeviceToHost, st));
    TCK(cudaStreamSynchronize(st));
    if (score) for (uint32_t k = 0; k < n_worldlines; k++) { unsigned long long fx; std::memcpy(&fx, &score[k], 8); score[k] = (double)fx / 65536.0; }
    if (flags && incongruent) for (uint32_t k = 0; k < n_worldlines; k++) flags[k] = incongruent[k] ? 1 : 0;
    else if (flags) for (uint32_t k0 = 0; k0 < n_worldlines; k0 += 256) {
        const uint32_t c = std::min<uint32_t>(256u, n_worldlines - k0);
        TCK(cudaMemcpy(local_count, d_count + k0, (size_t)c * 4, cudaMemcpyDeviceToHost));
        for (uint32_t k = 0; k < c; k++) flags[k0 + k] = local_count[k] ? 1 : 0;
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
        int sms = 148; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, w->device);
        const int blocks = (int)std::min<uint32_t>((n + 255) / 256, (uint32_t)(sms * 8));
        k_fork_age<<<blocks, 256, 0, st>>>(w->W.in.ts, n, (uint32_t)w->W.coherency, d);
        TCK(cudaGetLastError());
        w->launches += 1;
    }
    if (!n_promoted && !resolution_needed) return ICX_OK;      // asynchronous use: the promotions are all the caller wants
    uint32_t h[2] = {0, 0};
    TCK(cudaMemcpyAsync(h, d, 8, cudaMemcpyDeviceToHost, st));
    TCK(cudaStreamSynchronize(st));
    if (n_promoted) *n_promoted = h[0];
    if (resolution_needed) *resolution_needed = h[1] ? 1 : 0;
    return ICX_OK;
}

// The snapshot stream between adjacent timeslices (future -> past, server_network_dynamo.cpp:642-679
// push_past_timestream) when the two timeslices live on different GPUs / processes: the slot stamped
// `coherency` packed into a caller-owned DEVICE buffer (what a ncclSend / cudaMemcpyPeer moves) as
//   [10][n] float4 | [n] uint32 ids | {uint32 rows, 0, 0, 0} | rows x 176-byte collider rows
// and its inverse on the receiving side, which lands in the ring exactly like icx_record (push_back).
// Both are asynchronous on `stream` (cudaStream_t, NULL = the default stream): the export waits for the
// record on the recording world's stream through an event, a replay / scan on another stream waits for the
// import the same way -- no host synchronisation unless the number of collider rows is unknown to the host.
// ---- one recorded entity out of a slot (the components an ENTITY_CREATED / JUMP_ARRIVAL handler needs, see include/icx.h) ----
struct FetchOut { float4 v[TL_ARRAYS]; uint32_t found, has_row, pad[2]; IcxColliderComponent C; };
__global__ void __launch_bounds__(256) k_entity_fetch(const float4* __restrict__ state, const uint32_t* __restrict__ ids, uint32_t n, uint32_t n_max,
                                                      const DeltaRow* __restrict__ rows, const uint32_t* __restrict__ d_n, uint32_t cap, uint32_t id, FetchOut* out)
{
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        const uint32_t j = find_id(ids, n, id, ICX_NONE);
        out->found = j != ICX_NONE ? 1u : 0u;
        if (j != ICX_NONE) for (int k = 0; k < TL_ARRAYS; k++) out->v[k] = state[(size_t)k * n_max + j];
    }
    const uint32_t n_rows = rows ? min(*d_n, cap) : 0u;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n_rows; i += gridDim.x * blockDim.x) if (rows[i].id == id) { out->C = rows[i].C; out->has_row = 1u; }
}

// ---- event records ----
// the events block of slot s belongs to the tick stamped `coherency`: a block left by an older tick is emptied
static void ev_claim(IcxTimeline* tl, int s, uint16_t coherency)
{
    if (tl->ev_valid[s] && tl->ev_stamp[s] == coherency) return;
    tl->h_ev[s].n_events = 0; tl->ev_valid[s] = 0; tl->ev_pending[s] = 0; tl->ev_behind[s] = 0; tl->ev_stamp[s] = coherency;
}

int icx_timeline_event_push(IcxTimeline* tl, uint16_t coherency, const IcxTimelineEvent* events, uint32_t count)
{
    if (!tl || (!events && count)) return ICX_E_INVALID;
    if (!count) return ICX_OK;
    const int s = slot_of(tl, coherency);
    cudaSetDevice(tl->device);
    if (tl->ev_pending[s]) { TCK(cudaEventSynchronize(tl->slot_ev[s])); tl->ev_pending[s] = 0; }
    ev_claim(tl, s, coherency);
    IcxTimeline::EvBlock& b = tl->h_ev[s];
    if (b.n_events + count > ICX_TIMELINE_EVENTS_PER_TICK) { tl->error = "event_push: more than ICX_TIMELINE_EVENTS_PER_TICK records in one tick"; return ICX_E_CAPACITY; }
    for (uint32_t i = 0; i < count; i++) { IcxTimelineEvent e = events[i]; e.coherency = coherency; e.reserved = 0; b.ev[b.n_events++] = e; }
    if (!tl->ev_valid[s]) { tl->ev_valid[s] = 1; tl->ev_seq[s] = tl->next_seq++; tl->ev_behind[s] = 0; }   // (pushed in front of a breakpoint, like a record at it)
    TCK(cudaMemcpy(tl->d_ev + s, &b, sizeof(b), cudaMemcpyHostToDevice));   // rare and 1 KB: synchronous, so any later export sees it
    return ICX_OK;
}

int icx_timeline_events_pop(IcxTimeline* tl, uint16_t now, int mode, IcxTimelineEvent* out, size_t cap, size_t* n_out)
{
    if (!tl) return ICX_E_INVALID;
    if (n_out) *n_out = 0;
    if (mode != ICX_REPLAY_SERVER && mode != ICX_REPLAY_PARALLEL) return ICX_E_INVALID;
    std::vector<int> popped;
    for (uint32_t k = 0; k < tl->depth; k++) {
        if (!tl->ev_valid[k]) continue;
        const uint16_t a = now, b = tl->ev_stamp[k];
        const bool le_now = (a >= b && a - b <= 32768) || (a <= b && b - a > 32768);      // message.h:179-183
        if (!le_now) continue;
        const bool behind = tl->breakpoint_active && tl->ev_behind[k];
        if (mode == ICX_REPLAY_SERVER ? behind : !behind) continue;
        popped.push_back((int)k);
    }
    std::sort(popped.begin(), popped.end(), [&](int x, int y) { return tl->ev_seq[x] < tl->ev_seq[y]; });
    cudaSetDevice(tl->device);
    size_t n = 0;
    for (int k : popped) {
        if (tl->ev_pending[k]) { TCK(cudaEventSynchronize(tl->slot_ev[k])); tl->ev_pending[k] = 0; }
        const IcxTimeline::EvBlock& b = tl->h_ev[k];
        tl->ev_valid[k] = 0; tl->ev_behind[k] = 0;
        if (tl->ev_stamp[k] < now) continue;                                            // server_network_dynamo.cpp:391 (plain compare)
        for (uint32_t i = 0; i < std::min<uint32_t>(b.n_events, ICX_TIMELINE_EVENTS_PER_TICK); i++) { if (out && n < cap) out[n] = b.ev[i]; n++; }
    }
    if (n_out) *n_out = n;
    return ICX_OK;
}

int icx_timeline_entity_fetch(IcxTimeline* tl, uint16_t coherency, IcxEntity id, IcxTransform* T, IcxPhysics* P, uint8_t* has_physics,
                              IcxColliderComponent* C, int* found)
{
    if (!tl || !found) return ICX_E_INVALID;
    *found = 0;
    const int s = slot_of(tl, coherency);
    if (!(tl->valid[s] && tl->stamp[s] == coherency)) return ICX_OK;                // no record of that tick (any more)
    cudaSetDevice(tl->device);
    if (!tl->d_fetch) TCK(cudaMalloc(&tl->d_fetch, sizeof(FetchOut)));
    TCK(cudaEventSynchronize(tl->slot_ev[s]));
    TCK(cudaMemset(tl->d_fetch, 0, sizeof(FetchOut)));
    k_entity_fetch<<<tl->sms * 2, 256>>>(slot_state(tl, s), slot_ids(tl, s), tl->count[s], tl->n_max,
                               tl->delta_n[s] != 0 && tl->delta ? tl->delta + (size_t)s * tl->delta_cap : nullptr, tl->d_delta_n + s, tl->delta_cap, id, (FetchOut*)tl->d_fetch);
    FetchOut h;
    TCK(cudaMemcpy(&h, tl->d_fetch, sizeof(h), cudaMemcpyDeviceToHost));
    if (!h.found) return ICX_OK;
    *found = 1 | (h.has_row ? 2 : 0);
    // the slot's SoA words back into component rows (the inverse of k_import; same mapping as k_export, icx_convert.cu)
    const float4 pos = h.v[0], rot = h.v[1], vel = h.v[2], acc = h.v[3], av = h.v[4], aa = h.v[5], p0 = h.v[6], p1 = h.v[7], lp = h.v[8], lr = h.v[9];
    uint32_t ef; std::memcpy(&ef, &p1.w, 4);
    if (T) {
        std::memset(T, 0, sizeof(*T));
        T->origin[0] = pos.x; T->origin[1] = pos.y; T->origin[2] = pos.z;
        T->orientation[0] = rot.x; T->orientation[1] = rot.y; T->orientation[2] = rot.z; T->orientation[3] = rot.w;
        T->scale[0] = p0.x; T->scale[1] = p0.y; T->scale[2] = p0.z;
    }
    if (has_physics) *has_physics = (ef & EF_HAS_PHYSICS) ? 1 : 0;
    if (P) {
        std::memset(P, 0, sizeof(*P));
        P->velocity[0] = vel.x; P->velocity[1] = vel.y; P->velocity[2] = vel.z;
        P->acceleration[0] = acc.x; P->acceleration[1] = acc.y; P->acceleration[2] = acc.z;
        P->angularVelocity[0] = av.x; P->angularVelocity[1] = av.y; P->angularVelocity[2] = av.z;
        P->angularAcceleration[0] = aa.x; P->angularAcceleration[1] = aa.y; P->angularAcceleration[2] = aa.z;
        P->linearDrag = p1.x; P->angularDrag = p1.y; P->collisionAngularDrag = p1.z; P->collisionLinearDrag = lp.w; P->mass = p0.w;
        P->lockOrigin = (ef & EF_LOCK_ORIGIN) ? 1 : 0; P->lockOrientation = (ef & EF_LOCK_ORIENTATION) ? 1 : 0; P->isAsleep = (ef & EF_ASLEEP) ? 1 : 0;
        P->lockedOrigin[0] = lp.x; P->lockedOrigin[1] = lp.y; P->lockedOrigin[2] = lp.z;
        P->lockedOrientation[0] = lr.x; P->lockedOrientation[1] = lr.y; P->lockedOrientation[2] = lr.z; P->lockedOrientation[3] = lr.w;
    }
    if (C && h.has_row) *C = h.C;
    return ICX_OK;
}

int icx_timeline_slot_export(IcxTimeline* tl, uint16_t coherency, void* device_buffer, size_t capacity_bytes, uint32_t* n_out, void* stream)
{
    if (!tl || !device_buffer) return ICX_E_INVALID;
    const int s = slot_of(tl, coherency);
    if (!(tl->valid[s] && tl->stamp[s] == coherency)) { tl->error = "slot_export: no record with this coherency"; return ICX_E_STATE; }
    const uint32_t n = tl->count[s];
    if (n_out) *n_out = n;
    cudaSetDevice(tl->device);
    cudaStream_t st = (cudaStream_t)stream;
    TCK(cudaStreamWaitEvent(st, tl->slot_ev[s], 0));
    int64_t rows = tl->delta_n[s];
    if (rows < 0) {                                     // an imported slot: ask the device
        uint32_t r = 0;
        TCK(cudaMemcpyAsync(&r, tl->d_delta_n + s, 4, cudaMemcpyDeviceToHost, st)); TCK(cudaStreamSynchronize(st));
        rows = std::min<uint32_t>(r, tl->delta_cap);
    }
    const size_t dense = ICX_SNAPSHOT_DENSE_BYTES(n);
    if (dense + ICX_SNAPSHOT_HEADER_BYTES + (size_t)rows * sizeof(DeltaRow) > capacity_bytes) { tl->error = "slot_export: buffer too small"; return ICX_E_CAPACITY; }
    char* out = (char*)device_buffer;
    if (n) {
        TCK(cudaMemcpy2DAsync(out, (size_t)n * 16, slot_state(tl, s), (size_t)tl->n_max * 16, (size_t)n * 16, TL_ARRAYS, cudaMemcpyDeviceToDevice, st));
        TCK(cudaMemcpyAsync(out + (size_t)n * 16 * TL_ARRAYS, slot_ids(tl, s), (size_t)n * 4, cudaMemcpyDeviceToDevice, st));
    }
    if (tl->ev_valid[s] && tl->ev_stamp[s] == coherency) TCK(cudaMemcpyAsync(out + dense, tl->d_ev + s, ICX_SNAPSHOT_HEADER_BYTES, cudaMemcpyDeviceToDevice, st));
    else TCK(cudaMemsetAsync(out + dense, 0, ICX_SNAPSHOT_HEADER_BYTES, st));
    TCK(cudaMemcpyAsync(out + dense, tl->d_delta_n + s, 4, cudaMemcpyDeviceToDevice, st));
    if (rows) TCK(cudaMemcpyAsync(out + dense + ICX_SNAPSHOT_HEADER_BYTES, tl->delta + (size_t)s * tl->delta_cap, (size_t)rows * sizeof(DeltaRow), cudaMemcpyDeviceToDevice, st));
    return ICX_OK;
}

int icx_timeline_slot_import(IcxTimeline* tl, uint16_t coherency, uint32_t n, const void* device_buffer, size_t buffer_bytes, void* stream)
{
    if (!tl || (!device_buffer && n)) return ICX_E_INVALID;
    if (n > tl->n_max) { tl->error = "slot_import: more entities than the timeline holds"; return ICX_E_CAPACITY; }
    const size_t dense = ICX_SNAPSHOT_DENSE_BYTES(n);
    if (buffer_bytes < dense + ICX_SNAPSHOT_HEADER_BYTES) { tl->error = "slot_import: buffer shorter than the dense snapshot + header"; return ICX_E_INVALID; }
    const int s = slot_of(tl, coherency);
    cudaSetDevice(tl->device);
    cudaStream_t st = (cudaStream_t)stream;
    const char* in = (const char*)device_buffer;
    // the row count stays on the device; room for as many rows as the buffer could hold
    const uint32_t room = (uint32_t)std::min<size_t>((buffer_bytes - dense - ICX_SNAPSHOT_HEADER_BYTES) / sizeof(DeltaRow), tl->n_max);
    if (room) { int rc = ensure_delta_cap(tl, room); if (rc != ICX_OK) return rc; }
    if (n) {
        TCK(cudaMemcpy2DAsync(slot_state(tl, s), (size_t)tl->n_max * 16, in, (size_t)n * 16, (size_t)n * 16, TL_ARRAYS, cudaMemcpyDeviceToDevice, st));
        TCK(cudaMemcpyAsync(slot_ids(tl, s), in + (size_t)n * 16 * TL_ARRAYS, (size_t)n * 4, cudaMemcpyDeviceToDevice, st));
    }
    TCK(cudaMemcpyAsync(tl->d_delta_n + s, in + dense, 4, cudaMemcpyDeviceToDevice, st));
    if (tl->ev_pending[s]) { TCK(cudaEventSynchronize(tl->slot_ev[s])); tl->ev_pending[s] = 0; }
    TCK(cudaMemcpyAsync(tl->d_ev + s, in + dense, ICX_SNAPSHOT_HEADER_BYTES, cudaMemcpyDeviceToDevice, st));
    TCK(cudaMemcpyAsync(tl->h_ev + s, in + dense, ICX_SNAPSHOT_HEADER_BYTES, cudaMemcpyDeviceToHost, st));      // pinned: read by events_pop after slot_ev
    tl->ev_valid[s] = 1; tl->ev_pending[s] = 1; tl->ev_stamp[s] = coherency; tl->ev_seq[s] = tl->next_seq++;
    tl->ev_behind[s] = tl->breakpoint_active ? 1 : 0;
    if (room) k_copy_rows<<<std::min<int>(tl->sms * 2, (int)((room * 11u + 255u) / 256u)), 256, 0, st>>>(
                  (uint4*)(tl->delta + (size_t)s * tl->delta_cap), (const uint4*)(in + dense + ICX_SNAPSHOT_HEADER_BYTES), tl->d_delta_n + s, room);
    TCK(cudaEventRecord(tl->slot_ev[s], st));
    tl->delta_n[s] = room ? -1 : 0;
    tl->stamp[s] = coherency; tl->valid[s] = 1; tl->count[s] = n; tl->seq[s] = tl->next_seq++;
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
