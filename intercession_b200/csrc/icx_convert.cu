// AoS <-> SoA boundary kernels.  The caller's arrays are the reference's own component structs
// (TransformComponent 40 B, PhysicsComponent 108 B, ColliderComponent 168 B); a device-resident
// AoS mirror keeps their constant fields so a download can hand back complete structs.
#include "icx_world.cuh"
#include "icx_host.h"
#include "icx_math.cuh"
#include "icx_rows.cuh"

namespace icx {

// One thread per listed entity.  src_* (optional) are compact arrays indexed by i and are first
// copied into the mirror at e = index ? index[i] : i; the SoA parts fed by each given array are
// then rebuilt.  full != 0: every SoA part is rebuilt (first upload).
__global__ void __launch_bounds__(128) k_import(DeviceWorld W, Mirror M, const uint32_t* __restrict__ index,
                                                const uint32_t* src_ids, const IcxTransform* src_T, const IcxPhysics* src_P,
                                                const IcxColliderComponent* src_C, const uint8_t* src_hasP,
                                                const uint8_t* src_ts, const uint16_t* src_tc,
                                                uint32_t count, int full, int have_C, uint32_t epoch)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    const uint32_t e = index ? index[i] : i;
    if (src_ids && src_ids != M.ids) M.ids[e] = src_ids[i];
    if (src_T && src_T != M.T) M.T[e] = src_T[i];
    if (src_P && src_P != M.P) M.P[e] = src_P[i];
    if (src_C && src_C != M.C) M.C[e] = src_C[i];
    if (full) M.hasP[e] = src_hasP ? src_hasP[i] : (uint8_t)(M.P != nullptr);
    const bool hasP = M.hasP[e] != 0;
    const uint32_t id = M.ids[e];
    const IcxPhysics* P = hasP ? &M.P[e] : nullptr;
    const IcxColliderComponent* C = have_C ? &M.C[e] : nullptr;

    if (full) {
        W.entity_id[e] = id;
        const uint32_t st = src_ts ? src_ts[i] : 0u, tc = src_tc ? src_tc[i] : 0u;
        W.in.ts[e] = st | (tc << 16);
    }
    if (full || src_T) {
        const IcxTransform& t = M.T[e];
        W.in.pos[e] = make_float4(t.origin[0], t.origin[1], t.origin[2], 0.0f);
        W.in.rot[e] = make_float4(t.orientation[0], t.orientation[1], t.orientation[2], t.orientation[3]);
    }
    if (full || src_P) {
        if (hasP) {
            W.in.vel[e] = make_float4(P->velocity[0], P->velocity[1], P->velocity[2], 0.0f);
            W.in.acc[e] = make_float4(P->acceleration[0], P->acceleration[1], P->acceleration[2], 0.0f);
            W.in.angvel[e] = make_float4(P->angularVelocity[0], P->angularVelocity[1], P->angularVelocity[2], 0.0f);
            W.in.angacc[e] = make_float4(P->angularAcceleration[0], P->angularAcceleration[1], P->angularAcceleration[2], 0.0f);
            W.locked_pos[e] = make_float4(P->lockedOrigin[0], P->lockedOrigin[1], P->lockedOrigin[2], P->collisionLinearDrag);
            W.locked_rot[e] = make_float4(P->lockedOrientation[0], P->lockedOrientation[1], P->lockedOrientation[2], P->lockedOrientation[3]);
        } else {
            const float4 z = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
            W.in.vel[e] = z; W.in.acc[e] = z; W.in.angvel[e] = z; W.in.angacc[e] = z;
        }
    }
    // Constants: the SoA is the source of truth (a replay overwrites it without touching the mirror), so only the parts fed by
    // a GIVEN array are rebuilt; the flag word mixes all three components and takes the missing inputs from its old value.
    {
        const float4 old0 = full ? make_float4(1.0f, 1.0f, 1.0f, 0.0f) : W.prm0[e];
        const float4 old1 = full ? make_float4(0.0f, 0.0f, 0.0f, 0.0f) : W.prm1[e];
        const uint32_t oldf = __float_as_uint(old1.w);
        const bool newT = full || src_T, newP = full || src_P, newC = full || src_C;
        float4 p0 = old0;
        if (newT) { const IcxTransform& t = M.T[e]; p0.x = t.scale[0]; p0.y = t.scale[1]; p0.z = t.scale[2]; }
        if (newP) p0.w = hasP ? P->mass : 0.0f;
        W.prm0[e] = p0;
        const bool lockO = newP ? (hasP && P->lockOrigin) : (oldf & EF_LOCK_ORIGIN) != 0;
        const bool lockR = newP ? (hasP && P->lockOrientation) : (oldf & EF_LOCK_ORIENTATION) != 0;
        const bool asleep = newP ? (hasP && P->isAsleep) : (oldf & EF_ASLEEP) != 0;
        const uint32_t cbits = newC ? collider_flag_bits(C) : oldf;
        const uint32_t ef = entity_flags(id, hasP, lockO, lockR, asleep, p0.w, cbits);
        float4 p1 = old1;
        if (newP) { p1.x = hasP ? P->linearDrag : 0.0f; p1.y = hasP ? P->angularDrag : 0.0f; p1.z = hasP ? P->collisionAngularDrag : 0.0f; }
        p1.w = __uint_as_float(ef);
        W.prm1[e] = p1;
        if (newC) W.c_ver[e] = epoch;
        if (C && (newC || src_T)) import_collider_row(W, e, *C, p0.x, p0.y, p0.z, ef, newC);
    }
}

// Patch the mutable fields of the mirror from the current state, then the host copies it out.
__global__ void __launch_bounds__(128) k_export(DeviceWorld W, Mirror M, int have_T, int have_P, int have_C,
                                                uint8_t* ts_out, uint16_t* tc_out, uint32_t n)
{
    const uint32_t e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n) return;
    if (have_T) {
        const float4 p = W.in.pos[e], r = W.in.rot[e];
        IcxTransform& t = M.T[e];
        t.origin[0] = p.x; t.origin[1] = p.y; t.origin[2] = p.z;
        t.orientation[0] = r.x; t.orientation[1] = r.y; t.orientation[2] = r.z; t.orientation[3] = r.w;
        const float4 sm = W.prm0[e];                       // the scale is device state too (icx_projectiles_fixed_update pulses it)
        t.scale[0] = sm.x; t.scale[1] = sm.y; t.scale[2] = sm.z;
    }
    if (have_P && M.hasP[e]) {
        const float4 v = W.in.vel[e], a = W.in.acc[e], w = W.in.angvel[e], al = W.in.angacc[e];
        IcxPhysics& P = M.P[e];
        P.velocity[0] = v.x; P.velocity[1] = v.y; P.velocity[2] = v.z;
        P.acceleration[0] = a.x; P.acceleration[1] = a.y; P.acceleration[2] = a.z;
        P.angularVelocity[0] = w.x; P.angularVelocity[1] = w.y; P.angularVelocity[2] = w.z;
        P.angularAcceleration[0] = al.x; P.angularAcceleration[1] = al.y; P.angularAcceleration[2] = al.z;
        // the constants live in the SoA too (a replay overwrites them there, server_network_dynamo.cpp:408-412)
        const float4 p0 = W.prm0[e], p1 = W.prm1[e], lp = W.locked_pos[e], lr = W.locked_rot[e];
        const uint32_t ef = __float_as_uint(p1.w);
        P.linearDrag = p1.x; P.angularDrag = p1.y; P.collisionAngularDrag = p1.z; P.collisionLinearDrag = lp.w; P.mass = p0.w;
        P.lockOrigin = (ef & EF_LOCK_ORIGIN) ? 1 : 0; P.lockOrientation = (ef & EF_LOCK_ORIENTATION) ? 1 : 0; P.isAsleep = (ef & EF_ASLEEP) ? 1 : 0;
        P.lockedOrigin[0] = lp.x; P.lockedOrigin[1] = lp.y; P.lockedOrigin[2] = lp.z;
        P.lockedOrientation[0] = lr.x; P.lockedOrientation[1] = lr.y; P.lockedOrientation[2] = lr.z; P.lockedOrientation[3] = lr.w;
    }
    if (have_C) {
        const uint32_t fp = W.first_packet[e];
        const uint32_t npk = fp >> 30, p0 = fp & 0x3FFFFFFFu;
        uint32_t k = 0;
        for (int s = 0; s < ICX_COLLIDERS_PER_ENTITY && k < npk; s++) {
            IcxCollider& c = M.C[e].colliders[s];
            if (!c.isActive) continue;
            c.minParametricValue = W.pk_minparam[p0 + k]; k++;
        }
    }
    const uint32_t ts = W.in.ts[e];
    if (ts_out) ts_out[e] = (uint8_t)(ts & 0xFFu);
    if (tc_out) tc_out[e] = (uint16_t)(ts >> 16);
}

// The 76 bytes per entity a tick writes, as AoS rows.  A block stages its 256 rows (19 456 contiguous bytes) in shared memory and
// writes them out with full-width coalesced stores.
__global__ void __launch_bounds__(256) k_motion_export(DeviceWorld W, float* __restrict__ out, uint32_t n)
{
    __shared__ float rows[256 * 19];
    const uint32_t base = blockIdx.x * 256u, e = base + threadIdx.x;
    if (e < n) {
        const float4 p = W.in.pos[e], r = W.in.rot[e], v = W.in.vel[e], a = W.in.acc[e], w = W.in.angvel[e], al = W.in.angacc[e];
        float* o = rows + threadIdx.x * 19;                                         // stride 19 words: conflict-free
        o[0] = p.x; o[1] = p.y; o[2] = p.z; o[3] = r.x; o[4] = r.y; o[5] = r.z; o[6] = r.w;
        o[7] = v.x; o[8] = v.y; o[9] = v.z; o[10] = a.x; o[11] = a.y; o[12] = a.z;
        o[13] = w.x; o[14] = w.y; o[15] = w.z; o[16] = al.x; o[17] = al.y; o[18] = al.z;
    }
    __syncthreads();
    const uint32_t m = min(256u, n - base) * 19u;
    float* dst = out + (size_t)base * 19;
    for (uint32_t i = threadIdx.x; i < m; i += 256u) dst[i] = rows[i];
}
void launch_motion_export(const DeviceWorld& W, IcxMotionState* out, uint32_t n, cudaStream_t st)
{
    if (n) k_motion_export<<<(n + 255) / 256, 256, 0, st>>>(W, reinterpret_cast<float*>(out), n);
}

// ENTITY_UPDATE messages in the reference's wire layout (include/icx.h "wire / record format"): one thread per 4-byte word of
// a message, so that a warp writes 128 contiguous bytes.  The mirror rows are current (k_export ran just before).
__global__ void __launch_bounds__(256) k_wire_messages(DeviceWorld W, Mirror M, IcxWireLayout lay, uint32_t coherency, int have_C, uint32_t* __restrict__ out, uint32_t n)
{
    const uint32_t WPM = ICX_WIRE_MAX_BYTES / 4;                                   // words per message slot
    const size_t total = (size_t)n * WPM;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const uint32_t e = (uint32_t)(i / WPM), wd = (uint32_t)(i % WPM);
        const bool hasP = M.hasP[e] != 0;
        const uint32_t nC = have_C ? 42u : 0u, nP = hasP ? 27u : 0u, nT = 10u;      // words: 168 / 108 / 40 bytes
        const uint32_t body = nC + nP + nT + 6u;
        uint32_t v = 0u;
        if (wd == 0u) v = lay.event_id;
        else if (wd == 1u) v = coherency & 0xFFFFu;                                  // uint16 coherency + 2 bytes of padding
        else if (wd == 2u) v = body * 4u;                                            // header.size
        else {
            uint32_t b = wd - 3u;
            if (b < nC) v = reinterpret_cast<const uint32_t*>(&M.C[e])[b];
            else if ((b -= nC) < nP) v = reinterpret_cast<const uint32_t*>(&M.P[e])[b];
            else if ((b -= nP) < nT) v = reinterpret_cast<const uint32_t*>(&M.T[e])[b];
            else if ((b -= nT) < 6u) {
                // ENTITY_UPDATE_params: Entity (uint16) @0, Signature (bitset<32> = unsigned long) @8, ComponentCategory @16
                uint32_t id = W.entity_id[e];
                if (lay.increment_link && (id & 0xFu) + 1u < 15u) id += 1u;         // increment_causal_chain_link, ecs_types.h:118-136
                const uint32_t id16 = ((id >> 28) << 12) | (((id >> 4) & 0xFFu) << 4) | (id & 0xFu);
                const uint32_t sign = (1u << lay.type_transform) | (hasP ? 1u << lay.type_physics : 0u) | (have_C ? 1u << lay.type_collider : 0u);
                v = b == 0u ? id16 : (b == 2u ? sign : 0u);                          // category all = 0
            }
        }
        out[i] = v;
    }
}
void launch_wire_messages(const DeviceWorld& W, Mirror M, const IcxWireLayout& lay, uint32_t coherency, int have_C, uint32_t* out, uint32_t n, cudaStream_t st)
{
    if (n) k_wire_messages<<<148 * 8, 256, 0, st>>>(W, M, lay, coherency, have_C, out, n);
}

// ---- packet structure = ColliderSynchro::update's submission order (collider_synchro.cpp:28-42) ----
// One packet per ACTIVE collider, entity by entity.  Computed on the device from the uploaded
// components (count, exclusive scan, pack): walking 168 B per entity on the host costs more than the
// PCIe transfer of the whole world.
__global__ void __launch_bounds__(256) k_count_active(const IcxColliderComponent* __restrict__ C, uint32_t n, uint32_t* __restrict__ counts,
                                                      uint8_t* __restrict__ mask)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t m = 0;
    if (C) for (int s = 0; s < ICX_COLLIDERS_PER_ENTITY; s++) if (C[i].colliders[s].isActive) m |= 1u << s;
    counts[i] = __popc(m); mask[i] = (uint8_t)m;
}
__global__ void __launch_bounds__(256) k_make_first(const uint32_t* __restrict__ offs, const uint8_t* __restrict__ mask, uint32_t n,
                                                    uint32_t* __restrict__ first)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) first[i] = offs[i] | ((uint32_t)__popc((uint32_t)mask[i]) << 30);
}
void launch_exclusive_scan(uint32_t* data, uint32_t* partial, uint32_t n, cudaStream_t st);
// scan_buf: n + 1 words (scan_buf[n] = number of packets afterwards)
void launch_packet_structure(const IcxColliderComponent* C, uint32_t n, uint32_t* scan_buf, uint32_t* partial, uint8_t* mask, uint32_t* first,
                             cudaStream_t st)
{
    if (n) k_count_active<<<(n + 255) / 256, 256, 0, st>>>(C, n, scan_buf, mask);
    launch_exclusive_scan(scan_buf, partial, n, st);
    if (n) k_make_first<<<(n + 255) / 256, 256, 0, st>>>(scan_buf, mask, n, first);
}

// ---- ParallelCosmosContext::load_and_link / extract_entity_updates (a22) ------------------
// Gather the AoS rows of the entities listed in idx (ascending source indices) into compact
// staging arrays.  mode 0 = clone (parallel_cosmos_context.cpp:107-151): a FORKED source entity
// carries its state over (stamped with the shared coherency) and is re-merged in the source
// (:145-150); every other entity starts as {merged, 0} in the fresh cosmos.  mode 1 = extract
// (:206-262): the state travels unchanged (the destination decides what to keep).
__global__ void __launch_bounds__(128) k_gather_rows(DeviceWorld W, Mirror M, const uint32_t* __restrict__ idx, uint32_t count, int mode,
                                                     uint32_t* o_ids, IcxTransform* o_T, IcxPhysics* o_P, IcxColliderComponent* o_C,
                                                     uint8_t* o_hasP, uint8_t* o_ts, uint16_t* o_tc)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    const uint32_t e = idx[i];
    o_ids[i] = M.ids[e]; o_T[i] = M.T[e]; o_hasP[i] = M.hasP[e];
    if (o_P) o_P[i] = M.P[e];
    if (o_C) o_C[i] = M.C[e];
    const uint32_t ts = W.in.ts[e], state = ts & 0xFFu;
    if (mode == 0) {
        if (state == ICX_TS_FORKED) {
            o_ts[i] = ICX_TS_FORKED; o_tc[i] = W.coherency;
            W.in.ts[e] = ICX_TS_MERGED | ((uint32_t)W.coherency << 16);             // set_timestream_state stamps the coherency, cosmos.cpp:162
        } else { o_ts[i] = ICX_TS_MERGED; o_tc[i] = 0; }
    } else { o_ts[i] = (uint8_t)state; o_tc[i] = (uint16_t)(ts >> 16); }
}

// extract, destination side (:255 and :268-271): entity index[i] takes the extracted state if its
// chain-link is > 0; timeslice 0 re-merges every extracted entity.  Both stamp the coherency.
__global__ void __launch_bounds__(128) k_extract_ts(DeviceWorld W, const uint32_t* __restrict__ index, const uint8_t* __restrict__ ts, uint32_t count)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    const uint32_t e = index[i];
    uint32_t cur = W.in.ts[e];
    if ((W.entity_id[e] & 0xFu) > 0u) cur = (uint32_t)ts[i] | ((uint32_t)W.coherency << 16);
    if (W.host_timeslice == 0u) cur = ICX_TS_MERGED | ((uint32_t)W.coherency << 16);
    W.in.ts[e] = cur;
}

void launch_gather_rows(const DeviceWorld& W, Mirror M, const uint32_t* idx, uint32_t count, int mode, uint32_t* o_ids, IcxTransform* o_T,
                        IcxPhysics* o_P, IcxColliderComponent* o_C, uint8_t* o_hasP, uint8_t* o_ts, uint16_t* o_tc, cudaStream_t st)
{
    if (!count) return;
    k_gather_rows<<<(count + 127) / 128, 128, 0, st>>>(W, M, idx, count, mode, o_ids, o_T, o_P, o_C, o_hasP, o_ts, o_tc);
}
void launch_extract_ts(const DeviceWorld& W, const uint32_t* index, const uint8_t* ts, uint32_t count, cudaStream_t st)
{
    if (!count) return;
    k_extract_ts<<<(count + 127) / 128, 128, 0, st>>>(W, index, ts, count);
}

void launch_import_kernel(const DeviceWorld& W, Mirror M, const uint32_t* index, const uint32_t* ids, const IcxTransform* T,
                          const IcxPhysics* P, const IcxColliderComponent* C, const uint8_t* hasP, const uint8_t* ts,
                          const uint16_t* tc, uint32_t count, int full, int have_C, uint32_t epoch, cudaStream_t st)
{
    if (!count) return;
    k_import<<<(count + 127) / 128, 128, 0, st>>>(W, M, index, ids, T, P, C, hasP, ts, tc, count, full, have_C, epoch);
}
void launch_export_kernel(const DeviceWorld& W, Mirror M, int have_T, int have_P, int have_C, uint8_t* ts, uint16_t* tc,
                          uint32_t n, cudaStream_t st)
{
    if (!n) return;
    k_export<<<(n + 127) / 128, 128, 0, st>>>(W, M, have_T, have_P, have_C, ts, tc, n);
}

} // namespace icx
