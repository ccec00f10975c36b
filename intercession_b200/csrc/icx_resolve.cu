// Ordered narrowphase + response + interception (CollisionPhysicsRelay::engage,
// source/physics/collision_physics_relay.h:24-231) on the broadphase candidates.
//
// The reference resolves every hit immediately inside a sequential lexicographic pair loop, so
// later pair tests see moved bodies.  Two facts make that loop parallel without changing a bit:
//  (1) a pair test that does not hit changes nothing, and a body that is never hit keeps its
//      post-integration state for the whole pass.  So every candidate is first tested
//      SPECULATIVELY on the post-integration state (k_narrow, one thread per pair); bodies with a
//      speculative hit are "dirty".  Pairs between two clean bodies are final non-hits.
//  (2) pairs sharing no dirty body commute.  The pairs of the dirty bodies are replayed by
//      dependency ROUNDS: each dirty body keeps a cursor into its sorted candidate list and a pair
//      of two dirty bodies runs in the round in which it is at the cursor of both (the globally
//      smallest pending pair always is, so every round progresses).  Clean and immutable partners
//      impose no order.  If such a replayed test hits a body that was clean, the set was too small:
//      that body is added, the dirty bodies are re-integrated from the tick input (bit-identical)
//      and the replay restarts, all inside the one persistent kernel (k_resolve).
// Immutable entities (infinite mass, non-temporal, no ray collider: nothing in a response can
// change them) are never dirty and never written.
#include <cooperative_groups.h>
#include "icx_world.cuh"
#include "icx_response.cuh"
#include "icx_mesh.cuh"
#include "icx_device.cuh"

namespace cg = cooperative_groups;

namespace icx {

#ifdef ICX_PROFILE_PAIR
#define PROF_ADD(k, v) do { if ((threadIdx.x & 31u) == 0u) atomicAdd(&W.counters[CTR_PROF + (k)], (uint32_t)(v)); } while (0)
#else
#define PROF_ADD(k, v) do { } while (0)
#endif
// list entry: x = other packet | ICX_NB_IMMUTABLE, y = its entity index (saves one dependent load per walk step)
struct Cursor { uint32_t p0, npk, n0, n1, e; const uint2* l0; const uint2* l1; };

// slot_info[slot] = {first packet | count << 30, n0 | n1 << 16, pool offset, entity}: everything a walk needs in ONE load
__device__ __forceinline__ Cursor make_cursor(const DeviceWorld& W, uint32_t slot, uint4 info)
{
    Cursor c;
    c.npk = info.x >> 30; c.p0 = info.x & 0x3FFFFFFFu; c.n0 = info.y & 0xFFFFu; c.n1 = info.y >> 16; c.e = info.w;
    c.l0 = W.nbr + info.z; c.l1 = c.l0 + c.n0;
    return c;
}
__device__ __forceinline__ Cursor load_cursor(const DeviceWorld& W, uint32_t slot) { return make_cursor(W, slot, W.slot_info[slot]); }
// Next entry of the body's merged queue.  Order = the reference's lexicographic pair order seen
// from this body: first pairs whose other packet precedes mine, by (other, my slot); then all
// pairs of my slot-0 packet, then all of my slot-1 packet.
// entry.y = partner's entity index (24 bits) | quantised lower bound of the gap between the two shapes on the post-integration
// state << 24 (ICX_GAP_UNIT margins per step; 0 = touching / unknown): see gap_of / cannot_meet
#define ICX_NB_ENT_MASK 0x00FFFFFFu
#ifndef ICX_SWEEP_MIN_SHARE
#define ICX_SWEEP_MIN_SHARE 50u     // another list + prediction sweep only if more than 1 / this of the dirty bodies are new
#endif
#ifndef ICX_GAPS_MIN_XPAIRS
#define ICX_GAPS_MIN_XPAIRS 2048u   // extra pairs from which the flagged lists get their may_meet gaps (lists_extend_flagged)
#endif
#define ICX_GAP_UNIT 0.5f
__device__ __forceinline__ bool cursor_head(const Cursor& c, uint32_t hw, uint32_t& k, uint32_t& entry, uint32_t& eo)
{
    const uint32_t h0 = hw & 0xFFFFu, h1 = hw >> 16;
    const bool has0 = h0 < c.n0, has1 = h1 < c.n1;
    if (!has0 && !has1) return false;
    uint32_t a = 0, b = 0, ea = 0, eb = 0;
    if (has0) { const uint2 v = c.l0[h0]; a = v.x; ea = v.y; }
    if (has1) { const uint2 v = c.l1[h1]; b = v.x; eb = v.y; }
    if (has0 && has1) {
        const uint32_t ao = a & ICX_NB_INDEX_MASK, bo = b & ICX_NB_INDEX_MASK;
        const bool alow = ao < c.p0, blow = bo < c.p0;
        if (alow || blow) k = (alow && blow) ? (bo < ao ? 1u : 0u) : (alow ? 0u : 1u);
        else k = 0u;
    } else k = has0 ? 0u : 1u;
    entry = k ? b : a; eo = k ? eb : ea;                     // eo: raw .y (entity | gap << 24)
    return true;
}
__device__ __forceinline__ uint32_t cursor_advance(uint32_t hw, uint32_t k) { return k ? hw + 0x10000u : hw + 1u; }
__device__ __forceinline__ bool cursor_pending(const Cursor& c, uint32_t hw) { return (hw & 0xFFFFu) < c.n0 || (hw >> 16) < c.n1; }

__device__ __forceinline__ Shape load_shape(const float4* d, V3 origin, int type, V3& base)
{
    Shape s;
    const float4 d0 = d[0], d1 = d[1], d2 = d[2], d3 = d[3], d4 = d[4], d5 = d[5];
    s.M.c[0] = v3(d0); s.M.c[1] = v3(d1); s.M.c[2] = v3(d2);
    base = v3(d0.w, d1.w, d2.w);
    s.M.t = base + (origin + v3(0.0f, 0.0f, 0.0f));
    s.inv.c[0] = v3(d3); s.inv.c[1] = v3(d4); s.inv.c[2] = v3(d5);
    s.type = type;
    return s;
}

__device__ __forceinline__ bool divergent(uint32_t s) { return s == ICX_TS_FORKING || s == ICX_TS_FORKED; }   // timestream_state.h:41-44

// Conservative, cheap stand-in for the 15-axis SAT of box_box / sphere_box_intersect
// (collision_procedures.cpp:41-120): false only if some axis separates the two shapes by clearly
// more than rounding noise, in which case the reference's own (normalised, differently rounded)
// test on that axis reports penetration <= 0 as well.  True means "may hit": the exact procedure
// decides.  Used to pick the dirty set and to skip exact tests; never to produce a result.
// slack > 0 also keeps pairs closer than slack (world units): the speculative pass uses it to
// pre-mark bodies a neighbour may be pushed into, which saves replay restarts.
// Penetration on axis L is rA(L) + rB(L) - |(cA - cB).L| and its sign does not depend on |L|.
__device__ __forceinline__ float box_radius(V3 c0, V3 c1, V3 c2, V3 L) { return 0.5f * (fabsf(dot(c0, L)) + fabsf(dot(c1, L)) + fabsf(dot(c2, L))); }

__device__ __forceinline__ bool sat_may_hit(const float4* __restrict__ dA, const float4* __restrict__ dB, V3 posA, V3 posB, int tA, int tB, float slack)
{
    const float4 a0 = dA[0], a1 = dA[1], a2 = dA[2], a3 = dA[3], a4 = dA[4], a5 = dA[5];
    const float4 b0 = dB[0], b1 = dB[1], b2 = dB[2], b3 = dB[3], b4 = dB[4], b5 = dB[5];
    const V3 A0 = v3(a0), A1 = v3(a1), A2 = v3(a2), B0 = v3(b0), B1 = v3(b1), B2 = v3(b2);
    const V3 cA = v3(a0.w, a1.w, a2.w) + posA, cB = v3(b0.w, b1.w, b2.w) + posB;
    const V3 d = cA - cB;
    const float cmag = fabsf(cA.x) + fabsf(cA.y) + fabsf(cA.z) + fabsf(cB.x) + fabsf(cB.y) + fabsf(cB.z);
    const float sphA = tA == ICX_COLLIDER_SPHERE ? 0.5f * sqrtf(dot(A0, A0)) : -1.0f;       // project_sphere radius, :97
    const float sphB = tB == ICX_COLLIDER_SPHERE ? 0.5f * sqrtf(dot(B0, B0)) : -1.0f;
    // face normals: rows of the inverse linear part (columns of its transpose), :41-52
    V3 n[6];
    n[0] = v3(a3.x, a4.x, a5.x); n[1] = v3(a3.y, a4.y, a5.y); n[2] = v3(a3.z, a4.z, a5.z);
    n[3] = v3(b3.x, b4.x, b5.x); n[4] = v3(b3.y, b4.y, b5.y); n[5] = v3(b3.z, b4.z, b5.z);
#pragma unroll
    for (int a = 0; a < 15; a++) {
        V3 L; float rel = 1.0e-4f;
        if (a < 6) L = n[a];
        else {
            const V3 x = n[(a - 6) / 3], y = n[3 + (a - 6) % 3];
            L = cross(x, y);
            // nearly parallel normals: the direction of the cross product is noise, the axis proves nothing
            if (dot(L, L) < 1.0e-4f * dot(x, x) * dot(y, y)) continue;
            rel = 1.0e-3f;
        }
        const float l1 = fabsf(L.x) + fabsf(L.y) + fabsf(L.z);
        const float rA = sphA >= 0.0f ? sphA * sqrtf(dot(L, L)) : box_radius(A0, A1, A2, L);
        const float rB = sphB >= 0.0f ? sphB * sqrtf(dot(L, L)) : box_radius(B0, B1, B2, L);
        const float s = rA + rB, dist = fabsf(dot(d, L));
        if (dist > s + rel * (s + dist) + (4.0e-6f * cmag + slack) * l1) return false;
    }
    return true;
}
// Same filter, but it also measures: returns true ("may hit") with `value` = an upper bound of the
// penetration depth the exact procedure can report (how far static resolution may push the two
// bodies), or false with `value` = a lower bound of the gap between the shapes (how far a body
// would have to be pushed before the exact procedure could report a hit).  A gap above `far`
// returns at once.  Used only to PREDICT which bodies the ordered replay will touch.
__device__ __forceinline__ bool sat_measure(const float4* __restrict__ dA, const float4* __restrict__ dB, V3 posA, V3 posB, int tA, int tB,
                                            float far, float& value, V3* push_dir = nullptr)
{
    const float4 a0 = dA[0], a1 = dA[1], a2 = dA[2], a3 = dA[3], a4 = dA[4], a5 = dA[5];
    const float4 b0 = dB[0], b1 = dB[1], b2 = dB[2], b3 = dB[3], b4 = dB[4], b5 = dB[5];
    const V3 A0 = v3(a0), A1 = v3(a1), A2 = v3(a2), B0 = v3(b0), B1 = v3(b1), B2 = v3(b2);
    const V3 cA = v3(a0.w, a1.w, a2.w) + posA, cB = v3(b0.w, b1.w, b2.w) + posB;
    const V3 d = cA - cB;
    const float cmag = fabsf(cA.x) + fabsf(cA.y) + fabsf(cA.z) + fabsf(cB.x) + fabsf(cB.y) + fabsf(cB.z);
    const float sphA = tA == ICX_COLLIDER_SPHERE ? 0.5f * sqrtf(dot(A0, A0)) : -1.0f;
    const float sphB = tB == ICX_COLLIDER_SPHERE ? 0.5f * sqrtf(dot(B0, B0)) : -1.0f;
    V3 n[6];
    n[0] = v3(a3.x, a4.x, a5.x); n[1] = v3(a3.y, a4.y, a5.y); n[2] = v3(a3.z, a4.z, a5.z);
    n[3] = v3(b3.x, b4.x, b5.x); n[4] = v3(b3.y, b4.y, b5.y); n[5] = v3(b3.z, b4.z, b5.z);
    float minpen = INFINITY, maxsep = 0.0f; bool separated = false;
#pragma unroll
    for (int a = 0; a < 15; a++) {
        V3 L; float rel = 1.0e-4f;
        if (a < 6) L = n[a];
        else {
            const V3 x = n[(a - 6) / 3], y = n[3 + (a - 6) % 3];
            L = cross(x, y);
            if (dot(L, L) < 1.0e-4f * dot(x, x) * dot(y, y)) continue;
            rel = 1.0e-3f;
        }
        const float l2 = dot(L, L), len = sqrtf(l2), l1 = fabsf(L.x) + fabsf(L.y) + fabsf(L.z);
        const float rA = sphA >= 0.0f ? sphA * len : box_radius(A0, A1, A2, L);
        const float rB = sphB >= 0.0f ? sphB * len : box_radius(B0, B1, B2, L);
        const float s = rA + rB, dist = fabsf(dot(d, L));
        const float tol = rel * (s + dist) + 4.0e-6f * cmag * l1;
        if (dist > s + tol) {
            separated = true;
            maxsep = fmaxf(maxsep, (dist - s - tol) / len);
            if (maxsep > far) { value = maxsep; return false; }
        } else {
            const float pen = (s - dist + tol) / len;
            if (pen < minpen) { minpen = pen; if (push_dir) *push_dir = L * ((dot(d, L) >= 0.0f ? 1.0f : -1.0f) / len); }
        }
    }
    value = separated ? maxsep : (minpen < INFINITY ? minpen : 0.0f);
    return !separated;
}

__device__ __forceinline__ bool is_sat_pair(int tA, int tB) {
    return (tA == ICX_COLLIDER_BOX && (tB == ICX_COLLIDER_BOX || tB == ICX_COLLIDER_SPHERE)) || (tA == ICX_COLLIDER_SPHERE && tB == ICX_COLLIDER_BOX);
}
// current-state filter for packets pa, pb: false = the exact procedure cannot report a hit
// The current-state filter (sat_may_hit) by a whole warp for ONE pair (uniform arguments): lane a evaluates
// axis a, so the verdict costs one axis of latency instead of fifteen.  Same arithmetic per axis.
__device__ __forceinline__ bool may_hit_warp(const DeviceWorld& W, uint32_t pa, uint32_t pb, uint32_t ea, uint32_t eb)
{
    const uint32_t FULL = 0xFFFFFFFFu, lane = threadIdx.x & 31u;
    const float wA = W.pk_lpos[pa].w, wB = W.pk_lpos[pb].w;
    const float4 pA = W.out.pos[ea], pB = W.out.pos[eb];
    const float4* __restrict__ dA = W.pk_d + (size_t)pa * 6; const float4* __restrict__ dB = W.pk_d + (size_t)pb * 6;
    const float4 a0 = dA[0], a1 = dA[1], a2 = dA[2], a3 = dA[3], a4 = dA[4], a5 = dA[5];
    const float4 b0 = dB[0], b1 = dB[1], b2 = dB[2], b3 = dB[3], b4 = dB[4], b5 = dB[5];
    const int tA = __float_as_uint(wA) & PF_TYPE_MASK, tB = __float_as_uint(wB) & PF_TYPE_MASK;
    if (!is_sat_pair(tA, tB)) return true;
    const V3 A0 = v3(a0), A1 = v3(a1), A2 = v3(a2), B0 = v3(b0), B1 = v3(b1), B2 = v3(b2);
    const V3 cA = v3(a0.w, a1.w, a2.w) + v3(pA), cB = v3(b0.w, b1.w, b2.w) + v3(pB);
    const V3 d = cA - cB;
    const float cmag = fabsf(cA.x) + fabsf(cA.y) + fabsf(cA.z) + fabsf(cB.x) + fabsf(cB.y) + fabsf(cB.z);
    const float sphA = tA == ICX_COLLIDER_SPHERE ? 0.5f * sqrtf(dot(A0, A0)) : -1.0f;
    const float sphB = tB == ICX_COLLIDER_SPHERE ? 0.5f * sqrtf(dot(B0, B0)) : -1.0f;
    const V3 nA0 = v3(a3.x, a4.x, a5.x), nA1 = v3(a3.y, a4.y, a5.y), nA2 = v3(a3.z, a4.z, a5.z);
    const V3 nB0 = v3(b3.x, b4.x, b5.x), nB1 = v3(b3.y, b4.y, b5.y), nB2 = v3(b3.z, b4.z, b5.z);
    bool sep = false;
    if (lane < 15u) {
        V3 L; float rel = 1.0e-4f; bool skip = false;
        if (lane < 6u) L = lane == 0u ? nA0 : lane == 1u ? nA1 : lane == 2u ? nA2 : lane == 3u ? nB0 : lane == 4u ? nB1 : nB2;
        else {
            const uint32_t i = (lane - 6u) / 3u, j = (lane - 6u) % 3u;
            const V3 x = i == 0u ? nA0 : i == 1u ? nA1 : nA2, y = j == 0u ? nB0 : j == 1u ? nB1 : nB2;
            L = cross(x, y);
            skip = dot(L, L) < 1.0e-4f * dot(x, x) * dot(y, y);
            rel = 1.0e-3f;
        }
        const float l1 = fabsf(L.x) + fabsf(L.y) + fabsf(L.z);
        const float rA = sphA >= 0.0f ? sphA * sqrtf(dot(L, L)) : box_radius(A0, A1, A2, L);
        const float rB = sphB >= 0.0f ? sphB * sqrtf(dot(L, L)) : box_radius(B0, B1, B2, L);
        const float sum = rA + rB, dist = fabsf(dot(d, L));
        sep = !skip && dist > sum + rel * (sum + dist) + (4.0e-6f * cmag) * l1;
    }
    return __any_sync(FULL, sep) == 0;
}

// intersectProcedures[tA][tB], collision_procedures.h:411-417, for packets pa < pb on the current
// state.  mpA / mpB: Collider::minParametricValue of a ray side, updated on a hit.
static __device__ __noinline__ bool pair_intersect(const DeviceWorld& W, uint32_t pa, uint32_t pb, uint32_t ea, uint32_t eb,
                                                   int tA, int tB, V3 posA, V3 posB, Shape& A, Shape& B, V3& baseA, V3& baseB,
                                                   float& mpA, float& mpB, Hit& h)
{
    A = load_shape(W.pk_d + (size_t)pa * 6, posA, tA, baseA);
    B = load_shape(W.pk_d + (size_t)pb * 6, posB, tB, baseB);
    h.point = v3(0.0f, 0.0f, 0.0f); h.normal = v3(0.0f, 0.0f, 0.0f); h.depth = 0.0f;
    bool hit = false;
    if (tA == ICX_COLLIDER_BOX && tB == ICX_COLLIDER_BOX) hit = box_box_intersect(A, B, h);
    else if (tA == ICX_COLLIDER_SPHERE && tB == ICX_COLLIDER_BOX) hit = sphere_box_intersect(A, B, h);
    else if (tA == ICX_COLLIDER_BOX && tB == ICX_COLLIDER_SPHERE) {                 // box_sphere_intersect, .cpp:233-247
        hit = sphere_box_intersect(B, A, h);
        if (hit) { h.normal = h.normal * -1.0f; h.point = h.point + (h.normal * h.depth); }
    } else if (tA == ICX_COLLIDER_RAY && tB == ICX_COLLIDER_BOX) {
        hit = ray_box_intersect(A, posA, B, posB, mpA, h);
    } else if (tA == ICX_COLLIDER_BOX && tB == ICX_COLLIDER_RAY) {                  // box_ray_intersect, .cpp:249-263
        hit = ray_box_intersect(B, posB, A, posA, mpB, h);
        if (hit) { h.normal = h.normal * -1.0f; h.point = h.point + (h.normal * h.depth); }
    } else if ((W.cfg_flags & ICX_F_SPHERE_SPHERE) && tA == ICX_COLLIDER_SPHERE && tB == ICX_COLLIDER_SPHERE) {
        const float rA = ICX_UNIT_RADIUS * W.pk_lscale[pa].x * W.prm0[ea].x;         // .cpp:403-404
        const float rB = ICX_UNIT_RADIUS * W.pk_lscale[pb].x * W.prm0[eb].x;
        hit = sphere_sphere_intersect(A, rA, B, rB, h);
    }
    return hit;
}

// ---- warp-cooperative exact procedures ---------------------------------------------------
// The exact narrowphase + response of ONE pair is a single thread's worth of serial FP32 work
// (~6 K instructions): on a GPU that is pure latency.  The ordered resolve therefore gives every
// pair a whole warp.  All 32 lanes hold the same operands and execute the same instructions
// ("uniform" code: loads broadcast, results identical in every lane), except for the 15-axis SAT
// loop, where lane a evaluates axis a with exactly the arithmetic of the serial loop and the warp
// then reproduces the loop's decisions: any axis with penetration <= 0 -> no hit, otherwise the
// FIRST axis attaining the strict minimum.  Only lane 0 writes to memory.
struct SatPick { V3 normal; float depth; bool separated; };

__device__ __forceinline__ SatPick sat_pick_warp(const AxisResult r, bool have)
{
    const uint32_t FULL = 0xFFFFFFFFu, lane = threadIdx.x & 31u;
    SatPick p;
    p.separated = __any_sync(FULL, have && !r.skip && r.pen <= 0) != 0;            // :107 (any axis, order irrelevant)
    const bool cand = have && !r.skip && r.pen < INFINITY;                         // `pen < minD` can only ever hold for these
    float key = cand ? r.pen : INFINITY; uint32_t idx = lane;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float ok = __shfl_xor_sync(FULL, key, o); const uint32_t oi = __shfl_xor_sync(FULL, idx, o);
        if (ok < key || (ok == key && oi < idx)) { key = ok; idx = oi; }
    }
    const V3 n = r.flip ? -r.axis : r.axis;                                          // :113-117
    p.normal = v3(__shfl_sync(FULL, n.x, idx), __shfl_sync(FULL, n.y, idx), __shfl_sync(FULL, n.z, idx));
    p.depth = key;
    if (!(key < INFINITY)) p.normal = v3(0.0f, 0.0f, 0.0f);
    return p;
}

// the 15-axis loop of box_box_intersect (X, Y boxes) or sphere_box_intersect (X sphere, Y box), one axis per lane
__device__ __forceinline__ SatPick sat_warp(const Shape& X, const Shape& Y, bool x_is_sphere)
{
    const int lane = threadIdx.x & 31;
    const M3 nX = transpose(X.inv), nY = transpose(Y.inv);
    V3 cX[8], cY[8];
    box_corners(Y.M, cY);
    const V3 sO = xform(X.M, 0.0f, 0.0f, 0.0f);
    const float projRadius = length(xform(X.M, ICX_UNIT_RADIUS, 0.0f, 0.0f) - sO);
    if (!x_is_sphere) box_corners(X.M, cX);
    AxisResult r; r.axis = v3(0.0f, 0.0f, 0.0f); r.pen = 0.0f; r.flip = false; r.skip = true;
    if (lane < 15) r = x_is_sphere ? sphere_box_axis(nX, nY, sO, projRadius, cY, lane) : box_box_axis(nX, nY, cX, cY, lane);
    return sat_pick_warp(r, lane < 15);
}

// The same loop on ONE lane (lane-parallel items, see round_execute_lanes): the serial axis loop of box_box_intersect /
// sphere_box_intersect itself, collision_procedures.cpp:76-118 -- same arithmetic per axis, first strict minimum.
__device__ __forceinline__ SatPick sat_serial(const Shape& X, const Shape& Y, bool x_is_sphere)
{
    const M3 nX = transpose(X.inv), nY = transpose(Y.inv);
    V3 cX[8], cY[8];
    box_corners(Y.M, cY);
    const V3 sO = xform(X.M, 0.0f, 0.0f, 0.0f);
    const float projRadius = length(xform(X.M, ICX_UNIT_RADIUS, 0.0f, 0.0f) - sO);
    if (!x_is_sphere) box_corners(X.M, cX);
    SatPick p; p.separated = false; p.normal = v3(0.0f, 0.0f, 0.0f); p.depth = INFINITY;
#pragma unroll 1
    for (int a = 0; a < 15; a++) {
        const AxisResult r = x_is_sphere ? sphere_box_axis(nX, nY, sO, projRadius, cY, a) : box_box_axis(nX, nY, cX, cY, a);
        if (r.skip) continue;
        if (r.pen <= 0) { p.separated = true; break; }                              // :107
        if (r.pen < p.depth) { p.depth = r.pen; p.normal = r.flip ? -r.axis : r.axis; }   // :113-117
    }
    return p;
}

// everything one side of a pair contributes, fetched in ONE batch of independent loads
struct SideData { float4 pos, vel, angvel, acc, angacc, prm0, prm1, mat0, lscale, pos0, d[6]; float damping, minparam; uint32_t slot, ts, id; };

__device__ __forceinline__ SideData load_side(const DeviceWorld& W, uint32_t p, uint32_t e)
{
    SideData s;
    s.pos = W.out.pos[e]; s.vel = W.out.vel[e]; s.angvel = W.out.angvel[e]; s.acc = W.out.acc[e]; s.angacc = W.out.angacc[e];
    s.prm0 = W.prm0[e]; s.prm1 = W.prm1[e]; s.pos0 = W.pos0[e]; s.slot = W.dirty_slot[e]; s.ts = W.out.ts[e]; s.id = W.entity_id[e];
    s.mat0 = W.pk_mat0[p]; s.lscale = W.pk_lscale[p]; s.damping = W.pk_damping[p]; s.minparam = W.pk_minparam[p];
    const float4* d = W.pk_d + (size_t)p * 6;
#pragma unroll
    for (int k = 0; k < 6; k++) s.d[k] = d[k];
    return s;
}
__device__ __forceinline__ Shape shape_of(const SideData& s, int type, V3& base)
{
    Shape sh;
    sh.M.c[0] = v3(s.d[0]); sh.M.c[1] = v3(s.d[1]); sh.M.c[2] = v3(s.d[2]);
    base = v3(s.d[0].w, s.d[1].w, s.d[2].w);
    sh.M.t = base + (v3(s.pos) + v3(0.0f, 0.0f, 0.0f));
    sh.inv.c[0] = v3(s.d[3]); sh.inv.c[1] = v3(s.d[4]); sh.inv.c[2] = v3(s.d[5]);
    sh.type = type;
    return sh;
}
__device__ __forceinline__ Side side_of(const SideData& s, const Shape& sh, V3 base, uint32_t bits)
{
    Side S;
    S.origin = v3(s.pos); S.vel = v3(s.vel); S.angvel = v3(s.angvel); S.acc = v3(s.acc); S.angacc = v3(s.angacc);
    S.mass = s.prm0.w; S.colAngDrag = s.prm1.z;
    S.L = mat3_of(sh.M); S.base = base; S.inv = sh.inv; S.inertia = v3(s.d[3].w, s.d[4].w, s.d[5].w);
    S.staticFriction = s.mat0.x; S.dynamicFriction = s.mat0.y; S.restitution = s.mat0.z; S.stiffness = s.mat0.w;
    S.damping = s.damping; S.restLength = s.lscale.w;
    S.influenceOrientation = (bits & PF_INFLUENCE_ORIENT) != 0;
    return S;
}
__device__ __forceinline__ void store_side(const DeviceWorld& W, uint32_t e, const Side& S, const SideData& s, bool rr)
{
    W.out.vel[e] = make_float4(S.vel.x, S.vel.y, S.vel.z, 0.0f);
    W.out.angvel[e] = make_float4(S.angvel.x, S.angvel.y, S.angvel.z, 0.0f);
    if (rr) {
        W.out.pos[e] = make_float4(S.origin.x, S.origin.y, S.origin.z, 0.0f);
        // the candidate lists cover a displacement of bm[e] (DESIGN.md "Broadphase bound"): beyond it the replay restarts with
        // this body's budget raised (a NaN origin is the reference's problem as much as ours: it is carried along)
        const float lim = W.bm[e] * 0.98f;
        const float dmax = fmaxf(fmaxf(fabsf(S.origin.x - s.pos0.x), fabsf(S.origin.y - s.pos0.y)), fabsf(S.origin.z - s.pos0.z));
        if (dmax > lim && dmax < 3.0e38f) {
            atomicMax((int*)&W.bm_need[e], __float_as_int(dmax));
            atomicAdd(&W.counters[CTR_VIOLATORS], 1u);                             // the replay restarts with this body's budget raised
        }
    } else {
        W.out.acc[e] = make_float4(S.acc.x, S.acc.y, S.acc.z, 0.0f);
        W.out.angacc[e] = make_float4(S.angacc.x, S.angacc.y, S.angacc.z, 0.0f);
    }
}

// One iteration of the reference's inner loop body (collision_physics_relay.h:39-228) for packets
// pa < pb, executed by a whole warp (all lanes pass the same arguments).  n_dirty: bodies with
// dirty_slot < n_dirty may be written; a hit on any other mutable body aborts the attempt (see the
// file header).
// LANE = false: a whole warp runs ONE pair (all lanes pass the same arguments, lane 0 finishes and writes).
// LANE = true: every lane with `active` runs ITS OWN pair from start to end -- the plain SIMT form; the serial work
// that keeps 31 lanes idle in warp mode (manifold, clipping, response: ~35 k cycles) then runs for up to 32 pairs at once,
// with thread-local scratch (all lanes active: local memory lines are fully used).
template <bool LANE>
__device__ __forceinline__ void process_pair(const DeviceWorld& W, uint32_t pa, uint32_t pb, uint32_t n_dirty, uint32_t ctr_new, V3* scratch, bool active = true)
{
    if (LANE && !active) return;
    const bool writer = LANE ? true : (threadIdx.x & 31u) == 0u;
#ifdef ICX_PROFILE_PAIR
    const long long tp0 = clock64();
#endif
    const uint32_t ea = W.pk_entity[pa], eb = W.pk_entity[pb];
    const uint32_t bitsA = __float_as_uint(W.pk_lpos[pa].w), bitsB = __float_as_uint(W.pk_lpos[pb].w);
    const int tA = bitsA & PF_TYPE_MASK, tB = bitsB & PF_TYPE_MASK;
    const SideData a = load_side(W, pa, ea), b = load_side(W, pb, eb);
    V3 baseA, baseB;
    const Shape A = shape_of(a, tA, baseA), B = shape_of(b, tB, baseB);
    const V3 posA = v3(a.pos), posB = v3(b.pos);
    float mpA = (tA == ICX_COLLIDER_RAY) ? a.minparam : 1.0f, mpB = (tB == ICX_COLLIDER_RAY) ? b.minparam : 1.0f;

    // ***** intersection: intersectProcedures[tA][tB], collision_procedures.h:411-417 *****
    // SAT axis loop by the whole warp; everything after it is serial work with dynamically indexed
    // scratch arrays: lane 0 alone continues, so that only ONE stack frame per warp is live in L1
    Hit h; h.point = v3(0.0f, 0.0f, 0.0f); h.normal = v3(0.0f, 0.0f, 0.0f); h.depth = 0.0f;
    SatPick pick; pick.separated = true; pick.normal = h.normal; pick.depth = 0.0f;
    const bool bb = tA == ICX_COLLIDER_BOX && tB == ICX_COLLIDER_BOX, sb = tA == ICX_COLLIDER_SPHERE && tB == ICX_COLLIDER_BOX,
               bs = tA == ICX_COLLIDER_BOX && tB == ICX_COLLIDER_SPHERE;
    if (bb || sb || bs) pick = LANE ? sat_serial(bs ? B : A, bs ? A : B, !bb)
                                    : sat_warp(bs ? B : A, bs ? A : B, !bb);        // one instantiation per mode: the code has to stay small (instruction cache)
#ifdef ICX_PROFILE_PAIR
    if (writer) { atomicAdd(&W.counters[CTR_PROF + 5], (uint32_t)((clock64() - tp0) >> 4)); atomicAdd(&W.counters[CTR_PROF + 6], pick.separated ? 0u : 1u); }
#endif
    if (!writer) return;
    bool hit = false;
    if (bb) {
        if (!pick.separated) {
            h.normal = pick.normal; h.depth = pick.depth;
            if (LANE) hit = box_box_finish(A, B, h);
            else hit = box_box_finish_s(A, B, h, scratch, scratch + 8, scratch + 8 + ICX_MAXV);
        }
    } else if (sb) {
        if (!pick.separated) { h.normal = pick.normal; h.depth = pick.depth; hit = sphere_box_finish(xform(A.M, 0.0f, 0.0f, 0.0f), xform(A.M, ICX_UNIT_RADIUS, 0.0f, 0.0f), h); }
    } else if (bs) {                                                                // box_sphere_intersect, .cpp:233-247
        if (!pick.separated) {
            h.normal = pick.normal; h.depth = pick.depth; hit = sphere_box_finish(xform(B.M, 0.0f, 0.0f, 0.0f), xform(B.M, ICX_UNIT_RADIUS, 0.0f, 0.0f), h);
            h.normal = h.normal * -1.0f; h.point = h.point + (h.normal * h.depth);
        }
    } else if (tA == ICX_COLLIDER_RAY && tB == ICX_COLLIDER_BOX) {
        hit = ray_box_intersect(A, posA, B, posB, mpA, h);
    } else if (tA == ICX_COLLIDER_BOX && tB == ICX_COLLIDER_RAY) {                  // box_ray_intersect, .cpp:249-263
        hit = ray_box_intersect(B, posB, A, posA, mpB, h);
        if (hit) { h.normal = h.normal * -1.0f; h.point = h.point + (h.normal * h.depth); }
    } else if ((W.cfg_flags & ICX_F_SPHERE_SPHERE) && tA == ICX_COLLIDER_SPHERE && tB == ICX_COLLIDER_SPHERE) {
        hit = sphere_sphere_intersect(A, ICX_UNIT_RADIUS * a.lscale.x * a.prm0.x, B, ICX_UNIT_RADIUS * b.lscale.x * b.prm0.x, h);   // .cpp:403-404
    }
    if (!hit) return;
#ifdef ICX_PROFILE_PAIR
    const long long tp1 = clock64();
    atomicAdd(&W.counters[CTR_PROF + 7], (uint32_t)((tp1 - tp0) >> 4)); atomicAdd(&W.counters[CTR_PROF + 8], 1u);
    if (bb) { atomicAdd(&W.counters[CTR_PROF + 11], (uint32_t)((tp1 - tp0) >> 4)); atomicAdd(&W.counters[CTR_PROF + 12], 1u); }
#endif
    const uint32_t efA = __float_as_uint(a.prm1.w), efB = __float_as_uint(b.prm1.w);
    {
        const bool cleanA = !(efA & EF_IMMUTABLE) && a.slot >= n_dirty;
        const bool cleanB = !(efB & EF_IMMUTABLE) && b.slot >= n_dirty;
        if (cleanA || cleanB) {                     // the dirty set was too small: grow it, write nothing
            if (writer) {
                if (cleanA) mark_dirty(W, ea);
                if (cleanB) mark_dirty(W, eb);
                atomicAdd(&W.counters[ctr_new], 1u);
            }
            return;
        }
    }
    if (writer) {
        if (tA == ICX_COLLIDER_RAY) W.pk_minparam[pa] = mpA;
        if (tB == ICX_COLLIDER_RAY) W.pk_minparam[pb] = mpB;
    }

    // ***** physics response, :82-107 *****
    const uint32_t ctA = (bitsA >> PF_CTYPE_SHIFT) & 3u, ctB = (bitsB >> PF_CTYPE_SHIFT) & 3u;
    uint32_t cflags = 0;
    V3 originA = posA, originB = posB;
    if (ctA != ICX_COLLISION_NOOP && ctB != ICX_COLLISION_NOOP && (efA & EF_HAS_PHYSICS) && (efB & EF_HAS_PHYSICS)) {
        cflags |= ICX_CONTACT_RESPONDED;
        const bool rr = ctA == ICX_COLLISION_RIGID && ctB == ICX_COLLISION_RIGID;
        const bool sr = ctA == ICX_COLLISION_SPRING && ctB == ICX_COLLISION_RIGID;
        const bool rs = ctA == ICX_COLLISION_RIGID && ctB == ICX_COLLISION_SPRING;
        if (rr || sr || rs) {                                                       // other table entries are null_response
            Side SA = side_of(a, A, baseA, bitsA), SB = side_of(b, B, baseB, bitsB);
            if (rr) rigid_rigid_response_i(SA, SB, h);
            else {
                if (rs) {                                                           // rigid_spring_response, .cpp:1055-1065: swapped sides, inverted contact
                    const Side t = SA; SA = SB; SB = t;
                    Hit inv; inv.normal = -h.normal; inv.point = h.point - (h.normal * h.depth); inv.depth = h.depth;
                    spring_rigid_response_i(SA, SB, inv);
                    const Side u = SA; SA = SB; SB = u;
                } else spring_rigid_response_i(SA, SB, h);
            }
            if (writer) {
                if (!(efA & EF_IMMUTABLE)) store_side(W, ea, SA, a, rr);
                if (!(efB & EF_IMMUTABLE)) store_side(W, eb, SB, b, rr);
            }
            if (rr) { if (!(efA & EF_IMMUTABLE)) originA = SA.origin; if (!(efB & EF_IMMUTABLE)) originB = SB.origin; }     // what the callbacks below see
        }
    }

#ifdef ICX_PROFILE_PAIR
    const long long tp2 = clock64();
    atomicAdd(&W.counters[CTR_PROF + 9], (uint32_t)((tp2 - tp1) >> 4));
#endif
    // ***** behaviours response :114-155 -> contact record, replayed by the host adapter *****
    if (bitsA & PF_BEHAVIORS) cflags |= ICX_CONTACT_BEHAVIORS_A;
    if (bitsB & PF_BEHAVIORS) cflags |= ICX_CONTACT_BEHAVIORS_B;
    const uint32_t idA = a.id, idB = b.id;
    {
        const uint32_t ci = atomicAdd(&W.counters[CTR_CONTACTS], 1u);
        if (ci < W.max_contacts) {
            IcxContact c;
            c.entityA = idA; c.entityB = idB;
            c.slotA = (bitsA & PF_SLOT) ? 1 : 0; c.slotB = (bitsB & PF_SLOT) ? 1 : 0; c.flags = (uint16_t)cflags;
            c.point[0] = h.point.x; c.point[1] = h.point.y; c.point[2] = h.point.z;
            c.normal[0] = h.normal.x; c.normal[1] = h.normal.y; c.normal[2] = h.normal.z;
            c.depth = h.depth; c.packetA = pa; c.packetB = pb;
            W.contacts[ci] = c;
        } else atomicOr(W.flags, RF_CONTACT_OVERFLOW);
    }
    // The contact half of BipedBehaviors::on_collision (behaviors/biped_behaviors.h:336-366), which needs the CALLER's origin
    // at callback time (after this pair's static resolution): isGrounded, groundNormal = the normal as the caller is handed it,
    // groundDist = length(caller origin - collision point).  B is called with the inverted data (relay :137-146).  One record
    // per call; the last one per entity in pair order is what the component ends up with (icx_grounds_fetch).
    if (cflags & (ICX_CONTACT_BEHAVIORS_A | ICX_CONTACT_BEHAVIORS_B)) {
        for (int side = 0; side < 2; side++) {
            if (!(cflags & (side ? ICX_CONTACT_BEHAVIORS_B : ICX_CONTACT_BEHAVIORS_A))) continue;
            const V3 n = side ? -h.normal : h.normal;
            const V3 pt = side ? h.point - (h.normal * h.depth) : h.point;
            const V3 o = side ? originB : originA;
            const uint32_t gi = atomicAdd(&W.counters[CTR_GROUNDS], 1u);
            if (gi < W.max_contacts) {
                IcxGround g;
                g.entity = side ? idB : idA; g.normal[0] = n.x; g.normal[1] = n.y; g.normal[2] = n.z; g.dist = length(o - pt);
                g.packetA = pa; g.packetB = pb; g.side = (uint32_t)side;
                W.grounds[gi] = g;
            } else atomicOr(W.flags, RF_CONTACT_OVERFLOW);
        }
    }

    // ***** timestream response :157-228 *****
    const uint32_t thisLink = idA & 0xFu, otherLink = idB & 0xFu;                   // derive_causal_chain_link, ecs_types.h:92-95
    if (thisLink == ICX_NULL_CHAINLINK || otherLink == ICX_NULL_CHAINLINK) return;  // :185-188
    const bool dA = divergent(a.ts & 0xFFu), dB = divergent(b.ts & 0xFFu);
    if (!(((thisLink == 0) != (otherLink == 0))
          || (thisLink != 0 && otherLink != 0 && (dA != dB))
          || (thisLink != otherLink && (!dA || !dB) && W.host_timeslice == ICX_NULL_TIMESLICE)))
        return;                                                                     // :193-200
    uint32_t agent;
    if (thisLink != otherLink) agent = thisLink < otherLink ? idA : idB;            // :207-210
    else if (dA != dB) agent = dA ? idA : idB;
    else agent = ICX_NULL_ENTITY;
    const uint32_t recipient = agent == idA ? idB : idA;                            // :222
    const uint32_t recipientIndex = agent == idA ? eb : ea;
    const uint32_t ii = atomicAdd(&W.counters[CTR_INTERCEPTS], 1u);
    if (ii < W.max_contacts) {
        W.intercepts[ii].agent = agent; W.intercepts[ii].recipient = recipient;
        W.intercept_key[ii] = make_uint2(pa, pb);
    } else atomicOr(W.flags, RF_CONTACT_OVERFLOW);
    // handlers the synchronous event reaches (server_network_dynamo.cpp:813-838: forking, ignored
    // for link-0 recipients; parallel_network_dynamo.cpp:455-467: forked).  set_timestream_state
    // stamps the current coherency (cosmos.cpp:155-173).
    if (W.host_timeslice == ICX_NULL_TIMESLICE) W.out.ts[recipientIndex] = ICX_TS_FORKED | ((uint32_t)W.coherency << 16);
    else if ((recipient & 0xFu) > 0) W.out.ts[recipientIndex] = ICX_TS_FORKING | ((uint32_t)W.coherency << 16);
}

// ---------------------------------------------------------------------------------------
// K5: speculative narrowphase.  One thread per candidate pair whose un-inflated bounds overlap:
// a conservative separating-axis filter (exact procedure for ray / sphere-sphere pairs) on the
// post-integration state; writes nothing but the dirty marks -- the exact, ordered evaluation of
// every pair that may hit happens in k_resolve.  Hits between two
// immutable bodies change nothing and need no order; they are queued for their contact record.
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) k_narrow(const __grid_constant__ DeviceWorld W)
{
    const uint32_t n = min(W.counters[CTR_TIGHT], W.max_pairs);
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const uint2 pr = W.tight[i];
    const uint32_t pa = pr.x & ICX_NB_INDEX_MASK, pb = pr.y & ICX_NB_INDEX_MASK;
    const uint32_t ea = W.pk_entity[pa], eb = W.pk_entity[pb];
    const int tA = __float_as_uint(W.pk_lpos[pa].w) & PF_TYPE_MASK, tB = __float_as_uint(W.pk_lpos[pb].w) & PF_TYPE_MASK;
    float depth = 0.0f;
    V3 dir = v3(0.0f, 0.0f, 0.0f);                  // direction static resolution would push A (least-penetration axis, from B towards A)
    if (is_sat_pair(tA, tB)) {
        float value;
        if (!sat_measure(W.pk_d + (size_t)pa * 6, W.pk_d + (size_t)pb * 6, v3(W.out.pos[ea]), v3(W.out.pos[eb]), tA, tB, 0.0f, value, &dir)) continue;
        depth = value;
    } else {
        V3 baseA, baseB; Shape A, B; Hit h;
        float mpA = 1.0f, mpB = 1.0f;           // Collider::reset value; a ray that hits is dirty and replayed in order
        if (!pair_intersect(W, pa, pb, ea, eb, tA, tB, v3(W.out.pos[ea]), v3(W.out.pos[eb]), A, B, baseA, baseB, mpA, mpB, h)) continue;
        depth = (tA == ICX_COLLIDER_RAY || tB == ICX_COLLIDER_RAY) ? 0.0f : h.depth;      // springs do not push
    }
    if (pr.y & ICX_PAIR_STATIC) {
        const uint32_t k = atomicAdd(&W.counters[CTR_STATIC], 1u);
        if (k < W.max_static) W.static_pairs[k] = make_uint2(pa, pb); else atomicOr(W.flags, RF_PAIR_OVERFLOW);
        continue;
    }
    // Predicted push of each body.  Static resolution splits a penetration by inverse mass (collision_procedures.cpp:668-694)
    // and moves the body along the contact normal; contacts on opposite sides cancel.  Kept per body: the largest single
    // share (push) and the vector sum of all shares (pushv); budget_entity takes the larger of the two magnitudes.
    const bool ma = !(__float_as_uint(W.prm1[ea].w) & EF_IMMUTABLE), mb = !(__float_as_uint(W.prm1[eb].w) & EF_IMMUTABLE);
    const float massA = W.prm0[ea].w, massB = W.prm0[eb].w;                        // 0 = infinite
    float shareA = 1.0f, shareB = 1.0f;
    if (massA == 0.0f && massB == 0.0f) { shareA = 0.0f; shareB = 0.0f; }
    else if (massA == 0.0f) shareA = 0.0f;
    else if (massB == 0.0f) shareB = 0.0f;
    else { shareA = massB / (massA + massB); shareB = massA / (massA + massB); }
    if (ma) {
        mark_dirty(W, ea);
        const float d = depth * shareA;
        if (d > 0.0f) { atomicMax((int*)&W.push[ea], __float_as_int(d)); atomicAdd(&W.pushv[ea].x, dir.x * d); atomicAdd(&W.pushv[ea].y, dir.y * d); atomicAdd(&W.pushv[ea].z, dir.z * d); }
    }
    if (mb) {
        mark_dirty(W, eb);
        const float d = depth * shareB;
        if (d > 0.0f) { atomicMax((int*)&W.push[eb], __float_as_int(d)); atomicAdd(&W.pushv[eb].x, -dir.x * d); atomicAdd(&W.pushv[eb].y, -dir.y * d); atomicAdd(&W.pushv[eb].z, -dir.z * d); }
    }
    }
}

// ---- per-body displacement budgets, the movers' extra candidates, candidate lists -------------
// The stored bounds S are the shapes' bounds inflated by the uniform W.margin: enough for a body that static resolution
// moves by less than that during the pass.  A body the speculative pass predicts to be pushed further (push[e] = sum of
// the depths of its speculative contacts), or that an earlier attempt caught moving further (bm_keep), gets a budget
// bm[e] > margin.  Candidates are the pairs with overlap(S_i (+) x_i, S_j (+) x_j), x = bm - margin: for x_i = x_j = 0 that
// is the flat list k_pairs emitted; the rest (xpairs) is found by the movers themselves, one warp each, walking the grid
// cells their widened bound can reach.  Nobody else pays for a mover.
#ifndef ICX_PREDICT_CAP
#define ICX_PREDICT_CAP 8.0f
#endif
__device__ __forceinline__ float predicted_push(const DeviceWorld& W, uint32_t e)
{
    const float4 v = W.pushv[e];
    return fmaxf(*(volatile float*)&W.push[e], fmaxf(fmaxf(fabsf(v.x), fabsf(v.y)), fabsf(v.z)));
}

static __device__ __forceinline__ void budget_entity(const DeviceWorld& W, uint32_t e)
{
    const float m0 = W.margin;
    const float k = W.bm_keep[e];
    const float p = predicted_push(W, e);
    float b = fmaxf(m0, k);
    // the prediction is trusted up to ICX_PREDICT_CAP margins (static resolution rarely moves a body further: measured p99.9);
    // beyond that only a body that was CAUGHT moving further keeps a larger budget (bm_keep)
    if (p > 0.0f) b = fmaxf(b, fminf(1.25f * p + 0.25f * m0, ICX_PREDICT_CAP * m0));
    b = fminf(b, m0 + 6.0f * W.cell_size);                                          // a mover's walk stays bounded (beyond it: the attempt limit reports it)
    if (!(b == b) || (__float_as_uint(W.prm1[e].w) & EF_IMMUTABLE)) b = m0;
    W.bm[e] = b;
    if (b > m0 && (W.first_packet[e] >> 30)) {
        W.mover_list[atomicAdd(&W.counters[CTR_MOVERS], 1u)] = e;
        atomicMax((int*)&W.counters[CTR_XMAX], __float_as_int(b - m0));
    }
}

__device__ __forceinline__ bool overlap_x(float4 alo, float4 ahi, float xa, float4 blo, float4 bhi, float xb)
{
    return alo.x - xa <= bhi.x + xb && blo.x - xb <= ahi.x + xa && alo.y - xa <= bhi.y + xb && blo.y - xb <= ahi.y + xa &&
           alo.z - xa <= bhi.z + xb && blo.z - xb <= ahi.z + xa;
}

__device__ __forceinline__ void emit_xpair(const DeviceWorld& W, uint32_t p, uint32_t q, bool q_immut)
{
    const uint32_t at = atomicAdd(&W.counters[CTR_XPAIRS], 1u);
    if (at >= W.max_xpairs) { atomicOr(W.flags, RF_PAIR_OVERFLOW); return; }
    // p belongs to a mover (mutable)
    W.xpairs[at] = p < q ? make_uint2(p, q | (q_immut ? ICX_PAIR_B_IMMUT : 0u)) : make_uint2(q | (q_immut ? ICX_PAIR_A_IMMUT : 0u), p);
}

// One warp per mover entity: every packet whose widened bound overlaps the mover's widened bound but whose stored bound
// does not (those pairs are in the flat list already).  A partner is binned by the centre of its stored bound, which lies
// within half a cell of that bound: the walk covers the cells within x_i + x_max + cell / 2 of the mover's stored bound.
// A pair of two movers is emitted by the one with the lower packet index.
static __device__ __noinline__ void mover_walk_warp(const DeviceWorld& W, uint32_t e, float x_max)
{
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t fp = W.first_packet[e];
    const uint32_t npk = fp >> 30, p0 = fp & 0x3FFFFFFFu;
    const uint32_t wl = W.worldline_stride ? e / W.worldline_stride : 0u;
    const uint32_t n_large = min(W.counters[CTR_LARGE], W.max_large);
    const float m0 = W.margin, xi = W.bm[e] - m0;
    for (uint32_t k = 0; k < npk; k++) {
        const uint32_t p = p0 + k;
        if (W.pk_cell[p] >= 0xFFFFFFFEu) continue;                                  // inert (large colliders are immutable: never movers)
        const float4 lo = W.pk_lo[p], hi = W.pk_hi[p];
        const int type = __float_as_uint(W.pk_lpos[p].w) & PF_TYPE_MASK;
        const float reach = xi + x_max + 0.5005f * W.cell_size;
        const float sx = reach + 4.0e-6f * fmaxf(fabsf(lo.x), fabsf(hi.x)), sy = reach + 4.0e-6f * fmaxf(fabsf(lo.y), fabsf(hi.y)),
                    sz = reach + 4.0e-6f * fmaxf(fabsf(lo.z), fabsf(hi.z));
        const int cx0 = (int)floorf((lo.x - sx) * W.inv_cell), cx1 = (int)floorf((hi.x + sx) * W.inv_cell);
        const int cy0 = (int)floorf((lo.y - sy) * W.inv_cell), cy1 = (int)floorf((hi.y + sy) * W.inv_cell);
        const int cz0 = (int)floorf((lo.z - sz) * W.inv_cell), cz1 = (int)floorf((hi.z + sz) * W.inv_cell);
        // one grid row (x-run of up to 16 cells at fixed y, z) per lane: the walk is a chain of dependent loads per row, 32 rows hide it
        const int nxb = (cx1 >> ICX_XB) - (cx0 >> ICX_XB) + 1, ny = cy1 - cy0 + 1, nz = cz1 - cz0 + 1;
        const uint32_t n_rows = (uint32_t)nxb * (uint32_t)ny * (uint32_t)nz;
        for (uint32_t row = lane; row < n_rows; row += 32u) {
            const int xb = (cx0 >> ICX_XB) + (int)(row % (uint32_t)nxb), cy = cy0 + (int)((row / (uint32_t)nxb) % (uint32_t)ny), cz = cz0 + (int)(row / ((uint32_t)nxb * (uint32_t)ny));
            const uint32_t block = row_block(xb, cy, cz, wl, W.n_buckets);
            const int x_lo = max(cx0, xb << ICX_XB), x_hi = min(cx1, (xb << ICX_XB) + ICX_XMASK);
            const uint32_t beg = W.cell_count[(block << ICX_XB) | (uint32_t)(x_lo & ICX_XMASK)];
            const uint32_t end = W.cell_count[((block << ICX_XB) | (uint32_t)(x_hi & ICX_XMASK)) + 1u];
            for (uint32_t j = beg; j < end; j++) {
                const float4 blo = W.s_lo[j], bhi = W.s_hi[j];
                const uint32_t btag = __float_as_uint(blo.w), be = __float_as_uint(bhi.w), q = btag & ICX_NB_INDEX_MASK;
                if (be == e || (W.worldline_stride && be / W.worldline_stride != wl) || !can_intersect(type, (btag >> 28) & 3, W.cfg_flags)) continue;
                // rows that collide in the bucket hash share a block: a packet counts only in the row its own cell is in
                if (cell_coord(blo.y, bhi.y, W.inv_cell) != cy || cell_coord(blo.z, bhi.z, W.inv_cell) != cz ||
                    (cell_coord(blo.x, bhi.x, W.inv_cell) >> ICX_XB) != xb) continue;
                const bool q_immut = (btag & 0x40000000u) != 0;
                const float xj = q_immut ? 0.0f : W.bm[be] - m0;
                if (xj > 0.0f && q < p) continue;                                  // two movers: the lower packet's walk emits the pair
                if (!overlap_x(lo, hi, xi, blo, bhi, xj) || overlap(lo, hi, blo, bhi)) continue;
                emit_xpair(W, p, q, q_immut);
            }
        }
        __syncwarp();
        for (uint32_t i = lane; i < n_large; i += 32u) {                           // large colliders (immutable, not in the grid)
            const uint32_t q = W.large_list[i], qe = W.pk_entity[q];
            if (qe == e || (W.worldline_stride && qe / W.worldline_stride != wl) ||
                !can_intersect(type, (int)(__float_as_uint(W.pk_lpos[q].w) & PF_TYPE_MASK), W.cfg_flags)) continue;
            const float4 qlo = W.pk_lo[q], qhi = W.pk_hi[q];
            if (!overlap_x(lo, hi, xi, qlo, qhi, 0.0f) || overlap(lo, hi, qlo, qhi)) continue;
            emit_xpair(W, p, q, true);
        }
    }
}

// Candidate lists of the dirty bodies [first, n_dirty): the adjacency of the flat pair list (+ the movers' extra pairs),
// every list ascending by packet index = the order the reference's pair loop meets the partners
// (collision_physics_relay.h:29-37).  Streaming pass over the pairs with one atomic append per dirty endpoint, then one
// warp per body sorts (and drops the duplicates the movers' walks may have produced).
// Lists live in a pool that is bump-allocated per tick (CTR_POOL): count pass, allocation, fill pass, then one warp per body
// sorts (lists of up to 32 entries in registers, longer ones by ranking into a fresh pool region).
// mode of a list build: plain entries; entries with the gap between the shapes (may_meet); or the same for the flagged slots only
enum { LISTS_PLAIN = 0, LISTS_GAPS = 1, LISTS_FLAGGED = 2 };
__device__ __forceinline__ bool slot_wanted(const DeviceWorld& W, uint32_t s, uint32_t first, uint32_t n_dirty, int mode, uint32_t always_from)
{
    // s = ICX_NONE (clean) fails the range test; in LISTS_FLAGGED mode the slots from always_from on have no list yet and are wanted too
    return s >= first && s < n_dirty && (mode != LISTS_FLAGGED || s >= always_from || W.slot_flag[s] != 0u);
}
__device__ __forceinline__ void list_visit(const DeviceWorld& W, uint32_t s, uint32_t e, uint32_t p, uint32_t other, uint32_t eo, bool o_immut, bool fill, uint32_t gapq)
{
    const uint32_t k = p - (W.first_packet[e] & 0x3FFFFFFFu);
    if (!fill) { atomicAdd(&W.nbr_count[s * 2u + k], 1u); return; }
    const uint4 info = W.slot_info[s];                                              // {.., n0 | n1 << 16, offset, ..} written by the allocation pass
    if (info.z == ICX_NONE) return;                                                 // pool full: the tick is re-run with a larger pool
    const uint32_t at = atomicAdd(&W.nbr_fill[s * 2u + k], 1u);
    W.nbr[info.z + (k ? (info.y & 0xFFFFu) : 0u) + at] = make_uint2(other | (o_immut ? ICX_NB_IMMUTABLE : 0u), eo | (gapq << 24));
}

// Lower bound of the distance between the shapes of packets pa, pb on the post-integration state, in steps of
// ICX_GAP_UNIT * margin (0 = touching, not a separating-axis pair, or not worth knowing).
__device__ __forceinline__ uint32_t gap_of(const DeviceWorld& W, uint32_t pa, uint32_t pb, uint32_t ea, uint32_t eb)
{
    const int tA = __float_as_uint(W.pk_lpos[pa].w) & PF_TYPE_MASK, tB = __float_as_uint(W.pk_lpos[pb].w) & PF_TYPE_MASK;
    if (!is_sat_pair(tA, tB)) return 0u;
    float value;
    if (sat_measure(W.pk_d + (size_t)pa * 6, W.pk_d + (size_t)pb * 6, v3(W.pos0[ea]), v3(W.pos0[eb]), tA, tB, 3.0e38f, value)) return 0u;
    const float steps = value / (ICX_GAP_UNIT * W.margin);
    return steps >= 255.0f ? 255u : (uint32_t)steps;                                // truncation: a lower bound stays one
}

static __device__ __noinline__ void sort_lists_warp(const DeviceWorld& W, uint32_t slot)
{
    const uint32_t FULL = 0xFFFFFFFFu, lane = threadIdx.x & 31u;
    uint4 info = W.slot_info[slot];
    if (info.z == ICX_NONE) return;
    const uint32_t n0 = info.y & 0xFFFFu, n1 = info.y >> 16;
    uint32_t new_off = info.z;
    if (n0 > 32u || n1 > 32u) {                                                     // the ranked copy becomes the list
        if (lane == 0) { new_off = atomicAdd(&W.counters[CTR_POOL], n0 + n1); if (new_off + n0 + n1 > W.max_nbr) { atomicOr(W.flags, RF_NEIGHBOR_OVERFLOW); new_off = ICX_NONE; } }
        new_off = __shfl_sync(FULL, new_off, 0);
        if (new_off == ICX_NONE) return;
    }
    for (uint32_t k = 0; k < 2u; k++) {
        const uint32_t cnt = k ? n1 : n0;
        uint2* list = W.nbr + info.z + (k ? n0 : 0u);
        if (cnt <= 32u) {
            const uint2 mine = lane < cnt ? list[lane] : make_uint2(0xFFFFFFFFu, 0u);
            const uint32_t key = mine.x & ICX_NB_INDEX_MASK;
            uint32_t rank = 0;
            for (uint32_t q = 0; q < cnt; q++) { const uint32_t other = __shfl_sync(FULL, key, q); rank += other < key ? 1u : 0u; }
            __syncwarp();
            if (lane < cnt) (new_off == info.z ? list : W.nbr + new_off + (k ? n0 : 0u))[rank] = mine;      // keys are distinct packet indices
        } else {
            uint2* dst = W.nbr + new_off + (k ? n0 : 0u);
            for (uint32_t i = lane; i < cnt; i += 32u) {
                const uint2 mine = list[i]; const uint32_t key = mine.x & ICX_NB_INDEX_MASK;
                uint32_t rank = 0;
                for (uint32_t q = 0; q < cnt; q++) rank += (list[q].x & ICX_NB_INDEX_MASK) < key ? 1u : 0u;
                dst[rank] = mine;
            }
        }
        __syncwarp();
    }
    if (lane == 0 && new_off != info.z) { info.z = new_off; W.slot_info[slot] = info; }
}

// Budgets of the dirty bodies (a body an earlier attempt caught beyond its budget first gets its carried floor raised), then
// the movers' own candidates.  Called by every thread of the grid; ends behind a grid barrier.
template <typename Grid>
__device__ __forceinline__ void budgets_and_movers(const DeviceWorld& W, Grid& grid, uint32_t tid, uint32_t nth, uint32_t n_dirty)
{
    for (uint32_t s = tid; s < n_dirty; s += nth) {             // clean bodies keep the uniform margin (k_integrate_prep)
        const uint32_t e = W.dirty_list[s];
        const uint32_t need = W.bm_need[e];
        if (need) { W.bm_need[e] = 0u; W.bm_keep[e] = fmaxf(W.bm_keep[e], __uint_as_float(need) * 2.0f + 0.02f); }
        budget_entity(W, e);
    }
    grid.sync();
    const uint32_t n_mov = min(*(volatile uint32_t*)&W.counters[CTR_MOVERS], W.n_entities);
    const float x_max = __uint_as_float(*(volatile uint32_t*)&W.counters[CTR_XMAX]);
    for (uint32_t m = tid >> 5; m < n_mov; m += (nth >> 5)) mover_walk_warp(W, W.mover_list[m], x_max);
    grid.sync();
}

template <typename Grid>
__device__ __forceinline__ void lists_build(const DeviceWorld& W, uint32_t first, uint32_t n_dirty, Grid& grid, uint32_t tid, uint32_t nth, int mode, uint32_t always_from)
{
    if (first >= n_dirty) return;                                                   // uniform
    for (uint32_t s = first + tid; s < n_dirty; s += nth) if (slot_wanted(W, s, first, n_dirty, mode, always_from)) { W.nbr_count[s * 2u] = 0u; W.nbr_count[s * 2u + 1u] = 0u; }
    grid.sync();
    const uint32_t n_pairs = min(*(volatile uint32_t*)&W.counters[CTR_CANDIDATES], W.max_pairs);
    const uint32_t n_x = min(*(volatile uint32_t*)&W.counters[CTR_XPAIRS], W.max_xpairs);
    for (int fill = 0; fill < 2; fill++) {
        for (uint32_t i = tid; i < n_pairs + n_x; i += nth) {
            const uint2 pr = i < n_pairs ? W.pairs[i] : W.xpairs[i - n_pairs];
            const uint32_t pa = pr.x & ICX_NB_INDEX_MASK, pb = pr.y & ICX_NB_INDEX_MASK;
            const bool ia = (pr.x & ICX_PAIR_A_IMMUT) != 0, ib = (pr.y & ICX_PAIR_B_IMMUT) != 0;
            if (ia && ib) continue;
            const uint32_t ea = W.pk_entity[pa], eb = W.pk_entity[pb];
            const uint32_t sa = ia ? ICX_NONE : W.dirty_slot[ea], sb = ib ? ICX_NONE : W.dirty_slot[eb];
            const bool wa = slot_wanted(W, sa, first, n_dirty, mode, always_from), wb = slot_wanted(W, sb, first, n_dirty, mode, always_from);
            if (!wa && !wb) continue;
            uint32_t gapq = 0u;
            if (fill && mode != LISTS_PLAIN && !ia && !ib) gapq = gap_of(W, pa, pb, ea, eb);   // both can move: how far apart are they (may_meet)?
            if (wa) list_visit(W, sa, ea, pa, pb, eb, ib, fill != 0, gapq);
            if (wb) list_visit(W, sb, eb, pb, pa, ea, ia, fill != 0, gapq);
        }
        grid.sync();
        if (fill) break;
        for (uint32_t s = first + tid; s < n_dirty; s += nth) {                     // allocation
            if (!slot_wanted(W, s, first, n_dirty, mode, always_from)) continue;
            const uint32_t c0 = W.nbr_count[s * 2u], c1 = W.nbr_count[s * 2u + 1u];
            uint32_t off = ICX_NONE;
            if (c0 <= 0xFFFFu && c1 <= 0xFFFFu) { off = atomicAdd(&W.counters[CTR_POOL], c0 + c1); if (off + c0 + c1 > W.max_nbr) off = ICX_NONE; }
            if (off == ICX_NONE) atomicOr(W.flags, RF_NEIGHBOR_OVERFLOW);
            const uint32_t e = W.dirty_list[s];
            W.slot_info[s] = make_uint4(W.first_packet[e], off == ICX_NONE ? 0u : (c0 | (c1 << 16)), off, e);
            W.nbr_fill[s * 2u] = 0u; W.nbr_fill[s * 2u + 1u] = 0u;
        }
        grid.sync();
    }
    for (uint32_t s = first + (tid >> 5); s < n_dirty; s += (nth >> 5)) if (slot_wanted(W, s, first, n_dirty, mode, always_from)) sort_lists_warp(W, s);
}

// The lists of the flagged slots among the listed ones [0, n_listed) -- the movers and the bodies their extra pairs name -- get
// those extra pairs.  Their adjacency in the flat pair list is already there (sorted): the old entries are copied into a fresh pool
// region, the extra pairs appended by a pass over the extra pairs alone, and the result sorted again.  No pass over the tick's pair
// list, no gaps (0 = unknown: may_meet then decides on the bounds alone); the caller uses it while the extra pairs are few.
template <typename Grid>
__device__ __forceinline__ void lists_extend_flagged(const DeviceWorld& W, uint32_t n_listed, Grid& grid, uint32_t tid, uint32_t nth)
{
    const uint32_t FULL = 0xFFFFFFFFu, lane = tid & 31u;
    const uint32_t n_x = min(*(volatile uint32_t*)&W.counters[CTR_XPAIRS], W.max_xpairs);
    for (uint32_t s = tid; s < n_listed; s += nth) if (W.slot_flag[s]) { W.nbr_count[s * 2u] = 0u; W.nbr_count[s * 2u + 1u] = 0u; }
    grid.sync();
    for (int fill = 0; fill < 2; fill++) {
        for (uint32_t i = tid; i < n_x; i += nth) {
            const uint2 pr = W.xpairs[i];
            const uint32_t pa = pr.x & ICX_NB_INDEX_MASK, pb = pr.y & ICX_NB_INDEX_MASK;
            const bool ia = (pr.x & ICX_PAIR_A_IMMUT) != 0, ib = (pr.y & ICX_PAIR_B_IMMUT) != 0;
            if (ia && ib) continue;
            const uint32_t ea = W.pk_entity[pa], eb = W.pk_entity[pb];
            const uint32_t sa = ia ? ICX_NONE : W.dirty_slot[ea], sb = ib ? ICX_NONE : W.dirty_slot[eb];
            const bool wa = sa < n_listed && W.slot_flag[sa] != 0u, wb = sb < n_listed && W.slot_flag[sb] != 0u;
            if (!wa && !wb) continue;
            const uint32_t gapq = 0u;                                               // unknown: may_meet decides on the bounds alone
            if (wa) list_visit(W, sa, ea, pa, pb, eb, ib, fill != 0, gapq);
            if (wb) list_visit(W, sb, eb, pb, pa, ea, ia, fill != 0, gapq);
        }
        grid.sync();
        if (fill) break;
        for (uint32_t s = tid >> 5; s < n_listed; s += (nth >> 5)) {                // one warp per flagged slot: new region, old entries first
            if (!W.slot_flag[s]) continue;
            const uint4 old = W.slot_info[s];
            if (old.z == ICX_NONE) continue;                                        // (the pool overflowed earlier: the tick is re-run)
            const uint32_t n0 = old.y & 0xFFFFu, n1 = old.y >> 16;
            const uint32_t c0 = n0 + W.nbr_count[s * 2u], c1 = n1 + W.nbr_count[s * 2u + 1u];
            if (c0 == n0 && c1 == n1) continue;                                     // a mover whose walk found nothing new: its list stands
            uint32_t off = ICX_NONE;
            if (lane == 0u) {
                if (c0 <= 0xFFFFu && c1 <= 0xFFFFu) { off = atomicAdd(&W.counters[CTR_POOL], c0 + c1); if (off + c0 + c1 > W.max_nbr) off = ICX_NONE; }
                if (off == ICX_NONE) atomicOr(W.flags, RF_NEIGHBOR_OVERFLOW);
            }
            off = __shfl_sync(FULL, off, 0);
            if (off != ICX_NONE) {
                for (uint32_t i = lane; i < n0 + n1; i += 32u) {
                    const uint2 en = W.nbr[old.z + i];
                    W.nbr[off + (i < n0 ? i : c0 + (i - n0))] = en;
                }
            }
            if (lane == 0u) {
                W.slot_info[s] = make_uint4(old.x, off == ICX_NONE ? 0u : (c0 | (c1 << 16)), off, old.w);
                W.nbr_fill[s * 2u] = n0; W.nbr_fill[s * 2u + 1u] = n1;              // the extra pairs go behind the old entries
            }
        }
        grid.sync();
    }
    for (uint32_t s = tid >> 5; s < n_listed; s += (nth >> 5)) if (W.slot_flag[s]) sort_lists_warp(W, s);
}

// Put a dirty body back to its post-integration state (bit-identical recomputation from the tick input).
static __device__ __noinline__ void restore_body(const DeviceWorld& W, uint32_t e, double dt, int do_euler)
{
    const float4 p1 = W.prm1[e];
    const EntityRegs r = euler_entity(W, e, __float_as_uint(p1.w), p1, dt, do_euler);
    W.out.pos[e] = r.pos; W.out.rot[e] = r.rot; W.out.vel[e] = r.vel; W.out.acc[e] = r.acc;
    W.out.angvel[e] = r.av; W.out.angacc[e] = r.aa;
    W.out.ts[e] = W.in.ts[e];
    const uint32_t fp = W.first_packet[e];
    for (uint32_t k = 0; k < (fp >> 30); k++) W.pk_minparam[(fp & 0x3FFFFFFFu) + k] = 1.0f;
}

// One warp per dirty body: hand my predicted push on to the candidates it can reach (see resolve_body).  first_time:
// the body's list is new, so clean partners within reach are marked too; otherwise only pushes that matter are passed on.
static __device__ __noinline__ void predict_body_warp(const DeviceWorld& W, uint32_t sl, bool first_time)
{
    const uint32_t e = W.dirty_list[sl];
    const float pa = predicted_push(W, e);
    if (!(pa > 0.0f)) return;
    const float m0 = W.margin;
    if (!first_time && pa < 0.5f * m0) return;                                      // cannot lift anyone above the uniform margin (measured: skipping small pushes
                                                                                    // on the first pass too costs more in restarts than the pass itself)
    const uint4 info = W.slot_info[sl];
    if (info.z == ICX_NONE) return;
    const uint32_t p0 = info.x & 0x3FFFFFFFu, n0 = info.y & 0xFFFFu, n1 = info.y >> 16;
    const float ma = W.prm0[e].w;                                                   // 0 = infinite
    const float reach = 1.5f * pa + 0.02f * m0;
    for (uint32_t i = threadIdx.x & 31u; i < n0 + n1; i += 32u) {
        const uint32_t k = i < n0 ? 0u : 1u;
        const uint2 en = W.nbr[info.z + i];
        if (en.x & ICX_NB_IMMUTABLE) continue;
        const uint32_t o = en.x & ICX_NB_INDEX_MASK, eo = en.y & ICX_NB_ENT_MASK;
        const bool clean = W.dirty_slot[eo] == ICX_NONE;
        if (!clean && pa < 0.5f * m0) continue;
        const uint32_t p = p0 + k;
        const int tP = __float_as_uint(W.pk_lpos[p].w) & PF_TYPE_MASK, tO = __float_as_uint(W.pk_lpos[o].w) & PF_TYPE_MASK;
        if (!is_sat_pair(tP, tO)) continue;
        float value;
        const bool touching = sat_measure(W.pk_d + (size_t)p * 6, W.pk_d + (size_t)o * 6, v3(W.out.pos[e]), v3(W.out.pos[eo]), tP, tO, reach, value);
        if (!(touching || value < reach)) continue;
        if (clean) mark_dirty(W, eo);
        const float mb = W.prm0[eo].w;
        const float share = mb == 0.0f ? 0.0f : (ma == 0.0f ? 1.0f : ma / (ma + mb));     // the partner's part of a penetration between the two
        const float inherit = (touching ? pa : pa - value) * share;
        if (inherit > 0.0f) {
            const float before = __int_as_float(atomicMax((int*)&W.push[eo], __float_as_int(inherit)));
            if (inherit > before + 0.25f * m0) atomicAdd(&W.counters[CTR_PUSHED], 1u);          // worth another pass
        }
    }
}

// ---- dependency rounds -----------------------------------------------------------------
// A dirty body's cursor always rests on an entry that needs attention: a one-sided pair (clean or
// immutable partner) the filter could not reject, or a two-sided pair (dirty partner).  One round =
//   select   (one lane per active body)  one-sided pair -> work item; two-sided pair -> ready iff
//            the partner's cursor rests on the same pair (both read the round-start cursors); the
//            lower packet's lane then either consumes a certain miss or queues a work item
//   execute  exact narrowphase + response, then both cursors advance past the one-sided pairs the
//            filter rejects against the bodies' new state.  Thin rounds: one WARP per work item
//            (warp-cooperative SAT and cursor walk, lane 0 finishes).  Fat rounds (>= ICX_LANE_MIN
//            items per warp): one LANE per work item, 32 items per warp in plain SIMT form -- with
//            one lane active a warp spends its time streaming the ~85 KB procedure through the
//            instruction cache, which more warps per SM did not hide (measured)
// separated by grid barriers.  Cursors are double buffered (head[parity]); each active body has
// exactly one writer per round.


// A two-sided entry (dirty partner) normally waits until it is at the head of BOTH queues.  Not if the pair cannot hit
// wherever the partner may be: the partner's bound stays inside its stored bound widened by its budget for the whole pass
// (store_side enforces it), so if MY current bound -- translated by what static resolution moved me so far -- misses that
// region, or the gap between the two shapes after integration exceeds what I have moved plus what the partner can move, the
// exact test would miss too and nothing changes.  The partner learns it from my cursor having passed the
// entry (round_select), or finds the same on its own.  A mover's long list then costs a walk, not a chain of rounds.
__device__ __forceinline__ bool may_meet(const DeviceWorld& W, float4 alo, float4 ahi, float moved, float4 blo, float4 bhi, uint32_t ey)
{
    const float bmj = W.bm[ey & ICX_NB_ENT_MASK], m0 = W.margin;
    if (!overlap_x(alo, ahi, 0.0f, blo, bhi, bmj - m0)) return false;
    // the shapes were at least gap apart; I have moved `moved` (Euclidean), the partner at most sqrt(3) * its budget
    const float gap = (float)(ey >> 24) * (ICX_GAP_UNIT * m0);
    return !(gap > moved + 1.7320509f * bmj + 1.0e-4f);
}

// Advance a body's cursor past the leading one-sided entries that cannot hit on the current state
// (own state is final until the body's next exact pair, one-sided partners are never written).
// Done by a whole WARP (uniform arguments and result).  Pass 1, one entry per lane: two-sided
// entries (dirty partner) stop the walk; of the others only those whose un-inflated bounds overlap
// -- mine translated by what static resolution moved me so far -- are candidates (the test that
// defines ICX_PAIR_TIGHT).  Pass 2 replays the merged queue order over the 32 verdicts and gives
// each candidate it meets to the warp-cooperative SAT filter.  kind: 0 = queue exhausted,
// 1 = rests on a two-sided entry, 2 = rests on a one-sided pair to execute (list k_out, packet o_out).
__device__ __forceinline__ uint32_t skip_rejects_warp(const DeviceWorld& W, const Cursor& c, uint32_t hw, uint32_t n_dirty,
                                                      int& kind, uint32_t& k_out, uint32_t& o_out)
{
    const uint32_t FULL = 0xFFFFFFFFu, lane = threadIdx.x & 31u;
    const float4 pe = W.out.pos[c.e], pe0 = W.pos0[c.e];
    const float dx = pe.x - pe0.x, dy = pe.y - pe0.y, dz = pe.z - pe0.z;
    const float moved = sqrtf(dx * dx + dy * dy + dz * dz) * 1.000001f;
    kind = 0; k_out = 0; o_out = 0;
    for (;;) {
        const uint32_t h0 = hw & 0xFFFFu, h1 = hw >> 16;
        const uint32_t rem0 = h0 < c.n0 ? c.n0 - h0 : 0u, rem1 = h1 < c.n1 ? c.n1 - h1 : 0u;
        if (rem0 + rem1 == 0u) return hw;
        const uint32_t w0 = rem1 == 0u ? 32u : (rem0 == 0u ? 0u : 16u);        // lanes [0, w0) look at list 0, the rest at list 1
        const uint32_t k = lane < w0 ? 0u : 1u, idx = k ? lane - w0 : lane;
        const bool valid = idx < (k ? rem1 : rem0);
        uint32_t entry = 0, ent = 0; bool two = false, cand = false;
        if (valid) {
            PROF_ADD(3, 1);
            const uint2 en = k ? c.l1[h1 + idx] : c.l0[h0 + idx];
            entry = en.x; ent = en.y;
            const uint32_t p = c.p0 + k, o = entry & ICX_NB_INDEX_MASK;
            const uint32_t ds = (entry & ICX_NB_IMMUTABLE) ? ICX_NONE : W.dirty_slot[en.y & ICX_NB_ENT_MASK];
            float4 alo = W.pk_lo[p], ahi = W.pk_hi[p];
            const float4 blo = W.pk_lo[o], bhi = W.pk_hi[o];
            alo.x += dx; alo.y += dy; alo.z += dz; ahi.x += dx; ahi.y += dy; ahi.z += dz;
            two = ds < n_dirty && may_meet(W, alo, ahi, moved, blo, bhi, en.y);
            cand = !(ds < n_dirty) && tight_overlap(alo, ahi, blo, bhi, W.margin);
        }
        const uint32_t twos = __ballot_sync(FULL, two), cands = __ballot_sync(FULL, cand);
        uint32_t a0 = 0, a1 = 0; bool boundary = false;
        for (;;) {                                                              // merged order (cursor_head) over the chunk
            const bool has0 = a0 < rem0, has1 = a1 < rem1;
            if (!has0 && !has1) break;
            if ((has0 && a0 >= w0) || (has1 && a1 >= 32u - w0)) { boundary = true; break; }
            const uint32_t la = has0 ? a0 : 0u, lb = has1 ? w0 + a1 : 0u;
            const uint32_t ea = __shfl_sync(FULL, entry, la), eb = __shfl_sync(FULL, entry, lb);
            uint32_t pick;
            if (has0 && has1) {
                const uint32_t ao = ea & ICX_NB_INDEX_MASK, bo = eb & ICX_NB_INDEX_MASK;
                const bool alow = ao < c.p0, blow = bo < c.p0;
                if (alow || blow) pick = (alow && blow) ? (bo < ao ? 1u : 0u) : (alow ? 0u : 1u);
                else pick = 0u;
            } else pick = has0 ? 0u : 1u;
            const uint32_t bit = pick ? lb : la;
            if (twos >> bit & 1u) { kind = 1; break; }
            if (cands >> bit & 1u) {
                const uint32_t o = (pick ? eb : ea) & ICX_NB_INDEX_MASK, eo = __shfl_sync(FULL, ent, bit) & ICX_NB_ENT_MASK, p = c.p0 + pick;
                const bool mh = p < o ? may_hit_warp(W, p, o, c.e, eo) : may_hit_warp(W, o, p, eo, c.e);
                if (mh) { kind = 2; k_out = pick; o_out = o; break; }
            }
            if (pick) a1++; else a0++;
        }
        hw = (h0 + a0) | ((h1 + a1) << 16);
        if (!boundary) return hw;
    }
}


// Publish a body's cursor for the next round.  A body whose queue is exhausted leaves the work lists; its final cursor goes
// into BOTH buffers, because a partner that arrives at their common entry later must find it passed (round_select).
__device__ __forceinline__ void publish_cursor(const DeviceWorld& W, uint32_t s, const Cursor& c, uint32_t hw, int parity, uint32_t ctr_out)
{
    W.head[parity ^ 1][s] = hw;
    if (cursor_pending(c, hw)) W.active[parity ^ 1][atomicAdd(&W.counters[ctr_out], 1u)] = s;
    else W.head[parity][s] = hw;
}

// warp version of settle: lane 0 publishes the cursor
__device__ __forceinline__ void settle_warp(const DeviceWorld& W, uint32_t s, const Cursor& c, uint32_t hw, uint32_t n_dirty, int parity, uint32_t ctr_out)
{
    int kind; uint32_t kk, oo;
    hw = skip_rejects_warp(W, c, hw, n_dirty, kind, kk, oo);
    if ((threadIdx.x & 31u) == 0u) publish_cursor(W, s, c, hw, parity, ctr_out);
}

// The walk on ONE lane (lane-parallel items): the merged queue order is cursor_head's, the verdict per entry is the
// warp version's -- two-sided entries stop the walk, one-sided ones need overlapping un-inflated bounds (mine translated
// by what static resolution moved me) and the conservative filter, here sat_may_hit itself (same arithmetic per axis).
__device__ __forceinline__ uint32_t skip_rejects_lane(const DeviceWorld& W, const Cursor& c, uint32_t hw, uint32_t n_dirty,
                                                      int& kind, uint32_t& k_out, uint32_t& o_out)
{
    const float4 pe = W.out.pos[c.e], pe0 = W.pos0[c.e];
    const float dx = pe.x - pe0.x, dy = pe.y - pe0.y, dz = pe.z - pe0.z;
    const float moved = sqrtf(dx * dx + dy * dy + dz * dz) * 1.000001f;
    kind = 0; k_out = 0; o_out = 0;
#pragma unroll 1
    for (;;) {
        uint32_t k, entry, ey;
        if (!cursor_head(c, hw, k, entry, ey)) return hw;
        const uint32_t eo = ey & ICX_NB_ENT_MASK;
        const uint32_t p = c.p0 + k, o = entry & ICX_NB_INDEX_MASK;
        const uint32_t ds = (entry & ICX_NB_IMMUTABLE) ? ICX_NONE : W.dirty_slot[eo];
        float4 alo = W.pk_lo[p], ahi = W.pk_hi[p];
        const float4 blo = W.pk_lo[o], bhi = W.pk_hi[o];
        alo.x += dx; alo.y += dy; alo.z += dz; ahi.x += dx; ahi.y += dy; ahi.z += dz;
        if (ds < n_dirty) {
            if (may_meet(W, alo, ahi, moved, blo, bhi, ey)) { kind = 1; return hw; }
        } else if (tight_overlap(alo, ahi, blo, bhi, W.margin)) {
            const uint32_t qa = min(p, o), qb = max(p, o), ea = p < o ? c.e : eo, eb = p < o ? eo : c.e;
            const int tA = __float_as_uint(W.pk_lpos[qa].w) & PF_TYPE_MASK, tB = __float_as_uint(W.pk_lpos[qb].w) & PF_TYPE_MASK;
            const bool mh = !is_sat_pair(tA, tB)
                            || sat_may_hit(W.pk_d + (size_t)qa * 6, W.pk_d + (size_t)qb * 6, v3(W.out.pos[ea]), v3(W.out.pos[eb]), tA, tB, 0.0f);
            if (mh) { kind = 2; k_out = k; o_out = o; return hw; }
        }
        hw = cursor_advance(hw, k);
    }
}

__device__ __forceinline__ void round_select(const DeviceWorld& W, uint32_t s, uint32_t n_dirty, int parity, uint32_t ctr_out, uint32_t ctr_items)
{
    const volatile uint32_t* hin = W.head[parity];      // round-start cursors; a body that finishes during the round may already publish here (publish_cursor)
    const Cursor c = load_cursor(W, s);
    const uint32_t hw = hin[s];
    uint32_t k, entry, eo;
    if (!cursor_head(c, hw, k, entry, eo)) return;                                  // (not queued when nothing is pending)
    const uint32_t p = c.p0 + k, o = entry & ICX_NB_INDEX_MASK;
    uint32_t so = ICX_NONE;
    if (!(entry & ICX_NB_IMMUTABLE)) so = W.dirty_slot[eo & ICX_NB_ENT_MASK];
    if (so >= n_dirty) {                                                            // one-sided, not rejected by the filter
        W.items[atomicAdd(&W.counters[ctr_items], 1u)] = make_uint4(min(p, o), max(p, o), s, ICX_NONE);
        return;
    }
    const Cursor co = load_cursor(W, so);
    const uint32_t hwo = hin[so];
    uint32_t ko, entryo, eoo;
    const bool ready = cursor_head(co, hwo, ko, entryo, eoo) && co.p0 + ko == o && (entryo & ICX_NB_INDEX_MASK) == p;
    if (!ready) {
        // Has the partner's queue passed my entry?  Its lists are ascending by packet index and consumed front to back, so
        // the entry for my packet p in the list of its packet o is gone iff that list's head is beyond p.  Then the partner
        // found that the pair cannot hit wherever I am (may_meet): a miss for me too -- I walk on.
        const uint32_t kq = o - co.p0, hq = kq ? (hwo >> 16) : (hwo & 0xFFFFu), nq = kq ? co.n1 : co.n0;
        const bool passed = hq >= nq || p < ((kq ? co.l1 : co.l0)[hq].x & ICX_NB_INDEX_MASK);
        uint32_t hn = hw;
        if (passed) { int kind; uint32_t kk, oo; hn = skip_rejects_lane(W, c, cursor_advance(hw, k), n_dirty, kind, kk, oo); }
        publish_cursor(W, s, c, hn, parity, ctr_out);                               // (not passed: wait, the cursor is carried over)
        return;
    }
    if (p > o) return;                                                              // the lower packet's lane acts for both
    // the SAT filter is the executing warp's job (one axis per lane instead of fifteen in a row on this lane):
    // a pair it rejects is consumed on both sides without the exact procedure
    W.items[atomicAdd(&W.counters[ctr_items], 1u)] = make_uint4(p, o | ICX_ITEM_CHECK, s, so);
}

// One work item, by a whole warp: the exact pair (unless the filter already rejected it), then the
// cursors of its one or two dirty bodies move on.  The cursor records of both bodies are requested
// BEFORE the exact procedure and arrive while it runs.  One-sided pairs that follow need no
// partner and are executed at once -- through the same process_pair call site: the procedure is
// ~85 KB of code and must exist once.
__device__ __forceinline__ void round_execute(const DeviceWorld& W, uint4 item, uint32_t n_dirty, int parity, uint32_t ctr_out, uint32_t ctr_new, V3* scratch)
{
    const uint32_t FULL = 0xFFFFFFFFu, lane = threadIdx.x & 31u;
#ifdef ICX_PROFILE_PAIR
    const long long t0 = clock64();
    long long t_pair = 0;
#endif
    const uint32_t my_s = lane == 0u ? item.z : (lane == 1u ? item.w : ICX_NONE);
    uint4 my_info = make_uint4(0u, 0u, 0u, 0u); uint32_t my_hw = 0u;
    if (my_s != ICX_NONE) { my_info = W.slot_info[my_s]; my_hw = W.head[parity][my_s]; }

    uint32_t pa = item.x, pb = item.y & ICX_NB_INDEX_MASK;
    bool run = true;
    if (item.y & ICX_ITEM_CHECK) run = may_hit_warp(W, pa, pb, W.pk_entity[pa], W.pk_entity[pb]);
    int side = 0; bool entered = false;                                             // body being walked; its cursor is loaded
    Cursor c; uint32_t hw = 0, k_run = 0;
    c.p0 = c.npk = c.n0 = c.n1 = c.e = 0; c.l0 = c.l1 = nullptr;
    for (;;) {
        if (run) {
#ifdef ICX_PROFILE_PAIR
            const long long tq = clock64();
#endif
            process_pair<false>(W, pa, pb, n_dirty, ctr_new, scratch);
            __syncwarp();
#ifdef ICX_PROFILE_PAIR
            t_pair += clock64() - tq;
#endif
        }
        run = false;
        for (; side < 2; side++) {
            const uint32_t sb = side ? item.w : item.z;
            if (sb == ICX_NONE) continue;
            if (!entered) {                                                         // step over the entry just executed
                uint4 info; info.x = __shfl_sync(FULL, my_info.x, side); info.y = __shfl_sync(FULL, my_info.y, side);
                info.z = __shfl_sync(FULL, my_info.z, side); info.w = __shfl_sync(FULL, my_info.w, side);
                c = make_cursor(W, sb, info);
                hw = __shfl_sync(FULL, my_hw, side);
                uint32_t k, entry, eo;
                cursor_head(c, hw, k, entry, eo);
                hw = cursor_advance(hw, k);
                entered = true;
            } else hw = cursor_advance(hw, k_run);                                  // a follow-up pair was executed
            int kind; uint32_t kk, oo;
            hw = skip_rejects_warp(W, c, hw, n_dirty, kind, kk, oo);
            if (kind == 2) {                                                        // one-sided pair the filter keeps: run it now
                PROF_ADD(4, 1);
                const uint32_t p = c.p0 + kk;
                k_run = kk; pa = min(p, oo); pb = max(p, oo); run = true;
                break;
            }
            if (lane == 0u) publish_cursor(W, sb, c, hw, parity, ctr_out);
            entered = false;
        }
        if (!run) break;
    }
#ifdef ICX_PROFILE_PAIR
    PROF_ADD(0, 1); PROF_ADD(1, t_pair >> 4); PROF_ADD(2, (clock64() - t0 - t_pair) >> 4);
#endif
}

// Up to 32 work items by one warp, one per lane, from the exact pair to the end of its cursor walks (the scalar form of
// round_execute).  All lanes meet at ONE process_pair call site per iteration: a lane whose walk found a one-sided pair
// to execute brings it to the next iteration, the others idle there.
__device__ __forceinline__ void round_execute_lanes(const DeviceWorld& W, uint4 item, bool active, uint32_t n_dirty, int parity, uint32_t ctr_out, uint32_t ctr_new)
{
    const uint32_t FULL = 0xFFFFFFFFu;
    uint32_t pa = item.x, pb = item.y & ICX_NB_INDEX_MASK;
    bool run = active, done = !active, entered = false;
    int side = 0;
    Cursor c; uint32_t hw = 0, k_run = 0;
    c.p0 = c.npk = c.n0 = c.n1 = c.e = 0; c.l0 = c.l1 = nullptr;
#pragma unroll 1
    while (__any_sync(FULL, !done)) {
        process_pair<true>(W, pa, pb, n_dirty, ctr_new, nullptr, run && !done);
        if (done) continue;
        run = false;
#pragma unroll 1
        for (; side < 2; side++) {
            const uint32_t sb = side ? item.w : item.z;
            if (sb == ICX_NONE) continue;
            if (!entered) {                                                         // step over the entry just executed
                c = load_cursor(W, sb);
                hw = W.head[parity][sb];
                uint32_t k, entry, eo;
                cursor_head(c, hw, k, entry, eo);
                hw = cursor_advance(hw, k);
                entered = true;
            } else hw = cursor_advance(hw, k_run);                                  // a follow-up pair was executed
            int kind; uint32_t kk, oo;
            hw = skip_rejects_lane(W, c, hw, n_dirty, kind, kk, oo);
            if (kind == 2) {                                                        // one-sided pair the filter keeps: next iteration
                const uint32_t p = c.p0 + kk;
                k_run = kk; pa = min(p, oo); pb = max(p, oo); run = true;
                break;
            }
            publish_cursor(W, sb, c, hw, parity, ctr_out);
            entered = false;
        }
        if (!run) done = true;
    }
}

// ---------------------------------------------------------------------------------------
// K6: ordered resolve of the dirty bodies.  Persistent cooperative kernel (grid-wide barriers
// between dependency rounds, no host round trips); see the file header.
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ void trace_mark(const DeviceWorld& W, uint32_t& n, uint32_t tag)
{
    if (!W.trace || n + 2 > ICX_TRACE_CAP) return;
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    W.trace[n++] = t; W.trace[n++] = tag;
}

// The work before the rounds (candidate lists, push prediction, first cursor positions) is throughput-bound
// and light on registers; the rounds are latency-bound and want every register they can get (no spills in
// the exact procedures).  So the first attempt is split over two cooperative launches with different shapes:
// PHASE 0 = prepare (2 x 384 threads per SM), PHASE 1 = rounds (1 x 256 threads per SM, 251 registers) --
// which also carries the whole loop for the rare restarts.
#define ICX_MAX_ATTEMPTS 96u

#ifndef ICX_LANE_MIN
#define ICX_LANE_MIN 4u          // items per warp from which a round runs lane-parallel (break-even measured at ~4: the lane-parallel pass costs ~40 us flat, a warp-mode item ~19 us, a cursor walk ~9.5 us)
#endif
#ifndef RESOLVE_TPB
#define RESOLVE_TPB 256
#endif
#ifndef PREPARE_TPB
#define PREPARE_TPB 384
#define PREPARE_MINB 2
#endif
template <int PHASE>
__device__ __forceinline__ void resolve_body(const DeviceWorld& W, double dt, int do_euler, V3* scratch)
{
    uint32_t tn = 2;        // trace[0] = number of words used
    if (PHASE == 1 && W.trace) tn = (uint32_t)W.trace[0];
    cg::grid_group grid = cg::this_grid();
    const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x, nth = gridDim.x * blockDim.x;
    volatile uint32_t* ctr = W.counters;
    uint32_t attempts = 0, rounds = 0;
    bool rebudget = false;
    if (PHASE != 1 && tid == 0) trace_mark(W, tn, 0u);
    if (PHASE != 1) {
        // Predict how far static resolution may push each body, and with it which clean bodies the replay will reach.
        // push[e] starts as the sum of the depths of e's speculative contacts (k_narrow).  A body pushed by up to p reaches
        // the candidates whose gap is smaller and hands them the remainder, scaled by its share of the two masses (static
        // resolution splits a penetration by inverse mass, collision_procedures.cpp:668-694); clean ones become dirty.
        // Up to four passes over the candidate lists of the uniform margin.  Then budgets (bm = what the lists must cover),
        // the movers' own wider candidates, and -- only if there are any -- the lists once more with those included.
        // Only a prediction: a body it misses, or a budget that proves too small, restarts the replay further down.
        uint32_t done = 0;
        for (int sweep = 0; sweep < 3; sweep++) {
            const uint32_t n_dirty = ctr[CTR_DIRTY];
            const uint32_t changed = sweep ? ctr[CTR_PUSHED] : 1u;
            grid.sync();                                        // every thread has read the counters before thread 0 resets one
            // another pass only if it matters: many new bodies to list, or many pushes handed on (uniform)
            if (sweep && (n_dirty - done) * ICX_SWEEP_MIN_SHARE < n_dirty && changed * 100u < n_dirty) break;
            if (tid == 0) { ctr[CTR_PUSHED] = 0u; __threadfence(); }
            lists_build(W, done, n_dirty, grid, tid, nth, LISTS_PLAIN, 0u);
            if (tid == 0) trace_mark(W, tn, 0x60000u | min(n_dirty - done, 0xFFFFu));
            grid.sync();
            for (uint32_t sl = (tid >> 5); sl < n_dirty; sl += (nth >> 5)) predict_body_warp(W, sl, sl >= done);
            done = n_dirty;
            grid.sync();
            if (tid == 0) trace_mark(W, tn, 0x70000u | (uint32_t)sweep);                       // prediction sweep done
        }
        {
            const uint32_t n_dirty = ctr[CTR_DIRTY];
            budgets_and_movers(W, grid, tid, nth, n_dirty);
            if (tid == 0) trace_mark(W, tn, 0x80000u | min(ctr[CTR_MOVERS], 0xFFFFu));
            const uint32_t n_mov = min(ctr[CTR_MOVERS], W.n_entities), n_x = min(ctr[CTR_XPAIRS], W.max_xpairs);
            if (n_mov != 0u) {                                  // (uniform) the lists of the movers and of the bodies their extra candidates name
                for (uint32_t s = tid; s < n_dirty; s += nth) W.slot_flag[s] = 0u;      // again, now with those candidates and with the gaps may_meet prunes by
                grid.sync();
                for (uint32_t m = tid; m < n_mov; m += nth) { const uint32_t s = W.dirty_slot[W.mover_list[m]]; if (s < n_dirty) atomicMax(&W.slot_flag[s], 2u); }   // 2 = a mover: its list gets the gaps too
                for (uint32_t i = tid; i < n_x; i += nth) {
                    const uint2 pr = W.xpairs[i];
                    if (!(pr.x & ICX_PAIR_A_IMMUT)) { const uint32_t s = W.dirty_slot[W.pk_entity[pr.x & ICX_NB_INDEX_MASK]]; if (s < n_dirty) atomicMax(&W.slot_flag[s], 1u); }
                    if (!(pr.y & ICX_PAIR_B_IMMUT)) { const uint32_t s = W.dirty_slot[W.pk_entity[pr.y & ICX_NB_INDEX_MASK]]; if (s < n_dirty) atomicMax(&W.slot_flag[s], 1u); }
                }
                grid.sync();
                // Many extra pairs (a pile): the flagged lists are built again from the pair list, with the gap of every two-sided
                // pair measured once.  A handful (the sparse case, every tick of it): the extra pairs are appended to the lists that
                // exist and the gaps are left out -- measuring them would put ~100 us of cold-code latency on the critical path.
                const bool rebuild = n_x >= ICX_GAPS_MIN_XPAIRS;                    // (uniform)
                if (rebuild) lists_build(W, 0u, n_dirty, grid, tid, nth, LISTS_FLAGGED, done);      // (+ plain lists of the bodies the last pass marked: slots >= done)
                else lists_extend_flagged(W, done, grid, tid, nth);            // (those bodies get their lists -- extra pairs included -- with the attempt's)
                grid.sync();
                // ... and the clean bodies a mover can reach through them are replayed in order too
                const uint32_t listed = rebuild ? n_dirty : done;
                for (uint32_t m = (tid >> 5); m < n_mov; m += (nth >> 5)) {
                    const uint32_t sl = W.dirty_slot[W.mover_list[m]];
                    if (sl < listed) predict_body_warp(W, sl, true);
                }
                done = listed;
                grid.sync();
            }
        }
        if (tid == 0) ctr[CTR_LISTED] = done;
        grid.sync();
    }
    for (;;) {
        const uint32_t n_dirty = ctr[CTR_DIRTY], n_listed = ctr[CTR_LISTED], n_static = min(ctr[CTR_STATIC], W.max_static);
        if (PHASE == 1 && attempts == 0) {                  // the prepare launch did all of this for the first attempt
            if (*(volatile uint32_t*)W.flags & RF_NEIGHBOR_OVERFLOW) break;
        } else {
        if (rebudget) {                                     // (uniform) some body moved beyond its budget: new budgets, the movers walk again, all lists again
            if (tid == 0) { ctr[CTR_MOVERS] = 0u; ctr[CTR_XPAIRS] = 0u; ctr[CTR_XMAX] = 0u; ctr[CTR_VIOLATORS] = 0u; ctr[CTR_POOL] = 0u; __threadfence(); }
            grid.sync();
            budgets_and_movers(W, grid, tid, nth, n_dirty);
            lists_build(W, 0u, n_dirty, grid, tid, nth, LISTS_GAPS, 0u);
        } else lists_build(W, n_listed, n_dirty, grid, tid, nth, LISTS_PLAIN, 0u);
        for (uint32_t s = tid; s < n_dirty; s += nth)
            if (attempts > 0 && s < n_listed) restore_body(W, W.dirty_list[s], dt, do_euler);
        if (tid == 0) trace_mark(W, tn, 0x10000u | min(n_dirty, 0xFFFFu));
        grid.sync();
        if (tid == 0) trace_mark(W, tn, 0x20000u);
        if (*(volatile uint32_t*)W.flags & RF_NEIGHBOR_OVERFLOW) break;     // lists are incomplete: the host grows them and re-runs the tick
        if (tid == 0) {
            ctr[CTR_LISTED] = n_dirty; ctr[CTR_CONTACTS] = 0; ctr[CTR_INTERCEPTS] = 0; ctr[CTR_GROUNDS] = 0;
            ctr[CTR_NEWD] = 0; ctr[CTR_NEWD + 1] = 0; ctr[CTR_NEWD + 2] = 0;
            ctr[CTR_ACTIVE] = 0; ctr[CTR_ACTIVE + 1] = 0; ctr[CTR_ACTIVE + 2] = 0;
            ctr[CTR_ITEMS] = 0; ctr[CTR_ITEMS + 1] = 0; ctr[CTR_FETCH] = 0; ctr[CTR_FETCH + 1] = 0;
            __threadfence();
        }
        grid.sync();
        // static pairs (contact records only) join the first round's work items; every body's cursor goes to its first entry of interest
        for (uint32_t i = tid; i < n_static; i += nth) { const uint2 pr = W.static_pairs[i]; W.items[atomicAdd(&W.counters[CTR_ITEMS], 1u)] = make_uint4(pr.x, pr.y, ICX_NONE, ICX_NONE); }
        for (uint32_t s = tid; s < n_dirty; s += nth) {                              // one lane per body (the walk is a chain of dependent loads:
            const Cursor c = load_cursor(W, s);                                     // 32 bodies per warp hide it better than 32 entries of one body)
            int kind; uint32_t kk, oo;
            const uint32_t hw = skip_rejects_lane(W, c, 0u, n_dirty, kind, kk, oo);
            W.head[0][s] = hw; W.head[1][s] = hw;
            if (cursor_pending(c, hw)) W.active[0][atomicAdd(&W.counters[CTR_ACTIVE], 1u)] = s;   // writes head[0] / active[0]
        }
        grid.sync();
        if (tid == 0) trace_mark(W, tn, 0x50000u);
        }
        if (PHASE == 0) { if (tid == 0 && W.trace) W.trace[0] = tn; return; }
        int parity = 0;
        rounds = 0;
        for (;;) {
            const uint32_t n_act = ctr[CTR_ACTIVE + rounds % 3u];
            const uint32_t grew = rounds ? ctr[CTR_NEWD + (rounds + 2u) % 3u] : 0u;    // written in the previous round only
            if ((n_act == 0u && !(rounds == 0u && n_static)) || grew != 0u) break;       // done, or the dirty set grew: restart at once
            // (a body beyond its budget does not stop the attempt: finishing it catches the others of its kind in ONE restart)
            if (rounds >= ICX_MAX_ROUNDS) { if (tid == 0) atomicOr(W.flags, RF_ROUNDS_EXCEEDED); break; }
            const uint32_t ctr_out = CTR_ACTIVE + (rounds + 1u) % 3u, ctr_items = CTR_ITEMS + (rounds & 1u);
            if (tid == 0) { ctr[CTR_ACTIVE + (rounds + 2u) % 3u] = 0u; ctr[CTR_NEWD + (rounds + 1u) % 3u] = 0u; ctr[CTR_ITEMS + ((rounds + 1u) & 1u)] = 0u;
                            ctr[CTR_FETCH + ((rounds + 1u) & 1u)] = 0u; }
            for (uint32_t i = tid; i < n_act; i += nth) round_select(W, W.active[parity][i], n_dirty, parity, ctr_out, ctr_items);
            grid.sync();
            const uint32_t n_items = ctr[ctr_items];
            // items differ 3x in cost (filter miss ... box-box hit with clipping): warps fetch them one by one
            // A fat round (more than ICX_LANE_MIN items per warp) is throughput-bound on the single finishing lane: there a warp
            // takes a batch of items and runs them lane-parallel, one item per lane from the exact pair to the end of its cursor
            // walks (the items of a round touch disjoint bodies).  Thin rounds keep one item per warp.
            const uint32_t per_warp = (n_items + (nth >> 5) - 1u) / (nth >> 5);
            const bool force_lanes = (W.cfg_flags & ICX_F_LANE_ROUNDS) != 0u;
            const bool lane_mode = per_warp >= ICX_LANE_MIN || force_lanes;
            const uint32_t batch = lane_mode ? (force_lanes ? 32u : min(per_warp, 32u)) : 1u;
            for (;;) {
                uint32_t i = 0;
                if ((tid & 31u) == 0u) i = atomicAdd(&W.counters[CTR_FETCH + (rounds & 1u)], batch);
                i = __shfl_sync(0xFFFFFFFFu, i, 0);
                if (i >= n_items) break;
                if (!lane_mode) { round_execute(W, W.items[i], n_dirty, parity, ctr_out, CTR_NEWD + rounds % 3u, scratch); continue; }
                const uint32_t cnt = min(batch, n_items - i), lane = tid & 31u;
                uint4 mine = make_uint4(0u, 0u, ICX_NONE, ICX_NONE);
                if (lane < cnt) mine = W.items[i + lane];
                round_execute_lanes(W, mine, lane < cnt, n_dirty, parity, ctr_out, CTR_NEWD + rounds % 3u);
                __syncwarp();
            }
            grid.sync();
            if (tid == 0) trace_mark(W, tn, 0x30000u | min(n_items, 0xFFFFu));
            rounds++; parity ^= 1;
        }
        grid.sync();
        rebudget = ctr[CTR_VIOLATORS] != 0u;
        if (tid == 0) { if (rebudget) ctr[CTR_NBUDGET] += 1u; if (ctr[CTR_DIRTY] != n_dirty) ctr[CTR_NGREW] += 1u; }
        if ((ctr[CTR_DIRTY] == n_dirty && !rebudget) || (*(volatile uint32_t*)W.flags & RF_ROUNDS_EXCEEDED)) break;
        attempts++;
        if (attempts > ICX_MAX_ATTEMPTS) { if (tid == 0) atomicOr(W.flags, RF_MARGIN_VIOLATED); break; }      // budgets do not settle: the host reports it
        grid.sync();
    }
    if (tid == 0) { ctr[CTR_ROUNDS] = rounds; ctr[CTR_ATTEMPTS] = attempts; trace_mark(W, tn, 0x40000u); if (W.trace) W.trace[0] = tn; }
}

__global__ void __launch_bounds__(PREPARE_TPB, PREPARE_MINB) k_resolve_prepare(const __grid_constant__ DeviceWorld W, double dt, int do_euler)
{
    resolve_body<0>(W, dt, do_euler, nullptr);
}
__global__ void __launch_bounds__(RESOLVE_TPB) k_resolve(const __grid_constant__ DeviceWorld W, double dt, int do_euler)
{
    // per-warp scratch of the exact procedures (contact manifolds, clipping): see pseudo_clip_polyhedra_s
    __shared__ V3 scratch_all[RESOLVE_TPB / 32][8 + 2 * ICX_MAXV];
    resolve_body<1>(W, dt, do_euler, scratch_all[threadIdx.x >> 5]);
}

// ---- static triangle mesh pass (extension, no counterpart in the reference: include/icx.h, icx_mesh.cuh) ----------------
// One thread per entity, on the state the pair loop left in `in` (the tick has swapped): entities do not interact here, so
// the pass is embarrassingly parallel and still has ONE defined order per entity -- slot order, then ascending triangle
// index over the triangles whose bounds overlap the collider's bounds at the start of its turn.  The grid only narrows
// the search: a triangle is registered in every cell its (slightly inflated) bounds touch, and the exact bounds test below
// decides, so the visited set is the definition's whatever the cell size.
__device__ __forceinline__ bool tri_bounds_overlap(const Tri& t, V3 lo, V3 hi)
{
    const V3 tlo = v3(smin(smin(t.a.x, t.b.x), t.c.x), smin(smin(t.a.y, t.b.y), t.c.y), smin(smin(t.a.z, t.b.z), t.c.z));
    const V3 thi = v3(smax(smax(t.a.x, t.b.x), t.c.x), smax(smax(t.a.y, t.b.y), t.c.y), smax(smax(t.a.z, t.b.z), t.c.z));
    return tlo.x <= hi.x && lo.x <= thi.x && tlo.y <= hi.y && lo.y <= thi.y && tlo.z <= hi.z && lo.z <= thi.z;
}

// the mesh as the other side of a response: immovable, identity transform, its material
__device__ __forceinline__ Side mesh_side(const DeviceMesh& M)
{
    Side B;
    B.origin = v3(0.0f, 0.0f, 0.0f); B.vel = B.origin; B.angvel = B.origin; B.acc = B.origin; B.angacc = B.origin;
    B.mass = 0.0f; B.colAngDrag = 0.0f;
    B.L = diag3(v3(1.0f, 1.0f, 1.0f)); B.base = B.origin; B.inv = B.L; B.inertia = B.origin;
    B.staticFriction = M.staticFriction; B.dynamicFriction = M.dynamicFriction; B.restitution = M.restitution;
    B.stiffness = 0.0f; B.damping = 0.0f; B.restLength = 0.0f; B.influenceOrientation = false;
    return B;
}
__device__ __forceinline__ Tri mesh_tri(const DeviceMesh& M, uint32_t t) { Tri tr; tr.a = v3(M.tri[3 * t]); tr.b = v3(M.tri[3 * t + 1]); tr.c = v3(M.tri[3 * t + 2]); return tr; }
__device__ __forceinline__ bool mesh_cells(const DeviceMesh& M, V3 lo, V3 hi, int* c0, int* c1)
{
    if (!(lo.x == lo.x && hi.x == hi.x && lo.y == lo.y && hi.y == hi.y && lo.z == lo.z && hi.z == hi.z)) return false;      // NaN state: nothing overlaps
    const float l3[3] = {lo.x, lo.y, lo.z}, h3[3] = {hi.x, hi.y, hi.z};
    for (int a = 0; a < 3; a++) {
        const float f0 = floorf((l3[a] - M.lo[a]) * M.inv_cell), f1 = floorf((h3[a] - M.lo[a]) * M.inv_cell);
        if (f1 < 0.0f || f0 > (float)(M.dim[a] - 1)) return false;
        c0[a] = (int)fminf(fmaxf(f0, 0.0f), (float)(M.dim[a] - 1)); c1[a] = (int)fminf(fmaxf(f1, 0.0f), (float)(M.dim[a] - 1));
    }
    return true;
}
__device__ __forceinline__ void mesh_emit(const DeviceWorld& W, const DeviceMesh& M, uint32_t e, uint32_t bits, uint32_t t, const Hit& h)
{
    const uint32_t at = atomicAdd(&M.counter[0], 1u);
    if (at >= M.max_contacts) { M.counter[1] = 1u; return; }
    IcxMeshContact c;
    c.entity = W.entity_id[e]; c.row = e; c.slot = (bits & PF_SLOT) ? 1u : 0u; c.triangle = t;
    c.point[0] = h.point.x; c.point[1] = h.point.y; c.point[2] = h.point.z;
    c.normal[0] = h.normal.x; c.normal[1] = h.normal.y; c.normal[2] = h.normal.z; c.depth = h.depth;
    M.contacts[at] = c;
}

// Phase 1, one thread per entity: who can touch the mesh at all?  The packet bounds of this tick (post-integration, inflated by
// the margin) moved with the body by what the pair loop displaced it -- a superset of the exact bounds -- against the occupied
// grid cells.  Streams 70 B per entity; the few that pass go to phase 2.
__global__ void __launch_bounds__(256) k_mesh_select(const __grid_constant__ DeviceWorld W, const __grid_constant__ DeviceMesh M)
{
    const uint32_t e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= W.n_entities) return;
    if (!(__float_as_uint(W.prm1[e].w) & EF_HAS_PHYSICS)) return;
    if (W.prm0[e].w == 0.0f) return;                                                // infinite mass: not visited (the response could not change it)
    const uint32_t fp = W.first_packet[e], npk = fp >> 30, p0 = fp & 0x3FFFFFFFu;
    if (!npk) return;
    const float4 pos = W.in.pos[e], q0 = W.pos0[e];
    const V3 delta = v3(pos.x - q0.x, pos.y - q0.y, pos.z - q0.z);
    const float slack = 1.0e-3f * (fabsf(delta.x) + fabsf(delta.y) + fabsf(delta.z)) + 1.0e-4f;
    bool near = false;
    for (uint32_t k = 0; k < npk && !near; k++) {
        const float4 blo = W.pk_lo[p0 + k], bhi = W.pk_hi[p0 + k];
        int c0[3], c1[3];
        if (!mesh_cells(M, v3(blo.x + delta.x - slack, blo.y + delta.y - slack, blo.z + delta.z - slack),
                        v3(bhi.x + delta.x + slack, bhi.y + delta.y + slack, bhi.z + delta.z + slack), c0, c1)) continue;
        for (int cz = c0[2]; cz <= c1[2] && !near; cz++) for (int cy = c0[1]; cy <= c1[1] && !near; cy++) {
            const uint32_t row = ((uint32_t)cz * (uint32_t)M.dim[1] + (uint32_t)cy) * (uint32_t)M.dim[0];
            near = M.cell_start[row + (uint32_t)c1[0] + 1u] != M.cell_start[row + (uint32_t)c0[0]];      // cells of one row are contiguous in the CSR
        }
    }
    if (near) M.active[atomicAdd(&M.counter[2], 1u)] = e;
}

// Phase 2, one WARP per selected entity.  Per collider: the lanes gather the triangles of the overlapped cells whose bounds
// overlap the collider's (exact test), sort them ascending and drop duplicates in shared memory; then the ordered walk runs
// 32 triangles at a time -- every lane tests ITS triangle against the current state, the lowest lane that hits is the next
// contact of the sequential order (the misses before it saw the state they would have seen), all lanes apply its response
// (same inputs, same result), and the lanes behind it test again.
#define ICX_MESH_MAXC 256
__global__ void __launch_bounds__(128) k_mesh_pass(const __grid_constant__ DeviceWorld W, const __grid_constant__ DeviceMesh M)
{
    __shared__ uint32_t s_list[4][2][ICX_MESH_MAXC];
    const uint32_t FULL = 0xFFFFFFFFu, lane = threadIdx.x & 31u, wib = threadIdx.x >> 5;
    const uint32_t n_act = min(*(volatile uint32_t*)&M.counter[2], W.n_entities);
    uint32_t* raw = s_list[wib][0]; uint32_t* sorted = s_list[wib][1];
    for (uint32_t ai = blockIdx.x * 4u + wib; ai < n_act; ai += gridDim.x * 4u) {
        const uint32_t e = M.active[ai];
        const float4 prm1 = W.prm1[e];
        const uint32_t fp = W.first_packet[e], npk = fp >> 30, p0 = fp & 0x3FFFFFFFu;
        // the entity as the pair loop left it; the second collider sees what the first one's contacts did
        float4 pos = W.in.pos[e], vel = W.in.vel[e], angvel = W.in.angvel[e], acc = W.in.acc[e], angacc = W.in.angacc[e];
        bool touched = false;
        for (uint32_t k = 0; k < npk; k++) {
            const uint32_t p = p0 + k;
            const uint32_t bits = __float_as_uint(W.pk_lpos[p].w);
            const int type = (int)(bits & PF_TYPE_MASK);
            const uint32_t ctype = (bits >> PF_CTYPE_SHIFT) & 3u;
            const bool is_ray = type == ICX_COLLIDER_RAY;                            // a ray collider meets the mesh as a spring (the biped's legs), boxes and spheres as rigid bodies
            if (is_ray ? ctype != ICX_COLLISION_SPRING : ((type != ICX_COLLIDER_BOX && type != ICX_COLLIDER_SPHERE) || ctype != ICX_COLLISION_RIGID)) continue;
            SideData sd;
            sd.pos = pos; sd.vel = vel; sd.angvel = angvel; sd.acc = acc; sd.angacc = angacc;
            sd.prm0 = W.prm0[e]; sd.prm1 = prm1; sd.pos0 = sd.pos; sd.slot = ICX_NONE; sd.ts = 0u; sd.id = 0u;
            sd.mat0 = W.pk_mat0[p]; sd.lscale = W.pk_lscale[p]; sd.damping = W.pk_damping[p]; sd.minparam = W.pk_minparam[p];
            const float4* d = W.pk_d + (size_t)p * 6;
#pragma unroll
            for (int q = 0; q < 6; q++) sd.d[q] = d[q];
            V3 base; Shape sh = shape_of(sd, type, base);
            V3 half;                                                                // bounds of the collider when its turn starts
            V3 lo, hi;
            if (is_ray) {
                const V3 o = xform(sh.M, 0.0f, 0.0f, 0.0f), en = xform(sh.M, 0.0f, 0.0f, 1.0f);
                lo = v3(smin(o.x, en.x), smin(o.y, en.y), smin(o.z, en.z)); hi = v3(smax(o.x, en.x), smax(o.y, en.y), smax(o.z, en.z));
                half = v3(0.0f, 0.0f, 0.0f);
            } else if (type == ICX_COLLIDER_SPHERE) { const float r = length(xform(sh.M, ICX_UNIT_RADIUS, 0.0f, 0.0f) - xform(sh.M, 0.0f, 0.0f, 0.0f)); half = v3(r, r, r); }
            else half = v3(((fabsf(sh.M.c[0].x) + fabsf(sh.M.c[1].x)) + fabsf(sh.M.c[2].x)) * ICX_UNIT_RADIUS,
                           ((fabsf(sh.M.c[0].y) + fabsf(sh.M.c[1].y)) + fabsf(sh.M.c[2].y)) * ICX_UNIT_RADIUS,
                           ((fabsf(sh.M.c[0].z) + fabsf(sh.M.c[1].z)) + fabsf(sh.M.c[2].z)) * ICX_UNIT_RADIUS);
            if (!is_ray) { lo = sh.M.t - half; hi = sh.M.t + half; }
            int c0[3], c1[3];
            if (!mesh_cells(M, lo, hi, c0, c1)) continue;
            Side A = side_of(sd, sh, base, bits);
            const Side B0 = mesh_side(M);
            float min_param = sd.minparam;
            // Candidates are staged ICX_MESH_MAXC at a time over windows of the triangle index range (normally ONE window: the
            // whole range); a window that holds more is halved.  The bounds are fixed for the collider's turn, so the windows
            // partition one ascending sequence.
            uint32_t win_lo = 0u, span = M.n_tri;
            while (win_lo < M.n_tri) {
                const uint32_t win_hi = (span >= M.n_tri - win_lo) ? M.n_tri : win_lo + span;
                // ---- gather (all lanes), exact bounds test, append through a ballot ----
                uint32_t n_raw = 0;
                for (int cz = c0[2]; cz <= c1[2]; cz++) for (int cy = c0[1]; cy <= c1[1]; cy++) {
                    const uint32_t row = ((uint32_t)cz * (uint32_t)M.dim[1] + (uint32_t)cy) * (uint32_t)M.dim[0];
                    const uint32_t b = M.cell_start[row + (uint32_t)c0[0]], en = M.cell_start[row + (uint32_t)c1[0] + 1u];
                    for (uint32_t i0 = b; i0 < en; i0 += 32u) {
                        const uint32_t i = i0 + lane;
                        uint32_t t = ICX_NONE; bool want = false;
                        if (i < en) { t = M.cell_tri[i]; want = t >= win_lo && t < win_hi && tri_bounds_overlap(mesh_tri(M, t), lo, hi); }
                        const uint32_t m = __ballot_sync(FULL, want);
                        const uint32_t at = n_raw + __popc(m & ((1u << lane) - 1u));
                        if (want && at < ICX_MESH_MAXC) raw[at] = t;
                        n_raw += __popc(m);
                    }
                }
                if (n_raw > ICX_MESH_MAXC) {
                    if (span > 1u) { span = (span + 1u) / 2u; __syncwarp(); continue; }
                    n_raw = 1u;                                                     // one triangle registered in more cells than the staging holds: raw[0] is it
                }
                __syncwarp();
                uint32_t n_c = 0;
                if (n_raw) {
                    // ---- sort ascending (rank by counting; equal values keep their order), then drop duplicates ----
                    for (uint32_t i = lane; i < n_raw; i += 32u) {
                        const uint32_t x = raw[i]; uint32_t rank = 0;
                        for (uint32_t j = 0; j < n_raw; j++) { const uint32_t y = raw[j]; rank += (y < x || (y == x && j < i)) ? 1u : 0u; }
                        sorted[rank] = x;
                    }
                    __syncwarp();
                    for (uint32_t i0 = 0; i0 < n_raw; i0 += 32u) {
                        const uint32_t i = i0 + lane;
                        const bool keep = i < n_raw && (i == 0u || sorted[i] != sorted[i - 1u]);
                        const uint32_t x = i < n_raw ? sorted[i] : 0u;
                        const uint32_t m = __ballot_sync(FULL, keep);
                        __syncwarp();
                        if (keep) raw[n_c + __popc(m & ((1u << lane) - 1u))] = x;  // raw is free again: position <= i
                        n_c += __popc(m);
                        __syncwarp();
                    }
                }
                // ---- the ordered walk, 32 candidates at a time ----
                uint32_t at = 0;
                while (at < n_c) {
                    const uint32_t i = at + lane;
                    Hit h; h.point = v3(0.0f, 0.0f, 0.0f); h.normal = v3(0.0f, 0.0f, 0.0f); h.depth = 0.0f;
                    bool hit = false; uint32_t t = ICX_NONE; float h_param = 0.0f;
                    if (i < n_c) {
                        t = raw[i];
                        sh.M.t = A.base + (A.origin + v3(0.0f, 0.0f, 0.0f));
                        const Tri tr = mesh_tri(M, t);
                        float prm = 0.0f;
                        hit = is_ray ? ray_triangle_hit(sh.M, tr, min_param, h, prm) : (type == ICX_COLLIDER_SPHERE ? sphere_triangle_hit(sh.M, tr, h) : box_triangle_hit(sh.M, tr, h));
                        if (is_ray && hit) h_param = prm;
                    }
                    const uint32_t m = __ballot_sync(FULL, hit);
                    if (!m) { at += 32u; continue; }
                    const int f = __ffs(m) - 1;                                     // the next contact of the sequential order
                    h.point.x = __shfl_sync(FULL, h.point.x, f); h.point.y = __shfl_sync(FULL, h.point.y, f); h.point.z = __shfl_sync(FULL, h.point.z, f);
                    h.normal.x = __shfl_sync(FULL, h.normal.x, f); h.normal.y = __shfl_sync(FULL, h.normal.y, f); h.normal.z = __shfl_sync(FULL, h.normal.z, f);
                    h.depth = __shfl_sync(FULL, h.depth, f); t = __shfl_sync(FULL, t, f); h_param = __shfl_sync(FULL, h_param, f);
                    Side B = B0;
                    if (is_ray) { min_param = h_param; spring_rigid_response(A, B, h); }   // the nearest hit so far; the spring pushes along the normal
                    else rigid_rigid_response(A, B, h);                             // every lane: same inputs, same result
                    touched = true;
                    if (lane == 0u) mesh_emit(W, M, e, bits, t, h);
                    at += (uint32_t)f + 1u;
                }
                __syncwarp();
                win_lo = win_hi;
                span = span >= (M.n_tri >> 1) ? M.n_tri : span * 2u;
            }
            pos = make_float4(A.origin.x, A.origin.y, A.origin.z, 0.0f);
            vel = make_float4(A.vel.x, A.vel.y, A.vel.z, 0.0f);
            angvel = make_float4(A.angvel.x, A.angvel.y, A.angvel.z, 0.0f);
            acc = make_float4(A.acc.x, A.acc.y, A.acc.z, 0.0f);
            angacc = make_float4(A.angacc.x, A.angacc.y, A.angacc.z, 0.0f);
            if (is_ray && lane == 0u && min_param != sd.minparam) W.pk_minparam[p] = min_param;
            __syncwarp();
        }
        if (touched && lane == 0u) { W.in.pos[e] = pos; W.in.vel[e] = vel; W.in.angvel[e] = angvel; W.in.acc[e] = acc; W.in.angacc[e] = angacc; }
    }
}

void launch_mesh_pass(const DeviceWorld& W, const DeviceMesh& M, int sms, cudaStream_t st)
{
    if (!W.n_entities || !M.n_tri) return;
    k_mesh_select<<<(W.n_entities + 255) / 256, 256, 0, st>>>(W, M);
    k_mesh_pass<<<sms * 8, 128, 0, st>>>(W, M);
}

void launch_narrow(const DeviceWorld& W, int blocks, cudaStream_t st)
{
    k_narrow<<<blocks, 128, 0, st>>>(W);
}

int resolve_grid_blocks(int device, int phase)
{
    int per_sm = 0, sms = 0;
    if (phase == 0) cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_resolve_prepare, PREPARE_TPB, 0);
    else cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_resolve, RESOLVE_TPB, 0);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    return per_sm * sms;
}

cudaError_t launch_resolve(const DeviceWorld& W, double dt, int do_euler, int prepare_blocks, int blocks, cudaStream_t st)
{
    DeviceWorld Wc = W;
    void* args[] = {&Wc, &dt, &do_euler};
    cudaError_t e = cudaLaunchCooperativeKernel((const void*)k_resolve_prepare, dim3(prepare_blocks), dim3(PREPARE_TPB), args, 0, st);
    if (e != cudaSuccess) return e;
    return cudaLaunchCooperativeKernel((const void*)k_resolve, dim3(blocks), dim3(RESOLVE_TPB), args, 0, st);
}

} // namespace icx
