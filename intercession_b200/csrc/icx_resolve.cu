// Ordered narrowphase + response + interception (CollisionPhysicsRelay::engage,
// source/physics/collision_physics_relay.h:24-231) on the candidate lists.
//
// The reference resolves every hit immediately inside a sequential lexicographic pair loop, so
// later pair tests see moved bodies.  Two pairs commute exactly iff they share no entity, hence
// the loop is replayed by ROUNDS: every entity keeps a cursor into its (sorted) candidate list,
// and a pair runs in the round in which it is at the cursor of BOTH its entities.  The pair at
// the global minimum is always runnable, so every round makes progress; within a round all
// runnable pairs touch disjoint entities.  Immutable entities (infinite mass, non-temporal, no
// ray collider: nothing in a response can change them) impose no ordering and are never written.
#include "icx_world.cuh"
#include "icx_response.cuh"

namespace icx {

struct Cursor { uint32_t p0, npk, c0, c1, n0, n1; };

__device__ __forceinline__ Cursor load_cursor(const DeviceWorld& W, uint32_t e)
{
    Cursor c;
    const uint32_t fp = W.first_packet[e];
    c.npk = fp >> 30; c.p0 = fp & 0x3FFFFFFFu;
    c.c0 = c.npk > 0 ? W.pos_of[c.p0] : 0xFFFFFFFFu;
    c.c1 = c.npk > 1 ? W.pos_of[c.p0 + 1] : 0xFFFFFFFFu;
    c.n0 = c.c0 != 0xFFFFFFFFu ? W.nbr_count[c.c0] : 0u;
    c.n1 = c.c1 != 0xFFFFFFFFu ? W.nbr_count[c.c1] : 0u;
    return c;
}
// Next entry of the entity's merged queue.  Order = the reference's lexicographic pair order seen
// from this entity: first pairs whose other packet precedes mine, by (other, my slot); then all
// pairs of my slot-0 packet, then all of my slot-1 packet.
__device__ __forceinline__ bool cursor_head(const DeviceWorld& W, const Cursor& c, uint32_t hw, uint32_t& k, uint32_t& entry)
{
    const uint32_t h0 = hw & 0xFFFFu, h1 = hw >> 16;
    const bool has0 = h0 < c.n0, has1 = h1 < c.n1;
    if (!has0 && !has1) return false;
    uint32_t a = 0, b = 0;
    if (has0) a = W.nbr[(size_t)h0 * W.n_packets + c.c0];
    if (has1) b = W.nbr[(size_t)h1 * W.n_packets + c.c1];
    if (has0 && has1) {
        const uint32_t ao = a & ICX_NB_INDEX_MASK, bo = b & ICX_NB_INDEX_MASK;
        const bool alow = ao < c.p0, blow = bo < c.p0;
        if (alow || blow) k = (alow && blow) ? (bo < ao ? 1u : 0u) : (alow ? 0u : 1u);
        else k = 0u;
    } else k = has0 ? 0u : 1u;
    entry = k ? b : a;
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

// One iteration of the reference's inner loop body (collision_physics_relay.h:39-228) for
// packets pa < pb.
static __device__ __noinline__ void process_pair(const DeviceWorld& W, uint32_t pa, uint32_t pb)
{
    const uint32_t ea = W.pk_entity[pa], eb = W.pk_entity[pb];
    const float4 lposA = W.pk_lpos[pa], lposB = W.pk_lpos[pb];
    const uint32_t bitsA = __float_as_uint(lposA.w), bitsB = __float_as_uint(lposB.w);
    const int tA = bitsA & PF_TYPE_MASK, tB = bitsB & PF_TYPE_MASK;
    const float4 posA4 = W.out.pos[ea], posB4 = W.out.pos[eb];
    V3 baseA, baseB;
    const Shape A = load_shape(W.pk_d + (size_t)pa * 6, v3(posA4), tA, baseA);
    const Shape B = load_shape(W.pk_d + (size_t)pb * 6, v3(posB4), tB, baseB);

    // ***** intersection: intersectProcedures[tA][tB], collision_procedures.h:411-417 *****
    Hit h; h.point = v3(0.0f, 0.0f, 0.0f); h.normal = v3(0.0f, 0.0f, 0.0f); h.depth = 0.0f;
    bool hit = false;
    if (tA == ICX_COLLIDER_BOX && tB == ICX_COLLIDER_BOX) hit = box_box_intersect(A, B, h);
    else if (tA == ICX_COLLIDER_SPHERE && tB == ICX_COLLIDER_BOX) hit = sphere_box_intersect(A, B, h);
    else if (tA == ICX_COLLIDER_BOX && tB == ICX_COLLIDER_SPHERE) {                 // box_sphere_intersect, .cpp:233-247
        hit = sphere_box_intersect(B, A, h);
        if (hit) { h.normal = h.normal * -1.0f; h.point = h.point + (h.normal * h.depth); }
    } else if (tA == ICX_COLLIDER_RAY && tB == ICX_COLLIDER_BOX) {
        float mp = W.pk_minparam[pa];
        hit = ray_box_intersect(A, v3(posA4), B, v3(posB4), mp, h);
        if (hit) W.pk_minparam[pa] = mp;
    } else if (tA == ICX_COLLIDER_BOX && tB == ICX_COLLIDER_RAY) {                  // box_ray_intersect, .cpp:249-263
        float mp = W.pk_minparam[pb];
        hit = ray_box_intersect(B, v3(posB4), A, v3(posA4), mp, h);
        if (hit) { W.pk_minparam[pb] = mp; h.normal = h.normal * -1.0f; h.point = h.point + (h.normal * h.depth); }
    } else if ((W.cfg_flags & ICX_F_SPHERE_SPHERE) && tA == ICX_COLLIDER_SPHERE && tB == ICX_COLLIDER_SPHERE) {
        const float rA = ICX_UNIT_RADIUS * W.pk_lscale[pa].x * W.prm0[ea].x;         // .cpp:403-404
        const float rB = ICX_UNIT_RADIUS * W.pk_lscale[pb].x * W.prm0[eb].x;
        hit = sphere_sphere_intersect(A, rA, B, rB, h);
    }
    if (!hit) return;

    // ***** physics response, :82-107 *****
    const uint32_t efA = __float_as_uint(W.prm1[ea].w), efB = __float_as_uint(W.prm1[eb].w);
    const uint32_t ctA = (bitsA >> PF_CTYPE_SHIFT) & 3u, ctB = (bitsB >> PF_CTYPE_SHIFT) & 3u;
    uint32_t cflags = 0;
    if (ctA != ICX_COLLISION_NOOP && ctB != ICX_COLLISION_NOOP && (efA & EF_HAS_PHYSICS) && (efB & EF_HAS_PHYSICS)) {
        cflags |= ICX_CONTACT_RESPONDED;
        const bool rr = ctA == ICX_COLLISION_RIGID && ctB == ICX_COLLISION_RIGID;
        const bool sr = ctA == ICX_COLLISION_SPRING && ctB == ICX_COLLISION_RIGID;
        const bool rs = ctA == ICX_COLLISION_RIGID && ctB == ICX_COLLISION_SPRING;
        if (rr || sr || rs) {                                                       // other table entries are null_response
            Side SA, SB;
            {
                const float4 v = W.out.vel[ea], w = W.out.angvel[ea], p0 = W.prm0[ea], p1 = W.prm1[ea], m0 = W.pk_mat0[pa];
                SA.origin = v3(posA4); SA.vel = v3(v); SA.angvel = v3(w); SA.mass = p0.w; SA.colAngDrag = p1.z;
                SA.L = mat3_of(A.M); SA.base = baseA; SA.inv = A.inv;
                const float4* d = W.pk_d + (size_t)pa * 6; SA.inertia = v3(d[3].w, d[4].w, d[5].w);
                SA.staticFriction = m0.x; SA.dynamicFriction = m0.y; SA.restitution = m0.z; SA.stiffness = m0.w;
                SA.damping = W.pk_damping[pa]; SA.restLength = W.pk_lscale[pa].w;
                SA.influenceOrientation = (bitsA & PF_INFLUENCE_ORIENT) != 0;
                if (!rr) { SA.acc = v3(W.out.acc[ea]); SA.angacc = v3(W.out.angacc[ea]); }
            }
            {
                const float4 v = W.out.vel[eb], w = W.out.angvel[eb], p0 = W.prm0[eb], p1 = W.prm1[eb], m0 = W.pk_mat0[pb];
                SB.origin = v3(posB4); SB.vel = v3(v); SB.angvel = v3(w); SB.mass = p0.w; SB.colAngDrag = p1.z;
                SB.L = mat3_of(B.M); SB.base = baseB; SB.inv = B.inv;
                const float4* d = W.pk_d + (size_t)pb * 6; SB.inertia = v3(d[3].w, d[4].w, d[5].w);
                SB.staticFriction = m0.x; SB.dynamicFriction = m0.y; SB.restitution = m0.z; SB.stiffness = m0.w;
                SB.damping = W.pk_damping[pb]; SB.restLength = W.pk_lscale[pb].w;
                SB.influenceOrientation = (bitsB & PF_INFLUENCE_ORIENT) != 0;
                if (!rr) { SB.acc = v3(W.out.acc[eb]); SB.angacc = v3(W.out.angacc[eb]); }
            }
            if (rr) rigid_rigid_response(SA, SB, h);
            else if (sr) spring_rigid_response(SA, SB, h);
            else {                                                                  // rigid_spring_response, .cpp:1055-1065
                Hit inv; inv.normal = -h.normal; inv.point = h.point - (h.normal * h.depth); inv.depth = h.depth;
                spring_rigid_response(SB, SA, inv);
            }
            const float lim = W.margin * 0.98f;
            if (!(efA & EF_IMMUTABLE)) {
                W.out.vel[ea] = make_float4(SA.vel.x, SA.vel.y, SA.vel.z, 0.0f);
                W.out.angvel[ea] = make_float4(SA.angvel.x, SA.angvel.y, SA.angvel.z, 0.0f);
                if (rr) {
                    W.out.pos[ea] = make_float4(SA.origin.x, SA.origin.y, SA.origin.z, 0.0f);
                    const float4 o0 = W.pos0[ea];
                    if (!(fabsf(SA.origin.x - o0.x) <= lim && fabsf(SA.origin.y - o0.y) <= lim && fabsf(SA.origin.z - o0.z) <= lim))
                        atomicOr(W.flags, RF_MARGIN_VIOLATED);
                } else {
                    W.out.acc[ea] = make_float4(SA.acc.x, SA.acc.y, SA.acc.z, 0.0f);
                    W.out.angacc[ea] = make_float4(SA.angacc.x, SA.angacc.y, SA.angacc.z, 0.0f);
                }
            }
            if (!(efB & EF_IMMUTABLE)) {
                W.out.vel[eb] = make_float4(SB.vel.x, SB.vel.y, SB.vel.z, 0.0f);
                W.out.angvel[eb] = make_float4(SB.angvel.x, SB.angvel.y, SB.angvel.z, 0.0f);
                if (rr) {
                    W.out.pos[eb] = make_float4(SB.origin.x, SB.origin.y, SB.origin.z, 0.0f);
                    const float4 o0 = W.pos0[eb];
                    if (!(fabsf(SB.origin.x - o0.x) <= lim && fabsf(SB.origin.y - o0.y) <= lim && fabsf(SB.origin.z - o0.z) <= lim))
                        atomicOr(W.flags, RF_MARGIN_VIOLATED);
                } else {
                    W.out.acc[eb] = make_float4(SB.acc.x, SB.acc.y, SB.acc.z, 0.0f);
                    W.out.angacc[eb] = make_float4(SB.angacc.x, SB.angacc.y, SB.angacc.z, 0.0f);
                }
            }
        }
    }

    // ***** behaviours response :114-155 -> contact record, replayed by the host adapter *****
    if (bitsA & PF_BEHAVIORS) cflags |= ICX_CONTACT_BEHAVIORS_A;
    if (bitsB & PF_BEHAVIORS) cflags |= ICX_CONTACT_BEHAVIORS_B;
    const uint32_t idA = W.entity_id[ea], idB = W.entity_id[eb];
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

    // ***** timestream response :157-228 *****
    const uint32_t thisLink = idA & 0xFu, otherLink = idB & 0xFu;                   // derive_causal_chain_link, ecs_types.h:92-95
    if (thisLink == ICX_NULL_CHAINLINK || otherLink == ICX_NULL_CHAINLINK) return;  // :185-188
    const uint32_t tsA = W.out.ts[ea], tsB = W.out.ts[eb];
    const bool dA = divergent(tsA & 0xFFu), dB = divergent(tsB & 0xFFu);
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

// One dependency round.  One thread per still-active entity.
__global__ void __launch_bounds__(128) k_resolve_round(DeviceWorld W, int parity, uint32_t ctr_in, uint32_t ctr_out)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t n_active = W.counters[ctr_in];
    if (i >= n_active) return;
    const uint32_t* __restrict__ hin = W.head[parity];
    uint32_t* __restrict__ hout = W.head[parity ^ 1];
    const uint32_t e = W.active[parity][i];
    const Cursor c = load_cursor(W, e);
    const bool immut_me = (__float_as_uint(W.prm1[e].w) & EF_IMMUTABLE) != 0;
    uint32_t hw = hin[e];
    bool first = true;
    for (;;) {
        uint32_t k, entry;
        if (!cursor_head(W, c, hw, k, entry)) break;
        const uint32_t p = c.p0 + k;
        const uint32_t o = entry & ICX_NB_INDEX_MASK;
        if (immut_me) {
            // only immutable partners are listed; no state changes, no ordering: run the pairs I drive
            const bool ilarge = W.pk_cell[p] == 0xFFFFFFFEu, olarge = (entry & ICX_NB_LARGE) != 0;
            if (ilarge == olarge ? p < o : !ilarge) process_pair(W, min(p, o), max(p, o));
            hw = cursor_advance(hw, k); first = false;
            continue;
        }
        if (entry & ICX_NB_IMMUTABLE) {             // partner imposes no ordering and is never written
            process_pair(W, min(p, o), max(p, o));
            hw = cursor_advance(hw, k); first = false;
            continue;
        }
        if (!first) break;                          // a two-sided pair must be at my cursor at round start
        const uint32_t eo = W.pk_entity[o];
        const Cursor co = load_cursor(W, eo);
        const uint32_t hwo = hin[eo];
        uint32_t ko, entryo;
        if (!cursor_head(W, co, hwo, ko, entryo)) break;           // cannot happen for a consistent list
        if (co.p0 + ko != o || (entryo & ICX_NB_INDEX_MASK) != p) break;    // partner is not ready: wait
        if (p > o) return;                          // I am the B side: the A thread advances and re-queues me
        process_pair(W, p, o);
        hw = cursor_advance(hw, k); first = false;
        const uint32_t hwo2 = cursor_advance(hwo, ko);
        hout[eo] = hwo2;
        if (cursor_pending(co, hwo2)) W.active[parity ^ 1][atomicAdd(&W.counters[ctr_out], 1u)] = eo;
    }
    hout[e] = hw;
    if (cursor_pending(c, hw)) W.active[parity ^ 1][atomicAdd(&W.counters[ctr_out], 1u)] = e;
}

void launch_resolve_round(const DeviceWorld& W, int parity, uint32_t ctr_in, uint32_t ctr_out, uint32_t n_threads, cudaStream_t st)
{
    if (n_threads == 0) return;
    k_resolve_round<<<(n_threads + 127) / 128, 128, 0, st>>>(W, parity, ctr_in, ctr_out);
}

} // namespace icx
