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
#include "icx_device.cuh"

namespace cg = cooperative_groups;

namespace icx {

struct Cursor { uint32_t p0, npk, n0, n1; const uint32_t* l0; const uint32_t* l1; };

__device__ __forceinline__ Cursor load_cursor(const DeviceWorld& W, uint32_t slot, uint32_t e)
{
    Cursor c;
    const uint32_t fp = W.first_packet[e];
    c.npk = fp >> 30; c.p0 = fp & 0x3FFFFFFFu;
    c.l0 = W.nbr + (size_t)(slot * 2u) * W.max_neighbors; c.l1 = c.l0 + W.max_neighbors;
    c.n0 = W.nbr_count[slot * 2u]; c.n1 = W.nbr_count[slot * 2u + 1u];
    return c;
}
// Next entry of the body's merged queue.  Order = the reference's lexicographic pair order seen
// from this body: first pairs whose other packet precedes mine, by (other, my slot); then all
// pairs of my slot-0 packet, then all of my slot-1 packet.
__device__ __forceinline__ bool cursor_head(const Cursor& c, uint32_t hw, uint32_t& k, uint32_t& entry)
{
    const uint32_t h0 = hw & 0xFFFFu, h1 = hw >> 16;
    const bool has0 = h0 < c.n0, has1 = h1 < c.n1;
    if (!has0 && !has1) return false;
    uint32_t a = 0, b = 0;
    if (has0) a = c.l0[h0];
    if (has1) b = c.l1[h1];
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

// One iteration of the reference's inner loop body (collision_physics_relay.h:39-228) for packets
// pa < pb.  n_dirty: bodies with dirty_slot < n_dirty may be written; a hit on any other mutable
// body aborts the attempt (see file header).
static __device__ __noinline__ void process_pair(const DeviceWorld& W, uint32_t pa, uint32_t pb, uint32_t n_dirty)
{
    const uint32_t ea = W.pk_entity[pa], eb = W.pk_entity[pb];
    const float4 lposA = W.pk_lpos[pa], lposB = W.pk_lpos[pb];
    const uint32_t bitsA = __float_as_uint(lposA.w), bitsB = __float_as_uint(lposB.w);
    const int tA = bitsA & PF_TYPE_MASK, tB = bitsB & PF_TYPE_MASK;
    const float4 posA4 = W.out.pos[ea], posB4 = W.out.pos[eb];
    V3 baseA, baseB; Shape A, B; Hit h;
    float mpA = (tA == ICX_COLLIDER_RAY) ? W.pk_minparam[pa] : 1.0f, mpB = (tB == ICX_COLLIDER_RAY) ? W.pk_minparam[pb] : 1.0f;
    const bool hit = pair_intersect(W, pa, pb, ea, eb, tA, tB, v3(posA4), v3(posB4), A, B, baseA, baseB, mpA, mpB, h);
    if (!hit) return;
    const uint32_t efA = __float_as_uint(W.prm1[ea].w), efB = __float_as_uint(W.prm1[eb].w);
    {
        const bool cleanA = !(efA & EF_IMMUTABLE) && W.dirty_slot[ea] >= n_dirty;
        const bool cleanB = !(efB & EF_IMMUTABLE) && W.dirty_slot[eb] >= n_dirty;
        if (cleanA || cleanB) {                     // the dirty set was too small: grow it, write nothing
            if (cleanA) mark_dirty(W, ea);
            if (cleanB) mark_dirty(W, eb);
            atomicAdd(&W.counters[CTR_NEW_DIRTY], 1u);
            return;
        }
    }
    if (tA == ICX_COLLIDER_RAY) W.pk_minparam[pa] = mpA;
    if (tB == ICX_COLLIDER_RAY) W.pk_minparam[pb] = mpB;

    // ***** physics response, :82-107 *****
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

// ---------------------------------------------------------------------------------------
// K5: speculative narrowphase.  One thread per candidate pair whose un-inflated bounds overlap:
// test on the post-integration state, write nothing but the dirty marks.  Hits between two
// immutable bodies change nothing and need no order; they are queued for their contact record.
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) k_narrow(DeviceWorld W)
{
    const uint32_t n = min(W.counters[CTR_CANDIDATES], W.max_pairs);
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const uint2 pr = W.pairs[i];
    if (!(pr.y & ICX_PAIR_TIGHT)) continue;
    const uint32_t pa = pr.x, pb = pr.y & ICX_NB_INDEX_MASK;
    const uint32_t ea = W.pk_entity[pa], eb = W.pk_entity[pb];
    const int tA = __float_as_uint(W.pk_lpos[pa].w) & PF_TYPE_MASK, tB = __float_as_uint(W.pk_lpos[pb].w) & PF_TYPE_MASK;
    V3 baseA, baseB; Shape A, B; Hit h;
    float mpA = 1.0f, mpB = 1.0f;               // Collider::reset value; a ray that hits is dirty and replayed in order
    if (!pair_intersect(W, pa, pb, ea, eb, tA, tB, v3(W.out.pos[ea]), v3(W.out.pos[eb]), A, B, baseA, baseB, mpA, mpB, h)) continue;
    if (pr.y & ICX_PAIR_STATIC) {
        const uint32_t k = atomicAdd(&W.counters[CTR_STATIC], 1u);
        if (k < W.max_static) W.static_pairs[k] = make_uint2(pa, pb); else atomicOr(W.flags, RF_PAIR_OVERFLOW);
        continue;
    }
    if (!(__float_as_uint(W.prm1[ea].w) & EF_IMMUTABLE)) mark_dirty(W, ea);
    if (!(__float_as_uint(W.prm1[eb].w) & EF_IMMUTABLE)) mark_dirty(W, eb);
    }
}

// Sorted candidate lists of one dirty body (both collider slots): every packet of another entity
// whose inflated bound overlaps, ascending by packet index = the order the reference's pair loop
// meets them (collision_physics_relay.h:29-37).
__device__ __forceinline__ void list_insert(uint32_t* list, uint32_t& cnt, uint32_t cap, uint32_t entry, bool& overflow)
{
    if (cnt >= cap) { overflow = true; return; }
    uint32_t i = cnt;
    const uint32_t key = entry & ICX_NB_INDEX_MASK;
    while (i > 0 && (list[i - 1] & ICX_NB_INDEX_MASK) > key) { list[i] = list[i - 1]; i--; }
    list[i] = entry; cnt++;
}

static __device__ __noinline__ void build_lists(const DeviceWorld& W, uint32_t slot)
{
    const uint32_t e = W.dirty_list[slot];
    const uint32_t fp = W.first_packet[e];
    const uint32_t npk = fp >> 30, p0 = fp & 0x3FFFFFFFu;
    const uint32_t wl = W.worldline_stride ? e / W.worldline_stride : 0u;
    const uint32_t n_large = min(W.counters[CTR_LARGE], W.max_large);
    bool overflow = false;
    for (uint32_t k = 0; k < 2u; k++) {
        uint32_t* list = W.nbr + (size_t)(slot * 2u + k) * W.max_neighbors;
        uint32_t cnt = 0;
        if (k < npk && W.pk_cell[p0 + k] < 0xFFFFFFFEu) {
            const uint32_t p = p0 + k;
            const float4 lo = W.pk_lo[p], hi = W.pk_hi[p];
            const int type = __float_as_uint(W.pk_lpos[p].w) & PF_TYPE_MASK;
            for_each_neighbor(W, cell_coord(lo.x, hi.x, W.inv_cell), cell_coord(lo.y, hi.y, W.inv_cell), cell_coord(lo.z, hi.z, W.inv_cell), wl,
                [&](uint32_t j) {
                    const float4 blo = W.s_lo[j], bhi = W.s_hi[j];
                    if (!overlap(lo, hi, blo, bhi)) return;
                    const uint32_t btag = __float_as_uint(blo.w), be = __float_as_uint(bhi.w);
                    if (be == e) return;
                    if (W.worldline_stride && be / W.worldline_stride != wl) return;
                    if (!can_intersect(type, (btag >> 28) & 3, W.cfg_flags)) return;
                    list_insert(list, cnt, W.max_neighbors, (btag & ICX_NB_INDEX_MASK) | ((btag & 0x40000000u) ? ICX_NB_IMMUTABLE : 0u), overflow);
                });
            for (uint32_t q = 0; q < n_large; q++) {
                const uint32_t o = W.large_list[q];
                const uint32_t oe = W.pk_entity[o];
                if (oe == e) continue;
                if (W.worldline_stride && oe / W.worldline_stride != wl) continue;
                if (!can_intersect(type, (int)(__float_as_uint(W.pk_lpos[o].w) & PF_TYPE_MASK), W.cfg_flags)) continue;
                if (!overlap(lo, hi, W.pk_lo[o], W.pk_hi[o])) continue;
                list_insert(list, cnt, W.max_neighbors, o | ICX_NB_IMMUTABLE, overflow);
            }
        }
        W.nbr_count[slot * 2u + k] = cnt;
    }
    if (overflow) atomicOr(W.flags, RF_NEIGHBOR_OVERFLOW);
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

// One dependency round for dirty slot s.
__device__ __forceinline__ void round_step(const DeviceWorld& W, uint32_t s, uint32_t n_dirty, int parity, uint32_t ctr_out)
{
    const uint32_t* __restrict__ hin = W.head[parity];
    uint32_t* __restrict__ hout = W.head[parity ^ 1];
    const uint32_t e = W.dirty_list[s];
    const Cursor c = load_cursor(W, s, e);
    uint32_t hw = hin[s];
    bool first = true;
    for (;;) {
        uint32_t k, entry;
        if (!cursor_head(c, hw, k, entry)) break;
        const uint32_t p = c.p0 + k;
        const uint32_t o = entry & ICX_NB_INDEX_MASK;
        uint32_t eo = 0, so = ICX_NONE;
        if (!(entry & ICX_NB_IMMUTABLE)) { eo = W.pk_entity[o]; so = W.dirty_slot[eo]; }
        if (so >= n_dirty) {                        // immutable or clean partner: no ordering, never written
            process_pair(W, min(p, o), max(p, o), n_dirty);
            hw = cursor_advance(hw, k); first = false;
            continue;
        }
        if (!first) break;                          // a two-sided pair must be at my cursor at round start
        const Cursor co = load_cursor(W, so, eo);
        const uint32_t hwo = hin[so];
        uint32_t ko, entryo;
        if (!cursor_head(co, hwo, ko, entryo)) break;           // cannot happen for a consistent list
        if (co.p0 + ko != o || (entryo & ICX_NB_INDEX_MASK) != p) break;    // partner is not ready: wait
        if (p > o) return;                          // I am the B side: the A thread advances and re-queues me
        process_pair(W, p, o, n_dirty);
        hw = cursor_advance(hw, k); first = false;
        const uint32_t hwo2 = cursor_advance(hwo, ko);
        hout[so] = hwo2;
        if (cursor_pending(co, hwo2)) W.active[parity ^ 1][atomicAdd(&W.counters[ctr_out], 1u)] = so;
    }
    hout[s] = hw;
    if (cursor_pending(c, hw)) W.active[parity ^ 1][atomicAdd(&W.counters[ctr_out], 1u)] = s;
}

// ---------------------------------------------------------------------------------------
// K6: ordered resolve of the dirty bodies.  Persistent cooperative kernel (grid-wide barriers
// between dependency rounds, no host round trips); see the file header.
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) k_resolve(DeviceWorld W, double dt, int do_euler)
{
    cg::grid_group grid = cg::this_grid();
    const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x, nth = gridDim.x * blockDim.x;
    volatile uint32_t* ctr = W.counters;
    uint32_t attempts = 0, rounds = 0;
    for (;;) {
        const uint32_t n_dirty = ctr[CTR_DIRTY], n_listed = ctr[CTR_LISTED], n_static = min(ctr[CTR_STATIC], W.max_static);
        for (uint32_t s = n_listed + tid; s < n_dirty; s += nth) build_lists(W, s);
        for (uint32_t s = tid; s < n_dirty; s += nth) {
            if (attempts > 0 && s < n_listed) restore_body(W, W.dirty_list[s], dt, do_euler);
            W.head[0][s] = 0u; W.active[0][s] = s;
        }
        grid.sync();
        if (*(volatile uint32_t*)W.flags & RF_NEIGHBOR_OVERFLOW) break;     // lists are incomplete: the host grows them and re-runs the tick
        if (tid == 0) {
            ctr[CTR_LISTED] = n_dirty; ctr[CTR_CONTACTS] = 0; ctr[CTR_INTERCEPTS] = 0; ctr[CTR_NEW_DIRTY] = 0;
            ctr[CTR_ACTIVE] = n_dirty; ctr[CTR_ACTIVE + 1] = 0; ctr[CTR_ACTIVE + 2] = 0;
            __threadfence();
        }
        grid.sync();
        for (uint32_t i = tid; i < n_static; i += nth) { const uint2 pr = W.static_pairs[i]; process_pair(W, pr.x, pr.y, n_dirty); }
        int parity = 0;
        rounds = 0;
        for (;;) {
            const uint32_t n_act = ctr[CTR_ACTIVE + rounds % 3u];
            if (n_act == 0u) break;
            if (rounds >= ICX_MAX_ROUNDS) { if (tid == 0) atomicOr(W.flags, RF_ROUNDS_EXCEEDED); break; }
            if (tid == 0) ctr[CTR_ACTIVE + (rounds + 2u) % 3u] = 0u;
            for (uint32_t i = tid; i < n_act; i += nth) round_step(W, W.active[parity][i], n_dirty, parity, CTR_ACTIVE + (rounds + 1u) % 3u);
            grid.sync();
            rounds++; parity ^= 1;
        }
        grid.sync();
        if (ctr[CTR_NEW_DIRTY] == 0u || (*(volatile uint32_t*)W.flags & RF_ROUNDS_EXCEEDED)) break;
        attempts++;
        grid.sync();                                // everyone has read CTR_NEW_DIRTY before the next attempt resets it
    }
    if (tid == 0) { ctr[CTR_ROUNDS] = rounds; ctr[CTR_ATTEMPTS] = attempts; }
}

void launch_narrow(const DeviceWorld& W, int blocks, cudaStream_t st)
{
    k_narrow<<<blocks, 128, 0, st>>>(W);
}

int resolve_grid_blocks(int device)
{
    int per_sm = 0, sms = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_resolve, 128, 0);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    return per_sm * sms;
}

cudaError_t launch_resolve(const DeviceWorld& W, double dt, int do_euler, int blocks, cudaStream_t st)
{
    DeviceWorld Wc = W;
    void* args[] = {&Wc, &dt, &do_euler};
    return cudaLaunchCooperativeKernel((const void*)k_resolve, dim3(blocks), dim3(128), args, 0, st);
}

} // namespace icx
