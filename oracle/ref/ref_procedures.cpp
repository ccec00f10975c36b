// TEST INFRASTRUCTURE -- not product code (see ref_harness.cpp header).
//
// The reference's relay / procedure tables driven over plain arrays:
//   icxref_euler    -> the real EulerPhysicsRelay (euler_physics_relay.h:21-97)
//   icxref_collide  -> the pair loop of CollisionPhysicsRelay::engage
//                      (collision_physics_relay.h:24-231) restated around the
//                      real intersectProcedures / responseProcedures tables
//                      (collision_procedures.h:411-430), because the relay
//                      itself needs a Cosmos whose ids are uint16.
//                      Optional candidate list = oracle-side cull:
//                      the same loop visiting only the listed pairs, in order.
#include "../../include/icx.h"

#include "physics/transform_component.h"
#include "events/event_broker.h"
#include "physics/physics_component.h"
#include "physics/collider_component.h"
#include "physics/euler_physics_relay.h"
#include "physics/collision_procedures.h"
#include "spacetime/timestream_state.h"

#include <chrono>

using namespace pleep;

namespace {

struct Packet {
    uint32_t entityIndex;
    uint8_t  slot;
};

inline bool divergent(uint8_t s) { return is_divergent((TimestreamState)s); }

struct CollideCtx {
    uint32_t n; const IcxEntity* id;
    TransformComponent* T; PhysicsComponent* P; const uint8_t* hasP;
    ColliderComponent* C;
    uint8_t* tstate; uint16_t* tcoh; uint16_t coherency; uint32_t host; uint32_t flags;
    uint32_t stride;
    std::vector<Packet> packets;
    std::vector<ColliderPacket> cps;
    IcxContact* hits; size_t hcap; size_t nh;
    IcxInterception* ints; size_t icap; size_t ni;

    // one iteration of the inner loop body, collision_physics_relay.h:39-228
    void visit(uint32_t ia, uint32_t ib) {
        ColliderPacket& dataA = cps[ia];
        ColliderPacket& dataB = cps[ib];
        const Packet& pa = packets[ia];
        const Packet& pb = packets[ib];
        if (dataB.collider.colliderType == ColliderType::none || !dataB.collider.isActive) return;  // :47
        if (pa.entityIndex == pb.entityIndex) return;                                             // :51
        if (stride && pa.entityIndex / stride != pb.entityIndex / stride) return;   // world-line batch (extension, G7)

        glm::vec3 collisionNormal; float collisionDepth; glm::vec3 collisionPoint;
        size_t ta = (size_t)dataA.collider.colliderType, tb = (size_t)dataB.collider.colliderType;
        bool hit;
        if ((flags & ICX_F_SPHERE_SPHERE) && ta == (size_t)ColliderType::sphere && tb == (size_t)ColliderType::sphere)
            hit = sphere_sphere_intersect(dataA, dataB, collisionPoint, collisionNormal, collisionDepth);  // opt-in, G3
        else
            hit = intersectProcedures[ta][tb](dataA, dataB, collisionPoint, collisionNormal, collisionDepth); // :61-64
        if (!hit) return;

        uint16_t cflags = 0;
        if (dataA.collider.collisionType != CollisionType::noop && dataB.collider.collisionType != CollisionType::noop) {
            // :84-106 -- get_component<PhysicsComponent> throws when absent; swallowed
            bool haveA = P && (!hasP || hasP[pa.entityIndex]);
            bool haveB = P && (!hasP || hasP[pb.entityIndex]);
            if (haveA && haveB) {
                responseProcedures[(size_t)dataA.collider.collisionType][(size_t)dataB.collider.collisionType](
                    dataA, P[pa.entityIndex], dataB, P[pb.entityIndex], collisionPoint, collisionNormal, collisionDepth);
                cflags |= ICX_CONTACT_RESPONDED;
            }
        }
        if (dataA.collider.useBehaviorsResponse) cflags |= ICX_CONTACT_BEHAVIORS_A;   // :114
        if (dataB.collider.useBehaviorsResponse) cflags |= ICX_CONTACT_BEHAVIORS_B;   // :133
        if (nh < hcap) {
            IcxContact& c = hits[nh];
            c.entityA = id ? id[pa.entityIndex] : pa.entityIndex; c.entityB = id ? id[pb.entityIndex] : pb.entityIndex;
            c.slotA = pa.slot; c.slotB = pb.slot; c.flags = cflags;
            for (int k = 0; k < 3; k++) { c.point[k] = collisionPoint[k]; c.normal[k] = collisionNormal[k]; }
            c.depth = collisionDepth; c.packetA = ia; c.packetB = ib;
        }
        nh++;

        // ***** TIMESTREAM RESPONSE ***** :157-228 (wide ids: low nibble = chain-link)
        if (!id) return;
        IcxEntity eA = id[pa.entityIndex], eB = id[pb.entityIndex];
        uint32_t thisLink = eA & 0xFu, otherLink = eB & 0xFu;
        uint8_t thisState = tstate ? tstate[pa.entityIndex] : 0, otherState = tstate ? tstate[pb.entityIndex] : 0;
        if (thisLink == NULL_CAUSALCHAINLINK || otherLink == NULL_CAUSALCHAINLINK) return;
        if (!(((thisLink == 0) ^ (otherLink == 0))
              || (thisLink != 0 && otherLink != 0 && (divergent(thisState) ^ divergent(otherState)))
              || (thisLink != otherLink && (!divergent(thisState) || !divergent(otherState)) && host == NULL_TIMESLICEID)))
            return;
        IcxEntity agent;
        if (thisLink != otherLink) agent = thisLink < otherLink ? eA : eB;
        else if (divergent(thisState) ^ divergent(otherState)) agent = divergent(thisState) ? eA : eB;
        else agent = ICX_NULL_ENTITY;
        IcxEntity recipient = agent == eA ? eB : eA;
        uint32_t recipientIndex = agent == eA ? pb.entityIndex : pa.entityIndex;
        if (ni < icap) { ints[ni].agent = agent; ints[ni].recipient = recipient; }
        ni++;
        // handlers: server_network_dynamo.cpp:813-838 / parallel_network_dynamo.cpp:455-467
        if (tstate) {
            if (host == NULL_TIMESLICEID) { tstate[recipientIndex] = (uint8_t)TimestreamState::forked; if (tcoh) tcoh[recipientIndex] = coherency; }
            else if ((recipient & 0xFu) > 0) { tstate[recipientIndex] = (uint8_t)TimestreamState::forking; if (tcoh) tcoh[recipientIndex] = coherency; }
        }
    }
};

} // namespace

extern "C" {

// EulerPhysicsRelay::engage over entities that have a PhysicsComponent.
int icxref_euler(uint32_t n, IcxTransform* T, IcxPhysics* P, const uint8_t* has_physics, double dt, double* seconds)
{
    if (!T || !P) return 1;
    EulerPhysicsRelay relay(nullptr);
    TransformComponent* t = reinterpret_cast<TransformComponent*>(T);
    PhysicsComponent* p = reinterpret_cast<PhysicsComponent*>(P);
    for (uint32_t i = 0; i < n; i++)
        if (!has_physics || has_physics[i]) relay.submit(PhysicsPacket{t[i], p[i]});
    auto t0 = std::chrono::steady_clock::now();
    relay.engage(dt);
    auto t1 = std::chrono::steady_clock::now();
    if (seconds) *seconds = std::chrono::duration<double>(t1 - t0).count();
    relay.clear();
    return 0;
}

// Collision pass.  cand == NULL: all pairs (the reference loop).  Otherwise only
// the listed packet pairs, which must be sorted lexicographically.
int icxref_collide(uint32_t n, const IcxEntity* id,
                   IcxTransform* T, IcxPhysics* P, const uint8_t* has_physics,
                   IcxColliderComponent* C,
                   uint8_t* tstate, uint16_t* tcoh, uint16_t coherency,
                   uint32_t host_timeslice, uint32_t flags, uint32_t worldline_stride,
                   const IcxPair* cand, size_t n_cand,
                   IcxContact* hits, size_t hit_cap, size_t* n_hits,
                   IcxInterception* ints, size_t int_cap, size_t* n_ints,
                   double* seconds)
{
    if (!T || !C) return 1;
    CollideCtx cx;
    cx.n = n; cx.id = id;
    cx.T = reinterpret_cast<TransformComponent*>(T);
    cx.P = reinterpret_cast<PhysicsComponent*>(P);
    cx.hasP = has_physics;
    cx.C = reinterpret_cast<ColliderComponent*>(C);
    cx.tstate = tstate; cx.tcoh = tcoh; cx.coherency = coherency; cx.host = host_timeslice; cx.flags = flags;
    cx.stride = worldline_stride;
    cx.hits = hits; cx.hcap = hit_cap; cx.nh = 0; cx.ints = ints; cx.icap = int_cap; cx.ni = 0;
    // ColliderSynchro::update, collider_synchro.cpp:28-42
    for (uint32_t i = 0; i < n; i++)
        for (int s = 0; s < (int)COLLIDERS_PER_ENTITY; s++)
            if (cx.C[i].colliders[s].isActive) {
                cx.C[i].colliders[s].reset();
                cx.packets.push_back(Packet{i, (uint8_t)s});
                cx.cps.push_back(ColliderPacket{cx.T[i], cx.C[i].colliders[s], (Entity)0, std::weak_ptr<Cosmos>()});
            }
    const uint32_t np = (uint32_t)cx.packets.size();
    auto t0 = std::chrono::steady_clock::now();
    if (!cand) {
        for (uint32_t a = 0; a < np; a++) {
            if (cx.cps[a].collider.colliderType == ColliderType::none || !cx.cps[a].collider.isActive) continue;  // :32
            for (uint32_t b = a + 1; b < np; b++) cx.visit(a, b);
        }
    } else {
        for (size_t k = 0; k < n_cand; k++) {
            uint32_t a = cand[k].a, b = cand[k].b;
            if (a >= b || b >= np) return 1;
            if (k && !(cand[k - 1].a < a || (cand[k - 1].a == a && cand[k - 1].b < b))) return 1;
            if (cx.cps[a].collider.colliderType == ColliderType::none || !cx.cps[a].collider.isActive) continue;
            cx.visit(a, b);
        }
    }
    auto t1 = std::chrono::steady_clock::now();
    if (seconds) *seconds = std::chrono::duration<double>(t1 - t0).count();
    if (n_hits) *n_hits = cx.nh;
    if (n_ints) *n_ints = cx.ni;
    return 0;
}

// Single-pair probes of the reference procedures, for per-function KATs.
int icxref_intersect(const IcxTransform* tA, const IcxCollider* cA, const IcxTransform* tB, const IcxCollider* cB,
                     uint32_t flags, float* point, float* normal, float* depth, float* minParamA, float* minParamB)
{
    TransformComponent TA = *reinterpret_cast<const TransformComponent*>(tA), TB = *reinterpret_cast<const TransformComponent*>(tB);
    Collider CA = *reinterpret_cast<const Collider*>(cA), CB = *reinterpret_cast<const Collider*>(cB);
    ColliderPacket a{TA, CA, (Entity)0, std::weak_ptr<Cosmos>()}, b{TB, CB, (Entity)1, std::weak_ptr<Cosmos>()};
    glm::vec3 p(0.0f), n(0.0f); float d = 0.0f;
    size_t ta = (size_t)CA.colliderType, tb = (size_t)CB.colliderType;
    bool hit;
    if ((flags & ICX_F_SPHERE_SPHERE) && CA.colliderType == ColliderType::sphere && CB.colliderType == ColliderType::sphere)
        hit = sphere_sphere_intersect(a, b, p, n, d);
    else hit = intersectProcedures[ta][tb](a, b, p, n, d);
    for (int k = 0; k < 3; k++) { point[k] = p[k]; normal[k] = n[k]; }
    *depth = d;
    if (minParamA) *minParamA = CA.minParametricValue;
    if (minParamB) *minParamB = CB.minParametricValue;
    return hit ? 1 : 0;
}

int icxref_response(IcxTransform* tA, IcxPhysics* pA, IcxCollider* cA, IcxTransform* tB, IcxPhysics* pB, IcxCollider* cB,
                    float* point, float* normal, float* depth)
{
    ColliderPacket a{*reinterpret_cast<TransformComponent*>(tA), *reinterpret_cast<Collider*>(cA), (Entity)0, std::weak_ptr<Cosmos>()};
    ColliderPacket b{*reinterpret_cast<TransformComponent*>(tB), *reinterpret_cast<Collider*>(cB), (Entity)1, std::weak_ptr<Cosmos>()};
    glm::vec3 p(point[0], point[1], point[2]), n(normal[0], normal[1], normal[2]); float d = *depth;
    responseProcedures[(size_t)a.collider.collisionType][(size_t)b.collider.collisionType](
        a, *reinterpret_cast<PhysicsComponent*>(pA), b, *reinterpret_cast<PhysicsComponent*>(pB), p, n, d);
    for (int k = 0; k < 3; k++) { point[k] = p[k]; normal[k] = n[k]; }
    *depth = d;
    return 0;
}

} // extern "C"
