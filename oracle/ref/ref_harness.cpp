// TEST INFRASTRUCTURE -- not product code.  Only tests/, __graft_entry__.smoke()
// and bench.py's cpu_baseline / --impl reference leg may load the library
// built from this file (oracle/_ref/libicx_ref.so).
//
// Headless harness around the UNMODIFIED reference sources under
// /root/reference/source (compiled where they lie, see oracle/Makefile).
// Two entry families:
//   icxref_verbatim_run   -- the reference objects themselves: Cosmos +
//                            PhysicsSynchro/ColliderSynchro + PhysicsDynamo
//                            (EulerPhysicsRelay + CollisionPhysicsRelay), call
//                            sequence of SURVEY.md 3.1.  uint16 entity ids.
//   icxref_euler / icxref_collide -- the reference's relay / procedure tables
//                            (EulerPhysicsRelay::engage, intersectProcedures,
//                            responseProcedures) driven over plain arrays so
//                            more than 65 535 bodies and wide ids are possible.
//                            The pair loop and the interception predicate are
//                            restated from collision_physics_relay.h:24-231 and
//                            checked against the verbatim path in tests/.
#include "../../include/icx.h"

#include "core/cosmos.h"
#include "events/event_broker.h"
#include "physics/physics_dynamo.h"
#include "physics/physics_synchro.h"
#include "physics/collider_synchro.h"
#include "physics/physics_component.h"
#include "physics/collider_component.h"
#include "physics/transform_component.h"
#include "physics/euler_physics_relay.h"
#include "physics/collision_procedures.h"
#include "behaviors/behaviors_component.h"
#include "behaviors/i_behaviors_drivetrain.h"
#include "spacetime/timestream_state.h"

#include <chrono>
#include <map>

using namespace pleep;

static_assert(sizeof(TransformComponent) == sizeof(IcxTransform), "Transform layout");
static_assert(sizeof(PhysicsComponent) == sizeof(IcxPhysics), "Physics layout");
static_assert(sizeof(Collider) == sizeof(IcxCollider), "Collider layout");
static_assert(sizeof(ColliderComponent) == sizeof(IcxColliderComponent), "ColliderComponent layout");
static_assert(offsetof(PhysicsComponent, lockedOrientation) == offsetof(IcxPhysics, lockedOrientation), "Physics offsets");
static_assert(offsetof(PhysicsComponent, isAsleep) == offsetof(IcxPhysics, isAsleep), "Physics offsets");
static_assert(offsetof(Collider, localTransform) == offsetof(IcxCollider, localTransform), "Collider offsets");
static_assert(offsetof(Collider, minParametricValue) == offsetof(IcxCollider, minParametricValue), "Collider offsets");
static_assert(offsetof(Collider, restLength) == offsetof(IcxCollider, restLength), "Collider offsets");

// The real BehaviorsLibrary (behaviors/behaviors_library.cpp) drags in GL-bound
// staging headers; nothing on the physics path fetches a drivetrain from it.
std::shared_ptr<I_BehaviorsDrivetrain> BehaviorsLibrary::fetch_behaviors(BehaviorsType) { return nullptr; }

namespace {

inline IcxEntity widen(Entity e) { return ((uint32_t)(e & 0xF000u) << 16) | (uint32_t)(e & 0x0FFFu); }

struct HitSink {
    IcxContact* out = nullptr; size_t cap = 0; size_t n = 0;
    void push(const IcxContact& c) { if (n < cap) out[n] = c; n++; }
};
struct IntSink {
    IcxInterception* out = nullptr; size_t cap = 0; size_t n = 0;
    void push(IcxEntity a, IcxEntity r) { if (n < cap) { out[n].agent = a; out[n].recipient = r; } n++; }
};

// ---- verbatim path -------------------------------------------------------
struct Recorder : public I_BehaviorsDrivetrain {
    Cosmos* cosmos = nullptr;
    HitSink* sink = nullptr;
    std::unordered_map<Entity, uint32_t> firstPacket;   // entity -> packet index of its first active collider
    std::unordered_map<Entity, std::array<uint8_t, 2>> wantsBehaviors;  // caller's original flags
    std::map<IcxEntity, IcxGround> grounds;                             // last on_collision call per entity (cleared per tick)
    uint32_t packet_of(const ColliderPacket& p, uint8_t& slot) {
        ColliderComponent& cc = cosmos->get_component<ColliderComponent>(p.collidee);
        slot = (&p.collider == &cc.colliders[1]) ? 1 : 0;
        uint32_t base = firstPacket[p.collidee];
        return base + ((slot == 1 && cc.colliders[0].isActive) ? 1u : 0u);
    }
    void on_collision(ColliderPacket a, ColliderPacket b, glm::vec3 n, float depth, glm::vec3 point,
                      std::shared_ptr<EventBroker>) override {
        uint8_t sa, sb;
        uint32_t pa = packet_of(a, sa), pb = packet_of(b, sb);
        if (wantsBehaviors[a.collidee][sa]) {
            // what BipedBehaviors::on_collision does with these very arguments (behaviors/biped_behaviors.h:348-351), evaluated
            // INSIDE the reference's own callback: callerData.transform is the live transform at callback time
            IcxGround g{};
            g.entity = widen(a.collidee);
            for (int k = 0; k < 3; k++) g.normal[k] = n[k];
            g.dist = glm::length(a.transform.origin - point);
            g.packetA = pa < pb ? pa : pb; g.packetB = pa < pb ? pb : pa; g.side = pa < pb ? 0u : 1u;
            grounds[g.entity] = g;
        }
        if (pa > pb) return;   // B-side callback of the same hit (inverted data): recorded once, from A
        IcxContact c{};
        c.entityA = widen(a.collidee); c.entityB = widen(b.collidee);
        c.slotA = sa; c.slotB = sb; c.packetA = pa; c.packetB = pb;
        c.flags = 0;
        if (wantsBehaviors[a.collidee][sa]) c.flags |= ICX_CONTACT_BEHAVIORS_A;
        if (wantsBehaviors[b.collidee][sb]) c.flags |= ICX_CONTACT_BEHAVIORS_B;
        for (int k = 0; k < 3; k++) { c.point[k] = point[k]; c.normal[k] = n[k]; }
        c.depth = depth;
        sink->push(c);
    }
};

struct InterceptionListener {
    std::shared_ptr<Cosmos> cosmos; IntSink* sink; bool parallel;
    // Restates ServerNetworkDynamo::_timestream_interception_handler
    // (server/server_network_dynamo.cpp:813-838) resp. the parallel one
    // (spacetime/parallel_network_dynamo.cpp:455-467): neither file compiles
    // here (asio absent), and the state they set is observed by later pairs.
    void handle(EventMessage& evRef) {
        EventMessage ev = evRef;    // handlers pop a copy; later listeners still see the params
        events::cosmos::TIMESTREAM_INTERCEPTION_params info;
        ev >> info;
        sink->push(widen(info.agent), widen(info.recipient));
        if (parallel) {
            cosmos->set_timestream_state(info.recipient, TimestreamState::forked);
        } else {
            if (derive_causal_chain_link(info.recipient) <= 0) return;
            cosmos->set_timestream_state(info.recipient, TimestreamState::forking);
        }
    }
};

} // namespace

static std::map<IcxEntity, IcxGround> g_last_grounds;

extern "C" {

int icxref_abi_version(void) { return ICX_ABI_VERSION; }

// per entity: the arguments of the last on_collision call of the LAST tick of the last icxref_verbatim_run, as
// BipedBehaviors::on_collision would have stored them (ascending entity id)
int icxref_last_grounds(IcxGround* out, size_t cap, size_t* n)
{
    size_t k = 0;
    for (auto& kv : g_last_grounds) { if (out && k < cap) out[k] = kv.second; k++; }
    if (n) *n = k;
    return 0;
}

// Reference objects, reference call sequence:
//   cosmos->update(); dynamo->run_relays(dt); dynamo->reset_relays(); cosmos->increment_coherency();
// Contacts / interceptions returned are those of the LAST tick.  tstate / tcoh
// are in-out.  Returns 0, or 1 on bad arguments.
int icxref_verbatim_run(uint32_t n, const uint16_t* ids,
                        IcxTransform* T, IcxPhysics* P, const uint8_t* has_physics,
                        IcxColliderComponent* C, const uint8_t* has_collider,
                        uint8_t* tstate, uint16_t* tcoh, uint16_t* coherency,
                        uint32_t host_timeslice, double dt, uint32_t n_ticks,
                        IcxContact* hits, size_t hit_cap, size_t* n_hits,
                        IcxInterception* ints, size_t int_cap, size_t* n_ints,
                        double* seconds)
{
    if (!ids || !T) return 1;
    auto broker = std::make_shared<EventBroker>();
    auto cosmos = std::make_shared<Cosmos>(broker, (TimesliceId)host_timeslice);
    auto dynamo = std::make_shared<PhysicsDynamo>(broker);
    cosmos->register_component<TransformComponent>();
    cosmos->register_component<PhysicsComponent>();
    cosmos->register_component<ColliderComponent>();
    cosmos->register_component<BehaviorsComponent>();
    cosmos->register_synchro<PhysicsSynchro>()->attach_dynamo(dynamo);
    cosmos->register_synchro<ColliderSynchro>()->attach_dynamo(dynamo);

    HitSink hs; hs.out = hits; hs.cap = hit_cap;
    IntSink is; is.out = ints; is.cap = int_cap;
    auto rec = std::make_shared<Recorder>();
    rec->cosmos = cosmos.get(); rec->sink = &hs;

    InterceptionListener lis{cosmos, &is, host_timeslice == NULL_TIMESLICEID};
    std::function<void(EventMessage&)> cb = [&lis](EventMessage& ev) { lis.handle(ev); };
    broker->add_listener(events::cosmos::TIMESTREAM_INTERCEPTION, cb);

    // entities must be registered in ascending id order to mirror std::set iteration
    std::vector<uint32_t> order(n);
    for (uint32_t i = 0; i < n; i++) order[i] = i;
    std::sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return ids[a] < ids[b]; });
    for (uint32_t i = 0; i + 1 < n; i++) if (ids[i] >= ids[i + 1]) return 1;  // caller must pass ascending ids

    uint32_t packet = 0;
    for (uint32_t i = 0; i < n; i++) {
        Entity e = ids[i];
        if (!cosmos->register_entity(e)) return 1;
        cosmos->add_component(e, *reinterpret_cast<TransformComponent*>(&T[i]));
        if (P && (!has_physics || has_physics[i]))
            cosmos->add_component(e, *reinterpret_cast<PhysicsComponent*>(&P[i]));
        if (C && (!has_collider || has_collider[i])) {
            ColliderComponent cc = *reinterpret_cast<ColliderComponent*>(&C[i]);
            rec->firstPacket[e] = packet;
            std::array<uint8_t, 2> want{};
            for (int s = 0; s < 2; s++) {
                want[s] = cc.colliders[s].useBehaviorsResponse;
                if (cc.colliders[s].isActive) packet++;
                cc.colliders[s].useBehaviorsResponse = true;   // instrument: record every hit
            }
            rec->wantsBehaviors[e] = want;
            cosmos->add_component(e, cc);
            BehaviorsComponent bc; bc.drivetrain = rec;
            cosmos->add_component(e, bc);
        }
        if (tstate && (tstate[i] != 0 || (tcoh && tcoh[i] != 0))) {
            cosmos->set_coherency(tcoh ? tcoh[i] : 0);
            cosmos->set_timestream_state(e, (TimestreamState)tstate[i]);
        }
    }
    cosmos->set_coherency(coherency ? *coherency : 0);

    auto t0 = std::chrono::steady_clock::now();
    for (uint32_t t = 0; t < n_ticks; t++) {
        hs.n = 0; is.n = 0; rec->grounds.clear();
        cosmos->update();
        dynamo->run_relays(dt);
        dynamo->reset_relays();
        cosmos->increment_coherency();
    }
    auto t1 = std::chrono::steady_clock::now();
    if (seconds) *seconds = std::chrono::duration<double>(t1 - t0).count();

    for (uint32_t i = 0; i < n; i++) {
        Entity e = ids[i];
        T[i] = *reinterpret_cast<IcxTransform*>(&cosmos->get_component<TransformComponent>(e));
        if (P && (!has_physics || has_physics[i]))
            P[i] = *reinterpret_cast<IcxPhysics*>(&cosmos->get_component<PhysicsComponent>(e));
        if (C && (!has_collider || has_collider[i])) {
            ColliderComponent cc = cosmos->get_component<ColliderComponent>(e);
            for (int s = 0; s < 2; s++) cc.colliders[s].useBehaviorsResponse = rec->wantsBehaviors[e][s];
            C[i] = *reinterpret_cast<IcxColliderComponent*>(&cc);
        }
        if (tstate) {
            auto st = cosmos->get_timestream_state(e);
            tstate[i] = (uint8_t)st.first;
            if (tcoh) tcoh[i] = st.second;
        }
    }
    g_last_grounds = rec->grounds;
    if (coherency) *coherency = cosmos->get_coherency();
    if (n_hits) *n_hits = hs.n;
    if (n_ints) *n_ints = is.n;
    return 0;
}

} // extern "C"
