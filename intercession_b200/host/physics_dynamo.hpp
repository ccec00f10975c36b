// Host-side mirror of the reference's physics operator interface, forwarding to the C ABI.
//
//   reference                                   here (namespace icx_host)
//   pleep::PhysicsDynamo          physics_dynamo.h:18-44      PhysicsDynamo   (same public methods)
//   pleep::A_PhysicsRelay         a_physics_relay.h:13-33     A_PhysicsRelay  (engage / clear)
//   pleep::EulerPhysicsRelay      euler_physics_relay.h       EulerPhysicsRelay
//   pleep::CollisionPhysicsRelay  collision_physics_relay.h   CollisionPhysicsRelay
//   pleep::PhysicsPacket          physics_packet.h:11-15      PhysicsPacket  {Transform&, Physics&}
//   pleep::ColliderPacket         collider_packet.h:16-25     ColliderPacket {Transform&, Collider&, collidee}
//
// Packets hold NON-OWNING references into the caller's component storage, valid until
// reset_relays() -- exactly the reference's contract.  run_relays(dt) gathers them into the AoS
// layouts of include/icx.h (byte-identical to the reference structs), runs the tick on the GPU,
// writes the results back through the same references and then replays, in the reference's pair
// order, the two up-calls the reference makes from inside its pair loop:
//   on_collision(A, B, n, depth, point) for colliders with useBehaviorsResponse (A side, then B
//   side with inverted normal / point, collision_physics_relay.h:114-155) and
//   TIMESTREAM_INTERCEPTION{agent, recipient} (collision_physics_relay.h:208-228).
// Errors: the C ABI never throws; this adapter turns a non-zero status into std::runtime_error,
// which is what the reference's own code raises.  No CPU fallback exists.
//
// Header-only; link with libicx.so.  In the reference tree TransformComponent / PhysicsComponent /
// Collider are the glm-based structs; they can be passed here with reinterpret_cast because the
// layouts are byte-identical (static_asserts in oracle/ref/ref_harness.cpp check that against the
// real headers).
#pragma once
#include <cstdint>
#include <algorithm>
#include <cstring>
#include <functional>
#include <memory>
#include <stdexcept>
#include <string>
#include <unordered_map>
#include <vector>

#include "../../include/icx.h"
#include "component_pools.hpp"

namespace icx_host {

using Entity = IcxEntity;
using TransformComponent = IcxTransform;
using PhysicsComponent = IcxPhysics;
using Collider = IcxCollider;
using ColliderComponent = IcxColliderComponent;

// ---- entity id helpers on the widened id (source/ecs/ecs_types.h:80-155) ----
inline uint32_t derive_timeslice_id(Entity e) { return e >> 28; }
inline uint32_t derive_genesis_id(Entity e) { return (e >> 4) & 0xFFFFFFu; }
inline uint32_t derive_causal_chain_link(Entity e) { return e & 0xFu; }
inline Entity compose_entity(uint32_t t, uint32_t g, uint32_t c) { return (t << 28) | (g << 4) | c; }
inline bool increment_causal_chain_link(Entity& e)
{
    if (e == ICX_NULL_ENTITY) return false;
    const uint32_t link = derive_causal_chain_link(e) + 1;
    if (link >= ICX_NULL_CHAINLINK) return false;
    e = compose_entity(derive_timeslice_id(e), derive_genesis_id(e), link);
    return true;
}
inline bool decrement_causal_chain_link(Entity& e)
{
    if (e == ICX_NULL_ENTITY) return false;
    const uint32_t link = derive_causal_chain_link(e);
    if (link >= ICX_NULL_CHAINLINK) return false;
    if (link == 0) throw std::range_error("Called decrement on entity whose chainlink value is at minimum (0).");
    e = compose_entity(derive_timeslice_id(e), derive_genesis_id(e), link - 1);
    return true;
}

struct PhysicsPacket { TransformComponent& transform; PhysicsComponent& physics; };
struct ColliderPacket { TransformComponent& transform; Collider& collider; Entity collidee; };

// what I_BehaviorsDrivetrain::on_collision receives (i_behaviors_drivetrain.h:43)
using CollisionCallback = std::function<void(ColliderPacket collider, ColliderPacket collidee, const float normal[3], float depth, const float point[3])>;
// events::cosmos::TIMESTREAM_INTERCEPTION (event_types.h:133-139)
using InterceptionCallback = std::function<void(Entity agent, Entity recipient)>;
// Cosmos::get_timestream_state / set_timestream_state (cosmos.cpp:155-178)
using StateGetter = std::function<std::pair<uint8_t, uint16_t>(Entity)>;
using StateSetter = std::function<void(Entity, uint8_t state, uint16_t coherency)>;

inline void check(int rc, IcxWorld* w, const char* what)
{
    if (rc != ICX_OK) throw std::runtime_error(std::string(what) + ": " + icx_status_string(rc) + " -- " + (w ? icx_last_error(w) : ""));
}

class PhysicsDynamo;

class A_PhysicsRelay {
public:
    explicit A_PhysicsRelay(PhysicsDynamo* owner) : m_owner(owner) {}
    virtual ~A_PhysicsRelay() = default;
    virtual void engage(double deltaTime) = 0;
    virtual void clear() = 0;
protected:
    PhysicsDynamo* m_owner;
};

class PhysicsDynamo {
public:
    // host_timeslice = Cosmos::get_host_id(); ICX_NULL_TIMESLICE for a parallel cosmos
    explicit PhysicsDynamo(uint32_t max_entities, uint32_t host_timeslice = 0, int device = 0, uint32_t flags = 0)
    {
        IcxConfig cfg;
        icx_config_default(&cfg);
        cfg.max_entities = max_entities; cfg.host_timeslice = host_timeslice; cfg.device = device; cfg.flags = flags;
        check(icx_world_create(&cfg, &m_world), nullptr, "icx_world_create (is a CUDA device present? there is no CPU fallback)");
        m_motionStep.reset(new Euler(this));
        m_collisionStep.reset(new Collision(this));
    }
    ~PhysicsDynamo() { icx_world_destroy(m_world); }
    PhysicsDynamo(const PhysicsDynamo&) = delete;
    PhysicsDynamo& operator=(const PhysicsDynamo&) = delete;

    // ---- the reference's public interface (physics_dynamo.cpp:23-53) ----
    void submit(PhysicsPacket data) { m_physics.push_back(data); }
    void submit(ColliderPacket data) { m_colliders.push_back(data); }
    void run_relays(double deltaTime)
    {
        if (m_pools) m_pools->to_device(m_world); else gather();
        check(icx_coherency_set(m_world, m_coherency), m_world, "icx_coherency_set");
        m_motionStep->engage(deltaTime);        // Euler then collision, as physics_dynamo.cpp:38-46
        m_collisionStep->engage(deltaTime);
        if (m_pools) m_pools->from_device(m_world, m_collidersBack); else scatter();
        replay_upcalls();
    }
    // Device-shared component storage instead of per-entity packets (host/component_pools.hpp): the synchros have nothing to
    // submit, run_relays sends the rows the host edited and takes Transform + Physics back in place.  colliders_back: also the
    // ColliderComponents (minParametricValue of ray colliders is written by the tick).
    void bind(ComponentPools* pools, bool colliders_back = false) { m_pools = pools; m_collidersBack = colliders_back; }
    void reset_relays() { m_motionStep->clear(); m_collisionStep->clear(); }

    // ---- what the reference reaches through Cosmos / EventBroker / BehaviorsComponent ----
    void set_coherency(uint16_t c) { m_coherency = c; }                        // Cosmos::get_coherency
    void set_collision_callback(CollisionCallback cb) { m_onCollision = std::move(cb); }
    void set_interception_callback(InterceptionCallback cb) { m_onInterception = std::move(cb); }
    void set_timestream_state_access(StateGetter g, StateSetter s) { m_getState = std::move(g); m_setState = std::move(s); }
    IcxWorld* world() { return m_world; }
    const std::vector<IcxContact>& last_contacts() const { return m_contacts; }

private:
    struct Euler : A_PhysicsRelay {
        using A_PhysicsRelay::A_PhysicsRelay;
        void engage(double dt) override { check(icx_euler_engage(m_owner->m_world, dt), m_owner->m_world, "icx_euler_engage"); }
        void clear() override { m_owner->m_physics.clear(); }
    };
    struct Collision : A_PhysicsRelay {
        using A_PhysicsRelay::A_PhysicsRelay;
        void engage(double) override { check(icx_collision_engage(m_owner->m_world), m_owner->m_world, "icx_collision_engage"); }
        void clear() override { m_owner->m_colliders.clear(); }
    };

    // one row per entity, in submission (= ascending id) order; packets of one entity share a Transform
    void gather()
    {
        // The packets of one frame are the packets of the last one unless entities came or went: the row table (and the
        // device's packet structure with it) is rebuilt only when the submitted references differ.
        m_sig.clear();
        for (auto& p : m_colliders) { m_sig.push_back(&p.transform); m_sig.push_back(&p.collider); m_sig.push_back((const void*)(uintptr_t)p.collidee); }
        m_sig.push_back(nullptr);
        for (auto& p : m_physics) { m_sig.push_back(&p.transform); m_sig.push_back(&p.physics); }
        const bool same_rows = m_uploaded && m_sig == m_sigLast;
        if (!same_rows) {
            m_rows.clear(); m_index.clear();
            auto row_of = [&](TransformComponent* t) -> size_t {
                auto it = m_index.find(t);
                if (it != m_index.end()) return it->second;
                m_index.emplace(t, m_rows.size());
                m_rows.push_back(Row{t, nullptr, {nullptr, nullptr}, ICX_NULL_ENTITY});
                return m_rows.size() - 1;
            };
            for (auto& p : m_colliders) {                               // collider_synchro.cpp:28-42: slot 0 then slot 1
                Row& r = m_rows[row_of(&p.transform)];
                r.entity = p.collidee;
                r.col[r.col[0] ? 1 : 0] = &p.collider;
            }
            for (auto& p : m_physics) m_rows[row_of(&p.transform)].physics = &p.physics;
            m_rowOf.clear();
            for (size_t i = 0; i < m_rows.size(); i++) m_rowOf.emplace_back(m_rows[i].entity, i);
            std::sort(m_rowOf.begin(), m_rowOf.end());
            m_sigLast = m_sig;
        }
        const size_t n = m_rows.size();
        m_ids.resize(n); m_T.resize(n); m_P.resize(n); m_hasP.resize(n); m_Cnew.resize(n); m_ts.resize(n); m_tc.resize(n);
        bool state_changed = false;
        for (size_t i = 0; i < n; i++) {
            const Row& r = m_rows[i];
            m_ids[i] = r.entity; m_T[i] = *r.transform;
            m_hasP[i] = r.physics ? 1 : 0;
            if (r.physics) m_P[i] = *r.physics;
            for (int s = 0; s < ICX_COLLIDERS_PER_ENTITY; s++) {
                if (r.col[s]) m_Cnew[i].colliders[s] = *r.col[s];
                else { std::memset(&m_Cnew[i].colliders[s], 0, sizeof(Collider)); m_Cnew[i].colliders[s] = Collider{}; m_Cnew[i].colliders[s].isActive = 0; }
            }
            std::pair<uint8_t, uint16_t> st{ICX_TS_MERGED, 0};
            if (m_getState && r.entity != ICX_NULL_ENTITY) st = m_getState(r.entity);
            if (same_rows && (m_ts2[i] != st.first || m_tc2[i] != st.second)) state_changed = true;      // the host changed a timestream state
            m_ts[i] = st.first; m_tc[i] = st.second;
        }
        const bool same_colliders = same_rows && m_C.size() == n && std::memcmp(m_C.data(), m_Cnew.data(), n * sizeof(ColliderComponent)) == 0;
        m_C.swap(m_Cnew);
        if (same_rows && same_colliders && !state_changed) {            // Transform + Physics only; the ColliderComponents stay resident
            check(icx_bodies_update(m_world, (uint32_t)n, nullptr, m_T.data(), m_P.data(), nullptr), m_world, "icx_bodies_update");
            return;
        }
        check(icx_bodies_upload(m_world, (uint32_t)n, m_ids.data(), m_T.data(), m_P.data(), m_hasP.data(), m_C.data(), m_ts.data(), m_tc.data()),
              m_world, "icx_bodies_upload");
        m_uploaded = true;
    }
    void scatter()
    {
        const size_t n = m_rows.size();
        m_ts2 = m_ts; m_tc2 = m_tc;
        check(icx_bodies_download(m_world, (uint32_t)n, m_T.data(), m_P.data(), m_C.data(), m_ts2.data(), m_tc2.data()), m_world, "icx_bodies_download");
        for (size_t i = 0; i < n; i++) {
            const Row& r = m_rows[i];
            *r.transform = m_T[i];
            if (r.physics) *r.physics = m_P[i];
            for (int s = 0; s < ICX_COLLIDERS_PER_ENTITY; s++) if (r.col[s]) *r.col[s] = m_C[i].colliders[s];
            if (m_setState && (m_ts2[i] != m_ts[i] || m_tc2[i] != m_tc[i])) m_setState(r.entity, m_ts2[i], m_tc2[i]);
        }
    }
    void replay_upcalls()
    {
        size_t n = 0;
        check(icx_contacts_fetch(m_world, nullptr, 0, &n), m_world, "icx_contacts_fetch");
        m_contacts.resize(n);
        if (n) check(icx_contacts_fetch(m_world, m_contacts.data(), n, &n), m_world, "icx_contacts_fetch");
        if (m_onCollision && m_pools) {
            for (const IcxContact& c : m_contacts) {
                if (!(c.flags & (ICX_CONTACT_BEHAVIORS_A | ICX_CONTACT_BEHAVIORS_B))) continue;
                const size_t ra = m_pools->row_of(c.entityA), rb = m_pools->row_of(c.entityB);
                // the callee gets mutable references, as in the reference: the rows count as edited
                ColliderPacket A{m_pools->edit_transform(ra), m_pools->edit_colliders(ra).colliders[c.slotA], c.entityA};
                ColliderPacket B{m_pools->edit_transform(rb), m_pools->edit_colliders(rb).colliders[c.slotB], c.entityB};
                upcall_pair(A, B, c);
            }
        } else if (m_onCollision) {
            for (const IcxContact& c : m_contacts) {
                if (!(c.flags & (ICX_CONTACT_BEHAVIORS_A | ICX_CONTACT_BEHAVIORS_B))) continue;
                const Row& ra = m_rows[row_of_entity(c.entityA)]; const Row& rb = m_rows[row_of_entity(c.entityB)];
                ColliderPacket A{*ra.transform, *ra.col[packet_slot(ra, c.slotA)], c.entityA}, B{*rb.transform, *rb.col[packet_slot(rb, c.slotB)], c.entityB};
                upcall_pair(A, B, c);
            }
        }
        if (m_onInterception) {
            check(icx_interceptions_fetch(m_world, nullptr, 0, &n), m_world, "icx_interceptions_fetch");
            std::vector<IcxInterception> ev(n);
            if (n) check(icx_interceptions_fetch(m_world, ev.data(), n, &n), m_world, "icx_interceptions_fetch");
            for (const IcxInterception& e : ev) m_onInterception(e.agent, e.recipient);
        }
    }
    void upcall_pair(ColliderPacket& A, ColliderPacket& B, const IcxContact& c)
    {
        if (c.flags & ICX_CONTACT_BEHAVIORS_A) m_onCollision(A, B, c.normal, c.depth, c.point);                     // :114-131
        if (c.flags & ICX_CONTACT_BEHAVIORS_B) {                                                                    // :133-155
            const float inv_n[3] = {-c.normal[0], -c.normal[1], -c.normal[2]};
            const float inv_p[3] = {c.point[0] - c.normal[0] * c.depth, c.point[1] - c.normal[1] * c.depth, c.point[2] - c.normal[2] * c.depth};
            m_onCollision(B, A, inv_n, c.depth, inv_p);
        }
    }
    // entity -> row of the packet path: a sorted table rebuilt only when the submitted references change
    size_t row_of_entity(Entity e) const
    {
        auto it = std::lower_bound(m_rowOf.begin(), m_rowOf.end(), std::make_pair(e, (size_t)0));
        return it->second;
    }
    struct Row { TransformComponent* transform; PhysicsComponent* physics; Collider* col[2]; Entity entity; };
    // IcxContact::slotX is the ColliderComponent slot; Row::col is filled in submission order
    static int packet_slot(const Row& r, uint8_t slot) { return (slot == 1 && r.col[1]) ? 1 : 0; }

    IcxWorld* m_world = nullptr;
    uint16_t m_coherency = 0;
    std::unique_ptr<A_PhysicsRelay> m_motionStep, m_collisionStep;
    std::vector<PhysicsPacket> m_physics;
    std::vector<ColliderPacket> m_colliders;
    std::vector<Row> m_rows;
    std::unordered_map<TransformComponent*, size_t> m_index;       // used only while the row table is rebuilt
    std::vector<std::pair<Entity, size_t>> m_rowOf;
    std::vector<const void*> m_sig, m_sigLast;
    std::vector<ColliderComponent> m_Cnew;
    bool m_uploaded = false;
    ComponentPools* m_pools = nullptr; bool m_collidersBack = false;
    std::vector<Entity> m_ids; std::vector<TransformComponent> m_T; std::vector<PhysicsComponent> m_P; std::vector<uint8_t> m_hasP;
    std::vector<ColliderComponent> m_C; std::vector<uint8_t> m_ts, m_ts2; std::vector<uint16_t> m_tc, m_tc2;
    std::vector<IcxContact> m_contacts;
    CollisionCallback m_onCollision; InterceptionCallback m_onInterception; StateGetter m_getState; StateSetter m_setState;
};

} // namespace icx_host
