// Drives the host-side mirror of the reference's operator interface (intercession_b200/host/*.hpp)
// the way ServerCosmosContext::_on_fixed drives the reference: every tick the
// "synchros" submit packets that reference the caller's component storage, run_relays(dt) ticks,
// reset_relays() drops the references.  Scene = BASELINE.json configs[0] (case-study scene, push).
// Prints one line per checkpoint; tests/test_gpu_host.py compares them with the golden fixture.
#include <cmath>
#include <cstdio>
#include <cstring>
#include "../../intercession_b200/host/timeline_api.hpp"
#include "../../intercession_b200/host/projectile_behaviors.hpp"

using namespace icx_host;

static TransformComponent transform(float x, float y, float z)
{
    TransformComponent t{}; t.origin[0] = x; t.origin[1] = y; t.origin[2] = z; t.orientation[0] = 1; t.scale[0] = t.scale[1] = t.scale[2] = 1; return t;
}
static PhysicsComponent physics(float mass)
{
    PhysicsComponent p{}; p.linearDrag = 0.01f; p.angularDrag = 0.01f; p.collisionAngularDrag = 0.03f; p.mass = mass; p.lockedOrientation[0] = 1; return p;   // physics_component.h:26-54
}
static Collider collider(int type)
{
    Collider c{}; c.colliderType = type; c.collisionType = ICX_COLLISION_RIGID; c.isActive = 1;
    c.localTransform = transform(0, 0, 0); c.inheritOrientation = 1; c.influenceOrientation = 1; c.minParametricValue = 1.0f;
    c.staticFriction = 0.45f; c.dynamicFriction = 0.45f; c.restitution = 0.70f; c.stiffness = 100.0f; c.damping = 10.0f; c.restLength = 0.1f;   // collider.h:58-92
    return c;
}

int main(int argc, char** argv)
{
    const int ticks = argc > 1 ? atoi(argv[1]) : 100;
    const float h = std::sqrt(0.5f);
    // ---- the caller's ECS storage (ComponentArray<T> in the reference) ----
    Entity ids[4]; TransformComponent T[4]; PhysicsComponent P[4]; ColliderComponent C[4];
    const float floors[2][6] = {{12, -4, 0, 24, 24, 2}, {-12, -6, 0, 24, 24, 0.05f}};
    for (int k = 0; k < 2; k++) {                       // moon_cosmos.cpp:25-72 floors
        ids[k] = compose_entity(0, k, ICX_NULL_CHAINLINK);
        T[k] = transform(floors[k][0], floors[k][1], floors[k][2]);
        T[k].orientation[0] = h; T[k].orientation[1] = -h; T[k].scale[0] = floors[k][3]; T[k].scale[1] = floors[k][4]; T[k].scale[2] = floors[k][5];
        P[k] = physics(0.0f); P[k].lockOrigin = 1; P[k].lockOrientation = 1;
        std::memcpy(P[k].lockedOrigin, T[k].origin, 12); std::memcpy(P[k].lockedOrientation, T[k].orientation, 16);
        C[k].colliders[0] = collider(ICX_COLLIDER_BOX); C[k].colliders[0].localTransform.origin[2] = -0.499f;
        C[k].colliders[1] = collider(ICX_COLLIDER_BOX); C[k].colliders[1].isActive = 0;
    }
    ids[2] = compose_entity(0, 2, 0);                   // "point" sphere
    T[2] = transform(12, -1, 0); P[2] = physics(100.0f);
    C[2].colliders[0] = collider(ICX_COLLIDER_SPHERE); C[2].colliders[0].dynamicFriction = 0.9f; C[2].colliders[0].restitution = 0.1f;
    C[2].colliders[1] = collider(ICX_COLLIDER_BOX); C[2].colliders[1].isActive = 0;
    ids[3] = compose_entity(0, 3, 0);                   // focal biped: box + ray-spring legs (client_focal_entity.h:30-91)
    T[3] = transform(12, 3, 0); P[3] = physics(40.0f); P[3].lockOrientation = 1;
    C[3].colliders[0] = collider(ICX_COLLIDER_BOX);
    Collider& legs = C[3].colliders[1];
    legs = collider(ICX_COLLIDER_RAY); legs.collisionType = ICX_COLLISION_SPRING;
    legs.localTransform.orientation[0] = h; legs.localTransform.orientation[1] = h; legs.localTransform.scale[2] = 1.5f;
    legs.stiffness = 2500; legs.damping = 400; legs.restLength = 0.3f; legs.staticFriction = 0; legs.dynamicFriction = 0;
    legs.inheritOrientation = 0; legs.influenceOrientation = 0; legs.useBehaviorsResponse = 1;

    int n_callbacks = 0, n_interceptions = 0;
    float last_ground[3] = {0, 0, 0};
    try {
        PhysicsDynamo dynamo(16);
        TimelineApi timeline(16, 8);
        dynamo.set_collision_callback([&](ColliderPacket self, ColliderPacket, const float n[3], float, const float[3]) {
            if (self.collidee == ids[3]) { n_callbacks++; std::memcpy(last_ground, n, 12); }       // BipedBehaviors::on_collision stores the ground normal
        });
        dynamo.set_interception_callback([&](Entity, Entity) { n_interceptions++; });
        for (int t = 0; t < ticks; t++) {
            dynamo.set_coherency((uint16_t)t);
            for (int e = 0; e < 4; e++) {               // PhysicsSynchro::update / ColliderSynchro::update
                dynamo.submit(PhysicsPacket{T[e], P[e]});
                for (int s = 0; s < 2; s++) if (C[e].colliders[s].isActive) {
                    C[e].colliders[s].minParametricValue = 1.0f;                               // Collider::reset, collider.h:75-78
                    dynamo.submit(ColliderPacket{T[e], C[e].colliders[s], ids[e]});
                }
            }
            dynamo.run_relays(1.0 / 45);
            timeline.push_past_timestream(dynamo, (uint16_t)t);
            dynamo.reset_relays();
            if (t + 1 == 1 || t + 1 == 10 || t + 1 == 100 || t + 1 == 1000)
                std::printf("tick %d point %.9g %.9g %.9g biped %.9g %.9g %.9g vel %.9g %.9g %.9g\n", t + 1, T[2].origin[0], T[2].origin[1], T[2].origin[2],
                            T[3].origin[0], T[3].origin[1], T[3].origin[2], P[3].velocity[0], P[3].velocity[1], P[3].velocity[2]);
        }
        // congruence of the live cosmos against what was recorded one tick ago
        const bool congruent = timeline.is_congruent(dynamo, (uint16_t)(ticks - 1));
        std::printf("callbacks %d interceptions %d ground_normal %.6g %.6g %.6g congruent %d\n", n_callbacks, n_interceptions,
                    last_ground[0], last_ground[1], last_ground[2], congruent ? 1 : 0);
        // ProjectileBehaviors on the device: the "point" sphere (row 2) becomes a projectile about to expire
        ProjectileBehaviors projectiles(dynamo.world());
        ProjectileComponent pc{299.99, 300.0, 1.0, 2.0, 1.0};
        projectiles.attach({2u}, {pc});
        projectiles.on_fixed_update(1.0 / 60);
        std::vector<ProjectileComponent> back;
        projectiles.fetch(back);
        std::printf("projectile life %.17g condemned %zu\n", back[0].currLifetime, projectiles.condemned().size());
        // ---- event records + entity lifecycle through the C++ mirror: a second cosmos (the "past") receives ENTITY_CREATED for the
        // biped at chain-link 1 in the same tick's record, creates it from the record's components, follows it, then removes it ----
        {
            PhysicsDynamo past(16, 1);
            Entity pid[3]; TransformComponent pT[3]; PhysicsComponent pP[3]; ColliderComponent pC[3]; uint8_t hp[3] = {1, 1, 1};
            for (int e = 0; e < 3; e++) { pid[e] = ids[e]; increment_causal_chain_link(pid[e]); pT[e] = T[e]; pP[e] = P[e]; pC[e] = C[e]; }
            check(icx_bodies_upload(past.world(), 3, pid, pT, pP, hp, pC, nullptr, nullptr), past.world(), "upload past");
            const uint16_t c = (uint16_t)ticks;
            Entity born = ids[3]; increment_causal_chain_link(born);
            IcxTimelineEvent ev{}; ev.type = ICX_EVENT_ENTITY_CREATED; ev.entity = born; ev.source = ICX_NULL_ENTITY; ev.payload[0] = 0xC1;
            timeline.push_past_event(c, &ev, 1);
            timeline.push_past_timestream(dynamo, c);                                   // the tick's record: all four entities at link + 1
            int created = 0, followed = 0, removed = 0;
            for (const IcxTimelineEvent& r : timeline.pop_future_events(c)) {
                if (r.type != ICX_EVENT_ENTITY_CREATED) continue;
                TransformComponent t; PhysicsComponent ph; ColliderComponent cc; bool has_p = false, has_c = false;
                if (!timeline.recorded_entity(c, r.entity, t, ph, has_p, cc, &has_c)) continue;
                if (!has_c) cc = C[3];          // its collider has not changed since the previous record: the ring carries the row only in the tick an entity is new to it
                const uint8_t one = has_p ? 1 : 0;
                created += (int)TimelineApi::create_entities(past, 1, &r.entity, &t, &ph, &one, &cc);
                followed += std::memcmp(&t, &T[3], sizeof(t)) == 0 && std::memcmp(&ph, &P[3], sizeof(ph)) == 0 ? 1 : 0;
            }
            const bool applied = timeline.pop_future_timestream(past, c);
            removed = (int)TimelineApi::remove_entities(past, 1, &born);
            std::printf("lifecycle created %d record_matches %d replayed %d removed %d\n", created, followed, applied ? 1 : 0, removed);
        }
    } catch (const std::exception& e) {
        std::printf("error: %s\n", e.what());
        return 2;
    }
    return 0;
}
