// Device-shared component pools (intercession_b200/host/component_pools.hpp) against the packet path of
// the same PhysicsDynamo (references into caller storage, everything resubmitted every tick): the same scene, the same host
// edits (a few velocities per tick, one entity created, one removed), the states compared byte for byte after every tick.
// Then both paths timed through PhysicsDynamo::run_relays.
//   test_host_pools <bodies> <ticks> [time_only]
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include "../../intercession_b200/host/physics_dynamo.hpp"

using namespace icx_host;

static TransformComponent transform(float x, float y, float z, float s)
{
    TransformComponent t{}; t.origin[0] = x; t.origin[1] = y; t.origin[2] = z; t.orientation[0] = 1; t.scale[0] = t.scale[1] = t.scale[2] = s; return t;
}
static PhysicsComponent physics(float mass)
{
    PhysicsComponent p{}; p.linearDrag = 0.01f; p.angularDrag = 0.01f; p.collisionAngularDrag = 0.03f; p.mass = mass; p.lockedOrientation[0] = 1; return p;
}
static Collider collider(int type, bool active)
{
    Collider c{}; c.colliderType = type; c.collisionType = ICX_COLLISION_RIGID; c.isActive = active ? 1 : 0;
    c.localTransform = transform(0, 0, 0, 1); c.inheritOrientation = 1; c.influenceOrientation = 1; c.minParametricValue = 1.0f;
    c.staticFriction = 0.45f; c.dynamicFriction = 0.45f; c.restitution = 0.70f; c.stiffness = 100.0f; c.damping = 10.0f; c.restLength = 0.1f;
    return c;
}
static uint32_t rng_state = 12345u;
static float rnd() { rng_state = rng_state * 1664525u + 1013904223u; return (float)(rng_state >> 8) / 16777216.0f; }

struct Body { Entity id; TransformComponent T; PhysicsComponent P; ColliderComponent C; bool alive; };

static Body make_body(uint32_t g, float x, float y, float z)
{
    Body b; std::memset(&b, 0, sizeof(b));
    b.id = compose_entity(0, g, ICX_NULL_CHAINLINK); b.alive = true;
    b.T = transform(x, y, z, 0.5f + 0.4f * rnd());
    b.P = physics(1.0f + 4.0f * rnd());
    for (int k = 0; k < 3; k++) b.P.velocity[k] = 2.0f * rnd() - 1.0f;
    b.C.colliders[0] = collider(g % 3 == 0 ? ICX_COLLIDER_SPHERE : ICX_COLLIDER_BOX, true);
    b.C.colliders[1] = collider(ICX_COLLIDER_BOX, false);
    return b;
}

static double now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

int main(int argc, char** argv)
{
    const uint32_t n = argc > 1 ? (uint32_t)atoi(argv[1]) : 4096;
    const int ticks = argc > 2 ? atoi(argv[2]) : 12;
    const bool time_only = argc > 3;
    const uint32_t side = (uint32_t)std::ceil(std::cbrt((double)n));
    std::vector<Body> bodies; bodies.reserve(n + 8);
    {   // floor: an immutable box (no physics component)
        Body f = make_body(0, 0, -2.0f, 0);
        f.T = transform(0.5f * 1.2f * side, -1.5f, 0.5f * 1.2f * side, 1.0f); f.T.scale[0] = f.T.scale[2] = 1.3f * side + 4.0f; f.T.scale[1] = 1.0f;
        f.P = physics(0.0f); f.P.lockOrigin = 1; f.P.lockOrientation = 1; std::memcpy(f.P.lockedOrigin, f.T.origin, 12);
        f.C.colliders[0] = collider(ICX_COLLIDER_BOX, true);
        bodies.push_back(f);
    }
    for (uint32_t i = 0; i < n; i++)
        bodies.push_back(make_body(1 + i, 1.2f * (i % side) + 0.1f * rnd(), 1.2f * ((i / side) % side) + 0.1f * rnd(), 1.2f * (i / (side * side)) + 0.1f * rnd()));
    const Body extra = make_body(n + 100, -3.0f, 2.0f, -3.0f);
    try {
        // ---- path A: pools bound to the dynamo ----
        ComponentPools pools(bodies.size() + 8);
        for (const Body& b : bodies) pools.insert(b.id, b.T, &b.P, b.C);
        PhysicsDynamo da((uint32_t)bodies.size() + 8), db((uint32_t)bodies.size() + 8);
        da.bind(&pools);
        // ---- path B: the caller's own storage, packets every tick ----
        std::vector<Body> store = bodies; store.reserve(bodies.size() + 8);
        size_t sent_total = 0, full_uploads = 0; int mismatches = 0;
        double sec_a = 0, sec_b = 0;
        for (int t = 0; t < ticks; t++) {
            // the host's own work between ticks: a handful of velocity edits; an entity comes at tick 3 and one goes at tick 5
            if (!time_only) {
                for (int k = 0; k < 5; k++) {
                    const size_t r = 1 + (size_t)(rnd() * (float)(n - 1));
                    const Entity id = bodies[r].id;
                    const size_t row = pools.row_of(id);
                    if (row == ComponentPools::npos) continue;
                    const float kick = 3.0f * rnd();
                    pools.edit_physics(row).velocity[1] += kick;
                    for (Body& b : store) if (b.alive && b.id == id) b.P.velocity[1] += kick;
                }
                if (t == 3) { pools.insert(extra.id, extra.T, &extra.P, extra.C); store.push_back(extra); }
                if (t == 5) { const Entity gone = bodies[n / 2].id; pools.remove(gone); for (Body& b : store) if (b.id == gone) b.alive = false; }
            }
            double t0 = now();
            da.set_coherency((uint16_t)t);
            da.run_relays(1.0 / 45);
            da.reset_relays();
            sec_a += now() - t0;
            sent_total += pools.last_sync().rows_sent; full_uploads += pools.last_sync().full_upload ? 1 : 0;
            t0 = now();
            db.set_coherency((uint16_t)t);
            // PhysicsSynchro / ColliderSynchro::update in ascending id order (std::set<Entity>): `extra` sorts last, as in the pools
            for (Body& b : store) if (b.alive) {
                db.submit(PhysicsPacket{b.T, b.P});
                for (int s = 0; s < 2; s++) if (b.C.colliders[s].isActive) db.submit(ColliderPacket{b.T, b.C.colliders[s], b.id});
            }
            db.run_relays(1.0 / 45);
            db.reset_relays();
            sec_b += now() - t0;
            if (!time_only) for (const Body& b : store) if (b.alive) {
                const size_t row = pools.row_of(b.id);
                if (row == ComponentPools::npos || std::memcmp(&pools.transform(row), &b.T, sizeof(b.T)) || std::memcmp(&pools.physics(row), &b.P, sizeof(b.P))) mismatches++;
            }
        }
        size_t alive = 0; for (const Body& b : store) alive += b.alive ? 1 : 0;
        std::printf("pools entities %zu alive %zu mismatches %d full_uploads %zu rows_sent %zu\n", pools.size(), alive, mismatches, full_uploads, sent_total);
        std::printf("timing bodies %u ticks %d pools_ms_per_tick %.3f packets_ms_per_tick %.3f\n", n, ticks, 1e3 * sec_a / ticks, 1e3 * sec_b / ticks);
        const float y0 = pools.transform(pools.row_of(bodies[1].id)).origin[1];
        std::printf("moved %d\n", y0 != bodies[1].T.origin[1] ? 1 : 0);
        return mismatches ? 1 : 0;
    } catch (const std::exception& e) {
        std::printf("error: %s\n", e.what());
        return 2;
    }
}
