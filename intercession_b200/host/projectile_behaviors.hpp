// Host-side mirror of the reference's projectile drivetrain, forwarding to the C ABI.
//
//   reference                                                   here (namespace icx_host)
//   pleep::ProjectileComponent  behaviors/projectile_component.h:8-16   ProjectileComponent (= IcxProjectile, 40 B)
//   pleep::ProjectileBehaviors  behaviors/projectile_behaviors.h:13-55  ProjectileBehaviors
//
// The reference calls ProjectileBehaviors::on_fixed_update(dt, behaviors, entity, cosmos, broker) once per
// BehaviorsPacket and fixed step (behaviors_dynamo -> fixed relay), each call touching one ProjectileComponent
// and one TransformComponent.  Here the components live on the device: attach() hands them over once,
// on_fixed_update(dt) is ONE call per fixed step for all of them, and condemned() lists what the reference
// would have passed to Cosmos::condemn_entity (cosmos.h:339-345) so the caller's ECS can destroy them.
// The scales land where PhysicsDynamo::run_relays reads them; no per-tick icx_bodies_update is needed.
//
// Header-only; link with libicx.so.  Errors become std::runtime_error (what the reference raises).
#pragma once
#include <cstdint>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/icx.h"

namespace icx_host {

using ProjectileComponent = IcxProjectile;

class ProjectileBehaviors {
public:
    explicit ProjectileBehaviors(IcxWorld* world) : m_world(world) {}

    // rows[i] = position of projectile i's entity in the order the bodies were submitted (PhysicsDynamo::submit)
    void attach(const std::vector<uint32_t>& rows, const std::vector<ProjectileComponent>& components)
    {
        if (rows.size() != components.size()) throw std::runtime_error("ProjectileBehaviors::attach: size mismatch");
        check(icx_projectiles_upload(m_world, (uint32_t)rows.size(), rows.data(), components.data()), "icx_projectiles_upload");
        m_count = (uint32_t)rows.size();
    }

    // projectile_behaviors.h:17-54 for every attached projectile
    void on_fixed_update(double deltaTime) { check(icx_projectiles_fixed_update(m_world, deltaTime), "icx_projectiles_fixed_update"); }

    // projectile numbers (positions in attach()'s vectors, ascending) whose lifetime ran out in the last update
    std::vector<uint32_t> condemned()
    {
        std::vector<uint32_t> out(m_count ? m_count : 1);
        size_t n = 0;
        check(icx_projectiles_condemned(m_world, out.data(), out.size(), &n), "icx_projectiles_condemned");
        out.resize(n);
        return out;
    }

    // component state back into the caller's storage
    void fetch(std::vector<ProjectileComponent>& components)
    {
        components.resize(m_count);
        check(icx_projectiles_download(m_world, m_count, components.data()), "icx_projectiles_download");
    }

private:
    void check(int rc, const char* what)
    {
        if (rc != ICX_OK) throw std::runtime_error(std::string(what) + ": " + icx_status_string(rc) + " (" + icx_last_error(m_world) + ")");
    }
    IcxWorld* m_world;
    uint32_t m_count = 0;
};

} // namespace icx_host
