// TEST INFRASTRUCTURE -- not product code (see ref_harness.cpp for who may load this library).
//
// The reference's own ProjectileBehaviors::on_fixed_update
// (/root/reference/source/behaviors/projectile_behaviors.h:17-54, compiled unmodified from the
// header where it lies) driven through a real Cosmos.  The reference's uint16 ids cap a cosmos at
// 255 entities per timeslice, so the rows are pushed through in chunks of 200; every chunk is a
// fresh Cosmos.  "Condemned" = the entity is gone after the next Cosmos::update()
// (cosmos.cpp:29-36 deletes what condemn_entity queued).
#include "../../include/icx.h"

#include "core/cosmos.h"
#include "events/event_broker.h"
#include "physics/transform_component.h"
#include "behaviors/behaviors_component.h"
#include "behaviors/projectile_component.h"
#include "behaviors/projectile_behaviors.h"

using namespace pleep;

static_assert(sizeof(ProjectileComponent) == sizeof(IcxProjectile), "Projectile layout");
static_assert(offsetof(ProjectileComponent, scaleMin) == offsetof(IcxProjectile, scaleMin), "Projectile offsets");

extern "C" {

// n projectiles; T[i] is the transform of projectile i's entity (scale is written), proj[i] in-out,
// condemned[i] = 1 when the update condemned the entity.  Returns 0, 1 on a registration failure.
int icxref_projectile_update(uint32_t n, double dt, uint32_t n_updates, IcxTransform* T, IcxProjectile* proj, uint8_t* condemned)
{
    const uint32_t CHUNK = 200;
    ProjectileBehaviors drivetrain;
    for (uint32_t base = 0; base < n; base += CHUNK) {
        const uint32_t m = std::min(CHUNK, n - base);
        auto broker = std::make_shared<EventBroker>();
        auto cosmos = std::make_shared<Cosmos>(broker, (TimesliceId)0);
        cosmos->register_component<TransformComponent>();
        cosmos->register_component<ProjectileComponent>();
        std::vector<Entity> ids(m);
        for (uint32_t k = 0; k < m; k++) {
            const Entity e = (Entity)((1u << 12) | ((k + 1) << 4));     // timeslice 1 | genesis k + 1 | chain-link 0: hosted elsewhere, so
                                                                    // Cosmos::destroy_entity (cosmos.h:348-372) keeps no host count for it
            ids[k] = e;
            if (!cosmos->register_entity(e)) return 1;
            cosmos->add_component(e, *reinterpret_cast<TransformComponent*>(&T[base + k]));
            cosmos->add_component(e, *reinterpret_cast<ProjectileComponent*>(&proj[base + k]));
            condemned[base + k] = 0;
        }
        BehaviorsComponent bc;
        for (uint32_t u = 0; u < n_updates; u++) {
            for (uint32_t k = 0; k < m; k++) {
                if (condemned[base + k]) continue;                   // destroyed entities get no further updates
                drivetrain.on_fixed_update(dt, bc, ids[k], cosmos, broker);
                T[base + k] = *reinterpret_cast<IcxTransform*>(&cosmos->get_component<TransformComponent>(ids[k]));
                proj[base + k] = *reinterpret_cast<IcxProjectile*>(&cosmos->get_component<ProjectileComponent>(ids[k]));
            }
            cosmos->update();
            for (uint32_t k = 0; k < m; k++) if (!cosmos->entity_exists(ids[k])) condemned[base + k] = 1;
        }
    }
    return 0;
}

} // extern "C"
