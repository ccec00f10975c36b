// Component row <-> SoA helpers shared by the AoS boundary (icx_convert.cu) and the timeline replay
// (icx_timeline.cu): entity flag derivation and the per-packet constants of one ColliderComponent.
#pragma once
#include "icx_world.cuh"

namespace icx {

// immutable <=> nothing CollisionPhysicsRelay::engage does can change the entity:
// non-temporal (no interception), no ray collider (minParametricValue), and either no
// PhysicsComponent (responses are skipped, relay :84-106) or infinite mass with an angular
// velocity the response cannot touch.
__host__ __device__ __forceinline__ bool immutable_rule(uint32_t id, bool hasP, float mass, bool lockOrientation, bool has_ray, bool influence)
{
    if ((id & 0xFu) != ICX_NULL_CHAINLINK) return false;
    if (has_ray) return false;
    if (!hasP) return true;
    if (mass != 0.0f) return false;
    return lockOrientation || !influence;
}

// EF_HAS_RAY / EF_INFLUENCE of a collider component (the part of the entity flags that depends on C)
__host__ __device__ __forceinline__ uint32_t collider_flag_bits(const IcxColliderComponent* C)
{
    uint32_t f = 0;
    if (C) for (int s = 0; s < ICX_COLLIDERS_PER_ENTITY; s++) {
        const IcxCollider& c = C->colliders[s];
        if (!c.isActive) continue;
        if (c.colliderType == ICX_COLLIDER_RAY) f |= EF_HAS_RAY;
        if (c.influenceOrientation) f |= EF_INFLUENCE;
    }
    return f;
}

// entity flag word from the physics lock / sleep bits and the collider-derived bits
__host__ __device__ __forceinline__ uint32_t entity_flags(uint32_t id, bool hasP, bool lockOrigin, bool lockOrientation, bool asleep, float mass, uint32_t cbits)
{
    uint32_t ef = cbits & (EF_HAS_RAY | EF_INFLUENCE);
    if (hasP) {
        ef |= EF_HAS_PHYSICS;
        if (lockOrigin) ef |= EF_LOCK_ORIGIN;
        if (lockOrientation) ef |= EF_LOCK_ORIENTATION;
        if (asleep) ef |= EF_ASLEEP;
    }
    if (immutable_rule(id, hasP, mass, lockOrientation, (cbits & EF_HAS_RAY) != 0, (cbits & EF_INFLUENCE) != 0)) ef |= EF_IMMUTABLE;
    return ef;
}

// Per-packet constants of entity e's active colliders (submission order: slot 0 then 1) from its component row.
// scale = TransformComponent::scale (initial grid-cell estimate only); ef = the entity's flag word.
__device__ __forceinline__ void import_collider_row(const DeviceWorld& W, uint32_t e, const IcxColliderComponent& C, float sx0, float sy0, float sz0,
                                                    uint32_t ef, bool set_minparam)
{
    const uint32_t fp = W.first_packet[e];
    const uint32_t npk = fp >> 30, p0 = fp & 0x3FFFFFFFu;
    if (!npk) return;
    uint32_t k = 0;
    for (int s = 0; s < ICX_COLLIDERS_PER_ENTITY && k < npk; s++) {
        const IcxCollider& c = C.colliders[s];
        if (!c.isActive) continue;
        const uint32_t p = p0 + k; k++;
        uint32_t bits = ((uint32_t)c.colliderType & PF_TYPE_MASK) | (((uint32_t)c.collisionType & 3u) << PF_CTYPE_SHIFT);
        if (c.useBehaviorsResponse) bits |= PF_BEHAVIORS;
        if (c.inheritOrientation) bits |= PF_INHERIT_ORIENT;
        if (c.influenceOrientation) bits |= PF_INFLUENCE_ORIENT;
        if (s == 1) bits |= PF_SLOT;
        const IcxTransform& l = c.localTransform;
        W.pk_entity[p] = e;
        W.pk_lpos[p] = make_float4(l.origin[0], l.origin[1], l.origin[2], __uint_as_float(bits));
        W.pk_lrot[p] = make_float4(l.orientation[0], l.orientation[1], l.orientation[2], l.orientation[3]);
        W.pk_lscale[p] = make_float4(l.scale[0], l.scale[1], l.scale[2], c.restLength);
        W.pk_mat0[p] = make_float4(c.staticFriction, c.dynamicFriction, c.restitution, c.stiffness);
        W.pk_damping[p] = c.damping;
        if (set_minparam) W.pk_minparam[p] = c.minParametricValue;
        // rotation-invariant extent estimate of mutable colliders -> initial grid cell size
        if (!(ef & EF_IMMUTABLE) && c.colliderType != ICX_COLLIDER_NONE) {
            const float sx = fabsf(sx0 * l.scale[0]), sy = fabsf(sy0 * l.scale[1]), sz = fabsf(sz0 * l.scale[2]);
            float ext;
            if (c.colliderType == ICX_COLLIDER_SPHERE) ext = 1.7320509f * sx;
            else if (c.colliderType == ICX_COLLIDER_RAY) ext = sz * 1.000001f;
            else ext = sqrtf(sx * sx + sy * sy + sz * sz);
            if (ext == ext) atomicMax((int*)&W.counters[CTR_EXTENT], __float_as_int(ext));
        }
    }
}

} // namespace icx
