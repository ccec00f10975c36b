// TEST INFRASTRUCTURE: compiles the PRODUCT's device math headers (csrc/icx_narrow.cuh,
// icx_response.cuh) for the host so the `-m "not gpu"` suite can check them against the oracle
// without a GPU.  Not a fallback: nothing in the product loads this.
#include "../../intercession_b200/csrc/icx_response.cuh"
#include "../../include/icx.h"
using namespace icx;

static Shape make_shape(const IcxTransform& t, const IcxCollider& c, V3& base, V3& inertia)
{
    Q4 pq; pq.w = t.orientation[0]; pq.x = t.orientation[1]; pq.y = t.orientation[2]; pq.z = t.orientation[3];
    if (!c.inheritOrientation) { pq.w = 1; pq.x = pq.y = pq.z = 0; }
    const IcxTransform& l = c.localTransform;
    Q4 lq; lq.w = l.orientation[0]; lq.x = l.orientation[1]; lq.y = l.orientation[2]; lq.z = l.orientation[3];
    F4 P = model_transform(v3(t.origin[0], t.origin[1], t.origin[2]), pq, v3(t.scale[0], t.scale[1], t.scale[2]));
    F4 Lm = model_transform(v3(l.origin[0], l.origin[1], l.origin[2]), lq, v3(l.scale[0], l.scale[1], l.scale[2]));
    F4 f = mul(P, Lm);
    V4 b4 = P.c[0] * Lm.c[3].x + P.c[1] * Lm.c[3].y + P.c[2] * Lm.c[3].z;
    Shape s;
    s.M.c[0] = v3(f.c[0].x, f.c[0].y, f.c[0].z); s.M.c[1] = v3(f.c[1].x, f.c[1].y, f.c[1].z); s.M.c[2] = v3(f.c[2].x, f.c[2].y, f.c[2].z);
    base = v3(b4.x, b4.y, b4.z);
    s.M.t = base + (v3(t.origin[0], t.origin[1], t.origin[2]) + v3(0.0f, 0.0f, 0.0f));
    s.inv = inverse(mat3_of(s.M));
    s.type = c.colliderType;
    V3 sc = v3(t.scale[0], t.scale[1], t.scale[2]) * v3(l.scale[0], l.scale[1], l.scale[2]);
    if (s.type == ICX_COLLIDER_BOX) inertia = v3((sc.y * sc.y + sc.z * sc.z) / 4.0f, (sc.x * sc.x + sc.z * sc.z) / 4.0f, (sc.x * sc.x + sc.y * sc.y) / 4.0f);
    else if (s.type == ICX_COLLIDER_RAY) inertia = v3((sc.z * sc.z) / 3.0f, (sc.z * sc.z) / 3.0f, 0.0001f);
    else { float r = smin(smin(sc.x, sc.y), sc.z); float d = r * r * (2.0f / 5.0f); inertia = v3(d, d, d); }
    // the translation column must equal the full product's
    if (!(s.M.t.x == f.c[3].x && s.M.t.y == f.c[3].y && s.M.t.z == f.c[3].z)) s.type |= 0x100;
    return s;
}

extern "C" int icxh_intersect(const IcxTransform* tA, const IcxCollider* cA, const IcxTransform* tB, const IcxCollider* cB,
                              uint32_t flags, float* point, float* normal, float* depth, float* minParamA, float* minParamB)
{
    V3 bA, bB, iA, iB;
    Shape A = make_shape(*tA, *cA, bA, iA), B = make_shape(*tB, *cB, bB, iB);
    if ((A.type | B.type) & 0x100) return -1;
    V3 oA = v3(tA->origin[0], tA->origin[1], tA->origin[2]), oB = v3(tB->origin[0], tB->origin[1], tB->origin[2]);
    Hit h; h.point = v3(0, 0, 0); h.normal = v3(0, 0, 0); h.depth = 0;
    float mpA = cA->minParametricValue, mpB = cB->minParametricValue;
    bool hit = false;
    const int a = A.type, b = B.type;
    if (a == ICX_COLLIDER_BOX && b == ICX_COLLIDER_BOX) hit = box_box_intersect(A, B, h);
    else if (a == ICX_COLLIDER_SPHERE && b == ICX_COLLIDER_BOX) hit = sphere_box_intersect(A, B, h);
    else if (a == ICX_COLLIDER_BOX && b == ICX_COLLIDER_SPHERE) { hit = sphere_box_intersect(B, A, h); if (hit) { h.normal = h.normal * -1.0f; h.point = h.point + (h.normal * h.depth); } }
    else if (a == ICX_COLLIDER_RAY && b == ICX_COLLIDER_BOX) hit = ray_box_intersect(A, oA, B, oB, mpA, h);
    else if (a == ICX_COLLIDER_BOX && b == ICX_COLLIDER_RAY) { hit = ray_box_intersect(B, oB, A, oA, mpB, h); if (hit) { h.normal = h.normal * -1.0f; h.point = h.point + (h.normal * h.depth); } }
    else if ((flags & ICX_F_SPHERE_SPHERE) && a == ICX_COLLIDER_SPHERE && b == ICX_COLLIDER_SPHERE)
        hit = sphere_sphere_intersect(A, 0.5f * cA->localTransform.scale[0] * tA->scale[0], B, 0.5f * cB->localTransform.scale[0] * tB->scale[0], h);
    point[0] = h.point.x; point[1] = h.point.y; point[2] = h.point.z;
    normal[0] = h.normal.x; normal[1] = h.normal.y; normal[2] = h.normal.z; *depth = h.depth;
    if (minParamA) *minParamA = mpA;
    if (minParamB) *minParamB = mpB;
    return hit ? 1 : 0;
}

static Side make_side(const IcxTransform& t, const IcxPhysics& p, const IcxCollider& c)
{
    Side s; V3 base, inertia;
    Shape sh = make_shape(t, c, base, inertia);
    s.origin = v3(t.origin[0], t.origin[1], t.origin[2]);
    s.vel = v3(p.velocity[0], p.velocity[1], p.velocity[2]); s.angvel = v3(p.angularVelocity[0], p.angularVelocity[1], p.angularVelocity[2]);
    s.acc = v3(p.acceleration[0], p.acceleration[1], p.acceleration[2]); s.angacc = v3(p.angularAcceleration[0], p.angularAcceleration[1], p.angularAcceleration[2]);
    s.mass = p.mass; s.colAngDrag = p.collisionAngularDrag;
    s.L = mat3_of(sh.M); s.base = base; s.inv = sh.inv; s.inertia = inertia;
    s.staticFriction = c.staticFriction; s.dynamicFriction = c.dynamicFriction; s.restitution = c.restitution;
    s.stiffness = c.stiffness; s.damping = c.damping; s.restLength = c.restLength;
    s.influenceOrientation = c.influenceOrientation != 0;
    return s;
}
static void store_side(const Side& s, IcxTransform& t, IcxPhysics& p)
{
    t.origin[0] = s.origin.x; t.origin[1] = s.origin.y; t.origin[2] = s.origin.z;
    p.velocity[0] = s.vel.x; p.velocity[1] = s.vel.y; p.velocity[2] = s.vel.z;
    p.angularVelocity[0] = s.angvel.x; p.angularVelocity[1] = s.angvel.y; p.angularVelocity[2] = s.angvel.z;
    p.acceleration[0] = s.acc.x; p.acceleration[1] = s.acc.y; p.acceleration[2] = s.acc.z;
    p.angularAcceleration[0] = s.angacc.x; p.angularAcceleration[1] = s.angacc.y; p.angularAcceleration[2] = s.angacc.z;
}

extern "C" int icxh_response(IcxTransform* tA, IcxPhysics* pA, IcxCollider* cA, IcxTransform* tB, IcxPhysics* pB, IcxCollider* cB,
                             float* point, float* normal, float* depth)
{
    Side A = make_side(*tA, *pA, *cA), B = make_side(*tB, *pB, *cB);
    Hit h; h.point = v3(point[0], point[1], point[2]); h.normal = v3(normal[0], normal[1], normal[2]); h.depth = *depth;
    const int a = cA->collisionType, b = cB->collisionType;
    if (a == ICX_COLLISION_RIGID && b == ICX_COLLISION_RIGID) rigid_rigid_response(A, B, h);
    else if (a == ICX_COLLISION_SPRING && b == ICX_COLLISION_RIGID) spring_rigid_response(A, B, h);
    else if (a == ICX_COLLISION_RIGID && b == ICX_COLLISION_SPRING) {
        Hit inv; inv.normal = -h.normal; inv.point = h.point - (h.normal * h.depth); inv.depth = h.depth;
        spring_rigid_response(B, A, inv);
    }
    store_side(A, *tA, *pA); store_side(B, *tB, *pB);
    point[0] = h.point.x; point[1] = h.point.y; point[2] = h.point.z;
    normal[0] = h.normal.x; normal[1] = h.normal.y; normal[2] = h.normal.z; *depth = h.depth;
    return 0;
}
