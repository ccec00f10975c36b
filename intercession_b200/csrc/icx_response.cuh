// Response procedures (source/physics/collision_procedures.cpp:636-1065) as device functions.
#pragma once
#include "icx_narrow.cuh"

namespace icx {

// everything a response needs about one side of a pair
struct Side {
    // entity
    V3 origin, vel, angvel, acc, angacc;
    float mass, colAngDrag;
    // collider
    M3 L; V3 base; M3 inv; V3 inertia;      // cached composite (linear part, offset), inverse, inertia diagonal
    float staticFriction, dynamicFriction, restitution, stiffness, damping, restLength;
    bool influenceOrientation;
};

ICX_DI M4 side_model(const Side& s) {
    M4 m; m.c[0] = s.L.c[0]; m.c[1] = s.L.c[1]; m.c[2] = s.L.c[2];
    m.t = s.base + (s.origin + v3(0.0f, 0.0f, 0.0f));   // ((P0*lx + P1*ly) + P2*lz) + P3, P3 = translate column (glm keeps +0)
    return m;
}
// inverse world moment, .cpp:739-754 / :943-958
ICX_DI M3 inverse_moment(const Side& s, float invMass) {
    if (invMass == 0) return diag3(v3(0.0f, 0.0f, 0.0f));
    M3 I = mul(diag3(s.inertia), s.mass);
    return inverse(mul(mul(transpose(s.inv), I), s.inv));
}
// friction direction, .cpp:771-782 / :975-985
ICX_DI V3 collision_tangent(V3 rel, V3 n) {
    V3 tangentCross = cross(rel, n);
    V3 t = v3(0.0f, 0.0f, 0.0f);
    if (length2(tangentCross) != 0.0f) {
        t = cross(n, tangentCross);
        if (length2(t) != 0.0f) t = normalize(t);
    }
    return t;
}
ICX_DI float impulse_denominator(float invMassA, float invMassB, const M3& imA, const M3& imB, V3 leverA, V3 leverB, V3 dir) {
    return invMassB + invMassA +
           dot(cross(mul(imB, cross(leverB, dir)), leverB) + cross(mul(imA, cross(leverA, dir)), leverA), dir);
}

// rigid_rigid_response, .cpp:636-874.  Mutates origin/vel/angvel of both sides and h.point.
ICX_DI void rigid_rigid_response_i(Side& A, Side& B, Hit& h) {
    if (A.mass == 0 && B.mass == 0) return;                                         // :649
    float invMassA = 0, invMassB = 0;
    if (A.mass != 0) invMassA = 1.0f / A.mass;
    if (B.mass != 0) invMassB = 1.0f / B.mass;
    const float restitution = A.restitution * B.restitution;                        // :673
    const float staticFriction = A.staticFriction * B.staticFriction;
    const float dynamicFriction = A.dynamicFriction * B.dynamicFriction;
    float massRatio = invMassA / (invMassA + invMassB);                             // :688
    V3 n = h.normal;
    A.origin = A.origin + n * h.depth * massRatio;                                  // :692
    B.origin = B.origin - n * h.depth * (1 - massRatio);                            // :693
    h.point = h.point - n * h.depth * (1 - massRatio);                              // :694
    const M4 modelA = side_model(A), modelB = side_model(B);
    const V3 comA = xform(modelA, 0.0f, 0.0f, 0.0f), comB = xform(modelB, 0.0f, 0.0f, 0.0f);
    const V3 leverA = h.point - comA, leverB = h.point - comB;
    const V3 rel = (A.vel + cross(A.angvel, leverA)) - (B.vel + cross(B.angvel, leverB));   // :718
    if (dot(rel, n) > 0) return;                                                    // :722
    const M3 imA = inverse_moment(A, invMassA), imB = inverse_moment(B, invMassB);
    const float contactImpulse = (-1.0f * (1 + restitution) * dot(rel, n)) /
                                 impulse_denominator(invMassA, invMassB, imA, imB, leverA, leverB, n);   // :757-765
    V3 tangent = collision_tangent(rel, n);
    const float tangentImpulse = -1.0f * dot(rel, tangent) /
                                 impulse_denominator(invMassA, invMassB, imA, imB, leverA, leverB, tangent);  // :787-793
    const float frictionCone = staticFriction * contactImpulse;
    const float frictionImpulse = fabsf(tangentImpulse) < fabsf(frictionCone) ? tangentImpulse : tangentImpulse * dynamicFriction;
    A.vel = A.vel + invMassA * (contactImpulse * n);  B.vel = B.vel - invMassB * (contactImpulse * n);        // :827-828
    A.vel = A.vel + invMassA * (frictionImpulse * tangent);  B.vel = B.vel - invMassB * (frictionImpulse * tangent);
    if (A.influenceOrientation) A.angvel = A.angvel + mul(imA, cross(leverA, (contactImpulse * n)));          // :835-846
    if (B.influenceOrientation) B.angvel = B.angvel - mul(imB, cross(leverB, (contactImpulse * n)));
    if (A.influenceOrientation) A.angvel = A.angvel + mul(imA, cross(leverA, (frictionImpulse * tangent)));   // :850-861
    if (B.influenceOrientation) B.angvel = B.angvel - mul(imB, cross(leverB, (frictionImpulse * tangent)));
    if (A.influenceOrientation) A.angvel = A.angvel * (1.0f - A.colAngDrag);                                  // :866-873
    if (B.influenceOrientation) B.angvel = B.angvel * (1.0f - B.colAngDrag);
}

// spring_rigid_response, .cpp:876-1053 (A is the spring).  Mutates acc/angacc/vel/angvel and h.point.
ICX_DI void spring_rigid_response_i(Side& A, Side& B, Hit& h) {
    if (A.mass == 0 && B.mass == 0) return;
    float invMassA = 0, invMassB = 0;
    if (A.mass != 0) invMassA = 1.0f / A.mass;
    if (B.mass != 0) invMassB = 1.0f / B.mass;
    const float staticFriction = A.staticFriction * B.staticFriction;
    const float dynamicFriction = A.dynamicFriction * B.dynamicFriction;
    const M4 modelA = side_model(A), modelB = side_model(B);
    const V3 comA = xform(modelA, 0.0f, 0.0f, 0.0f), comB = xform(modelB, 0.0f, 0.0f, 0.0f);
    V3 n = h.normal;
    h.point = h.point - (n * h.depth) / 2.0f;                                       // :926
    const V3 leverA = h.point - comA, leverB = h.point - comB;
    const V3 rel = (A.vel + cross(A.angvel, leverA)) - (B.vel + cross(B.angvel, leverB));   // :934
    const M3 imA = inverse_moment(A, invMassA), imB = inverse_moment(B, invMassB);
    const V3 deltaDepth = dot(rel, n) * n;                                          // :963
    V3 rest = mul(A.L, v3(0.0f, 0.0f, A.restLength));                               // :970-971
    const float scaledRest = length(rest);
    const float springMag = (h.depth - scaledRest) * A.stiffness;
    const V3 springForce = springMag * n;
    const V3 damped = springForce - A.damping * deltaDepth;                         // :975
    V3 tangent = collision_tangent(rel, n);
    const float tangentImpulse = -1.0f * dot(rel, tangent) /
                                 impulse_denominator(invMassA, invMassB, imA, imB, leverA, leverB, tangent);
    const float frictionCone = staticFriction * springMag;
    const float frictionImpulse = fabsf(tangentImpulse) < fabsf(frictionCone) ? tangentImpulse : tangentImpulse * dynamicFriction;
    A.acc = A.acc + invMassA * damped;                                              // :1010-1011
    B.acc = B.acc - invMassB * damped;
    A.vel = A.vel + invMassA * (frictionImpulse * tangent);  B.vel = B.vel - invMassB * (frictionImpulse * tangent);
    if (A.influenceOrientation) A.angacc = A.angacc + mul(imA, cross(leverA, damped));
    if (B.influenceOrientation) B.angacc = B.angacc - mul(imB, cross(leverB, damped));
    if (A.influenceOrientation) A.angvel = A.angvel + mul(imA, cross(leverA, frictionImpulse * tangent));
    if (B.influenceOrientation) B.angvel = B.angvel - mul(imB, cross(leverB, frictionImpulse * tangent));
    if (A.influenceOrientation) A.angvel = A.angvel * (1.0f - (A.colAngDrag * 2.0f));                         // :1044-1051
    if (B.influenceOrientation) B.angvel = B.angvel * (1.0f - (B.colAngDrag * 2.0f));
}

// out-of-line versions for the serial callers (speculative narrowphase, host checks)
ICX_DN void rigid_rigid_response(Side& A, Side& B, Hit& h) { rigid_rigid_response_i(A, B, h); }
ICX_DN void spring_rigid_response(Side& A, Side& B, Hit& h) { spring_rigid_response_i(A, B, h); }

} // namespace icx
