// TEST INFRASTRUCTURE -- not product code.  Only tests/, __graft_entry__.smoke()
// and bench.py's cpu_baseline / --impl reference leg may load the library built
// from this file (oracle/libicx_oracle.so).  Nothing under intercession_b200/
// includes, links or calls it.
//
// CPU restatement, in plain scalar C++ without glm, of the reference's physics
// tick and world-line record / replay / divergence path.  Every function cites
// the reference file:line it follows (paths relative to the reference root).
// Arithmetic order follows glm 0.9.9.9 as vendored by the reference
// (external/glm), so results are bit-identical to the reference compiled with
// the same flags (-O2 -ffp-contract=off, no FMA) -- tests/test_oracle_vs_ref.py
// pins that against oracle/_ref (the reference's own sources) and against the
// golden vectors in tests/golden/ generated from it.
//
// Multiplications by the literal 0/1 entries of glm's translate / scale / toMat4
// products are kept (not elided): they decide the SIGN of zeros in contact normals.
#include "../include/icx.h"

#include <chrono>
#include <cmath>
#include <cstring>
#include <list>
#include <map>
#include <vector>
#include <algorithm>
#include <thread>

namespace {

struct V3 { float x, y, z; };
struct Q4 { float w, x, y, z; };
struct M3 { V3 c[3]; };          // column-major, c[col]
struct M4 { V3 c[3]; V3 t; };    // affine: [c0 c1 c2 | t], bottom row (0,0,0,1)

inline V3 v3(float x, float y, float z) { V3 r = {x, y, z}; return r; }
inline V3 ld3(const float* p) { return v3(p[0], p[1], p[2]); }
inline void st3(float* p, V3 v) { p[0] = v.x; p[1] = v.y; p[2] = v.z; }
inline V3 operator+(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline V3 operator-(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline V3 operator-(V3 a) { return v3(-a.x, -a.y, -a.z); }
inline V3 operator*(V3 a, float s) { return v3(a.x * s, a.y * s, a.z * s); }
inline V3 operator*(float s, V3 a) { return v3(s * a.x, s * a.y, s * a.z); }
inline V3 operator*(V3 a, V3 b) { return v3(a.x * b.x, a.y * b.y, a.z * b.z); }
inline V3 operator/(V3 a, float s) { return v3(a.x / s, a.y / s, a.z / s); }
// external/glm/glm/detail/func_geometric.inl:48-55
inline float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
// func_geometric.inl:68-79
inline V3 cross(V3 x, V3 y) { return v3(x.y * y.z - y.y * x.z, x.z * y.x - y.z * x.x, x.x * y.y - y.x * x.y); }
inline float length2(V3 a) { return dot(a, a); }
inline float length(V3 a) { return std::sqrt(dot(a, a)); }
// func_geometric.inl:82-90 + func_exponential.inl:134-139
inline V3 normalize(V3 a) { return a * (1.0f / std::sqrt(dot(a, a))); }
// std::min / std::max argument conventions (NaN behaviour)
inline float smin(float a, float b) { return (b < a) ? b : a; }
inline float smax(float a, float b) { return (a < b) ? b : a; }

// glm mat3 * vec3, detail/type_mat3x3.inl:468-474
inline V3 mul(const M3& m, V3 v) {
    return v3(m.c[0].x * v.x + m.c[1].x * v.y + m.c[2].x * v.z,
              m.c[0].y * v.x + m.c[1].y * v.y + m.c[2].y * v.z,
              m.c[0].z * v.x + m.c[1].z * v.y + m.c[2].z * v.z);
}
// glm mat3 * mat3, detail/type_mat3x3.inl:486-519
inline M3 mul(const M3& a, const M3& b) {
    M3 r;
    for (int c = 0; c < 3; c++) {
        V3 bc = b.c[c];
        r.c[c] = v3(a.c[0].x * bc.x + a.c[1].x * bc.y + a.c[2].x * bc.z,
                    a.c[0].y * bc.x + a.c[1].y * bc.y + a.c[2].y * bc.z,
                    a.c[0].z * bc.x + a.c[1].z * bc.y + a.c[2].z * bc.z);
    }
    return r;
}
inline M3 mul(const M3& a, float s) { M3 r; for (int c = 0; c < 3; c++) r.c[c] = a.c[c] * s; return r; }
inline M3 transpose(const M3& m) {
    M3 r;
    r.c[0] = v3(m.c[0].x, m.c[1].x, m.c[2].x);
    r.c[1] = v3(m.c[0].y, m.c[1].y, m.c[2].y);
    r.c[2] = v3(m.c[0].z, m.c[1].z, m.c[2].z);
    return r;
}
// glm::inverse(mat3), detail/func_matrix.inl:322-344.  m[c][r] <-> m.c[c].{x,y,z}
inline M3 inverse(const M3& m) {
    const float m00 = m.c[0].x, m01 = m.c[0].y, m02 = m.c[0].z;
    const float m10 = m.c[1].x, m11 = m.c[1].y, m12 = m.c[1].z;
    const float m20 = m.c[2].x, m21 = m.c[2].y, m22 = m.c[2].z;
    float ood = 1.0f / (+m00 * (m11 * m22 - m21 * m12) - m10 * (m01 * m22 - m21 * m02) + m20 * (m01 * m12 - m11 * m02));
    M3 r;
    r.c[0].x = +(m11 * m22 - m21 * m12) * ood;
    r.c[1].x = -(m10 * m22 - m20 * m12) * ood;
    r.c[2].x = +(m10 * m21 - m20 * m11) * ood;
    r.c[0].y = -(m01 * m22 - m21 * m02) * ood;
    r.c[1].y = +(m00 * m22 - m20 * m02) * ood;
    r.c[2].y = -(m00 * m21 - m20 * m01) * ood;
    r.c[0].z = +(m01 * m12 - m11 * m02) * ood;
    r.c[1].z = -(m00 * m12 - m10 * m02) * ood;
    r.c[2].z = +(m00 * m11 - m10 * m01) * ood;
    return r;
}
inline M3 mat3_of(const M4& m) { M3 r; r.c[0] = m.c[0]; r.c[1] = m.c[1]; r.c[2] = m.c[2]; return r; }
// glm mat4 * vec4(x,y,z,1), detail/type_mat4x4.inl:536-582: (m0*x + m1*y) + (m2*z + m3*1)
inline V3 xform(const M4& m, float x, float y, float z) { return (m.c[0] * x + m.c[1] * y) + (m.c[2] * z + m.t); }

// glm::mat3_cast, external/glm/glm/gtc/quaternion.inl:47-72
inline M3 mat3_cast(Q4 q) {
    float qxx = q.x * q.x, qyy = q.y * q.y, qzz = q.z * q.z, qxz = q.x * q.z, qxy = q.x * q.y, qyz = q.y * q.z;
    float qwx = q.w * q.x, qwy = q.w * q.y, qwz = q.w * q.z;
    M3 r;
    r.c[0] = v3(1.0f - 2.0f * (qyy + qzz), 2.0f * (qxy + qwz), 2.0f * (qxz - qwy));
    r.c[1] = v3(2.0f * (qxy - qwz), 1.0f - 2.0f * (qxx + qzz), 2.0f * (qyz + qwx));
    r.c[2] = v3(2.0f * (qxz + qwy), 2.0f * (qyz - qwx), 1.0f - 2.0f * (qxx + qyy));
    return r;
}
// quaternion product p*q, external/glm/glm/detail/type_quat.inl:286-296
inline Q4 qmul(Q4 p, Q4 q) {
    Q4 r;
    r.w = p.w * q.w - p.x * q.x - p.y * q.y - p.z * q.z;
    r.x = p.w * q.x + p.x * q.w + p.y * q.z - p.z * q.y;
    r.y = p.w * q.y + p.y * q.w + p.z * q.x - p.x * q.z;
    r.z = p.w * q.z + p.z * q.w + p.x * q.y - p.y * q.x;
    return r;
}
// glm::angleAxis, external/glm/glm/ext/quaternion_trigonometric.inl:30-36
inline Q4 angle_axis(float angle, V3 v) {
    float s = std::sin(angle * 0.5f);
    Q4 r; r.w = std::cos(angle * 0.5f); r.x = v.x * s; r.y = v.y * s; r.z = v.z * s;
    return r;
}

// Full glm 4x4 arithmetic (column-major), used only to build composite transforms so that even
// the signs of zeros match the reference.
struct V4 { float x, y, z, w; };
struct F4 { V4 c[4]; };
inline V4 v4(float x, float y, float z, float w) { V4 r = {x, y, z, w}; return r; }
inline V4 operator*(V4 a, float s) { return v4(a.x * s, a.y * s, a.z * s, a.w * s); }
inline V4 operator+(V4 a, V4 b) { return v4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
inline F4 f4_identity() { F4 m; m.c[0] = v4(1, 0, 0, 0); m.c[1] = v4(0, 1, 0, 0); m.c[2] = v4(0, 0, 1, 0); m.c[3] = v4(0, 0, 0, 1); return m; }
// glm mat4*mat4, detail/type_mat4x4.inl:630-650
inline F4 mul(const F4& a, const F4& b) {
    F4 r;
    for (int j = 0; j < 4; j++) r.c[j] = a.c[0] * b.c[j].x + a.c[1] * b.c[j].y + a.c[2] * b.c[j].z + a.c[3] * b.c[j].w;
    return r;
}
// TransformComponent::get_model_transform, source/physics/transform_component.h:28-37:
// glm::translate(mat4(1), origin) * glm::toMat4(orientation) * glm::scale(mat4(1), scale)
inline F4 model_transform(const IcxTransform& t, bool zeroOrientation) {
    Q4 q = {t.orientation[0], t.orientation[1], t.orientation[2], t.orientation[3]};
    if (zeroOrientation) { q.w = 1.0f; q.x = q.y = q.z = 0.0f; }    // glm::quat(vec3(0))
    F4 I = f4_identity();
    F4 tr = I;      // external/glm/glm/ext/matrix_transform.inl: Result[3] = m[0]*v[0] + m[1]*v[1] + m[2]*v[2] + m[3]
    tr.c[3] = I.c[0] * t.origin[0] + I.c[1] * t.origin[1] + I.c[2] * t.origin[2] + I.c[3];
    M3 r3 = mat3_cast(q);
    F4 rot;         // mat4(mat3)
    rot.c[0] = v4(r3.c[0].x, r3.c[0].y, r3.c[0].z, 0); rot.c[1] = v4(r3.c[1].x, r3.c[1].y, r3.c[1].z, 0);
    rot.c[2] = v4(r3.c[2].x, r3.c[2].y, r3.c[2].z, 0); rot.c[3] = v4(0, 0, 0, 1);
    F4 sc;          // glm::scale: Result[i] = m[i] * v[i]
    sc.c[0] = I.c[0] * t.scale[0]; sc.c[1] = I.c[1] * t.scale[1]; sc.c[2] = I.c[2] * t.scale[2]; sc.c[3] = I.c[3];
    return mul(mul(tr, rot), sc);
}
// Collider::compose_transform, source/physics/collider.h:146-156
inline M4 compose_transform(const IcxCollider& c, const IcxTransform& parent) {
    F4 f = mul(model_transform(parent, !c.inheritOrientation), model_transform(c.localTransform, false));
    M4 r;
    for (int j = 0; j < 3; j++) r.c[j] = v3(f.c[j].x, f.c[j].y, f.c[j].z);
    r.t = v3(f.c[3].x, f.c[3].y, f.c[3].z);
    return r;
}
// Collider::get_inertia_tensor, source/physics/collider.h:97-142 (diagonal)
inline V3 inertia_diag(const IcxCollider& c, const float* parentScale) {
    V3 s = ld3(parentScale) * ld3(c.localTransform.scale);
    switch (c.colliderType) {
    case ICX_COLLIDER_BOX:
        return v3((s.y * s.y + s.z * s.z) / 4.0f, (s.x * s.x + s.z * s.z) / 4.0f, (s.x * s.x + s.y * s.y) / 4.0f);
    case ICX_COLLIDER_RAY:
        return v3((s.z * s.z) / 3.0f, (s.z * s.z) / 3.0f, 0.0001f);
    default: {
        float radius = smin(smin(s.x, s.y), s.z);      // glm::min(a,b) = (b<a)?b:a
        float d = radius * radius * (2.0f / 5.0f);
        return v3(d, d, d);
    }
    }
}

// EulerPhysicsRelay::engage body, source/physics/euler_physics_relay.h:25-95
inline void euler_one(IcxTransform& T, IcxPhysics& P, double deltaTime, const float* gravity) {
    if (P.isAsleep) return;
    V3 vel = ld3(P.velocity), acc = ld3(P.acceleration), av = ld3(P.angularVelocity), aa = ld3(P.angularAcceleration);
    V3 org = ld3(T.origin);
    Q4 q = {T.orientation[0], T.orientation[1], T.orientation[2], T.orientation[3]};
    acc = acc + ld3(gravity);                                                       // :32
    if (P.lockOrigin) { vel = v3(0, 0, 0); acc = v3(0, 0, 0); org = ld3(P.lockedOrigin); }      // :36-46
    if (P.lockOrientation) {                                                        // :47-57
        av = v3(0, 0, 0); aa = v3(0, 0, 0);
        q.w = P.lockedOrientation[0]; q.x = P.lockedOrientation[1]; q.y = P.lockedOrientation[2]; q.z = P.lockedOrientation[3];
    }
    const float halfDt = (float)(deltaTime / 2.0f);                                 // :62 (double divide, narrowed)
    const float dt = (float)deltaTime;
    for (int half = 0; half < 2; half++) {
        org = org + vel * halfDt;                                                   // :62 / :80
        float angularSpeed = length(av);
        if (angularSpeed != 0.0f) q = qmul(angle_axis(angularSpeed * halfDt, av / angularSpeed), q);   // :65-70
        if (half == 0) { vel = vel + acc * dt; av = av + aa * dt; }                 // :74-75
    }
    acc = v3(0, 0, 0); aa = v3(0, 0, 0);                                            // :90-91
    vel = vel * (float)(1.0f - P.linearDrag * deltaTime);                           // :94 (double, narrowed by glm *=)
    av = av * (float)(1.0f - P.angularDrag * deltaTime);                            // :95
    st3(T.origin, org);
    T.orientation[0] = q.w; T.orientation[1] = q.x; T.orientation[2] = q.y; T.orientation[3] = q.z;
    st3(P.velocity, vel); st3(P.acceleration, acc); st3(P.angularVelocity, av); st3(P.angularAcceleration, aa);
}


// ---------------------------------------------------------------------------
// Narrowphase helpers, source/physics/collision_procedures.h
// ---------------------------------------------------------------------------
const float UNIT_RADIUS = 0.5f;            // collision_procedures.h:14
const float MANIFOLD_DEPTH = 0.04f;        // :19
const float MANIFOLD_MIN_WEIGHT = 0.80f;   // :22
const int MAXV = 32;                       // fixed capacity standing in for std::vector

struct Interval { float x, y; };

// project_box, collision_procedures.h:27-68
inline Interval project_box(const M4& m, V3 axis) {
    float mn = INFINITY, mx = -INFINITY;
    for (int i = -1; i < 2; i += 2) for (int j = -1; j < 2; j += 2) for (int k = -1; k < 2; k += 2) {
        V3 vertex = xform(m, UNIT_RADIUS * i, UNIT_RADIUS * j, UNIT_RADIUS * k);
        float coeff = dot(vertex, axis);
        mn = smin(mn, coeff); mx = smax(mx, coeff);
    }
    Interval r = {mn, mx}; return r;
}
// project_ray, :73-86  -> [origin, end]
inline Interval project_ray(const M4& m, V3 axis) {
    V3 o = xform(m, 0, 0, 0), e = xform(m, 0, 0, 1);
    Interval r = {dot(o, axis), dot(e, axis)}; return r;
}
// project_sphere, :88-101 (radius from x-scale only)
inline Interval project_sphere(const M4& m, V3 axis) {
    V3 o = xform(m, 0, 0, 0), s = xform(m, UNIT_RADIUS, 0, 0);
    float op = dot(o, axis);
    float rp = dot(length(s - o) * axis, axis);
    Interval r = {op - rp, op + rp}; return r;
}
// build_contact_manifold, :106-160
inline int build_contact_manifold(const M4& m, V3 axis, V3* dest) {
    float coeffs[8]; V3 verts[8]; int nAll = 0;
    float maxCoeff = -INFINITY;
    float dims[3] = {-1, -1, -1};
    const int order[8] = {0, 1, 0, 2, 0, 1, 0, 2};
    for (int o = 0; o < 8; o++) {
        V3 vertex = xform(m, UNIT_RADIUS * dims[0], UNIT_RADIUS * dims[1], UNIT_RADIUS * dims[2]);
        float coeff = dot(vertex, axis);
        maxCoeff = coeff > maxCoeff ? coeff : maxCoeff;
        if (coeff >= maxCoeff - MANIFOLD_DEPTH) { coeffs[nAll] = coeff; verts[nAll] = vertex; nAll++; }
        int index = order[o];
        dims[index] = dims[index] < 0.0f ? 1.0f : -1.0f;
    }
    int n = 0;
    for (int i = nAll - 1; i >= 0; i--) if (coeffs[i] >= maxCoeff - MANIFOLD_DEPTH) dest[n++] = verts[i];
    return n;
}
// glm::intersectRayPlane, external/glm/glm/gtx/intersect.inl:6-26
inline bool intersect_ray_plane(V3 orig, V3 dir, V3 planeOrig, V3 planeNormal, float& dist) {
    float d = dot(dir, planeNormal);
    if (std::fabs(d) > 1.1920928955078125e-07f) {
        float t = dot(planeOrig - orig, planeNormal) / d;
        if (t > 0.0f) { dist = t; return true; }
    }
    return false;
}
// pseudo_clip_polyhedra, collision_procedures.h:163-313
inline void pseudo_clip_polyhedra(const V3* clipper, int nc, V3* clippee, int& ne, V3 axis) {
    V3 centre = v3(0, 0, 0);
    for (int i = 0; i < nc; i++) centre = centre + clipper[i];
    centre = centre / (float)nc;
    for (int i = 0; i < nc; i++) {
        V3 tangent = cross(axis, clipper[(i + 1) % nc] - clipper[i]);
        if (length2(tangent) == 0.0f) continue;
        V3 pn = normalize(tangent);
        float pw = -1.0f * (pn.x * clipper[i].x + pn.y * clipper[i].y + pn.z * clipper[i].z);
        float insideCoeff = 0.0f;
        if (nc > 2) insideCoeff = pn.x * centre.x + pn.y * centre.y + pn.z * centre.z + pw;
        if (insideCoeff == 0.0f) insideCoeff = INFINITY;
        V3 clipped[MAXV]; int nk = 0;
        float prevCoeff = -INFINITY;
        if (ne == 0) break;     // reference: modulo by zero (UB); never reached by a valid manifold
        for (int j = 0; j < ne + 1; j++) {
            if (ne == 2 && j >= 2) break;
            int J = j % ne;
            int J_1 = J == 0 ? ne - 1 : J - 1;
            float coeff = pn.x * clippee[J].x + pn.y * clippee[J].y + pn.z * clippee[J].z + pw;
            const float e = 0.00005f;
            const bool prevInside = (prevCoeff >= 0 - e && insideCoeff >= 0 - e) || (prevCoeff <= 0 + e && insideCoeff <= 0 + e);
            const bool currInside = (coeff >= 0 - e && insideCoeff >= 0 - e) || (coeff <= 0 + e && insideCoeff <= 0 + e);
            prevCoeff = coeff;
            if ((currInside ^ prevInside) && j != 0) {
                V3 insideVertex = clippee[J_1], outsideVertex = clippee[J];
                if (currInside) { insideVertex = clippee[J]; outsideVertex = clippee[J_1]; }
                V3 dir = normalize(insideVertex - outsideVertex);
                float dist;
                if (intersect_ray_plane(outsideVertex, dir, clipper[i], pn, dist)) {
                    if (nk < MAXV) clipped[nk++] = outsideVertex + dir * dist;
                }
            }
            if (currInside && j < ne) { if (nk < MAXV) clipped[nk++] = clippee[J]; }
        }
        for (int k = 0; k < nk; k++) clippee[k] = clipped[k];
        ne = nk;
    }
}

struct Hit { V3 point, normal; float depth; };

// The 15 SAT axes shared by box_box / sphere_box, collision_procedures.cpp:41-67
inline void sat_axes15(const M4& A, const M4& B, V3* axes) {
    M3 nA = transpose(inverse(mat3_of(A)));     // :41
    M3 nB = transpose(inverse(mat3_of(B)));     // :43
    // mat3 * vec3(1,0,0) etc. keep the zero products: they decide signs of zeros (:52-58)
    axes[0] = mul(nA, v3(1, 0, 0)); axes[1] = mul(nA, v3(0, 1, 0)); axes[2] = mul(nA, v3(0, 0, 1));
    axes[3] = mul(nB, v3(1, 0, 0)); axes[4] = mul(nB, v3(0, 1, 0)); axes[5] = mul(nB, v3(0, 0, 1));
    for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) axes[6 + 3 * i + j] = cross(axes[i], axes[3 + j]);
}

// box_box_intersect, collision_procedures.cpp:14-231
inline bool box_box_intersect(const IcxTransform& tA, const IcxCollider& cA, const IcxTransform& tB, const IcxCollider& cB, Hit& h) {
    M4 A = compose_transform(cA, tA), B = compose_transform(cB, tB);
    V3 axes[15]; sat_axes15(A, B, axes);
    V3 minN = v3(0, 0, 0); float minD = INFINITY;
    for (int a = 0; a < 15; a++) {
        axes[a] = normalize(axes[a]);
        if (std::isnan(axes[a].x)) continue;                                        // :82
        Interval iA = project_box(A, axes[a]), iB = project_box(B, axes[a]);
        float pen; bool flip = false;
        if ((iA.x + iA.y) / 2.0f > (iB.x + iB.y) / 2.0f) pen = iB.y - iA.x;         // :95-98
        else { pen = iA.y - iB.x; flip = true; }
        if (pen <= 0) return false;                                                 // :107
        if (pen < minD) { minD = pen; minN = flip ? -axes[a] : axes[a]; }           // :113-117
    }
    h.normal = minN; h.depth = minD;
    V3 mA[MAXV], mB[MAXV];
    int nA = build_contact_manifold(A, -h.normal, mA);                              // :128-132
    int nB = build_contact_manifold(B, h.normal, mB);
    if (nB == 0 || nA == 0) return false;                                           // :135-139
    if (nB == 1) { h.point = mB[0]; return true; }                                  // :140-146
    if (nA == 1) { h.point = mA[0] + (h.normal * h.depth); return true; }           // :147-153
    pseudo_clip_polyhedra(mA, nA, mB, nB, h.normal);                                // :159
    if (nB == 1) { h.point = mB[0]; return true; }                                  // :168-172
    float maxCoeff = -INFINITY; V3 origin = v3(0, 0, 0); float contributors = 0;    // :181-199
    for (int i = 0; i < nB; i++) {
        float vc = dot(mB[i], h.normal);
        if (vc > maxCoeff) { origin = mB[i]; contributors = 1; maxCoeff = vc; }
        else if (vc == maxCoeff) { origin = origin + mB[i]; contributors++; }
    }
    origin = origin / contributors;                                                 // :203
    const float manifoldRange = 1.0f - MANIFOLD_MIN_WEIGHT;
    V3 pt = v3(0, 0, 0);
    for (int i = 0; i < nB; i++) {                                                  // :210-221
        float vc = dot(mB[i], h.normal);
        V3 rel = mB[i] - origin;
        float w = (vc - maxCoeff + MANIFOLD_DEPTH) / MANIFOLD_DEPTH;
        pt = pt + rel * (w * manifoldRange + MANIFOLD_MIN_WEIGHT);
    }
    pt = pt / (float)nB;                                                            // :222
    h.point = pt + origin;
    return true;
}

// sphere_box_intersect, collision_procedures.cpp:266-378
inline bool sphere_box_intersect(const IcxTransform& tA, const IcxCollider& cA, const IcxTransform& tB, const IcxCollider& cB, Hit& h) {
    M4 A = compose_transform(cA, tA), B = compose_transform(cB, tB);
    V3 axes[15]; sat_axes15(A, B, axes);
    V3 minN = v3(0, 0, 0); float minD = INFINITY;
    for (int a = 0; a < 15; a++) {
        axes[a] = normalize(axes[a]);
        if (std::isnan(axes[a].x)) continue;
        Interval iA = project_sphere(A, axes[a]), iB = project_box(B, axes[a]);
        float pen; bool flip = false;
        if ((iA.x + iA.y) / 2.0f > (iB.x + iB.y) / 2.0f) pen = iB.y - iA.x;
        else { pen = iA.y - iB.x; flip = true; }
        if (pen <= 0) return false;
        if (pen < minD) { minD = pen; minN = flip ? -axes[a] : axes[a]; }
    }
    h.normal = minN; h.depth = minD;
    V3 surface = xform(A, UNIT_RADIUS, 0, 0);                                       // :367
    float radius = length(surface);                                                 // :368 (includes translation: reference quirk)
    h.point = xform(A, 0, 0, 0) - (h.normal * radius) + (h.normal * h.depth);       // :370-373
    return true;
}

// sphere_sphere_intersect, collision_procedures.cpp:380-423 (unwired in the dispatch table; opt-in)
inline bool sphere_sphere_intersect(const IcxTransform& tA, const IcxCollider& cA, const IcxTransform& tB, const IcxCollider& cB, Hit& h) {
    M4 A = compose_transform(cA, tA), B = compose_transform(cB, tB);
    V3 oA = xform(A, 0, 0, 0), oB = xform(B, 0, 0, 0);
    float rA = UNIT_RADIUS * cA.localTransform.scale[0] * tA.scale[0];
    float rB = UNIT_RADIUS * cB.localTransform.scale[0] * tB.scale[0];
    V3 dir = oA - oB;
    float dist = length(dir);
    if (dist == 0) return false;
    float depth = (rA + rB) - dist;
    if (depth <= 0) return false;
    h.depth = depth; h.normal = dir / dist; h.point = oB + (h.normal * rB);
    return true;
}

// ray_box_intersect, collision_procedures.cpp:442-590.  Mutates cA.minParametricValue.
inline bool ray_box_intersect(const IcxTransform& tA, IcxCollider& cA, const IcxTransform& tB, const IcxCollider& cB, Hit& h) {
    M4 A = compose_transform(cA, tA), B = compose_transform(cB, tB);
    M3 nB = transpose(inverse(mat3_of(B)));
    V3 axes[6];
    axes[0] = mul(nB, v3(1, 0, 0)); axes[1] = mul(nB, v3(0, 1, 0)); axes[2] = mul(nB, v3(0, 0, 1));
    V3 rayToBox = ld3(tB.origin) - ld3(tA.origin);                                  // :463
    V3 rayOrigin = xform(A, 0, 0, 0), rayEnd = xform(A, 0, 0, 1);
    axes[3] = rayEnd - rayOrigin;
    axes[4] = cross(axes[3], rayToBox);
    axes[5] = cross(axes[4], axes[3]);
    V3 minN = v3(0, 0, 0); float minD = INFINITY; float maxParam = 0.0f;
    for (int a = 0; a < 6; a++) {
        axes[a] = normalize(axes[a]);
        Interval ray = project_ray(A, axes[a]), box = project_box(B, axes[a]);
        float pen = 0, param = 0; bool flip = false;
        if (ray.x < (box.x + box.y) / 2.0f) flip = true;                            // :492
        if (ray.x >= box.x && ray.x <= box.y) {                                     // :508 origin inside
            param = 0;
            if (!flip) pen = box.y - smin(ray.x, ray.y);
            else pen = smax(ray.x, ray.y) - box.x;
            if (pen < 0) return false;
        } else {                                                                    // :529 origin outside
            if (!flip) {
                float range = ray.x - smax(box.y, ray.y);
                float delta = (ray.x - ray.y);
                if (delta != 0.0f) param = range / delta;
                pen = box.y - ray.y;
            } else {
                float range = smin(box.x, ray.y) - ray.x;
                float delta = (ray.y - ray.x);
                if (delta != 0.0f) param = range / delta;
                pen = ray.y - box.x;
            }
            if (pen < 0) return false;
        }
        if (param >= maxParam) { maxParam = param; minD = pen; minN = flip ? -axes[a] : axes[a]; }   // :553-558
    }
    h.normal = minN; h.depth = minD;
    if (maxParam < 0 || maxParam > 1) return false;                                 // :566
    if (maxParam >= cA.minParametricValue) return false;                            // :574-578
    cA.minParametricValue = maxParam;
    h.point = rayOrigin + maxParam * (rayEnd - rayOrigin);                          // :583
    return true;
}

// dispatch = intersectProcedures, collision_procedures.h:411-417 (+ wrappers .cpp:233-263)
inline bool intersect(const IcxTransform& tA, IcxCollider& cA, const IcxTransform& tB, IcxCollider& cB, uint32_t flags, Hit& h) {
    const int a = cA.colliderType, b = cB.colliderType;
    if (a == ICX_COLLIDER_BOX && b == ICX_COLLIDER_BOX) return box_box_intersect(tA, cA, tB, cB, h);
    if (a == ICX_COLLIDER_SPHERE && b == ICX_COLLIDER_BOX) return sphere_box_intersect(tA, cA, tB, cB, h);
    if (a == ICX_COLLIDER_RAY && b == ICX_COLLIDER_BOX) return ray_box_intersect(tA, cA, tB, cB, h);
    if (a == ICX_COLLIDER_BOX && (b == ICX_COLLIDER_SPHERE || b == ICX_COLLIDER_RAY)) {
        bool hit = (b == ICX_COLLIDER_SPHERE) ? sphere_box_intersect(tB, cB, tA, cA, h) : ray_box_intersect(tB, cB, tA, cA, h);
        if (!hit) return false;
        h.normal = h.normal * -1.0f;                                               // .cpp:241-243
        h.point = h.point + (h.normal * h.depth);
        return true;
    }
    if ((flags & ICX_F_SPHERE_SPHERE) && a == ICX_COLLIDER_SPHERE && b == ICX_COLLIDER_SPHERE)
        return sphere_sphere_intersect(tA, cA, tB, cB, h);
    return false;   // null_intersect and the ray/sphere stubs (.cpp:425-439, :592-622)
}


// ---------------------------------------------------------------------------
// Response procedures, source/physics/collision_procedures.cpp:636-1065
// ---------------------------------------------------------------------------
inline M3 diag3(V3 d) { M3 m; m.c[0] = v3(d.x, 0, 0); m.c[1] = v3(0, d.y, 0); m.c[2] = v3(0, 0, d.z); return m; }
inline M3 zero3() { return diag3(v3(0, 0, 0)); }

// inverse world moment, .cpp:739-754 / :943-958
inline M3 inverse_moment(const M4& model, const IcxCollider& c, const IcxTransform& t, float mass, float invMass) {
    if (invMass == 0) return zero3();
    M3 invModel = inverse(mat3_of(model));
    M3 I = mul(diag3(inertia_diag(c, t.scale)), mass);
    return inverse(mul(mul(transpose(invModel), I), invModel));
}

struct Tangent { V3 dir; };
// friction direction, .cpp:771-782 / :975-985
inline V3 collision_tangent(V3 relVelocity, V3 n) {
    V3 tangentCross = cross(relVelocity, n);
    V3 t = v3(0, 0, 0);
    if (length2(tangentCross) != 0.0f) {
        t = cross(n, tangentCross);
        if (length2(t) != 0.0f) t = normalize(t);
    }
    return t;
}
inline float impulse_denominator(float invMassA, float invMassB, const M3& imA, const M3& imB, V3 leverA, V3 leverB, V3 dir) {
    return invMassB + invMassA +
           dot(cross(mul(imB, cross(leverB, dir)), leverB) + cross(mul(imA, cross(leverA, dir)), leverA), dir);
}

// rigid_rigid_response, .cpp:636-874
inline void rigid_rigid_response(IcxTransform& tA, IcxCollider& cA, IcxPhysics& pA,
                                 IcxTransform& tB, IcxCollider& cB, IcxPhysics& pB, Hit& h) {
    if (pA.mass == 0 && pB.mass == 0) return;                                       // :649
    float invMassA = 0, invMassB = 0;
    if (pA.mass != 0) invMassA = 1.0f / pA.mass;
    if (pB.mass != 0) invMassB = 1.0f / pB.mass;
    const float restitution = cA.restitution * cB.restitution;                      // :673
    const float staticFriction = cA.staticFriction * cB.staticFriction;
    const float dynamicFriction = cA.dynamicFriction * cB.dynamicFriction;
    float massRatio = invMassA / (invMassA + invMassB);                             // :688
    V3 n = h.normal;
    st3(tA.origin, ld3(tA.origin) + n * h.depth * massRatio);                       // :692
    st3(tB.origin, ld3(tB.origin) - n * h.depth * (1 - massRatio));                 // :693
    h.point = h.point - n * h.depth * (1 - massRatio);                              // :694
    const M4 modelA = compose_transform(cA, tA), modelB = compose_transform(cB, tB);
    const V3 comA = xform(modelA, 0, 0, 0), comB = xform(modelB, 0, 0, 0);
    const V3 leverA = h.point - comA, leverB = h.point - comB;
    V3 vA = ld3(pA.velocity), vB = ld3(pB.velocity), wA = ld3(pA.angularVelocity), wB = ld3(pB.angularVelocity);
    const V3 rel = (vA + cross(wA, leverA)) - (vB + cross(wB, leverB));             // :718
    if (dot(rel, n) > 0) return;                                                    // :722
    const M3 imA = inverse_moment(modelA, cA, tA, pA.mass, invMassA);
    const M3 imB = inverse_moment(modelB, cB, tB, pB.mass, invMassB);
    const float normalImpulse = (-1.0f * (1 + restitution) * dot(rel, n)) /
                                impulse_denominator(invMassA, invMassB, imA, imB, leverA, leverB, n);   // :757-765
    const float contactImpulse = normalImpulse;
    V3 tangent = collision_tangent(rel, n);
    const float tangentImpulse = -1.0f * dot(rel, tangent) /
                                 impulse_denominator(invMassA, invMassB, imA, imB, leverA, leverB, tangent);  // :787-793
    const float frictionCone = staticFriction * contactImpulse;
    const float frictionImpulse = std::fabs(tangentImpulse) < std::fabs(frictionCone) ? tangentImpulse : tangentImpulse * dynamicFriction;
    vA = vA + invMassA * (contactImpulse * n);  vB = vB - invMassB * (contactImpulse * n);          // :827-828
    vA = vA + invMassA * (frictionImpulse * tangent);  vB = vB - invMassB * (frictionImpulse * tangent);
    if (cA.influenceOrientation) wA = wA + mul(imA, cross(leverA, (contactImpulse * n)));           // :835-846
    if (cB.influenceOrientation) wB = wB - mul(imB, cross(leverB, (contactImpulse * n)));
    if (cA.influenceOrientation) wA = wA + mul(imA, cross(leverA, (frictionImpulse * tangent)));    // :850-861
    if (cB.influenceOrientation) wB = wB - mul(imB, cross(leverB, (frictionImpulse * tangent)));
    if (cA.influenceOrientation) wA = wA * (1.0f - pA.collisionAngularDrag);                        // :866-873
    if (cB.influenceOrientation) wB = wB * (1.0f - pB.collisionAngularDrag);
    st3(pA.velocity, vA); st3(pB.velocity, vB); st3(pA.angularVelocity, wA); st3(pB.angularVelocity, wB);
}

// spring_rigid_response, .cpp:876-1053 (A is the spring)
inline void spring_rigid_response(IcxTransform& tA, IcxCollider& cA, IcxPhysics& pA,
                                  IcxTransform& tB, IcxCollider& cB, IcxPhysics& pB, Hit& h) {
    if (pA.mass == 0 && pB.mass == 0) return;
    float invMassA = 0, invMassB = 0;
    if (pA.mass != 0) invMassA = 1.0f / pA.mass;
    if (pB.mass != 0) invMassB = 1.0f / pB.mass;
    const float staticFriction = cA.staticFriction * cB.staticFriction;
    const float dynamicFriction = cA.dynamicFriction * cB.dynamicFriction;
    const M4 modelA = compose_transform(cA, tA), modelB = compose_transform(cB, tB);
    const V3 comA = xform(modelA, 0, 0, 0), comB = xform(modelB, 0, 0, 0);
    V3 n = h.normal;
    h.point = h.point - (n * h.depth) / 2.0f;                                       // :926
    const V3 leverA = h.point - comA, leverB = h.point - comB;
    V3 vA = ld3(pA.velocity), vB = ld3(pB.velocity), wA = ld3(pA.angularVelocity), wB = ld3(pB.angularVelocity);
    const V3 rel = (vA + cross(wA, leverA)) - (vB + cross(wB, leverB));             // :934
    const M3 imA = inverse_moment(modelA, cA, tA, pA.mass, invMassA);
    const M3 imB = inverse_moment(modelB, cB, tB, pB.mass, invMassB);
    const V3 deltaDepth = dot(rel, n) * n;                                          // :963
    V3 rest = mul(mat3_of(modelA), v3(0.0f, 0.0f, cA.restLength));                  // :970-971
    const float scaledRest = length(rest);
    const float springMag = (h.depth - scaledRest) * cA.stiffness;
    const V3 springForce = springMag * n;
    const V3 damped = springForce - cA.damping * deltaDepth;                        // :975
    V3 tangent = collision_tangent(rel, n);
    const float tangentImpulse = -1.0f * dot(rel, tangent) /
                                 impulse_denominator(invMassA, invMassB, imA, imB, leverA, leverB, tangent);
    const float frictionCone = staticFriction * springMag;
    const float frictionImpulse = std::fabs(tangentImpulse) < std::fabs(frictionCone) ? tangentImpulse : tangentImpulse * dynamicFriction;
    st3(pA.acceleration, ld3(pA.acceleration) + invMassA * damped);                 // :1010-1011
    st3(pB.acceleration, ld3(pB.acceleration) - invMassB * damped);
    vA = vA + invMassA * (frictionImpulse * tangent);  vB = vB - invMassB * (frictionImpulse * tangent);
    if (cA.influenceOrientation) st3(pA.angularAcceleration, ld3(pA.angularAcceleration) + mul(imA, cross(leverA, damped)));
    if (cB.influenceOrientation) st3(pB.angularAcceleration, ld3(pB.angularAcceleration) - mul(imB, cross(leverB, damped)));
    if (cA.influenceOrientation) wA = wA + mul(imA, cross(leverA, frictionImpulse * tangent));
    if (cB.influenceOrientation) wB = wB - mul(imB, cross(leverB, frictionImpulse * tangent));
    if (cA.influenceOrientation) wA = wA * (1.0f - (pA.collisionAngularDrag * 2.0f));               // :1044-1051
    if (cB.influenceOrientation) wB = wB * (1.0f - (pB.collisionAngularDrag * 2.0f));
    st3(pA.velocity, vA); st3(pB.velocity, vB); st3(pA.angularVelocity, wA); st3(pB.angularVelocity, wB);
}

// responseProcedures dispatch, collision_procedures.h:424-430; returns whether a non-null procedure exists
inline void respond(IcxTransform& tA, IcxCollider& cA, IcxPhysics& pA, IcxTransform& tB, IcxCollider& cB, IcxPhysics& pB, Hit& h) {
    const int a = cA.collisionType, b = cB.collisionType;
    if (a == ICX_COLLISION_RIGID && b == ICX_COLLISION_RIGID) rigid_rigid_response(tA, cA, pA, tB, cB, pB, h);
    else if (a == ICX_COLLISION_SPRING && b == ICX_COLLISION_RIGID) spring_rigid_response(tA, cA, pA, tB, cB, pB, h);
    else if (a == ICX_COLLISION_RIGID && b == ICX_COLLISION_SPRING) {               // rigid_spring_response, .cpp:1055-1065
        Hit inv; inv.normal = -h.normal; inv.point = h.point - (h.normal * h.depth); inv.depth = h.depth;
        spring_rigid_response(tB, cB, pB, tA, cA, pA, inv);
    }
}

inline bool divergent(uint8_t s) { return s == ICX_TS_FORKING || s == ICX_TS_FORKED; }   // timestream_state.h:41-44


// ---------------------------------------------------------------------------
// Broadphase bound (NEW WORK: the reference is all-pairs, collision_physics_relay.h:29-37).
// A candidate list is correct iff it is conservative: it contains every pair the
// narrowphase above can report as a hit.  Bounds, per collider packet:
//   box    : centre +- 0.5*(|c0|+|c1|+|c2|)            (15-axis SAT is exact for boxes)
//   sphere : centre +- sqrt(3)*r   (the reference's sphere test is a SAT over box-frame
//            axes, i.e. it tests the sphere's box-aligned bounding cube; r from x-scale)
//   ray    : midpoint +- |end-origin|/2 (ball around the segment; the ray test only
//            projects on the box normals and ray-derived axes)
// all inflated by `margin`, which also covers the origin shifts static resolution
// applies while the pass runs.  DESIGN.md derives these.
// ---------------------------------------------------------------------------
struct Aabb { float lo[3], hi[3]; };

inline Aabb packet_aabb(const IcxTransform& t, const IcxCollider& c, float margin) {
    M4 m = compose_transform(c, t);
    V3 ctr, h;
    if (c.colliderType == ICX_COLLIDER_SPHERE) {
        ctr = m.t;
        float r = length(xform(m, UNIT_RADIUS, 0, 0) - xform(m, 0, 0, 0)) * 1.7320509f;
        h = v3(r, r, r);
    } else if (c.colliderType == ICX_COLLIDER_RAY) {
        V3 o = xform(m, 0, 0, 0), e = xform(m, 0, 0, 1);
        ctr = (o + e) * 0.5f;
        float r = length(e - o) * 0.5f * 1.000001f;
        h = v3(r, r, r);
    } else {
        ctr = m.t;
        h = v3((std::fabs(m.c[0].x) + std::fabs(m.c[1].x) + std::fabs(m.c[2].x)) * 0.5f,
               (std::fabs(m.c[0].y) + std::fabs(m.c[1].y) + std::fabs(m.c[2].y)) * 0.5f,
               (std::fabs(m.c[0].z) + std::fabs(m.c[1].z) + std::fabs(m.c[2].z)) * 0.5f);
    }
    Aabb b;
    b.lo[0] = ctr.x - h.x - margin; b.lo[1] = ctr.y - h.y - margin; b.lo[2] = ctr.z - h.z - margin;
    b.hi[0] = ctr.x + h.x + margin; b.hi[1] = ctr.y + h.y + margin; b.hi[2] = ctr.z + h.z + margin;
    return b;
}
inline bool overlap(const Aabb& a, const Aabb& b) {
    return a.lo[0] <= b.hi[0] && b.lo[0] <= a.hi[0] && a.lo[1] <= b.hi[1] && b.lo[1] <= a.hi[1] &&
           a.lo[2] <= b.hi[2] && b.lo[2] <= a.hi[2];
}
// does intersectProcedures[a][b] have a non-null entry (collision_procedures.h:411-417)?
inline bool can_intersect(int a, int b, uint32_t flags) {
    if (a == ICX_COLLIDER_BOX) return b == ICX_COLLIDER_BOX || b == ICX_COLLIDER_SPHERE || b == ICX_COLLIDER_RAY;
    if (b == ICX_COLLIDER_BOX) return a == ICX_COLLIDER_SPHERE || a == ICX_COLLIDER_RAY;
    return (flags & ICX_F_SPHERE_SPHERE) && a == ICX_COLLIDER_SPHERE && b == ICX_COLLIDER_SPHERE;
}

struct Packet { uint32_t entityIndex; uint8_t slot; };
thread_local std::map<IcxEntity, IcxGround> g_grounds;      // per entity: what the last on_collision call of the last collision pass left

// ColliderSynchro::update, source/physics/collider_synchro.cpp:28-42
inline void build_packets(uint32_t n, IcxColliderComponent* C, std::vector<Packet>& packets, bool reset) {
    for (uint32_t i = 0; i < n; i++)
        for (int s = 0; s < ICX_COLLIDERS_PER_ENTITY; s++)
            if (C[i].colliders[s].isActive) {
                if (reset) C[i].colliders[s].minParametricValue = 1.0f;     // Collider::reset, collider.h:75-78
                Packet p = {i, (uint8_t)s}; packets.push_back(p);
            }
}

struct Collide {
    const IcxEntity* id; IcxTransform* T; IcxPhysics* P; const uint8_t* hasP; IcxColliderComponent* C;
    uint8_t* tstate; uint16_t* tcoh; uint16_t coherency; uint32_t host, flags, stride;
    std::vector<Packet> packets;
    IcxContact* hits; size_t hcap, nh; IcxInterception* ints; size_t icap, ni;

    // inner-loop body of CollisionPhysicsRelay::engage, collision_physics_relay.h:39-228
    void visit(uint32_t ia, uint32_t ib) {
        const Packet pa = packets[ia], pb = packets[ib];
        IcxCollider& cA = C[pa.entityIndex].colliders[pa.slot];
        IcxCollider& cB = C[pb.entityIndex].colliders[pb.slot];
        if (cB.colliderType == ICX_COLLIDER_NONE || !cB.isActive) return;           // :47
        if (pa.entityIndex == pb.entityIndex) return;                               // :51
        if (stride && pa.entityIndex / stride != pb.entityIndex / stride) return;   // world-line batch (extension)
        IcxTransform& tA = T[pa.entityIndex]; IcxTransform& tB = T[pb.entityIndex];
        Hit h; h.point = v3(0, 0, 0); h.normal = v3(0, 0, 0); h.depth = 0;
        if (!intersect(tA, cA, tB, cB, flags, h)) return;                           // :61-68
        uint16_t cflags = 0;
        if (cA.collisionType != ICX_COLLISION_NOOP && cB.collisionType != ICX_COLLISION_NOOP) {     // :82
            bool haveA = P && (!hasP || hasP[pa.entityIndex]), haveB = P && (!hasP || hasP[pb.entityIndex]);
            if (haveA && haveB) {                                                   // :84-106
                respond(tA, cA, P[pa.entityIndex], tB, cB, P[pb.entityIndex], h);
                cflags |= ICX_CONTACT_RESPONDED;
            }
        }
        if (cA.useBehaviorsResponse) cflags |= ICX_CONTACT_BEHAVIORS_A;             // :114
        if (cB.useBehaviorsResponse) cflags |= ICX_CONTACT_BEHAVIORS_B;             // :133
        if (nh < hcap) {
            IcxContact& c = hits[nh];
            c.entityA = id ? id[pa.entityIndex] : pa.entityIndex; c.entityB = id ? id[pb.entityIndex] : pb.entityIndex;
            c.slotA = pa.slot; c.slotB = pb.slot; c.flags = cflags;
            st3(c.point, h.point); st3(c.normal, h.normal); c.depth = h.depth; c.packetA = ia; c.packetB = ib;
        }
        nh++;
        // BipedBehaviors::on_collision, behaviors/biped_behaviors.h:336-366, for each side the relay calls (:114-155): the caller's
        // origin is the LIVE transform (after this pair's static resolution); B gets the inverted normal / point (:137-146)
        for (int side = 0; side < 2; side++) {
            if (!(cflags & (side ? ICX_CONTACT_BEHAVIORS_B : ICX_CONTACT_BEHAVIORS_A))) continue;
            const V3 n = side ? h.normal * -1.0f : h.normal;
            const V3 pt = side ? h.point - (h.normal * h.depth) : h.point;
            const IcxTransform& tc = side ? tB : tA;
            IcxGround g; g.entity = id ? id[side ? pb.entityIndex : pa.entityIndex] : (side ? pb.entityIndex : pa.entityIndex);
            st3(g.normal, n); g.dist = length(ld3(tc.origin) - pt); g.packetA = ia; g.packetB = ib; g.side = (uint32_t)side;
            g_grounds[g.entity] = g;                                                // calls arrive in pair order: the last one stays
        }
        // timestream response, :157-228
        if (!id) return;
        const IcxEntity eA = id[pa.entityIndex], eB = id[pb.entityIndex];
        const uint32_t thisLink = eA & 0xFu, otherLink = eB & 0xFu;                 // derive_causal_chain_link, ecs_types.h:92-95
        const uint8_t thisState = tstate ? tstate[pa.entityIndex] : 0, otherState = tstate ? tstate[pb.entityIndex] : 0;
        if (thisLink == ICX_NULL_CHAINLINK || otherLink == ICX_NULL_CHAINLINK) return;      // :185-188
        if (!(((thisLink == 0) ^ (otherLink == 0))
              || (thisLink != 0 && otherLink != 0 && (divergent(thisState) ^ divergent(otherState)))
              || (thisLink != otherLink && (!divergent(thisState) || !divergent(otherState)) && host == ICX_NULL_TIMESLICE)))
            return;                                                                 // :193-200
        IcxEntity agent;
        if (thisLink != otherLink) agent = thisLink < otherLink ? eA : eB;          // :207-210
        else if (divergent(thisState) ^ divergent(otherState)) agent = divergent(thisState) ? eA : eB;
        else agent = ICX_NULL_ENTITY;
        const IcxEntity recipient = agent == eA ? eB : eA;                          // :222
        const uint32_t recipientIndex = agent == eA ? pb.entityIndex : pa.entityIndex;
        if (ni < icap) { ints[ni].agent = agent; ints[ni].recipient = recipient; }
        ni++;
        // the synchronous handlers the event reaches: server_network_dynamo.cpp:813-838 (forking,
        // ignored for link-0 recipients) / parallel_network_dynamo.cpp:455-467 (forked);
        // Cosmos::set_timestream_state stamps the current coherency (cosmos.cpp:155-173)
        if (tstate) {
            if (host == ICX_NULL_TIMESLICE) { tstate[recipientIndex] = ICX_TS_FORKED; if (tcoh) tcoh[recipientIndex] = coherency; }
            else if ((recipient & 0xFu) > 0) { tstate[recipientIndex] = ICX_TS_FORKING; if (tcoh) tcoh[recipientIndex] = coherency; }
        }
    }
};

} // namespace

extern "C" {

int icxo_abi_version(void) { return ICX_ABI_VERSION; }

// EulerPhysicsRelay::engage, euler_physics_relay.h:21-97
int icxo_euler(uint32_t n, IcxTransform* T, IcxPhysics* P, const uint8_t* has_physics, double dt, double* seconds)
{
    if (!T || !P) return 1;
    const float g[3] = {0.0f, -10.0f, 0.0f};
    auto t0 = std::chrono::steady_clock::now();
    for (uint32_t i = 0; i < n; i++) if (!has_physics || has_physics[i]) euler_one(T[i], P[i], dt, g);
    auto t1 = std::chrono::steady_clock::now();
    if (seconds) *seconds = std::chrono::duration<double>(t1 - t0).count();
    return 0;
}

// CollisionPhysicsRelay::engage, collision_physics_relay.h:24-231.  cand == NULL: all pairs.
// what BipedBehaviors::on_collision left per entity during the last icxo_collide of this thread (ascending entity id)
int icxo_last_grounds(IcxGround* out, size_t cap, size_t* n)
{
    size_t k = 0;
    for (auto& kv : g_grounds) { if (out && k < cap) out[k] = kv.second; k++; }
    if (n) *n = k;
    return 0;
}

int icxo_collide(uint32_t n, const IcxEntity* id, IcxTransform* T, IcxPhysics* P, const uint8_t* has_physics,
                 IcxColliderComponent* C, uint8_t* tstate, uint16_t* tcoh, uint16_t coherency,
                 uint32_t host_timeslice, uint32_t flags, uint32_t worldline_stride,
                 const IcxPair* cand, size_t n_cand,
                 IcxContact* hits, size_t hit_cap, size_t* n_hits,
                 IcxInterception* ints, size_t int_cap, size_t* n_ints, double* seconds)
{
    if (!T || !C) return 1;
    Collide cx;
    cx.id = id; cx.T = T; cx.P = P; cx.hasP = has_physics; cx.C = C; cx.tstate = tstate; cx.tcoh = tcoh;
    cx.coherency = coherency; cx.host = host_timeslice; cx.flags = flags; cx.stride = worldline_stride;
    cx.hits = hits; cx.hcap = hit_cap; cx.nh = 0; cx.ints = ints; cx.icap = int_cap; cx.ni = 0;
    g_grounds.clear();
    build_packets(n, C, cx.packets, true);
    const uint32_t np = (uint32_t)cx.packets.size();
    auto t0 = std::chrono::steady_clock::now();
    if (!cand) {
        for (uint32_t a = 0; a < np; a++) {
            const IcxCollider& cA = C[cx.packets[a].entityIndex].colliders[cx.packets[a].slot];
            if (cA.colliderType == ICX_COLLIDER_NONE || !cA.isActive) continue;      // :32
            for (uint32_t b = a + 1; b < np; b++) cx.visit(a, b);
        }
    } else {
        for (size_t k = 0; k < n_cand; k++) {
            const uint32_t a = cand[k].a, b = cand[k].b;
            if (a >= b || b >= np) return 1;
            if (k && !(cand[k - 1].a < a || (cand[k - 1].a == a && cand[k - 1].b < b))) return 1;
            const IcxCollider& cA = C[cx.packets[a].entityIndex].colliders[cx.packets[a].slot];
            if (cA.colliderType == ICX_COLLIDER_NONE || !cA.isActive) continue;
            cx.visit(a, b);
        }
    }
    auto t1 = std::chrono::steady_clock::now();
    if (seconds) *seconds = std::chrono::duration<double>(t1 - t0).count();
    if (n_hits) *n_hits = cx.nh;
    if (n_ints) *n_ints = cx.ni;
    return 0;
}

// single-pair probes (per-function known-answer tests)
int icxo_intersect(const IcxTransform* tA, const IcxCollider* cA, const IcxTransform* tB, const IcxCollider* cB,
                   uint32_t flags, float* point, float* normal, float* depth, float* minParamA, float* minParamB)
{
    IcxCollider a = *cA, b = *cB;
    Hit h; h.point = v3(0, 0, 0); h.normal = v3(0, 0, 0); h.depth = 0;
    bool hit = intersect(*tA, a, *tB, b, flags, h);
    st3(point, h.point); st3(normal, h.normal); *depth = h.depth;
    if (minParamA) *minParamA = a.minParametricValue;
    if (minParamB) *minParamB = b.minParametricValue;
    return hit ? 1 : 0;
}

int icxo_response(IcxTransform* tA, IcxPhysics* pA, IcxCollider* cA, IcxTransform* tB, IcxPhysics* pB, IcxCollider* cB,
                  float* point, float* normal, float* depth)
{
    Hit h; h.point = ld3(point); h.normal = ld3(normal); h.depth = *depth;
    respond(*tA, *cA, *pA, *tB, *cB, *pB, h);
    st3(point, h.point); st3(normal, h.normal); *depth = h.depth;
    return 0;
}

// Per-packet broadphase bounds (6 floats: lo xyz, hi xyz) in packet order; returns packet count.
int icxo_aabbs(uint32_t n, const IcxTransform* T, IcxColliderComponent* C, float margin, float* out, size_t cap, size_t* n_packets)
{
    std::vector<Packet> packets; build_packets(n, C, packets, false);
    if (n_packets) *n_packets = packets.size();
    for (size_t k = 0; k < packets.size() && k < cap; k++) {
        Aabb b = packet_aabb(T[packets[k].entityIndex], C[packets[k].entityIndex].colliders[packets[k].slot], margin);
        std::memcpy(out + 6 * k, &b, sizeof(b));
    }
    return 0;
}

// Candidate list = all packet pairs a<b of different entities (same world-line), both of a type
// with a non-null intersect procedure, whose inflated bounds overlap; lexicographic order.
// method 0: brute force over all pairs (the definition).  method 1: sort-and-sweep on x
// (same set; used where N^2 is infeasible, validated against method 0 in tests).
int icxo_pairs(uint32_t n, const IcxTransform* T, IcxColliderComponent* C, float margin, uint32_t flags,
               uint32_t worldline_stride, int method, IcxPair* out, size_t cap, size_t* n_pairs)
{
    std::vector<Packet> packets; build_packets(n, C, packets, false);
    const uint32_t np = (uint32_t)packets.size();
    std::vector<Aabb> box(np); std::vector<int> type(np);
    for (uint32_t k = 0; k < np; k++) {
        const IcxCollider& c = C[packets[k].entityIndex].colliders[packets[k].slot];
        box[k] = packet_aabb(T[packets[k].entityIndex], c, margin);
        type[k] = c.colliderType;
    }
    std::vector<IcxPair> pairs;
    auto test = [&](uint32_t a, uint32_t b) {
        if (packets[a].entityIndex == packets[b].entityIndex) return;
        if (worldline_stride && packets[a].entityIndex / worldline_stride != packets[b].entityIndex / worldline_stride) return;
        if (!can_intersect(type[a], type[b], flags)) return;
        if (!overlap(box[a], box[b])) return;
        IcxPair p = {a, b}; pairs.push_back(p);
    };
    if (method == 0) {
        for (uint32_t a = 0; a < np; a++) for (uint32_t b = a + 1; b < np; b++) test(a, b);
    } else {
        std::vector<uint32_t> order(np);
        for (uint32_t k = 0; k < np; k++) order[k] = k;
        std::sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return box[a].lo[0] < box[b].lo[0]; });
        for (uint32_t i = 0; i < np; i++) {
            const uint32_t a = order[i];
            for (uint32_t j = i + 1; j < np && box[order[j]].lo[0] <= box[a].hi[0]; j++) {
                const uint32_t b = order[j];
                if (a < b) test(a, b); else test(b, a);
            }
        }
        std::sort(pairs.begin(), pairs.end(), [](const IcxPair& x, const IcxPair& y) { return x.a < y.a || (x.a == y.a && x.b < y.b); });
    }
    if (n_pairs) *n_pairs = pairs.size();
    for (size_t k = 0; k < pairs.size() && k < cap; k++) out[k] = pairs[k];
    return 0;
}

} // extern "C"

// ---------------------------------------------------------------------------
// World-line record / replay / divergence scan
// ---------------------------------------------------------------------------
namespace {

// cyclic coherency compare, source/events/message.h:179-183
inline bool coherency_ge(uint16_t a, uint16_t b) { return (a >= b && a - b <= 32768) || (a <= b && b - a > 32768); }

struct Record { uint16_t coherency; IcxEntity entity; IcxTransform T; IcxPhysics P; IcxColliderComponent C; bool hasC; };

// TsBreakpointQueue, source/networking/ts_breakpoint_queue.h:16-193: a FIFO whose pop_front stops at a
// movable breakpoint; the parallel context pops AT the breakpoint and inserts BEFORE it.
struct Stream {
    std::list<Record> list; std::list<Record>::iterator bp; bool active;
    Stream() : bp(list.end()), active(false) {}
    void push_back(const Record& r) { list.push_back(r); if (active && bp == list.end()) --bp; }       // :44-52
    bool pop_front(Record& d) {                                                                       // :53-64
        if (list.empty()) return false;
        if (active && bp == list.begin()) return false;
        d = list.front(); list.pop_front(); return true;
    }
    void set_breakpoint() { bp = list.begin(); active = true; }                                       // :107-112
    void remove_breakpoint() { bp = list.end(); active = false; }                                     // :113-118
    bool push_at_breakpoint(const Record& r) { if (!active) return false; list.insert(bp, r); return true; }   // :130-141
    bool pop_at_breakpoint(Record& d) {                                                               // :143-156
        if (list.empty() || !active || bp == list.end()) return false;
        d = *bp; bp = list.erase(bp); return true;
    }
    bool available() const { return !(list.empty() || bp == list.begin()); }                          // :158-164 (bp == end when inactive & non-empty)
    bool available_at_breakpoint() const { return bp != list.end(); }                                 // :165-171
};

// EntityTimestreamMap, source/networking/entity_timestream_map.h:43-230
struct Timeline {
    std::map<IcxEntity, Stream> streams; bool breakpoints = false;
    Stream& at(IcxEntity e) {
        auto it = streams.find(e);
        if (it == streams.end()) { it = streams.emplace(e, Stream()).first; it->second.bp = it->second.list.end(); if (breakpoints) it->second.set_breakpoint(); }   // :63-67
        return it->second;
    }
};

inline IcxEntity inc_link(IcxEntity e) { return (e & 0xFu) + 1u >= 15u ? e : e + 1u; }   // increment_causal_chain_link, ecs_types.h:122-137

} // namespace

extern "C" {

void* icxo_tl_create(void) { return new Timeline(); }
void icxo_tl_destroy(void* h) { delete static_cast<Timeline*>(h); }
void icxo_tl_breakpoint(void* h, int set)                                                             // entity_timestream_map.h:181-199
{
    Timeline* tl = static_cast<Timeline*>(h);
    for (auto& kv : tl->streams) { if (set) kv.second.set_breakpoint(); else kv.second.remove_breakpoint(); }
    tl->breakpoints = set != 0;
}

// server_network_dynamo.cpp:642-679 (link_inc = 1) / parallel_network_dynamo.cpp:369-385 (at_breakpoint = 1)
// ComponentCategory::all (:666-668): every component of the entity travels -- Transform, Physics and (C != NULL) ColliderComponent
int icxo_tl_record(void* h, uint32_t n, const IcxEntity* ids, const IcxTransform* T, const IcxPhysics* P, const IcxColliderComponent* C,
                   uint16_t coherency, int at_breakpoint, int link_inc)
{
    Timeline* tl = static_cast<Timeline*>(h);
    for (uint32_t i = 0; i < n; i++) {
        Record r; r.coherency = coherency; r.entity = link_inc ? inc_link(ids[i]) : ids[i]; r.T = T[i]; r.P = P[i];
        r.hasC = C != nullptr; if (C) r.C = C[i]; else std::memset(&r.C, 0, sizeof(r.C));
        Stream& s = tl->at(r.entity);
        if (at_breakpoint) s.push_at_breakpoint(r); else s.push_back(r);
    }
    return 0;
}

// server_network_dynamo.cpp:377-421 / parallel_network_dynamo.cpp:51-89
int icxo_tl_replay(void* h, uint32_t n, const IcxEntity* ids, IcxTransform* T, IcxPhysics* P, IcxColliderComponent* C, const uint8_t* tstate,
                   uint16_t now, int mode)
{
    Timeline* tl = static_cast<Timeline*>(h);
    int applied = 0;
    for (uint32_t i = 0; i < n; i++) {
        auto it = tl->streams.find(ids[i]);
        if (it == tl->streams.end()) continue;
        Stream& s = it->second;
        Record r;
        for (;;) {
            bool got;
            if (mode == ICX_REPLAY_PARALLEL) got = s.available_at_breakpoint() && coherency_ge(now, s.bp->coherency) && s.pop_at_breakpoint(r);   // :141-149, :214-221
            else got = s.available() && coherency_ge(now, s.list.front().coherency) && s.pop_front(r);                                              // :130-139, :205-212
            if (!got) break;
            if (mode == ICX_REPLAY_SERVER && r.coherency < now) continue;           // stale, :391
            if (mode == ICX_REPLAY_PARALLEL) continue;                              // upstream components only
            if (!divergent(tstate ? tstate[i] : 0)) { T[i] = r.T; P[i] = r.P; if (C && r.hasC) C[i] = r.C; applied++; }   // :408-412 all components
        }
    }
    return applied;
}

// TimejumpConditions::operator==, source/spacetime/timejump_conditions.h:21-25
int icxo_congruent(uint32_t n, const IcxTransform* live, const IcxTransform* rec, uint8_t* out)
{
    for (uint32_t i = 0; i < n; i++) out[i] = length2(ld3(live[i].origin) - ld3(rec[i].origin)) <= 2.0f ? 1 : 0;
    return 0;
}
int icxo_coherency_ge(uint16_t a, uint16_t b) { return coherency_ge(a, b) ? 1 : 0; }

// Per world-line reduction of the congruence predicate (the build's densification):
// rec holds n_rec transforms; body i is compared with rec[i % n_rec].  score = sum of
// min(length2, 2^20) in 48.16 fixed point, returned as a double.
int icxo_divergence_scan(uint32_t n, const IcxTransform* live, uint32_t n_rec, const IcxTransform* rec, uint32_t stride,
                         uint8_t* flags, uint32_t* incongruent, double* score)
{
    const uint32_t nw = stride ? (n + stride - 1) / stride : 1;
    std::vector<uint64_t> acc(nw, 0); std::vector<uint32_t> cnt(nw, 0);
    for (uint32_t i = 0; i < n; i++) {
        const float len2 = length2(ld3(live[i].origin) - ld3(rec[i % n_rec].origin));
        const uint32_t w = stride ? i / stride : 0;
        if (!(len2 <= 2.0f)) cnt[w]++;
        const float c = (len2 <= 1048576.0f) ? len2 : 1048576.0f;
        acc[w] += (uint64_t)(c * 65536.0f);
    }
    for (uint32_t w = 0; w < nw; w++) {
        if (flags) flags[w] = cnt[w] ? 1 : 0;
        if (incongruent) incongruent[w] = cnt[w];
        if (score) score[w] = (double)acc[w] / 65536.0;
    }
    return 0;
}

// Fork ageing, server_network_dynamo.cpp:13-15, :681-705
int icxo_fork_age(uint32_t n, uint8_t* tstate, uint16_t* tcoh, uint16_t coherency, uint32_t* n_promoted, int* resolution_needed)
{
    uint32_t promoted = 0; int need = 0;
    for (uint32_t i = 0; i < n; i++) {
        const uint8_t state = tstate[i]; const uint16_t since = tcoh[i];
        if (state == ICX_TS_FORKING && since + 13 <= coherency) { tstate[i] = ICX_TS_FORKED; tcoh[i] = coherency; promoted++; }
        if (state == ICX_TS_FORKED && since + 13 <= coherency) need = 1;
    }
    if (n_promoted) *n_promoted = promoted;
    if (resolution_needed) *resolution_needed = need;
    return 0;
}

// ProjectileBehaviors::on_fixed_update, behaviors/projectile_behaviors.h:17-54, n_updates fixed steps.
// T[i] is the transform of projectile i's entity.  A condemned entity is destroyed by the next
// Cosmos::update (cosmos.cpp:29-36) and receives no further updates.
int icxo_projectile_update(uint32_t n, double dt, uint32_t n_updates, IcxTransform* T, IcxProjectile* proj, uint8_t* condemned)
{
    for (uint32_t i = 0; i < n; i++) condemned[i] = 0;
    for (uint32_t u = 0; u < n_updates; u++)
        for (uint32_t i = 0; i < n; i++) {
            if (condemned[i]) continue;
            IcxProjectile& p = proj[i];
            p.currLifetime += dt;                                                      // :33
            const double s = std::sin(p.currLifetime * p.hz);                         // :35
            const double amp = (p.scaleMax - p.scaleMin) / 2.0;                        // :38
            const float sc = static_cast<float>(s * amp + (p.scaleMin + 1.0));         // :39
            T[i].scale[0] = sc; T[i].scale[1] = sc; T[i].scale[2] = sc;
            if (p.currLifetime > p.maxLifetime) condemned[i] = 1;                      // :41-44
        }
    return 0;
}

} // extern "C"

// ---- static triangle mesh pass: SELF-ORACLE, PARITY UNPINNED ------------------------------------------------------------
// The reference has no mesh collider (source/physics/collider.h:18-29 comments `mesh` out), so there is nothing
// to pin this against.  What follows restates the DEFINITION in include/icx.h ("Static triangle mesh (extension)") on the
// reference-shaped component structs with the oracle's own arithmetic, brute force over all triangles: a second, independent
// statement of the same rules that the device pass (csrc/icx_mesh.cuh, k_mesh_pass) must reproduce bit for bit.
namespace {
struct MeshTri { V3 a, b, c; };

V3 mesh_closest_point(V3 p, const MeshTri& t) {
    const V3 ab = t.b - t.a, ac = t.c - t.a, ap = p - t.a;
    const float d1 = dot(ab, ap), d2 = dot(ac, ap);
    if (d1 <= 0.0f && d2 <= 0.0f) return t.a;                                       // vertex region a
    const V3 bp = p - t.b;
    const float d3 = dot(ab, bp), d4 = dot(ac, bp);
    if (d3 >= 0.0f && d4 <= d3) return t.b;                                         // vertex region b
    const float vc = d1 * d4 - d3 * d2;
    if (vc <= 0.0f && d1 >= 0.0f && d3 <= 0.0f) return t.a + ab * (d1 / (d1 - d3)); // edge ab
    const V3 cp = p - t.c;
    const float d5 = dot(ab, cp), d6 = dot(ac, cp);
    if (d6 >= 0.0f && d5 <= d6) return t.c;                                         // vertex region c
    const float vb = d5 * d2 - d1 * d6;
    if (vb <= 0.0f && d2 >= 0.0f && d6 <= 0.0f) return t.a + ac * (d2 / (d2 - d6)); // edge ac
    const float va = d3 * d6 - d5 * d4;
    if (va <= 0.0f && (d4 - d3) >= 0.0f && (d5 - d6) >= 0.0f) return t.b + (t.c - t.b) * ((d4 - d3) / ((d4 - d3) + (d5 - d6)));   // edge bc
    const float denom = 1.0f / (va + vb + vc);                                      // face
    return (t.a + ab * (vb * denom)) + ac * (vc * denom);
}

bool mesh_sphere_hit(const M4& m, const MeshTri& t, Hit& h) {
    const V3 centre = xform(m, 0, 0, 0);
    const float radius = length(xform(m, UNIT_RADIUS, 0, 0) - centre);              // project_sphere's radius
    const V3 q = mesh_closest_point(centre, t);
    const V3 d = centre - q;
    const float dist2 = dot(d, d);
    if (!(dist2 < radius * radius)) return false;
    const float dist = std::sqrt(dist2);
    V3 n;
    if (dist > 0.0f) n = d * (1.0f / dist);
    else { n = normalize(cross(t.b - t.a, t.c - t.a)); if (std::isnan(n.x)) return false; }
    h.normal = n; h.depth = radius - dist; h.point = q;
    return true;
}

bool mesh_box_hit(const M4& m, const MeshTri& t, Hit& h) {
    const V3 centre = m.t;
    const V3 edge[3] = {t.b - t.a, t.c - t.b, t.a - t.c};
    V3 axes[13];
    axes[0] = cross(edge[0], t.c - t.a);
    for (int i = 0; i < 3; i++) axes[1 + i] = m.c[i];
    for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) axes[4 + 3 * i + j] = cross(m.c[i], edge[j]);
    float best = INFINITY; V3 bestN = v3(0, 0, 0);
    for (int k = 0; k < 13; k++) {
        V3 L = axes[k];
        const float len2 = dot(L, L);
        if (!(len2 > 1.0e-12f)) continue;
        L = L * (1.0f / std::sqrt(len2));
        const float rb = UNIT_RADIUS * ((std::fabs(dot(m.c[0], L)) + std::fabs(dot(m.c[1], L))) + std::fabs(dot(m.c[2], L)));
        const float cb = dot(centre, L);
        const float pa = dot(t.a, L), pb = dot(t.b, L), pc = dot(t.c, L);
        const float tmin = smin(smin(pa, pb), pc), tmax = smax(smax(pa, pb), pc);
        float pen; bool flip = false;
        if (cb > (tmin + tmax) * 0.5f) pen = tmax - (cb - rb);
        else { pen = (cb + rb) - tmin; flip = true; }
        if (pen <= 0.0f) return false;
        if (pen < best) { best = pen; bestN = flip ? -L : L; }
    }
    if (best == INFINITY) return false;
    h.normal = bestN; h.depth = best; h.point = mesh_closest_point(centre, t);
    return true;
}
// ray (origin m * 0, end m * (0, 0, 1)) against one triangle, Moeller-Trumbore; only in front of what the ray has met so far
bool mesh_ray_hit(const M4& m, const MeshTri& t, float minParam, Hit& h, float& param) {
    const V3 o = xform(m, 0, 0, 0), end = xform(m, 0, 0, 1);
    const V3 d = end - o;
    const V3 e1 = t.b - t.a, e2 = t.c - t.a;
    const V3 pv = cross(d, e2);
    const float det = dot(e1, pv);
    if (!(std::fabs(det) > 1.0e-12f)) return false;
    const float inv = 1.0f / det;
    const V3 tv = o - t.a;
    const float u = dot(tv, pv) * inv;
    if (!(u >= 0.0f && u <= 1.0f)) return false;
    const V3 qv = cross(tv, e1);
    const float v = dot(d, qv) * inv;
    if (!(v >= 0.0f && u + v <= 1.0f)) return false;
    const float s = dot(e2, qv) * inv;
    if (!(s >= 0.0f && s <= 1.0f)) return false;
    if (s >= minParam) return false;
    V3 n = normalize(cross(e1, e2));
    if (dot(n, d) > 0.0f) n = -n;
    param = s;
    h.point = o + s * d; h.normal = n; h.depth = dot(n, h.point - end);
    return true;
}
} // namespace

extern "C" int icxo_mesh_pass(uint32_t n, const IcxEntity* id, IcxTransform* T, IcxPhysics* P, const uint8_t* has_physics, IcxColliderComponent* C,
                              const float* vertices, uint32_t n_vertices, const uint32_t* triangles, uint32_t n_triangles, const IcxMeshMaterial* material,
                              IcxMeshContact* out, size_t cap, size_t* n_out)
{
    if (!T || !P || !C || !vertices || !triangles) return 1;
    std::vector<MeshTri> tris(n_triangles);
    for (uint32_t t = 0; t < n_triangles; t++) {
        for (int k = 0; k < 3; k++) if (triangles[3 * t + k] >= n_vertices) return 1;
        tris[t].a = ld3(vertices + 3 * (size_t)triangles[3 * t]); tris[t].b = ld3(vertices + 3 * (size_t)triangles[3 * t + 1]); tris[t].c = ld3(vertices + 3 * (size_t)triangles[3 * t + 2]);
    }
    // the mesh as a body: immovable, identity transform, its material
    IcxTransform tM; std::memset(&tM, 0, sizeof(tM)); tM.orientation[0] = 1.0f; tM.scale[0] = tM.scale[1] = tM.scale[2] = 1.0f;
    IcxPhysics pM; std::memset(&pM, 0, sizeof(pM)); pM.lockedOrientation[0] = 1.0f;
    IcxCollider cM; std::memset(&cM, 0, sizeof(cM));
    cM.colliderType = ICX_COLLIDER_BOX; cM.collisionType = ICX_COLLISION_RIGID; cM.isActive = 1; cM.localTransform = tM; cM.inheritOrientation = 1;
    cM.staticFriction = material ? material->staticFriction : 0.45f; cM.dynamicFriction = material ? material->dynamicFriction : 0.45f;
    cM.restitution = material ? material->restitution : 0.70f;
    size_t k = 0;
    for (uint32_t e = 0; e < n; e++) {
        if (has_physics && !has_physics[e]) continue;
        if (P[e].mass == 0.0f) continue;                                            // infinite mass: not visited
        for (uint32_t s = 0; s < ICX_COLLIDERS_PER_ENTITY; s++) {
            IcxCollider& c = C[e].colliders[s];
            if (!c.isActive) continue;
            const bool isRay = c.colliderType == ICX_COLLIDER_RAY;          // a ray meets the mesh as a spring (the biped's legs), boxes and spheres as rigid bodies
            if (isRay ? c.collisionType != ICX_COLLISION_SPRING
                      : ((c.colliderType != ICX_COLLIDER_BOX && c.colliderType != ICX_COLLIDER_SPHERE) || c.collisionType != ICX_COLLISION_RIGID)) continue;
            const M4 m0 = compose_transform(c, T[e]);
            V3 half = v3(0, 0, 0);
            if (isRay) {}
            else if (c.colliderType == ICX_COLLIDER_SPHERE) { const float r = length(xform(m0, UNIT_RADIUS, 0, 0) - xform(m0, 0, 0, 0)); half = v3(r, r, r); }
            else half = v3(((std::fabs(m0.c[0].x) + std::fabs(m0.c[1].x)) + std::fabs(m0.c[2].x)) * UNIT_RADIUS,
                           ((std::fabs(m0.c[0].y) + std::fabs(m0.c[1].y)) + std::fabs(m0.c[2].y)) * UNIT_RADIUS,
                           ((std::fabs(m0.c[0].z) + std::fabs(m0.c[1].z)) + std::fabs(m0.c[2].z)) * UNIT_RADIUS);
            V3 lo = m0.t - half, hi = m0.t + half;                                  // bounds when the collider's turn starts
            if (isRay) {
                const V3 o = xform(m0, 0, 0, 0), en = xform(m0, 0, 0, 1);
                lo = v3(smin(o.x, en.x), smin(o.y, en.y), smin(o.z, en.z)); hi = v3(smax(o.x, en.x), smax(o.y, en.y), smax(o.z, en.z));
            }
            if (std::isnan(lo.x) || std::isnan(hi.x)) continue;
            for (uint32_t t = 0; t < n_triangles; t++) {                            // ascending triangle index
                const MeshTri& tr = tris[t];
                const V3 tlo = v3(smin(smin(tr.a.x, tr.b.x), tr.c.x), smin(smin(tr.a.y, tr.b.y), tr.c.y), smin(smin(tr.a.z, tr.b.z), tr.c.z));
                const V3 thi = v3(smax(smax(tr.a.x, tr.b.x), tr.c.x), smax(smax(tr.a.y, tr.b.y), tr.c.y), smax(smax(tr.a.z, tr.b.z), tr.c.z));
                if (!(tlo.x <= hi.x && lo.x <= thi.x && tlo.y <= hi.y && lo.y <= thi.y && tlo.z <= hi.z && lo.z <= thi.z)) continue;
                const M4 m = compose_transform(c, T[e]);                            // where the previous triangle left the body
                Hit h; h.point = v3(0, 0, 0); h.normal = v3(0, 0, 0); h.depth = 0;
                float param = 0;
                const bool hit = isRay ? mesh_ray_hit(m, tr, c.minParametricValue, h, param)
                                       : (c.colliderType == ICX_COLLIDER_SPHERE ? mesh_sphere_hit(m, tr, h) : mesh_box_hit(m, tr, h));
                if (!hit) continue;
                IcxTransform tB = tM; IcxPhysics pB = pM; IcxCollider cB = cM;
                if (isRay) { c.minParametricValue = param; spring_rigid_response(T[e], c, P[e], tB, cB, pB, h); }
                else rigid_rigid_response(T[e], c, P[e], tB, cB, pB, h);
                if (out && k < cap) {
                    IcxMeshContact& mc = out[k];
                    mc.entity = id ? id[e] : e; mc.row = e; mc.slot = s; mc.triangle = t;
                    st3(mc.point, h.point); st3(mc.normal, h.normal); mc.depth = h.depth;
                }
                k++;
            }
        }
    }
    if (n_out) *n_out = k;
    return 0;
}
