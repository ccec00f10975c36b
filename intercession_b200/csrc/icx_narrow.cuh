// Narrowphase: the reference's intersect procedures (source/physics/collision_procedures.{h,cpp})
// as device functions over cached composite transforms.  Fixed-capacity local arrays replace
// the reference's std::vector manifolds; iteration order is preserved.
#pragma once
#include "icx_math.cuh"

namespace icx {

#ifndef ICX_UNIT_RADIUS
#define ICX_UNIT_RADIUS 0.5f            // collision_procedures.h:14
#endif
#define ICX_MANIFOLD_DEPTH 0.04f        // :19
#define ICX_MANIFOLD_MIN_WEIGHT 0.80f   // :22
#define ICX_MAXV 24                     // clipped-manifold capacity (8 in, +1 per clip edge)

// A collider as the narrowphase sees it: composite transform (Collider::compose_transform,
// collider.h:146-156), its linear inverse, and the collider type.
struct Shape {
    M4 M;
    M3 inv;         // inverse(mat3(M))
    int type;
};

struct Hit { V3 point, normal; float depth; };
struct Interval { float x, y; };

// project_box, collision_procedures.h:27-68, over precomputed corners (same i,j,k order)
ICX_DI void box_corners(const M4& m, V3* c) {
    int n = 0;
#pragma unroll
    for (int i = -1; i < 2; i += 2)
#pragma unroll
        for (int j = -1; j < 2; j += 2)
#pragma unroll
            for (int k = -1; k < 2; k += 2)
                c[n++] = xform(m, ICX_UNIT_RADIUS * i, ICX_UNIT_RADIUS * j, ICX_UNIT_RADIUS * k);
}
ICX_DI Interval project_corners(const V3* c, V3 axis) {
    float mn = INFINITY, mx = -INFINITY;
#pragma unroll
    for (int n = 0; n < 8; n++) {
        float coeff = dot(c[n], axis);
        mn = smin(mn, coeff); mx = smax(mx, coeff);
    }
    Interval r; r.x = mn; r.y = mx; return r;
}
// project_ray, :73-86
ICX_DI Interval project_ray(const M4& m, V3 axis) {
    V3 o = xform(m, 0.0f, 0.0f, 0.0f), e = xform(m, 0.0f, 0.0f, 1.0f);
    Interval r; r.x = dot(o, axis); r.y = dot(e, axis); return r;
}
// project_sphere, :88-101
ICX_DI Interval project_sphere(V3 o, float radius, V3 axis) {
    float op = dot(o, axis);
    float rp = dot(radius * axis, axis);
    Interval r; r.x = op - rp; r.y = op + rp; return r;
}

// build_contact_manifold, :106-160.  The reference pushes a vertex when it is within
// MANIFOLD_DEPTH of the RUNNING maximum, then re-filters against the final maximum walking the
// kept vertices backwards; `kept` is that first filter as a bitmask (no dynamically indexed
// temporaries).
ICX_DI int build_contact_manifold(const M4& m, V3 axis, V3* dest) {
    float coeffs[8]; V3 verts[8];
    uint32_t kept = 0;
    float maxCoeff = -INFINITY;
    float dx = -1.0f, dy = -1.0f, dz = -1.0f;
#pragma unroll
    for (int o = 0; o < 8; o++) {
        const V3 vertex = xform(m, ICX_UNIT_RADIUS * dx, ICX_UNIT_RADIUS * dy, ICX_UNIT_RADIUS * dz);
        const float coeff = dot(vertex, axis);
        maxCoeff = coeff > maxCoeff ? coeff : maxCoeff;
        if (coeff >= maxCoeff - ICX_MANIFOLD_DEPTH) kept |= 1u << o;
        coeffs[o] = coeff; verts[o] = vertex;
        // winding order {0,1,0,2,0,1,0,2}: which dimension flips after vertex o
        if ((o & 1) == 0) dx = dx < 0.0f ? 1.0f : -1.0f;
        else if ((o & 3) == 1) dy = dy < 0.0f ? 1.0f : -1.0f;
        else dz = dz < 0.0f ? 1.0f : -1.0f;
    }
    int n = 0;
#pragma unroll
    for (int o = 7; o >= 0; o--)
        if ((kept >> o & 1u) && coeffs[o] >= maxCoeff - ICX_MANIFOLD_DEPTH) dest[n++] = verts[o];
    return n;
}
// The same with rolled loops over caller-provided scratch (tmpv: 8 vertices, tmpc: 8 floats): the
// ordered resolve runs this once per warp on cold code, where instruction fetch is the cost.
ICX_DI int build_contact_manifold_s(const M4& m, V3 axis, V3* dest, V3* tmpv, float* tmpc) {
    uint32_t kept = 0;
    float maxCoeff = -INFINITY;
    float dx = -1.0f, dy = -1.0f, dz = -1.0f;
#pragma unroll 1
    for (int o = 0; o < 8; o++) {
        const V3 vertex = xform(m, ICX_UNIT_RADIUS * dx, ICX_UNIT_RADIUS * dy, ICX_UNIT_RADIUS * dz);
        const float coeff = dot(vertex, axis);
        maxCoeff = coeff > maxCoeff ? coeff : maxCoeff;
        if (coeff >= maxCoeff - ICX_MANIFOLD_DEPTH) kept |= 1u << o;
        tmpc[o] = coeff; tmpv[o] = vertex;
        if ((o & 1) == 0) dx = dx < 0.0f ? 1.0f : -1.0f;
        else if ((o & 3) == 1) dy = dy < 0.0f ? 1.0f : -1.0f;
        else dz = dz < 0.0f ? 1.0f : -1.0f;
    }
    int n = 0;
#pragma unroll 1
    for (int o = 7; o >= 0; o--)
        if ((kept >> o & 1u) && tmpc[o] >= maxCoeff - ICX_MANIFOLD_DEPTH) dest[n++] = tmpv[o];
    return n;
}
// glm::intersectRayPlane, gtx/intersect.inl:6-26
ICX_DI bool intersect_ray_plane(V3 orig, V3 dir, V3 planeOrig, V3 planeNormal, float& dist) {
    float d = dot(dir, planeNormal);
    if (fabsf(d) > 1.1920928955078125e-07f) {
        float t = dot(planeOrig - orig, planeNormal) / d;
        if (t > 0.0f) { dist = t; return true; }
    }
    return false;
}
// pseudo_clip_polyhedra, collision_procedures.h:163-313.  `clipped` = caller-provided scratch of
// ICX_MAXV vertices: the ordered resolve hands in shared memory, because a single active lane
// working on thread-local arrays touches one 128-byte line per word and thrashes L1.
ICX_DI void pseudo_clip_polyhedra_s(const V3* clipper, int nc, V3* clippee, int& ne, V3 axis, V3* clipped) {
    V3 centre = v3(0.0f, 0.0f, 0.0f);
#pragma unroll 1
    for (int i = 0; i < nc; i++) centre = centre + clipper[i];
    centre = centre / (float)nc;
#pragma unroll 1
    for (int i = 0; i < nc; i++) {
        V3 tangent = cross(axis, clipper[(i + 1) % nc] - clipper[i]);
        if (length2(tangent) == 0.0f) continue;
        V3 pn = normalize(tangent);
        float pw = -1.0f * (pn.x * clipper[i].x + pn.y * clipper[i].y + pn.z * clipper[i].z);
        float insideCoeff = 0.0f;
        if (nc > 2) insideCoeff = pn.x * centre.x + pn.y * centre.y + pn.z * centre.z + pw;
        if (insideCoeff == 0.0f) insideCoeff = INFINITY;
        int nk = 0;
        float prevCoeff = -INFINITY;
        if (ne == 0) break;
#pragma unroll 1
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
                    if (nk < ICX_MAXV) clipped[nk++] = outsideVertex + dir * dist;
                }
            }
            if (currInside && j < ne) { if (nk < ICX_MAXV) clipped[nk++] = clippee[J]; }
        }
#pragma unroll 1
        for (int k = 0; k < nk; k++) clippee[k] = clipped[k];
        ne = nk;
    }
}

ICX_DN void pseudo_clip_polyhedra(const V3* clipper, int nc, V3* clippee, int& ne, V3 axis) {
    V3 clipped[ICX_MAXV];
    pseudo_clip_polyhedra_s(clipper, nc, clippee, ne, axis, clipped);
}

// the 15 SAT axes of box_box / sphere_box, collision_procedures.cpp:41-67; a = 0..14
ICX_DI V3 sat_axis(const M3& nA, const M3& nB, int a) {
    // normalTransform * vec4(e_i, 0): the zero products are kept, they decide signs of zeros
    if (a < 3) return mul(nA, v3(a == 0 ? 1.0f : 0.0f, a == 1 ? 1.0f : 0.0f, a == 2 ? 1.0f : 0.0f));
    if (a < 6) return mul(nB, v3(a == 3 ? 1.0f : 0.0f, a == 4 ? 1.0f : 0.0f, a == 5 ? 1.0f : 0.0f));
    int i = (a - 6) / 3, j = (a - 6) % 3;
    V3 ai = mul(nA, v3(i == 0 ? 1.0f : 0.0f, i == 1 ? 1.0f : 0.0f, i == 2 ? 1.0f : 0.0f));
    V3 bj = mul(nB, v3(j == 0 ? 1.0f : 0.0f, j == 1 ? 1.0f : 0.0f, j == 2 ? 1.0f : 0.0f));
    return cross(ai, bj);
}

// One SAT axis of box_box / sphere_box_intersect (collision_procedures.cpp:76-118): the serial
// procedures below loop over a = 0..14, the warp-cooperative ones (icx_resolve.cu) give one axis to
// each lane; both run exactly this arithmetic.
struct AxisResult { V3 axis; float pen; bool flip, skip; };

ICX_DI AxisResult sat_axis_result(V3 axis, Interval iA, Interval iB) {
    AxisResult r; r.axis = axis; r.skip = false; r.flip = false;
    if ((iA.x + iA.y) / 2.0f > (iB.x + iB.y) / 2.0f) r.pen = iB.y - iA.x;           // :95-98
    else { r.pen = iA.y - iB.x; r.flip = true; }
    return r;
}
ICX_DI AxisResult box_box_axis(const M3& nA, const M3& nB, const V3* cA, const V3* cB, int a) {
    const V3 axis = normalize(sat_axis(nA, nB, a));
    if (isnan(axis.x)) { AxisResult r; r.axis = axis; r.pen = 0.0f; r.flip = false; r.skip = true; return r; }   // :82
    return sat_axis_result(axis, project_corners(cA, axis), project_corners(cB, axis));
}
ICX_DI AxisResult sphere_box_axis(const M3& nA, const M3& nB, V3 sO, float projRadius, const V3* cB, int a) {
    const V3 axis = normalize(sat_axis(nA, nB, a));
    if (isnan(axis.x)) { AxisResult r; r.axis = axis; r.pen = 0.0f; r.flip = false; r.skip = true; return r; }
    return sat_axis_result(axis, project_sphere(sO, projRadius, axis), project_corners(cB, axis));
}

// box_box_intersect after the axis loop, collision_procedures.cpp:120-231 (h.normal / h.depth are set)
// mA (8), mB (ICX_MAXV), clipped (ICX_MAXV): caller-provided scratch, see pseudo_clip_polyhedra_s
ICX_DI bool box_box_finish_s(const Shape& A, const Shape& B, Hit& h, V3* mA, V3* mB, V3* clipped) {
    int nMA = build_contact_manifold_s(A.M, -h.normal, mA, clipped, (float*)(clipped + 8));      // :128-132 (clipped is free until the clip)
    int nMB = build_contact_manifold_s(B.M, h.normal, mB, clipped, (float*)(clipped + 8));
    if (nMB == 0 || nMA == 0) return false;                                         // :135-139
    if (nMB == 1) { h.point = mB[0]; return true; }                                 // :140-146
    if (nMA == 1) { h.point = mA[0] + (h.normal * h.depth); return true; }          // :147-153
    pseudo_clip_polyhedra_s(mA, nMA, mB, nMB, h.normal, clipped);                   // :159
    if (nMB == 1) { h.point = mB[0]; return true; }                                 // :168-172
    float maxCoeff = -INFINITY; V3 origin = v3(0.0f, 0.0f, 0.0f); float contributors = 0.0f;   // :181-199
#pragma unroll 1
    for (int i = 0; i < nMB; i++) {
        float vc = dot(mB[i], h.normal);
        if (vc > maxCoeff) { origin = mB[i]; contributors = 1.0f; maxCoeff = vc; }
        else if (vc == maxCoeff) { origin = origin + mB[i]; contributors += 1.0f; }
    }
    origin = origin / contributors;                                                 // :203
    const float manifoldRange = 1.0f - ICX_MANIFOLD_MIN_WEIGHT;
    V3 pt = v3(0.0f, 0.0f, 0.0f);
#pragma unroll 1
    for (int i = 0; i < nMB; i++) {                                                 // :210-221
        float vc = dot(mB[i], h.normal);
        V3 rel = mB[i] - origin;
        float w = (vc - maxCoeff + ICX_MANIFOLD_DEPTH) / ICX_MANIFOLD_DEPTH;
        pt = pt + rel * (w * manifoldRange + ICX_MANIFOLD_MIN_WEIGHT);
    }
    pt = pt / (float)nMB;                                                           // :222
    h.point = pt + origin;
    return true;
}

ICX_DN bool box_box_finish(const Shape& A, const Shape& B, Hit& h) {
    V3 mA[8], mB[ICX_MAXV], clipped[ICX_MAXV];
    return box_box_finish_s(A, B, h, mA, mB, clipped);
}

// box_box_intersect, collision_procedures.cpp:14-231
ICX_DN bool box_box_intersect(const Shape& A, const Shape& B, Hit& h) {
    const M3 nA = transpose(A.inv), nB = transpose(B.inv);      // :41-44
    V3 cA[8], cB[8];
    box_corners(A.M, cA); box_corners(B.M, cB);
    V3 minN = v3(0.0f, 0.0f, 0.0f); float minD = INFINITY;
    for (int a = 0; a < 15; a++) {
        const AxisResult r = box_box_axis(nA, nB, cA, cB, a);
        if (r.skip) continue;
        if (r.pen <= 0) return false;                                               // :107
        if (r.pen < minD) { minD = r.pen; minN = r.flip ? -r.axis : r.axis; }       // :113-117
    }
    h.normal = minN; h.depth = minD;
    return box_box_finish(A, B, h);
}

// sphere_box_intersect after the axis loop, collision_procedures.cpp:362-378
ICX_DI bool sphere_box_finish(V3 sO, V3 sS, Hit& h) {
    const float radius = length(sS);                                                // :367-368 (reference quirk: includes translation)
    h.point = sO - (h.normal * radius) + (h.normal * h.depth);                      // :370-373
    return true;
}

// sphere_box_intersect, collision_procedures.cpp:266-378 (A sphere, B box)
ICX_DN bool sphere_box_intersect(const Shape& A, const Shape& B, Hit& h) {
    const M3 nA = transpose(A.inv), nB = transpose(B.inv);
    V3 cB[8]; box_corners(B.M, cB);
    const V3 sO = xform(A.M, 0.0f, 0.0f, 0.0f);
    const V3 sS = xform(A.M, ICX_UNIT_RADIUS, 0.0f, 0.0f);
    const float projRadius = length(sS - sO);                                       // project_sphere :97
    V3 minN = v3(0.0f, 0.0f, 0.0f); float minD = INFINITY;
    for (int a = 0; a < 15; a++) {
        const AxisResult r = sphere_box_axis(nA, nB, sO, projRadius, cB, a);
        if (r.skip) continue;
        if (r.pen <= 0) return false;
        if (r.pen < minD) { minD = r.pen; minN = r.flip ? -r.axis : r.axis; }
    }
    h.normal = minN; h.depth = minD;
    return sphere_box_finish(sO, sS, h);
}

// sphere_sphere_intersect, collision_procedures.cpp:380-423 (opt-in, ICX_F_SPHERE_SPHERE)
ICX_DI bool sphere_sphere_intersect(const Shape& A, float rA, const Shape& B, float rB, Hit& h) {
    V3 oA = xform(A.M, 0.0f, 0.0f, 0.0f), oB = xform(B.M, 0.0f, 0.0f, 0.0f);
    V3 dir = oA - oB;
    float dist = length(dir);
    if (dist == 0) return false;
    float depth = (rA + rB) - dist;
    if (depth <= 0) return false;
    h.depth = depth; h.normal = dir / dist; h.point = oB + (h.normal * rB);
    return true;
}

// ray_box_intersect, collision_procedures.cpp:442-590 (A ray, B box); minParam = A.collider.minParametricValue
ICX_DN bool ray_box_intersect(const Shape& A, V3 entityOriginA, const Shape& B, V3 entityOriginB,
                                               float& minParam, Hit& h) {
    const M3 nB = transpose(B.inv);
    V3 cB[8]; box_corners(B.M, cB);
    V3 axes[6];
    axes[0] = mul(nB, v3(1.0f, 0.0f, 0.0f)); axes[1] = mul(nB, v3(0.0f, 1.0f, 0.0f)); axes[2] = mul(nB, v3(0.0f, 0.0f, 1.0f));
    V3 rayToBox = entityOriginB - entityOriginA;                                    // :463
    V3 rayOrigin = xform(A.M, 0.0f, 0.0f, 0.0f), rayEnd = xform(A.M, 0.0f, 0.0f, 1.0f);
    axes[3] = rayEnd - rayOrigin;
    axes[4] = cross(axes[3], rayToBox);
    axes[5] = cross(axes[4], axes[3]);
    V3 minN = v3(0.0f, 0.0f, 0.0f); float minD = INFINITY; float maxParam = 0.0f;
    for (int a = 0; a < 6; a++) {
        V3 axis = normalize(axes[a]);
        Interval ray = project_ray(A.M, axis), box = project_corners(cB, axis);
        float pen = 0, param = 0; bool flip = false;
        if (ray.x < (box.x + box.y) / 2.0f) flip = true;                            // :492
        if (ray.x >= box.x && ray.x <= box.y) {                                     // :508
            param = 0;
            if (!flip) pen = box.y - smin(ray.x, ray.y);
            else pen = smax(ray.x, ray.y) - box.x;
            if (pen < 0) return false;
        } else {                                                                    // :529
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
        if (param >= maxParam) { maxParam = param; minD = pen; minN = flip ? -axis : axis; }   // :553-558
    }
    h.normal = minN; h.depth = minD;
    if (maxParam < 0 || maxParam > 1) return false;                                 // :566
    if (maxParam >= minParam) return false;                                         // :574-578
    minParam = maxParam;
    h.point = rayOrigin + maxParam * (rayEnd - rayOrigin);                          // :583
    return true;
}

} // namespace icx
