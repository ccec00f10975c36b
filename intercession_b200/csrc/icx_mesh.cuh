// Triangle-mesh narrowphase: sphere - triangle and box - triangle.
//
// NO COUNTERPART IN THE REFERENCE: ColliderType::mesh is commented out there (source/physics/collider.h:18-29) and no
// procedure exists.  BASELINE.json's north_star and configs[2] ("plus torus.obj triangle-mesh colliders")
// name it, so this is an extension with a definition of its own, checked against an independent CPU
// restatement of that definition (oracle/icx_oracle.cpp icxo_mesh_pass: "self-oracle", parity unpinned by construction).
// Conventions follow the reference's procedures so that the reference's response applies unchanged: the normal points from
// the other collider (here: the triangle) towards the caller, depth > 0 is the penetration along it, point lies on the
// other collider's surface (collision_procedures.cpp:668-694).
#pragma once
#include "icx_narrow.cuh"

namespace icx {

struct Tri { V3 a, b, c; };

// Closest point of triangle (a, b, c) to p: Voronoi-region walk (vertex regions, edge regions, face).
ICX_DI V3 closest_point_triangle(V3 p, const Tri& t)
{
    const V3 ab = t.b - t.a, ac = t.c - t.a, ap = p - t.a;
    const float d1 = dot(ab, ap), d2 = dot(ac, ap);
    if (d1 <= 0.0f && d2 <= 0.0f) return t.a;
    const V3 bp = p - t.b;
    const float d3 = dot(ab, bp), d4 = dot(ac, bp);
    if (d3 >= 0.0f && d4 <= d3) return t.b;
    const float vc = d1 * d4 - d3 * d2;
    if (vc <= 0.0f && d1 >= 0.0f && d3 <= 0.0f) return t.a + ab * (d1 / (d1 - d3));
    const V3 cp = p - t.c;
    const float d5 = dot(ab, cp), d6 = dot(ac, cp);
    if (d6 >= 0.0f && d5 <= d6) return t.c;
    const float vb = d5 * d2 - d1 * d6;
    if (vb <= 0.0f && d2 >= 0.0f && d6 <= 0.0f) return t.a + ac * (d2 / (d2 - d6));
    const float va = d3 * d6 - d5 * d4;
    if (va <= 0.0f && (d4 - d3) >= 0.0f && (d5 - d6) >= 0.0f) return t.b + (t.c - t.b) * ((d4 - d3) / ((d4 - d3) + (d5 - d6)));
    const float denom = 1.0f / (va + vb + vc);
    return (t.a + ab * (vb * denom)) + ac * (vc * denom);
}

// Sphere of the composite transform M (centre M.t, radius as project_sphere takes it: |M * (0.5, 0, 0) - M * 0|,
// collision_procedures.h:88-101) against one triangle.
ICX_DI bool sphere_triangle_hit(const M4& M, const Tri& t, Hit& h)
{
    const V3 centre = xform(M, 0.0f, 0.0f, 0.0f);
    const float radius = length(xform(M, ICX_UNIT_RADIUS, 0.0f, 0.0f) - centre);
    const V3 q = closest_point_triangle(centre, t);
    const V3 d = centre - q;
    const float dist2 = dot(d, d);
    if (!(dist2 < radius * radius)) return false;
    const float dist = sqrtf(dist2);
    V3 n;
    if (dist > 0.0f) n = d * (1.0f / dist);
    else { n = normalize(cross(t.b - t.a, t.c - t.a)); if (!(n.x == n.x)) return false; }       // centre on a degenerate triangle
    h.normal = n; h.depth = radius - dist; h.point = q;
    return true;
}

// Box of the composite transform M (unit cube, half-extent 0.5, collision_procedures.h:27-68) against one triangle:
// separating-axis test over the triangle normal, the three box axes and the nine edge cross products, in that order;
// the axis of least penetration gives normal and depth (ties: the first).  point = the triangle's closest point to the box centre.
ICX_DI bool box_triangle_hit(const M4& M, const Tri& t, Hit& h)
{
    const V3 centre = M.t;
    const V3 e0 = t.b - t.a, e1 = t.c - t.b, e2 = t.a - t.c;
    float best = INFINITY; V3 best_n = v3(0.0f, 0.0f, 0.0f);
    for (int k = 0; k < 13; k++) {
        V3 L;
        if (k == 0) L = cross(e0, t.c - t.a);
        else if (k < 4) L = M.c[k - 1];
        else { const int i = (k - 4) / 3, j = (k - 4) % 3; L = cross(M.c[i], j == 0 ? e0 : (j == 1 ? e1 : e2)); }
        const float len2 = dot(L, L);
        if (!(len2 > 1.0e-12f)) continue;                                           // parallel edges / degenerate triangle: no axis
        L = L * (1.0f / sqrtf(len2));
        const float rb = ICX_UNIT_RADIUS * ((fabsf(dot(M.c[0], L)) + fabsf(dot(M.c[1], L))) + fabsf(dot(M.c[2], L)));
        const float cb = dot(centre, L);
        const float pa = dot(t.a, L), pb = dot(t.b, L), pc = dot(t.c, L);
        const float tmin = smin(smin(pa, pb), pc), tmax = smax(smax(pa, pb), pc);
        float pen; bool flip = false;
        if (cb > (tmin + tmax) * 0.5f) pen = tmax - (cb - rb);
        else { pen = (cb + rb) - tmin; flip = true; }
        if (pen <= 0.0f) return false;
        if (pen < best) { best = pen; best_n = flip ? -L : L; }
    }
    if (best == INFINITY) return false;
    h.normal = best_n; h.depth = best; h.point = closest_point_triangle(centre, t);
    return true;
}

// Ray of the composite transform M (origin M * 0, end M * (0, 0, 1), collision_procedures.h:73-86) against one triangle:
// Moeller-Trumbore.  A hit counts only in front of what the ray has already met (min_param, Collider::minParametricValue:
// ray_box_intersect keeps the nearest hit the same way, collision_procedures.cpp:566-578).  normal = the triangle's normal on the
// ray origin's side, depth = how far the ray's end lies behind the triangle's plane, point = where the ray pierces it.
ICX_DI bool ray_triangle_hit(const M4& M, const Tri& t, float min_param, Hit& h, float& param)
{
    const V3 o = xform(M, 0.0f, 0.0f, 0.0f), end = xform(M, 0.0f, 0.0f, 1.0f);
    const V3 d = end - o;
    const V3 e1 = t.b - t.a, e2 = t.c - t.a;
    const V3 pv = cross(d, e2);
    const float det = dot(e1, pv);
    if (!(fabsf(det) > 1.0e-12f)) return false;                                     // parallel to the plane, or a degenerate triangle
    const float inv = 1.0f / det;
    const V3 tv = o - t.a;
    const float u = dot(tv, pv) * inv;
    if (!(u >= 0.0f && u <= 1.0f)) return false;
    const V3 qv = cross(tv, e1);
    const float v = dot(d, qv) * inv;
    if (!(v >= 0.0f && u + v <= 1.0f)) return false;
    const float s = dot(e2, qv) * inv;
    if (!(s >= 0.0f && s <= 1.0f)) return false;
    if (s >= min_param) return false;
    V3 n = normalize(cross(e1, e2));
    if (dot(n, d) > 0.0f) n = -n;
    param = s;
    h.point = o + s * d;
    h.normal = n;
    h.depth = dot(n, h.point - end);
    return true;
}

} // namespace icx
