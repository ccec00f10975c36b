// Device helpers shared by the tick kernels: grid bucket function, neighbourhood walk, the Euler
// step of one entity (used by k_integrate_prep and by the ordered resolve's restore).
#pragma once
#include "icx_world.cuh"
#include "icx_math.cuh"

namespace icx {

__device__ __forceinline__ int cell_coord(float lo, float hi, float inv_cell) { return (int)floorf((lo + hi) * 0.5f * inv_cell); }

// Hashed block of 2^ICX_XB buckets holding the x-run [xb << ICX_XB, ...] of grid row (cy, cz) of world-line wl.
__device__ __forceinline__ uint32_t row_block(int xb, int cy, int cz, uint32_t wl, uint32_t n_buckets) {
    uint32_t h = (uint32_t)xb * 73856093u ^ (uint32_t)cy * 19349663u ^ (uint32_t)cz * 83492791u ^ wl * 2654435761u;
    h ^= h >> 15; h *= 2246822519u; h ^= h >> 13; h *= 3266489917u; h ^= h >> 16;
    return h & ((n_buckets >> ICX_XB) - 1u);
}
__device__ __forceinline__ uint32_t bucket_of(int cx, int cy, int cz, uint32_t wl, uint32_t n_buckets) {
    return (row_block(cx >> ICX_XB, cy, cz, wl, n_buckets) << ICX_XB) | (uint32_t)(cx & ICX_XMASK);
}

// Visit every bucket-sorted position j whose cell is one of the 27 cells around (cx, cy, cz):
// 9 rows x (1 or 2) contiguous ranges.  Ranges reached twice through hash collisions are visited once.
template <typename F>
__device__ __forceinline__ void for_each_neighbor(const DeviceWorld& W, int cx, int cy, int cz, uint32_t wl, F&& f)
{
    uint32_t seen[18]; int n_seen = 0;
    const int xb0 = (cx - 1) >> ICX_XB, xb1 = (cx + 1) >> ICX_XB;
    for (int dz = -1; dz <= 1; dz++)
        for (int dy = -1; dy <= 1; dy++)
            for (int xb = xb0; xb <= xb1; xb++) {
                const uint32_t block = row_block(xb, cy + dy, cz + dz, wl, W.n_buckets);
                const uint32_t key = block * 2u + (uint32_t)(xb - xb0);
                bool dup = false;
                for (int s = 0; s < n_seen; s++) dup |= (seen[s] == key);
                if (dup) continue;
                seen[n_seen++] = key;
                const int x_lo = max(cx - 1, xb << ICX_XB), x_hi = min(cx + 1, (xb << ICX_XB) + ICX_XMASK);
                const uint32_t b0 = (block << ICX_XB) | (uint32_t)(x_lo & ICX_XMASK), b1 = (block << ICX_XB) | (uint32_t)(x_hi & ICX_XMASK);
                const uint32_t beg = W.cell_count[b0], end = W.cell_count[b1 + 1];
                for (uint32_t j = beg; j < end; j++) f(j);
            }
}

__device__ __forceinline__ bool can_intersect(int a, int b, uint32_t cfg_flags) {
    // non-null entries of intersectProcedures, collision_procedures.h:411-417
    if (a == ICX_COLLIDER_BOX) return b != ICX_COLLIDER_NONE;
    if (b == ICX_COLLIDER_BOX) return a != ICX_COLLIDER_NONE;
    return (cfg_flags & ICX_F_SPHERE_SPHERE) && a == ICX_COLLIDER_SPHERE && b == ICX_COLLIDER_SPHERE;
}

__device__ __forceinline__ bool overlap(float4 alo, float4 ahi, float4 blo, float4 bhi) {
    return alo.x <= bhi.x && blo.x <= ahi.x && alo.y <= bhi.y && blo.y <= ahi.y && alo.z <= bhi.z && blo.z <= ahi.z;
}

// un-inflated bounds overlap (with 10 % of the margin kept as slack against rounding): only then can
// the narrowphase report a hit on the post-integration state (DESIGN.md "Broadphase bound")
__device__ __forceinline__ bool tight_le(float a, float b, float m2) {
    const float s = m2 - 2.0e-6f * (fabsf(a) + fabsf(b));
    return a + fmaxf(s, 0.0f) <= b;
}
__device__ __forceinline__ bool tight_overlap(float4 alo, float4 ahi, float4 blo, float4 bhi, float margin) {
    const float m2 = 1.8f * margin;
    return tight_le(alo.x, bhi.x, m2) && tight_le(blo.x, ahi.x, m2) && tight_le(alo.y, bhi.y, m2) && tight_le(blo.y, ahi.y, m2) &&
           tight_le(alo.z, bhi.z, m2) && tight_le(blo.z, ahi.z, m2);
}

struct EntityRegs { float4 pos, rot, vel, acc, av, aa; };

// EulerPhysicsRelay::engage for one entity, euler_physics_relay.h:25-95.  Reads W.in, returns the
// post-integration state in registers (deterministic: the restore path recomputes it bit for bit).
__device__ __forceinline__ EntityRegs euler_entity(const DeviceWorld& W, uint32_t e, uint32_t ef, float4 p1, double deltaTime, int do_euler)
{
    EntityRegs r;
    r.pos = W.in.pos[e]; r.rot = W.in.rot[e]; r.vel = W.in.vel[e]; r.acc = W.in.acc[e];
    r.av = W.in.angvel[e]; r.aa = W.in.angacc[e];
    if (do_euler && (ef & EF_HAS_PHYSICS) && !(ef & EF_ASLEEP)) {
        V3 org = v3(r.pos), v = v3(r.vel), a = v3(r.acc), w = v3(r.av), al = v3(r.aa);
        Q4 q; q.w = r.rot.x; q.x = r.rot.y; q.y = r.rot.z; q.z = r.rot.w;
        a = a + v3(W.gx, W.gy, W.gz);                                               // :32
        if (ef & EF_LOCK_ORIGIN) { v = v3(0.0f, 0.0f, 0.0f); a = v3(0.0f, 0.0f, 0.0f); org = v3(W.locked_pos[e]); }   // :36-46
        if (ef & EF_LOCK_ORIENTATION) {                                             // :47-57
            w = v3(0.0f, 0.0f, 0.0f); al = v3(0.0f, 0.0f, 0.0f);
            float4 lq = W.locked_rot[e]; q.w = lq.x; q.x = lq.y; q.y = lq.z; q.z = lq.w;
        }
        const float halfDt = (float)(deltaTime / 2.0f);                             // :62 double divide, then narrowed
        const float dt = (float)deltaTime;
#pragma unroll
        for (int half = 0; half < 2; half++) {
            org = org + v * halfDt;                                                 // :62 / :80
            float angularSpeed = length(w);
            if (angularSpeed != 0.0f) q = qmul(angle_axis(angularSpeed * halfDt, w / angularSpeed), q);   // :65-70
            if (half == 0) { v = v + a * dt; w = w + al * dt; }                     // :74-75
        }
        a = v3(0.0f, 0.0f, 0.0f); al = v3(0.0f, 0.0f, 0.0f);                         // :90-91
        v = v * (float)(1.0f - p1.x * deltaTime);                                   // :94 evaluated in double, narrowed by glm's *=
        w = w * (float)(1.0f - p1.y * deltaTime);                                   // :95
        r.pos = make_float4(org.x, org.y, org.z, 0.0f);
        r.rot = make_float4(q.w, q.x, q.y, q.z);
        r.vel = make_float4(v.x, v.y, v.z, 0.0f); r.acc = make_float4(a.x, a.y, a.z, 0.0f);
        r.av = make_float4(w.x, w.y, w.z, 0.0f); r.aa = make_float4(al.x, al.y, al.z, 0.0f);
    }
    return r;
}

// Mark entity e dirty (idempotent, any thread); returns true for the thread that appended it.
__device__ __forceinline__ bool mark_dirty(const DeviceWorld& W, uint32_t e)
{
    if (W.dirty_slot[e] != ICX_NONE) return false;
    if (atomicCAS(&W.dirty_slot[e], ICX_NONE, 0xFFFFFFFEu) != ICX_NONE) return false;
    const uint32_t s = atomicAdd(&W.counters[CTR_DIRTY], 1u);
    W.dirty_list[s] = e;
    __threadfence();
    atomicExch(&W.dirty_slot[e], s);
    return true;
}

} // namespace icx
