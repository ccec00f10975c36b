// Device-resident world state (SoA) and the parameter block every kernel receives by value.
// Layout rationale: DESIGN.md "Data layout in HBM".  All per-entity / per-packet arrays are
// float4/uint4-granular so each thread moves 128-bit words and a warp moves 512 contiguous bytes.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/icx.h"

namespace icx {

// entity flag bits (EntityConst::flags)
enum : uint32_t {
    EF_HAS_PHYSICS      = 1u << 0,
    EF_LOCK_ORIGIN      = 1u << 1,
    EF_LOCK_ORIENTATION = 1u << 2,
    EF_ASLEEP           = 1u << 3,
    EF_IMMUTABLE        = 1u << 4,   // no collision response can change it (see is_immutable)
    EF_HAS_RAY          = 1u << 5,
    EF_INFLUENCE        = 1u << 6,   // some active collider has influenceOrientation (is_immutable needs it when a replay changes mass / locks)
};
// packet flag bits (packed in PacketDerived::bits together with the two type enums)
enum : uint32_t {
    PF_TYPE_MASK        = 0x3u,          // ICX_COLLIDER_*
    PF_CTYPE_SHIFT      = 2,             // ICX_COLLISION_* in bits 2..3
    PF_BEHAVIORS        = 1u << 4,
    PF_INHERIT_ORIENT   = 1u << 5,
    PF_INFLUENCE_ORIENT = 1u << 6,
    PF_LARGE            = 1u << 7,
    PF_IMMUTABLE        = 1u << 8,
    PF_SLOT             = 1u << 9,
};

// neighbour-list entry of a dirty body: other packet index in the low 28 bits, then
#define ICX_NB_INDEX_MASK 0x0FFFFFFFu
#define ICX_NB_IMMUTABLE  0x10000000u     // other packet's entity is immutable (never written, no ordering)
// flat broadphase pair entry (uint2: x = lower packet, y = higher packet | flags)
#define ICX_PAIR_TIGHT    0x80000000u     // un-inflated bounds overlap too: the narrowphase can hit
#define ICX_PAIR_STATIC   0x40000000u     // both bodies immutable
#define ICX_PAIR_B_IMMUT  0x20000000u     // (in .y) the higher packet's body is immutable
#define ICX_PAIR_A_IMMUT  0x80000000u     // (in .x) the lower packet's body is immutable
#define ICX_NONE          0xFFFFFFFFu
#define ICX_ITEM_CHECK    0x80000000u     // work item (uint4.y): two-sided pair, the SAT filter has not seen it yet
// grid: 2^ICX_XB x-adjacent cells of one (y,z,world-line) row share a hashed block of buckets, so
// the 3 x-neighbours of a cell are one contiguous range of the bucket-sorted array
#define ICX_XB 4
#define ICX_XMASK ((1 << ICX_XB) - 1)

// error / retry flags raised by kernels (World::d_flags)
enum : uint32_t {
    RF_NEIGHBOR_OVERFLOW = 1u << 0,   // the candidate-list pool is full (or one list has more than 65535 entries)
    RF_MARGIN_VIOLATED   = 1u << 1,   // static resolution moved a body further than the margin
    RF_CELL_TOO_SMALL    = 1u << 2,   // a mutable collider is larger than a grid cell
    RF_CONTACT_OVERFLOW  = 1u << 3,
    RF_LARGE_OVERFLOW    = 1u << 4,
    RF_CLIP_OVERFLOW     = 1u << 5,
    RF_PAIR_OVERFLOW     = 1u << 6,   // the flat pair list / static-pair list exceeded its capacity
    RF_ROUNDS_EXCEEDED   = 1u << 7,
};

// mutable entity state: one buffer is the tick input, the other the tick output
struct EntityState {
    float4* pos;     // origin.xyz, -
    float4* rot;     // orientation w,x,y,z
    float4* vel;     // velocity.xyz, -
    float4* acc;     // acceleration.xyz, -
    float4* angvel;  // angularVelocity.xyz, -
    float4* angacc;  // angularAcceleration.xyz, -
    uint32_t* ts;    // timestream state (low 8 bits) | coherency stamp << 16   (cosmos.h:247)
};

struct DeviceWorld {
    uint32_t n_entities, n_packets;
    EntityState in, out;            // Euler reads `in`, writes `out`; the collision pass mutates `out`
    // per-entity constants
    float4* prm0;                   // scale.xyz, mass
    float4* prm1;                   // linearDrag, angularDrag, collisionAngularDrag, bits(flags)
    float4* locked_pos;             // lockedOrigin.xyz, collisionLinearDrag (unused by the tick, carried so that P is complete in SoA)
    float4* locked_rot;             // lockedOrientation wxyz
    uint32_t* entity_id;            // IcxEntity
    uint32_t* first_packet;         // packet index of the entity's first active collider | count << 30
    uint32_t* c_ver;                // per entity: world epoch (IcxWorld::c_epoch) at which its ColliderComponent last changed (record: sparse collider deltas)
    float4* pos0;                   // origin right after integration (margin check)
    // per-packet constants (uploaded AoS -> SoA)
    uint32_t* pk_entity;            // entity index
    float4* pk_lpos;                // localTransform.origin.xyz, bits
    float4* pk_lrot;                // localTransform.orientation wxyz
    float4* pk_lscale;              // localTransform.scale.xyz, restLength
    float4* pk_mat0;                // staticFriction, dynamicFriction, restitution, stiffness
    float* pk_damping;
    float* pk_minparam;             // Collider::minParametricValue (mutable, reset to 1 every tick)
    // per-packet, derived every tick by the prep kernel
    float4* pk_d;                   // 6 x float4 per packet, contiguous (96 B): d[p * 6 + k]
                                    //   k=0..2: composite linear column k .xyz, base[k]
                                    //   k=3..5: inverse linear column .xyz, inertia diag [k-3]
    float4* pk_lo;                  // inflated bound lo.xyz, -
    float4* pk_hi;                  // inflated bound hi.xyz, -
    uint32_t* pk_cell;              // hashed grid bucket (or 0xFFFFFFFF for large / inert)
    // grid
    uint32_t n_buckets;             // power of two
    uint32_t* cell_count;           // [n_buckets + 1] -> exclusive scan in place = cell_start
    uint32_t* cell_fill;
    uint32_t* scan_partial;
    float4* s_lo;                   // sorted by bucket: lo.xyz, asfloat(packet | type << 28 | immutable << 30)
    float4* s_hi;                   // hi.xyz, asfloat(entity index)
    uint32_t* large_list;           // packets larger than a cell (immutable only)
    // flat candidate pair list (the broadphase output, icx_pairs_fetch)
    uint2* pairs; uint32_t max_pairs;
    float* push;                    // per entity: largest single predicted push by static resolution this tick (own contacts, or handed on by a neighbour)
    float4* pushv;                  // per entity: vector sum of the predicted pushes of its own speculative contacts
    // Per-body displacement budget (DESIGN.md "Broadphase bound"): the stored bounds are inflated by the uniform `margin`;
    // a body the speculative pass predicts to be pushed further, or that was caught moving further, gets bm[e] > margin and is
    // a MOVER: its extra candidates (xpairs) come from a wide grid walk of its own, nobody else pays for it.
    float* bm;                      // per entity: budget of this tick, >= margin
    float* bm_keep;                 // per entity: floor carried over ticks (raised by a violation, decays)
    uint32_t* bm_need;              // per entity: float bits of the largest displacement seen beyond bm (0 = none)
    uint32_t* mover_list;           // entities with bm > margin (mutable ones)
    uint2* xpairs; uint32_t max_xpairs;     // candidate pairs only the movers' budgets create (same entry format as pairs)
    uint2* tight;                   // the subset with ICX_PAIR_TIGHT (input of the speculative narrowphase)
    uint2* static_pairs; uint32_t max_static;    // hits between two immutable bodies (contact records only)
    // bodies touched by at least one hit ("dirty"): only these need the reference's sequential order
    uint32_t* dirty_slot;           // per entity: index into dirty_list, ICX_NONE = clean
    uint32_t* dirty_list;           // entity indices, arbitrary order
    uint2* nbr;                     // pool of sorted candidate lists of the dirty bodies (bump-allocated per tick, CTR_POOL): entry = {other packet | flags, its entity}
    uint32_t* nbr_count;            // [slot * 2 + k] entries of the list being built
    uint32_t* nbr_fill;             // [slot * 2 + k] fill cursor while the list is written
    uint32_t* slot_flag;            // per dirty slot: list to be rebuilt with the movers' candidates
    uint4* slot_info;               // per dirty slot: {first packet | count << 30, n0 | n1 << 16, pool offset of list 0 (list 1 follows), entity}
    uint32_t max_neighbors;         // pool capacity per entity (pool = max_neighbors * capacity in entities)
    uint32_t max_nbr;               // pool capacity in entries
    uint32_t* head[2];              // per dirty slot: consumed entries of the slot-0 list | slot-1 list << 16
    uint32_t* active[2];            // dirty-slot work lists of the dependency rounds
    uint4* items;                   // per round: pairs that need the exact procedure {pa, pb, slot a, slot b | ICX_NONE}
    // outputs
    IcxContact* contacts; IcxInterception* intercepts; uint2* intercept_key; uint32_t max_contacts;
    IcxGround* grounds;             // one record per on_collision call a BipedBehaviors drivetrain would receive (CTR_GROUNDS)
    // counters, one cache line each would be nicer; they are cold
    uint32_t* counters;             // see CTR_*
    uint32_t* flags;                // RF_*
    unsigned long long* trace;      // optional device timeline of k_resolve (globaltimer ns), ICX_TRACE_CAP entries; may be null
    // parameters
    float margin, cell_size, inv_cell;
    float gx, gy, gz;
    uint32_t host_timeslice, cfg_flags, worldline_stride, max_large;
    uint16_t coherency;
};

enum { CTR_LARGE = 0, CTR_DIRTY = 1, CTR_CONTACTS = 2, CTR_INTERCEPTS = 3, CTR_CANDIDATES = 4, CTR_STATIC = 5,
       CTR_LISTED = 6,          // dirty bodies whose candidate lists are built
       CTR_ROUNDS = 7, CTR_ATTEMPTS = 8, CTR_NEAR = 9, CTR_TIGHT = 10,
       CTR_NEED_CELL = 11,      // float bits: cell size a too-large mutable collider needs
       CTR_EXTENT = 12,         // float bits: largest collider extent seen at import
       CTR_ACTIVE = 13,         // 3 rotating work-list counters: CTR_ACTIVE + (round % 3)
       CTR_ITEMS = 18,          // 2 alternating work-item counters
       CTR_NEWD = 20,           // 3 rotating per-round counts of bodies that turned dirty during the replay
       CTR_FETCH = 23,          // 2 alternating work-item fetch counters
       CTR_MOVERS = 25, CTR_XPAIRS = 26,
       CTR_XMAX = 27,           // float bits: largest (bm - margin) of this tick
       CTR_VIOLATORS = 28,      // bodies the current attempt moved beyond their budget
       CTR_POOL = 29,           // entries of the list pool in use
       CTR_NGREW = 30, CTR_NBUDGET = 31,   // restarts of the replay by reason (diagnostic)
       CTR_GROUNDS = 17,
       CTR_PUSHED = 16,         // predicted pushes raised by the current prediction pass
       CTR_PROF = 32,           // ICX_PROFILE_PAIR builds: cycle accumulators
       CTR_TOTAL = 48 };
#define ICX_MAX_ROUNDS 65535u
#define ICX_TRACE_CAP 512

// static triangle mesh (icx_mesh.cuh): triangles as 3 float4 each, a uniform grid over them (CSR, ascending indices per cell)
struct DeviceMesh {
    const float4* tri; const uint32_t* cell_start; const uint32_t* cell_tri;
    uint32_t n_tri; int dim[3]; float lo[3]; float inv_cell;
    float staticFriction, dynamicFriction, restitution;
    IcxMeshContact* contacts; uint32_t max_contacts; uint32_t* counter;      // counter[0] contacts, [1] overflow, [2] entities selected by phase 1
    uint32_t* active;               // [n_entities] rows phase 1 selected
};

} // namespace icx
