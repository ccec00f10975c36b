/*
 * icx.h -- C ABI of intercession-b200: the physics tick and the world-line
 * record / replay / divergence scan of meta-engineer/Intercession, on B200.
 *
 * Every entry point is `extern "C"`, takes plain pointers and sizes, returns an
 * int status (never throws / aborts across the boundary) and replaces the
 * reference interface cited next to it (paths relative to the reference
 * repository root).  The reference-side binding is shown in INTEGRATION.md.
 *
 * Threading: an IcxWorld is single-thread-affine (one CUDA stream); distinct
 * worlds are independent.  No CPU fallback exists: without a CUDA device
 * icx_world_create() fails with ICX_E_CUDA.
 */
#ifndef ICX_H
#define ICX_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ICX_ABI_VERSION 2

#if defined(__GNUC__)
#define ICX_API __attribute__((visibility("default")))
#else
#define ICX_API
#endif

/* ---- status codes (reference: C++ exceptions, SURVEY 8b "Errors") ---- */
#define ICX_OK           0
#define ICX_E_INVALID    1   /* bad argument / bad call order                */
#define ICX_E_CUDA       2   /* CUDA runtime error (see icx_last_error)      */
#define ICX_E_CAPACITY   3   /* a fixed-capacity device buffer overflowed    */
#define ICX_E_NOMEM      4
#define ICX_E_STATE      5   /* object is in the wrong state for this call   */

/* ---- enums mirroring source/physics/collider.h:18-37 ---- */
#define ICX_COLLIDER_NONE   0
#define ICX_COLLIDER_BOX    1
#define ICX_COLLIDER_SPHERE 2
#define ICX_COLLIDER_RAY    3

#define ICX_COLLISION_NOOP   0
#define ICX_COLLISION_RIGID  1
#define ICX_COLLISION_SPRING 2
#define ICX_COLLISION_FORCE  3

/* ---- TimestreamState, source/spacetime/timestream_state.h:11-18 ---- */
#define ICX_TS_MERGED        0
#define ICX_TS_FORKED        1
#define ICX_TS_FORKING       2
#define ICX_TS_SUPERPOSITION 3
#define ICX_TS_GHOST         4

/*
 * Entity id.  The reference packs a uint16 as 4b timeslice | 8b genesis | 4b
 * causal chain-link (source/ecs/ecs_types.h:16-58), which caps a cosmos at 255
 * hosted entities.  Here the same three fields live in a uint32:
 * 4b timeslice [31:28] | 24b genesis [27:4] | 4b chain-link [3:0].
 * A reference id `e` maps to ((e & 0xF000) << 16) | (e & 0x0FFF).
 */
typedef uint32_t IcxEntity;
#define ICX_NULL_ENTITY        0xFFFFFFFFu
#define ICX_NULL_TIMESLICE     14u          /* ecs_types.h: NULL_TIMESLICEID       */
#define ICX_NULL_CHAINLINK     15u          /* ecs_types.h: NULL_CAUSALCHAINLINK   */
#define ICX_COLLIDERS_PER_ENTITY 2          /* source/physics/collider_component.h:9 */

/* ---- AoS component layouts, byte-identical to the reference structs ---- */

/* source/physics/transform_component.h:14-20 (40 B; quaternion stored w,x,y,z) */
typedef struct IcxTransform {
    float origin[3];
    float orientation[4];   /* w, x, y, z */
    float scale[3];
} IcxTransform;

/* source/physics/physics_component.h:24-55 (108 B) */
typedef struct IcxPhysics {
    float   velocity[3];
    float   acceleration[3];
    float   angularVelocity[3];
    float   angularAcceleration[3];
    float   linearDrag;
    float   angularDrag;
    float   collisionLinearDrag;
    float   collisionAngularDrag;
    float   mass;                   /* 0 == INFINITE_MASS */
    uint8_t lockOrigin;      uint8_t _pad0[3];
    float   lockedOrigin[3];
    uint8_t lockOrientation; uint8_t _pad1[3];
    float   lockedOrientation[4];   /* w, x, y, z */
    uint8_t isAsleep;        uint8_t _pad2[3];
} IcxPhysics;

/* source/physics/collider.h:42-92 (84 B) */
typedef struct IcxCollider {
    int32_t colliderType;           /* ICX_COLLIDER_*  */
    int32_t collisionType;          /* ICX_COLLISION_* */
    uint8_t isActive;
    uint8_t useBehaviorsResponse;
    uint8_t _pad0[2];
    IcxTransform localTransform;
    uint8_t inheritOrientation;
    uint8_t influenceOrientation;
    uint8_t _pad1[2];
    float   minParametricValue;
    float   staticFriction;
    float   dynamicFriction;
    float   restitution;
    float   stiffness;
    float   damping;
    float   restLength;
} IcxCollider;

/* source/physics/collider_component.h:11-14 (168 B) */
typedef struct IcxColliderComponent {
    IcxCollider colliders[ICX_COLLIDERS_PER_ENTITY];
} IcxColliderComponent;

/*
 * One narrowphase hit, in the values the reference hands to
 * I_BehaviorsDrivetrain::on_collision (collision_physics_relay.h:114-155):
 * i.e. AFTER the response procedure adjusted collisionPoint.  Records are
 * returned in the reference's pair order (lexicographic on collider packets).
 * flags: bit0 = a physics response ran, bit1/bit2 = collider A/B has
 * useBehaviorsResponse (the adapter calls on_collision for those).
 */
typedef struct IcxContact {
    IcxEntity entityA;
    IcxEntity entityB;
    uint8_t   slotA;
    uint8_t   slotB;
    uint16_t  flags;
    float     point[3];
    float     normal[3];
    float     depth;
    uint32_t  packetA;      /* collider packet indices (submission order) */
    uint32_t  packetB;
} IcxContact;

#define ICX_CONTACT_RESPONDED   0x1u
#define ICX_CONTACT_BEHAVIORS_A 0x2u
#define ICX_CONTACT_BEHAVIORS_B 0x4u

/* What BipedBehaviors::on_collision (source/behaviors/biped_behaviors.h:336-366) writes into the caller's BipedComponent:
 * isGrounded = true, groundNormal = the collision normal as handed to the caller, groundDist = length(caller's origin at
 * callback time - collision point).  icx_grounds_fetch returns, per entity that received at least one such call in the last
 * tick (a collider with useBehaviorsResponse was hit), the values of the LAST call in the reference's pair order -- what the
 * component holds after the tick.  (packetA, packetB, side) identify that call. */
typedef struct IcxGround {
    IcxEntity entity;
    float     normal[3];
    float     dist;
    uint32_t  packetA, packetB;
    uint32_t  side;                 /* 0: the entity was the pair's A (lower packet), 1: its B */
} IcxGround;

/* events::cosmos::TIMESTREAM_INTERCEPTION_params, source/events/event_types.h:133-139 */
typedef struct IcxInterception {
    IcxEntity agent;
    IcxEntity recipient;
} IcxInterception;

/* A broadphase candidate: collider packet indices, a < b. */
typedef struct IcxPair {
    uint32_t a;
    uint32_t b;
} IcxPair;

#define ICX_F_SPHERE_SPHERE   0x1u  /* wire the reference's unwired sphere_sphere_intersect
                                       (collision_procedures.cpp:380-423); default off = parity
                                       with the dispatch table collision_procedures.h:411-417 */
#define ICX_F_NO_GRAVITY      0x2u  /* skip the hard-coded gravity (euler_physics_relay.h:32)  */
#define ICX_F_LANE_ROUNDS     0x4u  /* diagnostic: run EVERY dependency round of the ordered resolve
                                       lane-parallel (one work item per lane; normally only rounds with
                                       >= 4 items per warp do).  Results are identical either way. */

typedef struct IcxConfig {
    uint32_t struct_size;           /* = sizeof(IcxConfig)                                      */
    int32_t  device;                /* CUDA device ordinal                                      */
    uint32_t max_entities;
    uint32_t host_timeslice;        /* Cosmos::get_host_id(); ICX_NULL_TIMESLICE = parallel     */
    uint32_t flags;                 /* ICX_F_*                                                  */
    float    gravity[3];            /* default (0,-10,0), euler_physics_relay.h:32              */
    float    broadphase_margin;     /* AABB inflation; 0 = default 0.05                         */
    float    cell_size;             /* uniform-grid cell; 0 = chosen at upload                  */
    uint32_t max_neighbors;         /* candidate list capacity per dirty collider; 0 = 24 (grows) */
    uint32_t max_large;             /* colliders larger than a cell; 0 = default 1024           */
    uint32_t max_contacts;          /* contact / interception ring; 0 = 2 * max_entities        */
    uint32_t worldline_stride;      /* 0 = one world; else entities [k*stride,(k+1)*stride) form
                                       independent world-line k (pairs never cross world-lines) */
    uint32_t max_pairs;             /* broadphase pair list capacity; 0 = 8 * max_entities (grows)    */
} IcxConfig;

typedef struct IcxStats {
    uint64_t ticks;
    uint32_t n_entities;
    uint32_t n_packets;             /* active colliders submitted this tick                     */
    uint32_t n_large;
    uint32_t n_candidates;          /* broadphase pairs of the last tick                        */
    uint32_t n_contacts;
    uint32_t n_interceptions;
    uint32_t n_rounds;              /* dependency rounds of the ordered resolve (last attempt)  */
    uint32_t n_retries;             /* ticks re-run with a larger margin / capacity (cumulative)*/
    float    margin;                /* uniform part of the broadphase margin (per-body budgets may exceed it) */
    float    cell_size;
    uint32_t max_neighbors;
    uint32_t n_dirty;               /* bodies touched by a hit: the only ones resolved in order  */
    uint32_t n_attempts;            /* ordered-resolve attempts (1 + times the dirty set grew)   */
    uint32_t n_movers;              /* bodies whose displacement budget exceeded the uniform margin (own candidate walk) */
} IcxStats;

typedef struct IcxWorld IcxWorld;
typedef struct IcxTimeline IcxTimeline;

/* ---- library ---- */
ICX_API int         icx_abi_version(void);
ICX_API int         icx_device_count(void);             /* 0 when no CUDA device is usable */
ICX_API const char* icx_status_string(int status);
ICX_API void        icx_config_default(IcxConfig* cfg);

/* ---- world lifetime (replaces PhysicsDynamo ctor/dtor, physics_dynamo.cpp:7-21) ---- */
ICX_API int         icx_world_create(const IcxConfig* cfg, IcxWorld** out);
ICX_API void        icx_world_destroy(IcxWorld* w);
ICX_API const char* icx_last_error(const IcxWorld* w);

/*
 * Upload = PhysicsSynchro::update + ColliderSynchro::update
 * (physics_synchro.cpp:12-36, collider_synchro.cpp:11-43): entities in ascending
 * submission order; entity i has transform[i], optional physics (has_physics[i],
 * NULL = all have one; collision_physics_relay.h:84-106 skips the response
 * otherwise), optional collider component (colliders NULL = none), timestream
 * state (Cosmos::get_timestream_state, cosmos.cpp:175-178; NULL = merged, 0).
 */
ICX_API int icx_bodies_upload(IcxWorld* w, uint32_t n_entities,
                      const IcxEntity* entity,
                      const IcxTransform* transform,
                      const IcxPhysics* physics,
                      const uint8_t* has_physics,
                      const IcxColliderComponent* colliders,
                      const uint8_t* tstate,
                      const uint16_t* tstate_coherency);

/* Write device state back into the caller's AoS arrays (any pointer may be NULL). */
ICX_API int icx_bodies_download(IcxWorld* w, uint32_t n_entities,
                        IcxTransform* transform,
                        IcxPhysics* physics,
                        IcxColliderComponent* colliders,
                        uint8_t* tstate,
                        uint16_t* tstate_coherency);

/* What a tick writes: origin, orientation, velocity, acceleration, angularVelocity,
 * angularAcceleration -- 76 B per entity instead of the 148 B of the two whole components.  The adapter scatters the rows
 * back through its packet references. */
typedef struct IcxMotionState {
    float origin[3];
    float orientation[4];   /* w, x, y, z */
    float velocity[3];
    float acceleration[3];
    float angularVelocity[3];
    float angularAcceleration[3];
} IcxMotionState;
ICX_API int icx_motion_download(IcxWorld* w, uint32_t n_entities, IcxMotionState* out);

/* Host-side edits between ticks (behaviours, replay): overwrite `count` entities
 * listed in `index` (indices into the upload order; NULL = the first `count` entities) from the given AoS arrays,
 * which are indexed by position in `index`.  NULL arrays are left untouched. */
ICX_API int icx_bodies_update(IcxWorld* w, uint32_t count, const uint32_t* index,
                      const IcxTransform* transform,
                      const IcxPhysics* physics,
                      const IcxColliderComponent* colliders);

/*
 * ParallelCosmosContext::load_and_link (source/spacetime/parallel_cosmos_context.cpp:60-169):
 * make `dst` a copy of `src` as the parallel cosmos sees it.  Every entity is copied with all its
 * components except those with causal chain-link 0 and no future timestream (has_future[i] == 0;
 * NULL = every entity has one, :110-116).  dst takes src's coherency (:91).  A FORKED entity stays
 * forked in dst and is re-merged in src (:145-150); every other entity starts {merged, 0} in dst.
 * The two worlds may live on different GPUs (peer copy).  dst keeps its own IcxConfig
 * (host_timeslice = ICX_NULL_TIMESLICE for the parallel cosmos).  Linking the timestreams
 * (:157) is the caller's icx_timeline_* handle sharing.
 */
ICX_API int icx_world_clone(IcxWorld* src, IcxWorld* dst, const uint8_t* has_future, uint32_t* n_cloned);

/*
 * Static triangle mesh (extension).  NO COUNTERPART IN THE REFERENCE: ColliderType::mesh is commented out there
 * (source/physics/collider.h:18-29); BASELINE.json's north_star and configs[2] name triangle-mesh colliders, so the
 * definition is this library's own and its checker is a CPU restatement of that definition
 * (oracle icxo_mesh_pass: self-oracle, parity unpinned).
 * A world may hold one mesh: world-space triangles, immutable, infinite mass.  After the pair loop of every collision pass
 * (icx_collision_engage / icx_step) each entity that has a PhysicsComponent of finite mass (mass != 0) meets the mesh: entities in row order, their
 * active box / sphere colliders of collision type rigid and ray colliders of collision type spring in slot order, and for each collider the triangles whose bounds
 * overlap the collider's bounds (taken when the collider's turn starts) in ascending triangle index.  A hit (sphere -
 * triangle: closest point; box - triangle: 13-axis separating-axis test) is answered with the reference's
 * rigid_rigid_response (collision_procedures.cpp:636-874) against an immovable body of the mesh's material, so later
 * triangles see the moved body.  (The mesh belongs to the world it was uploaded to: icx_world_clone copies entities, not the
 * mesh -- a parallel cosmos that should meet it gets its own icx_mesh_upload.)  A ray (Moeller-Trumbore) only counts in front of what it has met so far this tick
 * (Collider::minParametricValue, as ray_box_intersect keeps it) and is answered with spring_rigid_response (:876-1053).  Entities do not interact in this pass.  Contacts: icx_mesh_contacts_fetch, in that order.
 */
typedef struct IcxMeshMaterial { float staticFriction, dynamicFriction, restitution; } IcxMeshMaterial;
typedef struct IcxMeshContact {
    IcxEntity entity;
    uint32_t row;           /* the entity's row */
    uint32_t slot;          /* collider slot */
    uint32_t triangle;
    float point[3];         /* on the triangle */
    float normal[3];        /* from the triangle towards the entity */
    float depth;
} IcxMeshContact;           /* 44 bytes */
/* vertices [n_vertices][3], triangles [n_triangles][3] vertex indices; material NULL = {0.45, 0.45, 0.70} (collider.h:58-92
 * defaults).  n_triangles == 0 removes the mesh.  Synchronous. */
ICX_API int icx_mesh_upload(IcxWorld* w, const float* vertices, uint32_t n_vertices, const uint32_t* triangles, uint32_t n_triangles,
                            const IcxMeshMaterial* material);
ICX_API int icx_mesh_contacts_fetch(IcxWorld* w, IcxMeshContact* out, size_t cap, size_t* n_out);
/* device time of the last mesh pass in microseconds (0 if none ran) */
ICX_API int icx_mesh_kernel_us(IcxWorld* w, float* us);

/* Several worlds on ONE device, ticked concurrently from different host threads (each world has its own stream): the two
 * persistent cooperative kernels of the ordered resolve normally fill the device, so a second world's could only start when the
 * first one's have finished.  With parts = k this world launches them over 1 / k of the co-resident blocks, and k such worlds
 * overlap -- which pays when a world is small enough for its dependency rounds to be latency-bound (config 4 sharded over many
 * GPUs: DESIGN.md 9).  Results do not depend on it.  parts = 1 restores the default. */
ICX_API int icx_device_share_set(IcxWorld* w, uint32_t parts);

/* Page-locked host memory for component storage that is handed to icx_bodies_upload / _update / _download every tick (the
 * copies then run at full PCIe rate and asynchronously; with pageable memory the driver stages them).  For hosts that do
 * not link the CUDA runtime themselves (icx_host::ComponentPools, host/component_pools.hpp, keeps its rows in it).
 * NULL when the allocation fails or no device is present. */
ICX_API void* icx_host_alloc(size_t bytes);
ICX_API void icx_host_free(void* p);

/*
 * Entity lifecycle on the device: Cosmos::register_entity / condemn_entity as the timeline's ENTITY_CREATED / ENTITY_REMOVED
 * handlers (source/server/server_network_dynamo.cpp:422-463) and the arrival of a time traveller (:748-812) call them.
 * The new rows come from the host; the entities the world already holds never leave the device: they are gathered into the
 * new submission order (ascending ids when the live ids are ascending, else the new rows go last) and re-imported there, so
 * that everything derived from the row order is what icx_bodies_upload of the merged arrays would have built.  An id that
 * exists already is left alone (:431-437), removing an id that is not there does nothing.  The coherency is kept.
 * Row indices of entities behind an inserted / removed row change: host-side row tables must be refreshed
 * (icx_entity_ids_fetch), and projectile rows (icx_projectiles_upload) are detached like after an upload.  Synchronous.
 */
ICX_API int icx_entities_create(IcxWorld* w, uint32_t count, const IcxEntity* entity, const IcxTransform* T, const IcxPhysics* P,
                                const uint8_t* has_physics, const IcxColliderComponent* C, const uint8_t* tstate, const uint16_t* tcoh,
                                uint32_t* n_created);
ICX_API int icx_entities_remove(IcxWorld* w, uint32_t count, const IcxEntity* entity, uint32_t* n_removed);
/* The ids of the world's entities in row order (n_out = their count; at most `cap` are written). */
ICX_API int icx_entity_ids_fetch(IcxWorld* w, IcxEntity* out, size_t cap, size_t* n_out);

/*
 * ParallelCosmosContext::extract_entity_updates (:171-423, entity loop :206-272): write every
 * temporal entity of the parallel world `par` back into `dst` under its id with the causal
 * chain-link decremented (:212), all components (:252); the entity's timestream state follows if
 * the decremented link is > 0 (:255) and is reset to merged when dst is timeslice 0 (:268-271).
 * Skipped like the reference: non-temporal ids (link 15, :213) and ids created during the parallel
 * run (timeslice 14, :223-229).  A link-0 id makes the reference throw (ecs_types.h:148):
 * ICX_E_INVALID.  Ids absent from dst are counted in *n_missing and left to the host (the
 * reference registers them, :241-245: entity lifecycle is host-side).
 */
ICX_API int icx_world_extract(IcxWorld* par, IcxWorld* dst, uint32_t* n_extracted, uint32_t* n_missing);
/* The same with flags.  ICX_EXTRACT_REGISTER: ids absent from dst are registered there first, from the parallel world's rows, as
 * the reference does (:241-245), and then extracted like the others (*n_missing = 0 afterwards unless dst is full:
 * ICX_E_CAPACITY).  dst's row indices change when something was registered (icx_entity_ids_fetch). */
#define ICX_EXTRACT_REGISTER 1u
ICX_API int icx_world_extract_ex(IcxWorld* par, IcxWorld* dst, uint32_t flags, uint32_t* n_extracted, uint32_t* n_missing);

/* Cosmos::get/set_coherency (cosmos.cpp:136-147) */
ICX_API int icx_coherency_set(IcxWorld* w, uint16_t coherency);
ICX_API int icx_coherency_get(IcxWorld* w, uint16_t* coherency);

/*
 * One tick = PhysicsDynamo::run_relays(dt) (physics_dynamo.cpp:38-46):
 * EulerPhysicsRelay::engage then CollisionPhysicsRelay::engage, followed by
 * Cosmos::increment_coherency (i_cosmos_context.cpp:64).  Asynchronous on the
 * world's stream; icx_sync() waits and reports deferred errors.
 */
ICX_API int icx_step(IcxWorld* w, double dt);
ICX_API int icx_step_n(IcxWorld* w, double dt, uint32_t n_ticks);
ICX_API int icx_sync(IcxWorld* w);

/* Individual relays, for per-stage parity tests (A_PhysicsRelay::engage). */
ICX_API int icx_euler_engage(IcxWorld* w, double dt);       /* euler_physics_relay.h:21-97     */
ICX_API int icx_collision_engage(IcxWorld* w);              /* collision_physics_relay.h:24-231 */

/* Results of the last tick.  *n receives the total available even if > cap. */
ICX_API int icx_pairs_fetch(IcxWorld* w, IcxPair* out, size_t cap, size_t* n);
ICX_API int icx_contacts_fetch(IcxWorld* w, IcxContact* out, size_t cap, size_t* n);
/* Broadphase, second part.  icx_pairs_fetch returns every pair whose bounds, inflated by the uniform margin, overlap on the
 * post-integration state.  A body that static resolution is predicted (or was caught) to move further than the margin gets a
 * larger per-body budget; the additional candidates this creates -- overlap(bound_i (+) budget_i, bound_j (+) budget_j) but not
 * in the uniform list -- are found by that body alone and returned here.  The ordered pass runs over the union. */
ICX_API int icx_extra_pairs_fetch(IcxWorld* w, IcxPair* out, size_t cap, size_t* n);
ICX_API int icx_budgets_fetch(IcxWorld* w, float* out, size_t cap);
ICX_API int icx_interceptions_fetch(IcxWorld* w, IcxInterception* out, size_t cap, size_t* n);
ICX_API int icx_grounds_fetch(IcxWorld* w, IcxGround* out, size_t cap, size_t* n);         /* ascending entity id */
ICX_API int icx_stats_fetch(IcxWorld* w, IcxStats* out);

/* Device-timed duration (ms) of the last icx_step/icx_step_n, CUDA events on
 * the world's stream. */
ICX_API int icx_last_step_ms(IcxWorld* w, float* ms);

/* Per-phase device timing of icx_step* (CUDA events on the world's stream, accumulated since
 * enabled): ms[ICX_N_PHASES] = {integrate+prep, grid sort, pair generation, speculative narrowphase,
 * ordered resolve}; n_launches = kernels launched by this world since creation.  Costs six extra
 * event records per tick. */
#define ICX_N_PHASES 5
ICX_API int icx_profile_enable(IcxWorld* w, int on);
ICX_API int icx_phase_ms(IcxWorld* w, double* ms, uint64_t* n_ticks, uint64_t* n_launches);
/* Sums over the ticks since the last icx_profile_enable call: out6 = {ticks, candidate pairs, contact records, dependency rounds,
 * replay attempts, dirty bodies} -- the per-tick quantities the roofline's algorithmic bytes are made of. */
ICX_API int icx_totals_fetch(IcxWorld* w, uint64_t* out6);

/* Device-side timeline of the last ordered resolve (pairs of {globaltimer ns, tag}); tags:
 * 0 start, 0x1nnnn lists built / bodies reset (nnnn = dirty bodies), 0x2.... barrier passed,
 * 0x3nnnn round finished (nnnn = bodies active in it), 0x4.... end.  enable != 0 turns recording on. */
ICX_API int icx_debug_trace(IcxWorld* w, int enable, uint64_t* out, size_t cap, size_t* n);

/* Raw device counters of the last tick (32 words; diagnostic, layout not stable). */
ICX_API int icx_debug_counters(IcxWorld* w, uint32_t* out32);

/* Per-entity diagnostic arrays of the last tick: which = 0 predicted push, 1 displacement budget, 2 largest component of the
 * displacement static resolution produced (layout not stable). */
ICX_API int icx_debug_array(IcxWorld* w, int which, float* out, size_t cap);

/* CUDA stream of the world as a void* (cudaStream_t), for callers that time it. */
ICX_API void* icx_world_stream(IcxWorld* w);

/* ==== behaviours on the device ======================================
 * ProjectileBehaviors::on_fixed_update (source/behaviors/projectile_behaviors.h:17-54) is the
 * per-tick host work between replay and physics for every projectile: it advances the lifetime,
 * pulses TransformComponent::scale and condemns expired entities.  Run on the device it needs no
 * per-tick icx_bodies_update of the scales.  IcxProjectile = ProjectileComponent
 * (source/behaviors/projectile_component.h:8-16, 40 B, doubles). */
typedef struct IcxProjectile {
    double currLifetime;
    double maxLifetime;
    double hz;
    double scaleMax;
    double scaleMin;
} IcxProjectile;

/* Projectile i belongs to the entity at upload position index[i] (each entity at most once).
 * Replaces any earlier set; n = 0 drops it.  Call again after icx_bodies_upload. */
ICX_API int icx_projectiles_upload(IcxWorld* w, uint32_t n, const uint32_t* index, const IcxProjectile* proj);
/* on_fixed_update(dt) for every uploaded projectile, asynchronously on the world's stream:
 *   currLifetime += dt;  scale = vec3((float)(sin(currLifetime * hz) * (scaleMax - scaleMin) / 2 + (scaleMin + 1)));
 *   currLifetime > maxLifetime -> condemned (Cosmos::condemn_entity, cosmos.h:339-345).
 * The scale lands where the next tick and icx_bodies_download read it.  Condemned projectiles keep
 * being reported until the caller removes them (entity lifecycle stays with the host ECS). */
ICX_API int icx_projectiles_fixed_update(IcxWorld* w, double dt);
/* Component state back to the host, projectile order of the upload. */
ICX_API int icx_projectiles_download(IcxWorld* w, uint32_t n, IcxProjectile* out);
/* Projectile numbers (positions in the uploaded set, ascending) condemned by the LAST update. */
ICX_API int icx_projectiles_condemned(IcxWorld* w, uint32_t* out, size_t cap, size_t* n);
/* Device-timed duration (ms) of the last icx_projectiles_fixed_update kernel (CUDA events). */
ICX_API int icx_projectiles_kernel_ms(IcxWorld* w, float* ms);

/* ==== world-line record / replay / divergence scan ========================================
 * The dense ENTITY_UPDATE stream between adjacent timeslices (EntityTimestreamMap,
 * source/networking/entity_timestream_map.h:43-230) as a device-resident ring of SoA snapshots:
 * slot = coherency % depth_ticks.  A record carries what the reference's does (ComponentCategory::all,
 * server_network_dynamo.cpp:666-677; layout component_registry.h:238-279): per entity the whole
 * TransformComponent and PhysicsComponent and the id the receiving cosmos knows it by (causal chain-link + 1,
 * :673); ColliderComponents travel as a sparse list of the rows that changed since the previous record into
 * the ring (every record of the reference repeats them unchanged).  Records are addressed by ID, as the reference's
 * per-entity streams are: the receiving cosmos may hold other entities than the recording one (a time traveller has no
 * future stream, a parallel cosmos omits it); an entity finds its record by its id -- at its own row when both hold the
 * same entities, else by binary search, which needs ids in ascending submission order (std::set<Entity>, ecs/i_synchro.h:48)
 * -- and an entity without a record is left alone and counted (icx_timeline_id_mismatches).
 * create / remove / jump events stay on the host TimelineApi queues. */
#define ICX_REPLAY_SERVER   0   /* ServerNetworkDynamo::run_relays phase 2, server_network_dynamo.cpp:377-421  */
#define ICX_REPLAY_PARALLEL 1   /* ParallelNetworkDynamo::run_relays,     parallel_network_dynamo.cpp:51-89 */

ICX_API int  icx_timeline_create(int device, uint32_t depth_ticks, uint32_t n_max, IcxTimeline** out);
ICX_API void icx_timeline_destroy(IcxTimeline* tl);
ICX_API const char* icx_timeline_last_error(const IcxTimeline* tl);

/* Measurement hook (no reference counterpart): CUDA events around the record / replay / divergence-scan kernel launches on
 * the world's stream.  which: 0 = k_record, 1 = k_replay, 2 = k_divergence_scan; *ms = duration of the LAST such launch
 * (waits for it).  ICX_E_STATE if that kernel has not run since profiling was switched on. */
ICX_API int icx_timeline_profile(IcxTimeline* tl, int enable);
ICX_API int icx_timeline_kernel_ms(IcxTimeline* tl, int which, float* ms);

/* Record = serialise every entity at `coherency` and push it.  Asynchronous on the world's stream.
 *   icx_record                TimelineApi::push_past_timestream, server_network_dynamo.cpp:642-679: the id travels with
 *                             its causal chain-link incremented (:673, refused at link >= 14, ecs_types.h:118-136)
 *   icx_record_at_breakpoint  push_timestream_at_breakpoint, parallel_network_dynamo.cpp:369-385: rewrites the slot in
 *                             front of the breakpoint, ids unchanged
 *   icx_record_ex             both spelled out (ICX_RECORD_*) */
#define ICX_RECORD_AT_BREAKPOINT 0x1u
#define ICX_RECORD_KEEP_LINK     0x2u   /* do not increment the chain-link (the record goes back to the cosmos it came from) */
ICX_API int icx_record(IcxWorld* w, IcxTimeline* tl, uint16_t coherency);
ICX_API int icx_record_at_breakpoint(IcxWorld* w, IcxTimeline* tl, uint16_t coherency);
ICX_API int icx_record_ex(IcxWorld* w, IcxTimeline* tl, uint16_t coherency, uint32_t flags);

/* Breakpoint of the timestreams (TsBreakpointQueue::set_breakpoint_at_begin / remove_breakpoint,
 * source/networking/ts_breakpoint_queue.h:107-121, set by link_timestreams when a parallel cosmos is
 * linked): records present when it is set, and later plain records, lie BEHIND it -- ICX_REPLAY_SERVER
 * cannot pop them, ICX_REPLAY_PARALLEL pops only them; icx_record_at_breakpoint inserts in front of it. */
ICX_API int icx_breakpoint_set(IcxTimeline* tl);
ICX_API int icx_breakpoint_clear(IcxTimeline* tl);

/* Replay = pop every record with coherency <= `coherency` (cyclic compare, events/message.h:179-183), drop those the
 * reference drops (`record.coherency < now`, a plain integer compare: server_network_dynamo.cpp:391) and apply the rest
 * oldest first: ICX_REPLAY_SERVER overwrites transform + physics (+ changed colliders) of every entity that is not
 * forking / forked; ICX_REPLAY_PARALLEL only consumes (a parallel cosmos takes upstream components only, of which the
 * physics path has none).  *applied = 1 if a record was applied.  Asynchronous on the world's stream. */
ICX_API int icx_replay(IcxWorld* w, IcxTimeline* tl, uint16_t coherency, int mode, int* applied);

/* Entities replays from this ring have left alone so far because the record held nothing addressed to their id, plus collider
 * rows whose packet structure was not the receiving entity's.  (A recorded row whose entity does not exist in the receiving
 * cosmos is a stream nobody pops: not counted.)  Synchronises the device. */
ICX_API int icx_timeline_id_mismatches(IcxTimeline* tl, uint64_t* total);

/* Divergence scan: live origins against the record stamped `coherency`, with the reference's
 * congruence predicate length2(a - b) <= 2.0f (TimejumpConditions::operator==,
 * source/spacetime/timejump_conditions.h:21-25).  Per world-line k (cfg.worldline_stride, or one):
 * flags[k] = any body incongruent, incongruent[k] = how many, score[k] = sum of min(length2, 2^20)
 * accumulated in 48.16 fixed point (the score is this build's definition, the flag is the
 * reference's).  The record may cover the whole world or ONE world-line (a base history shared by
 * all candidates).  icx_divergence_scan is synchronous; _enqueue only launches (results stay on the device, and a
 * per-world-line "flagged by some scan since the last reset" byte is kept), _fetch waits and reads the LAST scan's
 * results plus those bytes (any pointer may be NULL) -- a per-tick scan then costs no host round trip. */
ICX_API int icx_divergence_scan(IcxWorld* w, IcxTimeline* tl, uint16_t coherency, uint32_t n_worldlines,
                        uint8_t* flags, uint32_t* incongruent, double* score);
ICX_API int icx_divergence_scan_enqueue(IcxWorld* w, IcxTimeline* tl, uint16_t coherency, uint32_t n_worldlines);
ICX_API int icx_divergence_scan_fetch(IcxTimeline* tl, uint32_t n_worldlines, uint8_t* flags, uint8_t* flags_ever,
                              uint32_t* incongruent, double* score, int reset_ever);

/* icx_divergence_scan without any allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same arguments, same results as icx_divergence_scan (synchronous; flags / incongruent / score may each be NULL), but the
 * per-world-line accumulators live in a file-scope `static __device__` array of ICX_SCAN_STATIC_MAX_WORLDLINES entries
 * (96 KB, held by every GPU that loads the library until exit) instead of the timeline's lazily grown buffers: the call
 * performs no cudaMalloc / cudaFree and no host allocation.  More world-lines than the maximum: ICX_E_CAPACITY.  The array is
 * shared by every call on that GPU and cleared at the start of each: calls that overlap on one GPU (two worlds on different
 * streams, two host threads) race -- serialise them.  It does not touch the "ever flagged" bytes of _enqueue / _fetch. */
#define ICX_SCAN_STATIC_MAX_WORLDLINES 8192
ICX_API int icx_divergence_scan_static(IcxWorld* w, IcxTimeline* tl, uint16_t coherency, uint32_t n_worldlines,
                                       uint8_t* flags, uint32_t* incongruent, double* score);

/* Fork ageing (server_network_dynamo.cpp:13-15, :681-705): forking -> forked after 13 ticks;
 * *resolution_needed = some entity has been forked for >= 13 ticks (parallel_notify_divergence).
 * With both out-pointers NULL the call only enqueues the promotion kernel (no host round trip). */
ICX_API int icx_fork_age(IcxWorld* w, uint32_t* n_promoted, int* resolution_needed);

/*
 * Event records in the timeline: what a timeslice pushes into its past besides ENTITY_UPDATE -- ENTITY_CREATED / ENTITY_REMOVED
 * (_entity_created_handler / _entity_removed_handler, server_network_dynamo.cpp:748-811) and JUMP_REQUEST / JUMP_DEPARTURE /
 * JUMP_ARRIVAL (:878-976, :466-498).  A record is stamped with the coherency it is pushed at, lives in the ring slot of that
 * tick ON THE DEVICE next to the tick's snapshot, travels inside the same transport buffer (slot_export / slot_import: one
 * ncclSend per tick carries states, collider rows and events), and comes out of icx_timeline_events_pop in the order the
 * reference's per-entity streams hand messages out: every record with coherency <= now (cyclic compare), oldest first, those
 * with coherency < now (plain compare, :391) dropped.  `entity` / `source` are what the RECEIVING cosmos knows them by (the
 * sender increments the causal chain-link, :767-768).  payload is opaque (JUMP_params: tripId, timesliceDelta, isFocal, ...).
 * The handlers stay with the host (they call Cosmos::register_entity / condemn_entity = icx_entities_create / _remove).
 */
#define ICX_EVENT_ENTITY_CREATED 1
#define ICX_EVENT_ENTITY_REMOVED 2
#define ICX_EVENT_JUMP_REQUEST 3
#define ICX_EVENT_JUMP_DEPARTURE 4
#define ICX_EVENT_JUMP_ARRIVAL 5
typedef struct IcxTimelineEvent {
    uint32_t type;
    IcxEntity entity;
    IcxEntity source;
    uint16_t coherency;     /* written by the push */
    uint16_t reserved;
    uint32_t payload[4];
} IcxTimelineEvent;         /* 32 bytes */
/* Push `count` records stamped `coherency` (at most ICX_TIMELINE_EVENTS_PER_TICK per tick: ICX_E_CAPACITY beyond).  May be
 * called before or after the icx_record of the same tick. */
ICX_API int icx_timeline_event_push(IcxTimeline* tl, uint16_t coherency, const IcxTimelineEvent* events, uint32_t count);
/* Pop the records that are due at `now` (see above) WITHOUT consuming the ticks' snapshots: call it before icx_replay of
 * the same tick (ENTITY_CREATED must run before the ENTITY_UPDATE that fills the new entity's components).  At most `cap`
 * records are written, *n_out = how many were due.  mode as icx_replay (ICX_REPLAY_PARALLEL: behind the breakpoint). */
ICX_API int icx_timeline_events_pop(IcxTimeline* tl, uint16_t now, int mode, IcxTimelineEvent* out, size_t cap, size_t* n_out);

/* The components of ONE entity as the record of tick `coherency` holds them -- what the ENTITY_CREATED / JUMP_ARRIVAL
 * handlers deserialise out of their message in the reference (server_network_dynamo.cpp:940-976: the arrival carries the
 * traveller's serialised components; here the tick's ENTITY_UPDATE record already does).  id = the id the RECEIVING cosmos
 * knows the entity by.  *found: bit 0 = Transform / Physics found, bit 1 = the record also carries its ColliderComponent (it
 * does in the tick an entity is created: the row is new to the ring).  Call it after icx_timeline_events_pop and before the
 * icx_replay that consumes the tick; then icx_entities_create.  Synchronous, one small copy. */
ICX_API int icx_timeline_entity_fetch(IcxTimeline* tl, uint16_t coherency, IcxEntity id, IcxTransform* T, IcxPhysics* P,
                                      uint8_t* has_physics, IcxColliderComponent* C, int* found);

/* Snapshot transport between timeslices on different GPUs / processes (the future -> past ENTITY_UPDATE
 * stream, server_network_dynamo.cpp:642-679): pack the slot stamped `coherency` into / restore it from a
 * caller-owned DEVICE buffer, e.g. the buffer a ncclSend / ncclRecv moves:
 *   [10][n] float4 state | [n] uint32 id (padded to 16 B) | header {collider rows, events, 0, 0} + 32 event records of 32 B |
 *   rows x 176 B {sender row, id, ColliderComponent}
 * ICX_SNAPSHOT_CAPACITY(n, rows) is the buffer size for n entities and up to `rows` changed colliders.  `stream` is a
 * cudaStream_t (NULL = default stream); both calls are ASYNCHRONOUS on it and ordered against the recording / replaying
 * worlds' streams by events, so a timeslice needs no host synchronisation per tick. */
#define ICX_SNAPSHOT_BYTES 164
#define ICX_TIMELINE_EVENTS_PER_TICK 32
#define ICX_SNAPSHOT_HEADER_BYTES (16 + 32 * ICX_TIMELINE_EVENTS_PER_TICK)
#define ICX_SNAPSHOT_DELTA_BYTES 176
#define ICX_SNAPSHOT_DENSE_BYTES(n) ((size_t)(n) * 160 + (((size_t)(n) * 4 + 15) & ~(size_t)15))     /* the id array is padded to 16 bytes */
#define ICX_SNAPSHOT_CAPACITY(n, rows) (ICX_SNAPSHOT_DENSE_BYTES(n) + ICX_SNAPSHOT_HEADER_BYTES + (size_t)(rows) * ICX_SNAPSHOT_DELTA_BYTES)
ICX_API int icx_timeline_slot_export(IcxTimeline* tl, uint16_t coherency, void* device_buffer, size_t capacity_bytes,
                                     uint32_t* n_entities, void* stream);
ICX_API int icx_timeline_slot_import(IcxTimeline* tl, uint16_t coherency, uint32_t n_entities, const void* device_buffer,
                                     size_t buffer_bytes, void* stream);

ICX_API int icx_timeline_slot_info(const IcxTimeline* tl, uint16_t coherency, int* valid, uint32_t* n_recorded);

/* ==== wire / record format ==========================================
 * The reference's ENTITY_UPDATE message for every entity of the world, byte for byte as ServerNetworkDynamo's record phase
 * builds it for the child timestream (server_network_dynamo.cpp:666-677): MessageHeader<EventId> {id, coherency, size}
 * (events/message.h:22-28, 12 B), then the body stack (message.h:57-83) -- the components in DESCENDING component-type
 * order (ComponentRegistry::serialize_entity_components, ecs/component_registry.h:238-253), then ENTITY_UPDATE_params
 * {Entity, Signature, ComponentCategory::all} (events/event_types.h:102-111; 24 B under libstdc++, where std::bitset<32> is
 * one unsigned long) -- so that an un-accelerated timeslice or client can push it through its own
 * EventMessage >> params / deserialize_entity_components.  Only the physics segments exist on the device
 * (TransformComponent, PhysicsComponent, ColliderComponent); their ComponentType indices and the event id come from the
 * caller (hard_config_cosmos.h:47-58 registers them as 0, 6, 7; ENTITY_UPDATE is 102).  The entity id is the reference's
 * 16-bit one: ids whose genesis does not fit 8 bits make the call fail with ICX_E_INVALID.
 * Messages are written at a fixed stride (*stride_bytes = ICX_WIRE_MAX_BYTES); message k is entity k of the upload order,
 * its length is 12 + header.size. */
typedef struct IcxWireLayout {
    uint32_t event_id;              /* events::cosmos::ENTITY_UPDATE                              */
    uint8_t  type_transform, type_physics, type_collider;   /* ComponentType of each registered component */
    uint8_t  increment_link;        /* 1: the id travels with chain-link + 1 (push_past_timestream, :673) */
} IcxWireLayout;
#define ICX_WIRE_HEADER_BYTES 12
#define ICX_WIRE_PARAMS_BYTES 24
#define ICX_WIRE_MAX_BYTES (ICX_WIRE_HEADER_BYTES + 168 + 108 + 40 + ICX_WIRE_PARAMS_BYTES)
ICX_API void icx_wire_layout_default(IcxWireLayout* layout);
ICX_API int icx_entity_update_messages(IcxWorld* w, const IcxWireLayout* layout, uint16_t coherency,
                                       void* out, size_t capacity_bytes, size_t* stride_bytes, uint32_t* n_messages);

#ifdef __cplusplus
}
#endif
#endif /* ICX_H */
