"""ctypes binding of the C ABI in include/icx.h plus numpy views of its AoS structs.

The structs are byte-identical to the reference components
(source/physics/transform_component.h:14, physics_component.h:24, collider.h:42,
collider_component.h:11), so a numpy array of these dtypes can be handed to the
library -- or to the reference compiled in oracle/_ref -- without conversion.

There is no CPU fallback: load() raises if the CUDA library was not built.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("ICX_LIB", os.path.join(_HERE, "libicx.so"))      # ICX_LIB: tuning variants (tools/), never a fallback

# ---- numpy dtypes -----------------------------------------------------------
transform_dt = np.dtype([("origin", "<f4", 3), ("orientation", "<f4", 4), ("scale", "<f4", 3)])
assert transform_dt.itemsize == 40
# ProjectileComponent, source/behaviors/projectile_component.h:8-16
projectile_dt = np.dtype([("currLifetime", "<f8"), ("maxLifetime", "<f8"), ("hz", "<f8"), ("scaleMax", "<f8"), ("scaleMin", "<f8")])
assert projectile_dt.itemsize == 40

physics_dt = np.dtype({
    "names": ["velocity", "acceleration", "angularVelocity", "angularAcceleration",
              "linearDrag", "angularDrag", "collisionLinearDrag", "collisionAngularDrag", "mass",
              "lockOrigin", "lockedOrigin", "lockOrientation", "lockedOrientation", "isAsleep"],
    "formats": [("<f4", 3), ("<f4", 3), ("<f4", 3), ("<f4", 3),
                "<f4", "<f4", "<f4", "<f4", "<f4",
                "u1", ("<f4", 3), "u1", ("<f4", 4), "u1"],
    "offsets": [0, 12, 24, 36, 48, 52, 56, 60, 64, 68, 72, 84, 88, 104],
    "itemsize": 108})

collider_dt = np.dtype({
    "names": ["colliderType", "collisionType", "isActive", "useBehaviorsResponse", "localTransform",
              "inheritOrientation", "influenceOrientation", "minParametricValue", "staticFriction",
              "dynamicFriction", "restitution", "stiffness", "damping", "restLength"],
    "formats": ["<i4", "<i4", "u1", "u1", transform_dt, "u1", "u1",
                "<f4", "<f4", "<f4", "<f4", "<f4", "<f4", "<f4"],
    "offsets": [0, 4, 8, 9, 12, 52, 53, 56, 60, 64, 68, 72, 76, 80],
    "itemsize": 84})

collider_component_dt = np.dtype([("colliders", collider_dt, 2)])
assert collider_component_dt.itemsize == 168

motion_dt = np.dtype([("origin", "<f4", 3), ("orientation", "<f4", 4), ("velocity", "<f4", 3), ("acceleration", "<f4", 3),
                      ("angularVelocity", "<f4", 3), ("angularAcceleration", "<f4", 3)])
assert motion_dt.itemsize == 76
contact_dt = np.dtype([("entityA", "<u4"), ("entityB", "<u4"), ("slotA", "u1"), ("slotB", "u1"),
                       ("flags", "<u2"), ("point", "<f4", 3), ("normal", "<f4", 3), ("depth", "<f4"),
                       ("packetA", "<u4"), ("packetB", "<u4")])
assert contact_dt.itemsize == 48
ground_dt = np.dtype([("entity", "<u4"), ("normal", "<f4", 3), ("dist", "<f4"), ("packetA", "<u4"), ("packetB", "<u4"), ("side", "<u4")])
assert ground_dt.itemsize == 32
interception_dt = np.dtype([("agent", "<u4"), ("recipient", "<u4")])
pair_dt = np.dtype([("a", "<u4"), ("b", "<u4")])
mesh_contact_dt = np.dtype([("entity", "<u4"), ("row", "<u4"), ("slot", "<u4"), ("triangle", "<u4"), ("point", "<f4", 3), ("normal", "<f4", 3), ("depth", "<f4")])
assert mesh_contact_dt.itemsize == 44
mesh_material_dt = np.dtype([("staticFriction", "<f4"), ("dynamicFriction", "<f4"), ("restitution", "<f4")])
event_dt = np.dtype([("type", "<u4"), ("entity", "<u4"), ("source", "<u4"), ("coherency", "<u2"), ("reserved", "<u2"), ("payload", "<u4", 4)])
assert event_dt.itemsize == 32
EVENT_ENTITY_CREATED, EVENT_ENTITY_REMOVED, EVENT_JUMP_REQUEST, EVENT_JUMP_DEPARTURE, EVENT_JUMP_ARRIVAL = 1, 2, 3, 4, 5

COLLIDER_NONE, COLLIDER_BOX, COLLIDER_SPHERE, COLLIDER_RAY = 0, 1, 2, 3
COLLISION_NOOP, COLLISION_RIGID, COLLISION_SPRING, COLLISION_FORCE = 0, 1, 2, 3
TS_MERGED, TS_FORKED, TS_FORKING, TS_SUPERPOSITION, TS_GHOST = 0, 1, 2, 3, 4
NULL_TIMESLICE, NULL_CHAINLINK, NULL_ENTITY = 14, 15, 0xFFFFFFFF
F_SPHERE_SPHERE, F_NO_GRAVITY, F_LANE_ROUNDS = 1, 2, 4
CONTACT_RESPONDED, CONTACT_BEHAVIORS_A, CONTACT_BEHAVIORS_B = 1, 2, 4
OK, E_INVALID, E_CUDA, E_CAPACITY, E_NOMEM, E_STATE = 0, 1, 2, 3, 4, 5
REPLAY_SERVER, REPLAY_PARALLEL = 0, 1
TIMELINE_EVENTS_PER_TICK = 32
SNAPSHOT_BYTES, SNAPSHOT_HEADER_BYTES, SNAPSHOT_DELTA_BYTES = 164, 16 + 32 * TIMELINE_EVENTS_PER_TICK, 176
RECORD_AT_BREAKPOINT, RECORD_KEEP_LINK = 1, 2


def snapshot_capacity(n, rows=0):
    """ICX_SNAPSHOT_CAPACITY: transport buffer for n entities and up to `rows` changed ColliderComponents."""
    return n * 160 + ((n * 4 + 15) & ~15) + SNAPSHOT_HEADER_BYTES + rows * SNAPSHOT_DELTA_BYTES


def compose_entity(timeslice, genesis, link):
    """compose_entity (source/ecs/ecs_types.h:112-120) on the widened id."""
    return (int(timeslice) << 28) | (int(genesis) << 4) | int(link)


def widen_entity(e16):
    """Reference uint16 id -> IcxEntity (include/icx.h)."""
    e16 = int(e16)
    return ((e16 & 0xF000) << 16) | (e16 & 0x0FFF)


def default_transforms(n):
    t = np.zeros(n, transform_dt)
    t["orientation"][:, 0] = 1.0
    t["scale"] = 1.0
    return t


def default_physics(n):
    """PhysicsComponent defaults, physics_component.h:26-54."""
    p = np.zeros(n, physics_dt)
    p["linearDrag"] = 0.01
    p["angularDrag"] = 0.01
    p["collisionAngularDrag"] = 0.03
    p["mass"] = 1.0
    p["lockedOrientation"][:, 0] = 1.0
    return p


def default_colliders(n):
    """ColliderComponent of default-constructed Colliders, collider.h:58-92 (inactive boxes)."""
    c = np.zeros(n, collider_component_dt)
    cc = c["colliders"]
    cc["colliderType"] = COLLIDER_BOX
    cc["collisionType"] = COLLISION_RIGID
    cc["localTransform"]["orientation"][..., 0] = 1.0
    cc["localTransform"]["scale"] = 1.0
    cc["inheritOrientation"] = 1
    cc["influenceOrientation"] = 1
    cc["minParametricValue"] = 1.0
    cc["staticFriction"] = 0.45
    cc["dynamicFriction"] = 0.45
    cc["restitution"] = 0.70
    cc["stiffness"] = 100.0
    cc["damping"] = 10.0
    cc["restLength"] = 0.1
    return c


# ---- ctypes structs ---------------------------------------------------------
class IcxConfig(C.Structure):
    _fields_ = [("struct_size", C.c_uint32), ("device", C.c_int32), ("max_entities", C.c_uint32),
                ("host_timeslice", C.c_uint32), ("flags", C.c_uint32), ("gravity", C.c_float * 3),
                ("broadphase_margin", C.c_float), ("cell_size", C.c_float),
                ("max_neighbors", C.c_uint32), ("max_large", C.c_uint32), ("max_contacts", C.c_uint32),
                ("worldline_stride", C.c_uint32), ("max_pairs", C.c_uint32)]


class IcxWireLayout(C.Structure):
    _fields_ = [("event_id", C.c_uint32), ("type_transform", C.c_uint8), ("type_physics", C.c_uint8), ("type_collider", C.c_uint8),
                ("increment_link", C.c_uint8)]


WIRE_MAX_BYTES = 12 + 168 + 108 + 40 + 24


class IcxStats(C.Structure):
    _fields_ = [("ticks", C.c_uint64), ("n_entities", C.c_uint32), ("n_packets", C.c_uint32),
                ("n_large", C.c_uint32), ("n_candidates", C.c_uint32), ("n_contacts", C.c_uint32),
                ("n_interceptions", C.c_uint32), ("n_rounds", C.c_uint32), ("n_retries", C.c_uint32),
                ("margin", C.c_float), ("cell_size", C.c_float), ("max_neighbors", C.c_uint32),
                ("n_dirty", C.c_uint32), ("n_attempts", C.c_uint32), ("n_movers", C.c_uint32)]


_vp, _u32, _sz = C.c_void_p, C.c_uint32, C.c_size_t
_psz = C.POINTER(C.c_size_t)

# name -> (restype, argtypes); every symbol include/icx.h declares
SIGNATURES = {
    "icx_abi_version": (C.c_int, []),
    "icx_device_count": (C.c_int, []),
    "icx_status_string": (C.c_char_p, [C.c_int]),
    "icx_config_default": (None, [C.POINTER(IcxConfig)]),
    "icx_world_create": (C.c_int, [C.POINTER(IcxConfig), C.POINTER(_vp)]),
    "icx_world_destroy": (None, [_vp]),
    "icx_last_error": (C.c_char_p, [_vp]),
    "icx_bodies_upload": (C.c_int, [_vp, _u32, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "icx_bodies_download": (C.c_int, [_vp, _u32, _vp, _vp, _vp, _vp, _vp]),
    "icx_bodies_update": (C.c_int, [_vp, _u32, _vp, _vp, _vp, _vp]),
    "icx_motion_download": (C.c_int, [_vp, _u32, _vp]),
    "icx_coherency_set": (C.c_int, [_vp, C.c_uint16]),
    "icx_coherency_get": (C.c_int, [_vp, C.POINTER(C.c_uint16)]),
    "icx_step": (C.c_int, [_vp, C.c_double]),
    "icx_step_n": (C.c_int, [_vp, C.c_double, _u32]),
    "icx_sync": (C.c_int, [_vp]),
    "icx_euler_engage": (C.c_int, [_vp, C.c_double]),
    "icx_collision_engage": (C.c_int, [_vp]),
    "icx_pairs_fetch": (C.c_int, [_vp, _vp, _sz, _psz]),
    "icx_contacts_fetch": (C.c_int, [_vp, _vp, _sz, _psz]),
    "icx_extra_pairs_fetch": (C.c_int, [_vp, _vp, _sz, _psz]),
    "icx_budgets_fetch": (C.c_int, [_vp, _vp, _sz]),
    "icx_interceptions_fetch": (C.c_int, [_vp, _vp, _sz, _psz]),
    "icx_grounds_fetch": (C.c_int, [_vp, _vp, _sz, _psz]),
    "icx_stats_fetch": (C.c_int, [_vp, C.POINTER(IcxStats)]),
    "icx_last_step_ms": (C.c_int, [_vp, C.POINTER(C.c_float)]),
    "icx_world_stream": (_vp, [_vp]),
    "icx_profile_enable": (C.c_int, [_vp, C.c_int]),
    "icx_totals_fetch": (C.c_int, [_vp, C.POINTER(C.c_uint64)]),
    "icx_timeline_create": (C.c_int, [C.c_int, _u32, _u32, C.POINTER(_vp)]),
    "icx_timeline_destroy": (None, [_vp]),
    "icx_timeline_last_error": (C.c_char_p, [_vp]),
    "icx_timeline_profile": (C.c_int, [_vp, C.c_int]),
    "icx_timeline_kernel_ms": (C.c_int, [_vp, C.c_int, C.POINTER(C.c_float)]),
    "icx_record": (C.c_int, [_vp, _vp, C.c_uint16]),
    "icx_record_at_breakpoint": (C.c_int, [_vp, _vp, C.c_uint16]),
    "icx_record_ex": (C.c_int, [_vp, _vp, C.c_uint16, _u32]),
    "icx_timeline_id_mismatches": (C.c_int, [_vp, C.POINTER(C.c_uint64)]),
    "icx_timeline_event_push": (C.c_int, [_vp, C.c_uint16, _vp, _u32]),
    "icx_timeline_entity_fetch": (C.c_int, [_vp, C.c_uint16, _u32, _vp, _vp, _vp, _vp, C.POINTER(C.c_int)]),
    "icx_timeline_events_pop": (C.c_int, [_vp, C.c_uint16, C.c_int, _vp, _sz, _psz]),
    "icx_divergence_scan_enqueue": (C.c_int, [_vp, _vp, C.c_uint16, _u32]),
    "icx_divergence_scan_fetch": (C.c_int, [_vp, _u32, _vp, _vp, _vp, _vp, C.c_int]),
    "icx_breakpoint_set": (C.c_int, [_vp]),
    "icx_breakpoint_clear": (C.c_int, [_vp]),
    "icx_replay": (C.c_int, [_vp, _vp, C.c_uint16, C.c_int, C.POINTER(C.c_int)]),
    "icx_divergence_scan": (C.c_int, [_vp, _vp, C.c_uint16, _u32, _vp, _vp, _vp]),
    "icx_divergence_scan_static": (C.c_int, [_vp, _vp, C.c_uint16, _u32, _vp, _vp, _vp]),
    "icx_fork_age": (C.c_int, [_vp, C.POINTER(_u32), C.POINTER(C.c_int)]),
    "icx_timeline_slot_export": (C.c_int, [_vp, C.c_uint16, _vp, _sz, C.POINTER(_u32), _vp]),
    "icx_timeline_slot_import": (C.c_int, [_vp, C.c_uint16, _u32, _vp, _sz, _vp]),
    "icx_world_clone": (C.c_int, [_vp, _vp, _vp, C.POINTER(_u32)]),
    "icx_world_extract": (C.c_int, [_vp, _vp, C.POINTER(_u32), C.POINTER(_u32)]),
    "icx_world_extract_ex": (C.c_int, [_vp, _vp, _u32, C.POINTER(_u32), C.POINTER(_u32)]),
    "icx_mesh_upload": (C.c_int, [_vp, _vp, _u32, _vp, _u32, _vp]),
    "icx_mesh_contacts_fetch": (C.c_int, [_vp, _vp, _sz, _psz]),
    "icx_mesh_kernel_us": (C.c_int, [_vp, C.POINTER(C.c_float)]),
    "icx_device_share_set": (C.c_int, [_vp, _u32]),
    "icx_host_alloc": (C.c_void_p, [_sz]),
    "icx_host_free": (None, [_vp]),
    "icx_entities_create": (C.c_int, [_vp, _u32, _vp, _vp, _vp, _vp, _vp, _vp, _vp, C.POINTER(_u32)]),
    "icx_entities_remove": (C.c_int, [_vp, _u32, _vp, C.POINTER(_u32)]),
    "icx_entity_ids_fetch": (C.c_int, [_vp, _vp, C.c_size_t, C.POINTER(C.c_size_t)]),
    "icx_timeline_slot_info": (C.c_int, [_vp, C.c_uint16, C.POINTER(C.c_int), C.POINTER(_u32)]),
    "icx_debug_counters": (C.c_int, [_vp, _vp]),
    "icx_debug_array": (C.c_int, [_vp, C.c_int, _vp, _sz]),
    "icx_debug_trace": (C.c_int, [_vp, C.c_int, _vp, _sz, _psz]),
    "icx_phase_ms": (C.c_int, [_vp, C.POINTER(C.c_double), C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]),
    "icx_projectiles_upload": (C.c_int, [_vp, _u32, _vp, _vp]),
    "icx_projectiles_fixed_update": (C.c_int, [_vp, C.c_double]),
    "icx_projectiles_download": (C.c_int, [_vp, _u32, _vp]),
    "icx_projectiles_condemned": (C.c_int, [_vp, _vp, _sz, _psz]),
    "icx_projectiles_kernel_ms": (C.c_int, [_vp, C.POINTER(C.c_float)]),
    "icx_wire_layout_default": (None, [_vp]),
    "icx_entity_update_messages": (C.c_int, [_vp, _vp, C.c_uint16, _vp, _sz, _psz, C.POINTER(_u32)]),
    # include/icx_nccl.h
    "icx_nccl_available": (C.c_int, []),
    "icx_scores_allgather": (C.c_int, [_vp, _u32, _u32, C.c_int, _vp, _vp, _vp, _vp]),
    "icx_timeline_slot_send": (C.c_int, [_vp, C.c_uint16, _vp, C.c_int, _vp, _sz, _vp]),
    "icx_timeline_slot_recv": (C.c_int, [_vp, C.c_uint16, _u32, _vp, C.c_int, _vp, _sz, _vp]),
    "icx_nccl_group_start": (C.c_int, []),
    "icx_nccl_group_end": (C.c_int, []),
}

_lib = None


def load(path=None):
    """dlopen libicx.so and type every entry point.  Raises if it is missing."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    p = path or LIB_PATH
    if not os.path.exists(p):
        raise RuntimeError(
            f"intercession_b200: {p} not built -- run `python -c 'import __graft_entry__ as g; g.build()'`. "
            "There is no CPU fallback.")
    lib = C.CDLL(p, mode=C.RTLD_GLOBAL)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)     # AttributeError if the library lacks a declared symbol
        fn.restype = res
        fn.argtypes = args
    if path is None:
        _lib = lib
    return lib


def ptr(a):
    """void* of a C-contiguous numpy array (None -> NULL)."""
    if a is None:
        return None
    assert a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(C.c_void_p)
