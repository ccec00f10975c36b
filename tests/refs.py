"""Loaders for the CHECKERS (test infrastructure): oracle/_ref/libicx_ref.so = the reference's
own sources compiled headlessly, oracle/libicx_oracle.so = the CPU restatement."""
import ctypes as C
import os

import numpy as np

from intercession_b200 import abi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_PATH = os.path.join(ROOT, "oracle", "_ref", "libicx_ref.so")
ORACLE_PATH = os.path.join(ROOT, "oracle", "libicx_oracle.so")

_vp, _u32, _sz, _dbl = C.c_void_p, C.c_uint32, C.c_size_t, C.c_double
_psz = C.POINTER(C.c_size_t)
_pd = C.POINTER(C.c_double)


def _collide_sig():
    return (C.c_int, [_u32, _vp, _vp, _vp, _vp, _vp, _vp, _vp, C.c_uint16, _u32, _u32, _u32,
                      _vp, _sz, _vp, _sz, _psz, _vp, _sz, _psz, _pd])


REF_SIGS = {
    "icxref_verbatim_run": (C.c_int, [_u32, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, C.POINTER(C.c_uint16),
                                      _u32, _dbl, _u32, _vp, _sz, _psz, _vp, _sz, _psz, _pd]),
    "icxref_euler": (C.c_int, [_u32, _vp, _vp, _vp, _dbl, _pd]),
    "icxref_collide": _collide_sig(),
    "icxref_intersect": (C.c_int, [_vp, _vp, _vp, _vp, _u32, _vp, _vp, _vp, _vp, _vp]),
    "icxref_response": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "icxref_projectile_update": (C.c_int, [_u32, _dbl, _u32, _vp, _vp, _vp]),
}


def _load(path, sigs):
    if not os.path.exists(path):
        return None
    lib = C.CDLL(path)
    for name, (res, args) in sigs.items():
        fn = getattr(lib, name)
        fn.restype, fn.argtypes = res, args
    return lib


ORACLE_SIGS = {
    "icxo_euler": REF_SIGS["icxref_euler"],
    "icxo_collide": _collide_sig(),
    "icxo_intersect": REF_SIGS["icxref_intersect"],
    "icxo_response": REF_SIGS["icxref_response"],
    "icxo_aabbs": (C.c_int, [_u32, _vp, _vp, C.c_float, _vp, _sz, _psz]),
    "icxo_pairs": (C.c_int, [_u32, _vp, _vp, C.c_float, _u32, _u32, C.c_int, _vp, _sz, _psz]),
    "icxo_projectile_update": REF_SIGS["icxref_projectile_update"],
}

_ref = None
_oracle = None


def oracle():
    global _oracle
    if _oracle is None:
        _oracle = _load(ORACLE_PATH, ORACLE_SIGS)
    return _oracle


def oracle_pairs(scene, margin, flags=0, stride=0, method=1, cap=None):
    """Broadphase candidate list per the oracle's definition (method 0 = brute force)."""
    cap = cap or max(64 * scene.n, 4096)
    out = np.zeros(cap, abi.pair_dt)
    n = C.c_size_t(0)
    Cc = scene.C.copy()
    rc = oracle().icxo_pairs(scene.n, abi.ptr(scene.T), abi.ptr(Cc), margin, flags, stride, method,
                             abi.ptr(out), cap, C.byref(n))
    assert rc == 0 and n.value <= cap, (rc, n.value, cap)
    return out[:n.value]


def oracle_aabbs(scene, margin):
    out = np.zeros((2 * scene.n, 6), np.float32)
    n = C.c_size_t(0)
    Cc = scene.C.copy()
    oracle().icxo_aabbs(scene.n, abi.ptr(scene.T), abi.ptr(Cc), margin, abi.ptr(out), len(out), C.byref(n))
    return out[:n.value]


def ref():
    global _ref
    if _ref is None:
        _ref = _load(REF_PATH, REF_SIGS)
    return _ref


def narrow16(ids32):
    """IcxEntity -> reference uint16 id (genesis must fit 8 bits)."""
    ids32 = np.asarray(ids32, np.uint32)
    g = (ids32 >> 4) & 0xFFFFFF
    assert (g < 255).all()
    return (((ids32 >> 28) << 12) | (g << 4) | (ids32 & 0xF)).astype(np.uint16)


def verbatim_run(scene, n_ticks=1, dt=1.0 / 45, host=0, coherency=0, cap=1 << 16, ids16=None):
    """Run the reference objects (Cosmos + PhysicsDynamo) on a copy of `scene`."""
    s = scene.copy()
    ids = narrow16(s.ids) if ids16 is None else np.asarray(ids16, np.uint16)
    hits = np.zeros(cap, abi.contact_dt)
    ints = np.zeros(cap, abi.interception_dt)
    nh, ni, sec = C.c_size_t(0), C.c_size_t(0), C.c_double(0)
    coh = C.c_uint16(coherency)
    has_col = np.ascontiguousarray(s.C["colliders"]["isActive"].any(axis=1).astype(np.uint8))
    rc = ref().icxref_verbatim_run(s.n, abi.ptr(ids), abi.ptr(s.T), abi.ptr(s.P), abi.ptr(s.hasP),
                                   abi.ptr(s.C), abi.ptr(has_col), abi.ptr(s.tstate), abi.ptr(s.tcoh),
                                   C.byref(coh), host, dt, n_ticks, abi.ptr(hits), cap, C.byref(nh),
                                   abi.ptr(ints), cap, C.byref(ni), C.byref(sec))
    assert rc == 0, rc
    s["hits"], s["ints"], s["seconds"], s["coherency"] = hits[:nh.value], ints[:ni.value], sec.value, coh.value
    return s


def procedures_tick(lib, prefix, scene, dt=1.0 / 45, host=0, coherency=0, flags=0, stride=0, cand=None,
                    cap=None, euler=True, collide=True):
    """One tick through `<prefix>_euler` + `<prefix>_collide` (reference procedures or the oracle)."""
    s = scene.copy()
    cap = cap or max(4 * s.n, 1024)
    hits = np.zeros(cap, abi.contact_dt)
    ints = np.zeros(cap, abi.interception_dt)
    nh, ni, sec1, sec2 = C.c_size_t(0), C.c_size_t(0), C.c_double(0), C.c_double(0)
    if euler:
        rc = getattr(lib, prefix + "_euler")(s.n, abi.ptr(s.T), abi.ptr(s.P), abi.ptr(s.hasP), dt, C.byref(sec1))
        assert rc == 0
    if collide:
        if cand is not None:
            cand = np.ascontiguousarray(cand, abi.pair_dt)
        rc = getattr(lib, prefix + "_collide")(
            s.n, abi.ptr(s.ids), abi.ptr(s.T), abi.ptr(s.P), abi.ptr(s.hasP), abi.ptr(s.C),
            abi.ptr(s.tstate), abi.ptr(s.tcoh), coherency, host, flags, stride,
            abi.ptr(cand), 0 if cand is None else len(cand),
            abi.ptr(hits), cap, C.byref(nh), abi.ptr(ints), cap, C.byref(ni), C.byref(sec2))
        assert rc == 0, rc
    assert nh.value <= cap and ni.value <= cap
    s["hits"], s["ints"], s["seconds"] = hits[:nh.value], ints[:ni.value], sec1.value + sec2.value
    return s


# ---- timeline checkers (oracle/ref/ref_timeline.cpp: the reference's real containers; oracle: restatement) ----
_TL_SIGS = {
    "{p}_tl_create": (_vp, []),
    "{p}_tl_destroy": (None, [_vp]),
    "{p}_tl_breakpoint": (None, [_vp, C.c_int]),
    "{p}_tl_record": (C.c_int, [_vp, _u32, _vp, _vp, _vp, _vp, C.c_uint16, C.c_int, C.c_int]),
    "{p}_tl_replay": (C.c_int, [_vp, _u32, _vp, _vp, _vp, _vp, _vp, C.c_uint16, C.c_int]),
    "{p}_congruent": (C.c_int, [_u32, _vp, _vp, _vp]),
    "{p}_coherency_ge": (C.c_int, [C.c_uint16, C.c_uint16]),
}
REF_SIGS["icxref_tl_event_push"] = (C.c_int, [_vp, C.c_uint16, _vp, _u32, C.c_int])
REF_SIGS["icxref_tl_events_pop"] = (C.c_int, [_vp, C.c_uint16, C.c_int, _vp, _sz, C.POINTER(_sz)])
REF_SIGS.update({k.format(p="icxref"): v for k, v in _TL_SIGS.items()})
ORACLE_SIGS.update({k.format(p="icxo"): v for k, v in _TL_SIGS.items()})
ORACLE_SIGS["icxo_divergence_scan"] = (C.c_int, [_u32, _vp, _u32, _vp, _u32, _vp, _vp, _vp])
ORACLE_SIGS["icxo_mesh_pass"] = (C.c_int, [_u32, _vp, _vp, _vp, _vp, _vp, _vp, _u32, _vp, _u32, _vp, _vp, _sz, C.POINTER(_sz)])
ORACLE_SIGS["icxo_fork_age"] = (C.c_int, [_u32, _vp, _vp, C.c_uint16, C.POINTER(_u32), C.POINTER(C.c_int)])


class CpuTimeline:
    """EntityTimestreamMap driven through `<prefix>_tl_*` (reference containers or the oracle)."""

    def __init__(self, lib, prefix):
        self.lib, self.p = lib, prefix
        self.h = getattr(lib, prefix + "_tl_create")()

    def close(self):
        getattr(self.lib, self.p + "_tl_destroy")(self.h)

    def breakpoint(self, on):
        getattr(self.lib, self.p + "_tl_breakpoint")(self.h, 1 if on else 0)
        self._bp = bool(on)
        for ev in getattr(self, "_events", []):
            ev["behind"] = bool(on)

    def push_events(self, coherency, events, at_breakpoint=False):
        """Event records.  Reference containers: REAL EventMessages with the reference's params structs in the real per-entity
        streams.  The oracle has no event streams (its ids are wide, the reference's are 16 bits): a list with the same rules
        (due: cyclic <=; stale: plain <; old history hidden behind the breakpoint), pinned to the real ones by
        tests/test_gpu_events.py through the device."""
        events = np.ascontiguousarray(events, abi.event_dt).copy()
        if self.p == "icxref":
            assert getattr(self.lib, self.p + "_tl_event_push")(self.h, coherency, abi.ptr(events), len(events), 1 if at_breakpoint else 0) == 0
            return
        events["coherency"] = coherency
        self._events = getattr(self, "_events", [])
        for e in events:
            self._events.append({"rec": e.copy(), "behind": False})

    def pop_events(self, now, mode=0):
        if self.p == "icxref":
            out = np.zeros(4096, abi.event_dt)
            n = C.c_size_t(0)
            assert getattr(self.lib, self.p + "_tl_events_pop")(self.h, now, mode, abi.ptr(out), len(out), C.byref(n)) == 0
            return out[:n.value]
        keep, out = [], []
        for ev in getattr(self, "_events", []):
            c = int(ev["rec"]["coherency"])
            due = (now >= c and now - c <= 32768) or (now <= c and c - now > 32768)
            behind = getattr(self, "_bp", False) and ev["behind"]
            if not due or (behind if mode == 0 else not behind):
                keep.append(ev)
            elif not c < now:
                out.append(ev["rec"])
        self._events = keep
        return np.array(out, abi.event_dt) if out else np.zeros(0, abi.event_dt)

    def record(self, s, coherency, at_breakpoint=False, link_inc=True, colliders=True):
        """ENTITY_UPDATE of every entity with ALL its components (Transform, Physics, ColliderComponent)."""
        getattr(self.lib, self.p + "_tl_record")(self.h, s.n, abi.ptr(s.ids), abi.ptr(s.T), abi.ptr(s.P), abi.ptr(s.C) if colliders else None,
                                                 coherency, 1 if at_breakpoint else 0, 1 if link_inc else 0)

    def replay(self, s, now, mode=0):
        """Returns (scene after replay, entities overwritten)."""
        out = s.copy()
        n = getattr(self.lib, self.p + "_tl_replay")(self.h, out.n, abi.ptr(out.ids), abi.ptr(out.T), abi.ptr(out.P), abi.ptr(out.C),
                                                     abi.ptr(out.tstate), now, mode)
        return out, n


def oracle_scan(live, rec, stride=0):
    nw = (live.n + stride - 1) // stride if stride else 1
    flags, cnt, score = np.zeros(nw, np.uint8), np.zeros(nw, np.uint32), np.zeros(nw, np.float64)
    oracle().icxo_divergence_scan(live.n, abi.ptr(live.T), rec.n, abi.ptr(rec.T), stride, abi.ptr(flags), abi.ptr(cnt), abi.ptr(score))
    return flags, cnt, score


def oracle_mesh_pass(s, vertices, triangles, material=None, cap=None):
    """The static-mesh pass on a copy of `s` through the self-oracle (oracle/icx_oracle.cpp icxo_mesh_pass): (scene, contacts)."""
    out = s.copy()
    v = np.ascontiguousarray(vertices, np.float32).reshape(-1, 3)
    t = np.ascontiguousarray(triangles, np.uint32).reshape(-1, 3)
    m = None
    if material is not None:
        m = np.zeros(1, abi.mesh_material_dt)
        m["staticFriction"], m["dynamicFriction"], m["restitution"] = material
    cap = cap or max(16 * s.n, 1024)
    hits = np.zeros(cap, abi.mesh_contact_dt)
    k = C.c_size_t(0)
    rc = oracle().icxo_mesh_pass(out.n, abi.ptr(out.ids), abi.ptr(out.T), abi.ptr(out.P), abi.ptr(out.hasP), abi.ptr(out.C),
                                 abi.ptr(v), len(v), abi.ptr(t), len(t), abi.ptr(m), abi.ptr(hits), cap, C.byref(k))
    assert rc == 0 and k.value <= cap, (rc, k.value)
    return out, hits[:k.value]


def oracle_fork_age(s, coherency):
    out = s.copy()
    n, need = C.c_uint32(0), C.c_int(0)
    oracle().icxo_fork_age(out.n, abi.ptr(out.tstate), abi.ptr(out.tcoh), coherency, C.byref(n), C.byref(need))
    return out, n.value, bool(need.value)


def projectile_update(lib, prefix, T, proj, dt, n_updates=1):
    """ProjectileBehaviors::on_fixed_update through a checker: returns (scale per projectile, components, condemned)."""
    T, proj = T.copy(), proj.copy()
    condemned = np.zeros(len(proj), np.uint8)
    rc = getattr(lib, prefix + "_projectile_update")(len(proj), dt, n_updates, abi.ptr(T), abi.ptr(proj), abi.ptr(condemned))
    assert rc == 0
    return T["scale"].copy(), proj, condemned


# ---- a22: the reference's load_and_link / extract entity loops lifted around the real Cosmos (oracle/ref/ref_clone.cpp) ----
_u16p = C.POINTER(C.c_uint16)
REF_SIGS["icxref_load_and_link"] = (C.c_int, [_u32, _vp, _vp, _vp, _vp, _vp, _vp, _vp, C.c_uint16, _u32, _vp,
                                             _u32, C.POINTER(_u32), _vp, _vp, _vp, _vp, _vp, _vp, _vp, _u16p])
REF_SIGS["icxref_extract"] = (C.c_int, [_u32, _vp, _vp, _vp, _vp, _vp, _vp, _vp, C.c_uint16,
                                       _u32, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _u32, C.POINTER(_u32), C.POINTER(_u32)])


def ref_load_and_link(src, src_coherency, has_future=None, src_host=2):
    """Returns (parallel scene, parallel coherency); mutates src.tstate / src.tcoh like the reference (real Cosmos objects)."""
    from intercession_b200 import scenes
    ids16 = narrow16(src.ids)
    hf = None if has_future is None else np.ascontiguousarray(has_future, np.uint8)
    cap = src.n
    out = scenes._empty(cap)
    o_ids = np.zeros(cap, np.uint16)
    n_out, coh = _u32(0), C.c_uint16(0)
    rc = ref().icxref_load_and_link(src.n, abi.ptr(ids16), abi.ptr(src.T), abi.ptr(src.P), abi.ptr(src.hasP), abi.ptr(src.C),
                                    abi.ptr(src.tstate), abi.ptr(src.tcoh), src_coherency, src_host, abi.ptr(hf),
                                    cap, C.byref(n_out), abi.ptr(o_ids), abi.ptr(out.T), abi.ptr(out.P), abi.ptr(out.hasP), abi.ptr(out.C),
                                    abi.ptr(out.tstate), abi.ptr(out.tcoh), C.byref(coh))
    assert rc == 0, rc
    k = n_out.value
    par = scenes.Scene({key: (v[:k].copy() if isinstance(v, np.ndarray) else v) for key, v in out.items()})
    par["ids"] = np.array([abi.widen_entity(e) for e in o_ids[:k]], np.uint32)
    return par, coh.value


def ref_extract(par, dst, coherency, dst_host_id):
    """Mutates dst like the reference; returns (extracted, ids the reference had to register in dst) or raises on chain-link 0."""
    n_ex, n_reg = _u32(0), _u32(0)
    rc = ref().icxref_extract(par.n, abi.ptr(narrow16(par.ids)), abi.ptr(par.T), abi.ptr(par.P), abi.ptr(par.hasP), abi.ptr(par.C),
                              abi.ptr(par.tstate), abi.ptr(par.tcoh), coherency,
                              dst.n, abi.ptr(narrow16(dst.ids)), abi.ptr(dst.T), abi.ptr(dst.P), abi.ptr(dst.hasP), abi.ptr(dst.C),
                              abi.ptr(dst.tstate), abi.ptr(dst.tcoh), dst_host_id, C.byref(n_ex), C.byref(n_reg))
    if rc == 2:
        raise ValueError("chain-link 0 in the parallel cosmos (std::range_error in the reference)")
    assert rc == 0, rc
    return n_ex.value, n_reg.value


REF_SIGS["icxref_entity_update_messages"] = (C.c_int, [_u32, _vp, _vp, _vp, _vp, _vp, C.c_uint16, C.c_int, C.c_int, _vp, _sz])


def ref_entity_update_messages(s, coherency, hard_config_types=True, increment_link=True):
    """The reference's own ENTITY_UPDATE messages (real Cosmos::serialize_entity_components + EventMessage << params)."""
    stride = abi.WIRE_MAX_BYTES
    buf = np.zeros(s.n * stride, np.uint8)
    rc = ref().icxref_entity_update_messages(s.n, abi.ptr(narrow16(s.ids)), abi.ptr(s.T), abi.ptr(s.P), abi.ptr(s.hasP), abi.ptr(s.C), coherency,
                                             1 if hard_config_types else 0, 1 if increment_link else 0, abi.ptr(buf), stride)
    assert rc == 0, rc
    out = []
    for k in range(s.n):
        m = buf[k * stride:(k + 1) * stride]
        size = int(m[8:12].view(np.uint32)[0])
        out.append(bytes(m[:12 + size]))
    return out


REF_SIGS["icxref_last_grounds"] = (C.c_int, [_vp, _sz, _psz])
ORACLE_SIGS["icxo_last_grounds"] = (C.c_int, [_vp, _sz, _psz])


def last_grounds(lib, prefix):
    """What BipedBehaviors::on_collision left per entity during the checker's last collision pass."""
    n = C.c_size_t(0)
    fn = getattr(lib, prefix + "_last_grounds")
    fn(None, 0, C.byref(n))
    out = np.zeros(n.value, abi.ground_dt)
    if n.value:
        fn(abi.ptr(out), n.value, C.byref(n))
    return out
