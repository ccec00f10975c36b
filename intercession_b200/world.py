"""Thin Python driver over the C ABI (include/icx.h) for tests and bench.py.

Mirrors the reference's PhysicsDynamo call sequence (source/physics/physics_dynamo.cpp:23-53):
upload == the synchros' submit(), step == run_relays(dt) (+ Cosmos::increment_coherency),
download == the in-place component mutation the reference performs.  The production host side
is C++ (intercession_b200/host/physics_dynamo.hpp); this class exists so Python tests read
like the reference's tick.
"""
import ctypes as C

import numpy as np

from . import abi


class IcxError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"icx status {code}: {msg}")
        self.code = code


class World:
    def __init__(self, max_entities, host_timeslice=0, flags=0, margin=0.0, cell_size=0.0, max_neighbors=0,
                 worldline_stride=0, device=0, max_contacts=0, gravity=None):
        self.lib = abi.load()
        cfg = abi.IcxConfig()
        self.lib.icx_config_default(C.byref(cfg))
        cfg.device = device
        cfg.max_entities = max(int(max_entities), 1)
        cfg.host_timeslice = host_timeslice
        cfg.flags = flags
        cfg.broadphase_margin = margin
        cfg.cell_size = cell_size
        cfg.max_neighbors = max_neighbors
        cfg.worldline_stride = worldline_stride
        cfg.max_contacts = max_contacts
        if gravity is not None:
            cfg.gravity[:] = gravity
        self.handle = C.c_void_p()
        rc = self.lib.icx_world_create(C.byref(cfg), C.byref(self.handle))
        if rc != abi.OK:
            raise IcxError(rc, self.lib.icx_status_string(rc).decode() + " (icx_world_create; is a CUDA device present?)")
        self.n = 0

    def close(self):
        if self.handle:
            self.lib.icx_world_destroy(self.handle)
            self.handle = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _ck(self, rc):
        if rc != abi.OK:
            raise IcxError(rc, self.lib.icx_last_error(self.handle).decode())

    # ---- PhysicsSynchro / ColliderSynchro -> submit ----
    def upload(self, s):
        self.n = s.n
        self._ck(self.lib.icx_bodies_upload(self.handle, s.n, abi.ptr(s.ids), abi.ptr(s.T), abi.ptr(s.P),
                                            abi.ptr(s.hasP), abi.ptr(s.C), abi.ptr(s.tstate), abi.ptr(s.tcoh)))

    def download(self, like):
        out = like.copy()
        self._ck(self.lib.icx_bodies_download(self.handle, out.n, abi.ptr(out.T), abi.ptr(out.P), abi.ptr(out.C),
                                              abi.ptr(out.tstate), abi.ptr(out.tcoh)))
        return out

    def download_into(self, T=None, P=None, Cc=None, tstate=None, tcoh=None):
        self._ck(self.lib.icx_bodies_download(self.handle, self.n, abi.ptr(T), abi.ptr(P), abi.ptr(Cc),
                                              abi.ptr(tstate), abi.ptr(tcoh)))

    def update(self, index, T=None, P=None, Cc=None):
        index = np.ascontiguousarray(index, np.uint32)
        self._ck(self.lib.icx_bodies_update(self.handle, len(index), abi.ptr(index), abi.ptr(T), abi.ptr(P), abi.ptr(Cc)))

    def update_all(self, T=None, P=None):
        """Transform / physics of every entity (what the synchros resubmit each frame); colliders stay resident."""
        self._ck(self.lib.icx_bodies_update(self.handle, self.n, None, abi.ptr(T), abi.ptr(P), None))

    def motion_download(self, out=None):
        """The 76 B per entity a tick writes (origin, orientation, velocity, acceleration, angular velocity / acceleration)."""
        out = np.zeros(self.n, abi.motion_dt) if out is None else out
        self._ck(self.lib.icx_motion_download(self.handle, self.n, abi.ptr(out)))
        return out

    # ---- PhysicsDynamo::run_relays ----
    def step(self, dt=1.0 / 45, n=1):
        self._ck(self.lib.icx_step_n(self.handle, dt, n) if n != 1 else self.lib.icx_step(self.handle, dt))

    def euler_engage(self, dt=1.0 / 45):
        self._ck(self.lib.icx_euler_engage(self.handle, dt))

    def collision_engage(self):
        self._ck(self.lib.icx_collision_engage(self.handle))

    def share_device(self, parts):
        """This world's persistent kernels use 1 / parts of the device, so that `parts` worlds tick concurrently (icx_device_share_set)."""
        self._ck(self.lib.icx_device_share_set(self.handle, parts))

    def sync(self):
        self._ck(self.lib.icx_sync(self.handle))

    def stream(self):
        """cudaStream_t of the world as an int."""
        return self.lib.icx_world_stream(self.handle) or 0

    def set_coherency(self, c):
        self._ck(self.lib.icx_coherency_set(self.handle, c))

    def coherency(self):
        c = C.c_uint16(0)
        self._ck(self.lib.icx_coherency_get(self.handle, C.byref(c)))
        return c.value

    # ---- results of the last tick ----
    def _fetch(self, fn, dt):
        n = C.c_size_t(0)
        self._ck(fn(self.handle, None, 0, C.byref(n)))
        out = np.zeros(n.value, dt)
        if n.value:
            self._ck(fn(self.handle, abi.ptr(out), n.value, C.byref(n)))
        return out

    def contacts(self):
        return self._fetch(self.lib.icx_contacts_fetch, abi.contact_dt)

    def interceptions(self):
        return self._fetch(self.lib.icx_interceptions_fetch, abi.interception_dt)

    def pairs(self):
        return self._fetch(self.lib.icx_pairs_fetch, abi.pair_dt)

    def grounds(self):
        """Per entity whose BipedBehaviors::on_collision would have run: what it leaves in the BipedComponent."""
        return self._fetch(self.lib.icx_grounds_fetch, abi.ground_dt)

    def extra_pairs(self):
        return self._fetch(self.lib.icx_extra_pairs_fetch, abi.pair_dt)

    def all_pairs(self):
        """Uniform-margin pairs + the movers' extra pairs, canonical order: the candidate set of the ordered pass."""
        p = np.concatenate([self.pairs(), self.extra_pairs()])
        return p[np.lexsort((p["b"], p["a"]))]

    def budgets(self):
        out = np.zeros(self.n, np.float32)
        self._ck(self.lib.icx_budgets_fetch(self.handle, abi.ptr(out), len(out)))
        return out

    def entity_update_messages(self, coherency, types=(0, 6, 7), event_id=102, increment_link=True):
        """ENTITY_UPDATE messages of every entity in the reference's wire layout: list of bytes objects (header + body)."""
        lay = abi.IcxWireLayout(event_id, types[0], types[1], types[2], 1 if increment_link else 0)
        buf = np.zeros(max(self.n, 1) * abi.WIRE_MAX_BYTES, np.uint8)
        stride, n = C.c_size_t(0), C.c_uint32(0)
        self._ck(self.lib.icx_entity_update_messages(self.handle, C.byref(lay), coherency, abi.ptr(buf), buf.nbytes, C.byref(stride), C.byref(n)))
        out = []
        for k in range(n.value):
            m = buf[k * stride.value:(k + 1) * stride.value]
            size = int(m[8:12].view(np.uint32)[0])
            out.append(bytes(m[:12 + size]))
        return out

    def stats(self):
        st = abi.IcxStats()
        self._ck(self.lib.icx_stats_fetch(self.handle, C.byref(st)))
        return st

    # ---- ProjectileBehaviors::on_fixed_update on the device (behaviors/projectile_behaviors.h:17-54) ----
    def projectiles_upload(self, index, proj):
        index = np.ascontiguousarray(index, np.uint32)
        proj = np.ascontiguousarray(proj, abi.projectile_dt)
        assert len(index) == len(proj)
        self._n_proj = len(proj)
        self._ck(self.lib.icx_projectiles_upload(self.handle, len(proj), abi.ptr(index), abi.ptr(proj)))

    def projectiles_fixed_update(self, dt):
        self._ck(self.lib.icx_projectiles_fixed_update(self.handle, dt))

    def projectiles_download(self):
        out = np.zeros(self._n_proj, abi.projectile_dt)
        self._ck(self.lib.icx_projectiles_download(self.handle, len(out), abi.ptr(out)))
        return out

    def projectiles_condemned(self):
        out = np.zeros(max(self._n_proj, 1), np.uint32)
        n = C.c_size_t(0)
        self._ck(self.lib.icx_projectiles_condemned(self.handle, abi.ptr(out), len(out), C.byref(n)))
        return out[:n.value]

    def projectiles_kernel_ms(self):
        ms = C.c_float(0)
        self._ck(self.lib.icx_projectiles_kernel_ms(self.handle, C.byref(ms)))
        return ms.value

    def last_step_ms(self):
        ms = C.c_float(0)
        self._ck(self.lib.icx_last_step_ms(self.handle, C.byref(ms)))
        return ms.value

    def profile(self, on=True):
        self._ck(self.lib.icx_profile_enable(self.handle, 1 if on else 0))

    def phases(self):
        """({phase: accumulated ms}, ticks profiled, kernels launched since creation)."""
        ms = (C.c_double * 5)()
        nt, nl = C.c_uint64(0), C.c_uint64(0)
        self._ck(self.lib.icx_phase_ms(self.handle, ms, C.byref(nt), C.byref(nl)))
        names = ("integrate_prep", "grid_sort", "pairs", "narrow", "resolve")
        return dict(zip(names, list(ms))), nt.value, nl.value

    def totals(self):
        """Sums over the ticks since profile(): ticks, candidate pairs, contacts, rounds, attempts, dirty bodies."""
        out = (C.c_uint64 * 6)()
        self._ck(self.lib.icx_totals_fetch(self.handle, out))
        return dict(zip(("ticks", "candidates", "contacts", "rounds", "attempts", "dirty"), [int(v) for v in out]))

    def trace(self, enable=True):
        """Device timeline of the last ordered resolve: list of (ns since its start, tag)."""
        buf = np.zeros(512, np.uint64)
        n = C.c_size_t(0)
        self._ck(self.lib.icx_debug_trace(self.handle, 1 if enable else 0, abi.ptr(buf), len(buf), C.byref(n)))
        v = buf[:n.value].reshape(-1, 2)
        return [(int(t - v[0, 0]), int(tag)) for t, tag in v] if len(v) else []

    # ---- TimelineApi hooks ----
    def record(self, tl, coherency=None, at_breakpoint=False, keep_link=None):
        """push_past_timestream (chain-link + 1 on the way into the past) / push_timestream_at_breakpoint (ids unchanged)."""
        keep = at_breakpoint if keep_link is None else keep_link
        flags = (abi.RECORD_AT_BREAKPOINT if at_breakpoint else 0) | (abi.RECORD_KEEP_LINK if keep else 0)
        self._ck_tl(tl, self.lib.icx_record_ex(self.handle, tl.handle, self.coherency() if coherency is None else coherency, flags))

    def replay(self, tl, coherency=None, mode=abi.REPLAY_SERVER):
        applied = C.c_int(0)
        self._ck_tl(tl, self.lib.icx_replay(self.handle, tl.handle, self.coherency() if coherency is None else coherency, mode, C.byref(applied)))
        return bool(applied.value)

    def divergence_scan(self, tl, coherency, n_worldlines=1):
        flags = np.zeros(n_worldlines, np.uint8)
        count = np.zeros(n_worldlines, np.uint32)
        score = np.zeros(n_worldlines, np.float64)
        self._ck_tl(tl, self.lib.icx_divergence_scan(self.handle, tl.handle, coherency, n_worldlines, abi.ptr(flags), abi.ptr(count), abi.ptr(score)))
        return flags, count, score

    def divergence_scan_enqueue(self, tl, coherency, n_worldlines=1):
        self._ck_tl(tl, self.lib.icx_divergence_scan_enqueue(self.handle, tl.handle, coherency, n_worldlines))

    def fork_age_async(self):
        self._ck(self.lib.icx_fork_age(self.handle, None, None))

    def fork_age(self):
        n, need = C.c_uint32(0), C.c_int(0)
        self._ck(self.lib.icx_fork_age(self.handle, C.byref(n), C.byref(need)))
        return n.value, bool(need.value)

    # ---- ParallelCosmosContext::load_and_link / extract_entity_updates ----
    def clone_into(self, dst, has_future=None):
        """dst becomes the parallel-cosmos copy of this world; returns the number of entities copied."""
        hf = None if has_future is None else np.ascontiguousarray(has_future, np.uint8)
        n = C.c_uint32(0)
        rc = self.lib.icx_world_clone(self.handle, dst.handle, abi.ptr(hf), C.byref(n))
        if rc != abi.OK:
            raise IcxError(rc, self.lib.icx_last_error(dst.handle).decode())
        dst.n = n.value
        return n.value

    def extract_into(self, dst, register=False):
        """Write this (parallel) world's temporal entities back into dst at chain-link - 1: (extracted, missing).
        register: ids absent from dst are registered there first (parallel_cosmos_context.cpp:241-245)."""
        n, miss = C.c_uint32(0), C.c_uint32(0)
        rc = self.lib.icx_world_extract_ex(self.handle, dst.handle, 1 if register else 0, C.byref(n), C.byref(miss))
        if rc != abi.OK:
            raise IcxError(rc, self.lib.icx_last_error(dst.handle).decode())
        if register:
            dst.n = len(dst.entity_ids())
        return n.value, miss.value

    # ---- static triangle mesh (extension without a counterpart in the reference) ----
    def upload_mesh(self, vertices, triangles, material=None):
        v = np.ascontiguousarray(vertices, np.float32).reshape(-1, 3)
        t = np.ascontiguousarray(triangles, np.uint32).reshape(-1, 3)
        m = None
        if material is not None:
            m = np.zeros(1, abi.mesh_material_dt)
            m["staticFriction"], m["dynamicFriction"], m["restitution"] = material
        self._ck(self.lib.icx_mesh_upload(self.handle, abi.ptr(v), len(v), abi.ptr(t), len(t), abi.ptr(m)))

    def remove_mesh(self):
        self._ck(self.lib.icx_mesh_upload(self.handle, None, 0, None, 0, None))

    def mesh_contacts(self):
        return self._fetch(self.lib.icx_mesh_contacts_fetch, abi.mesh_contact_dt)

    def mesh_kernel_us(self):
        us = C.c_float(0)
        self._ck(self.lib.icx_mesh_kernel_us(self.handle, C.byref(us)))
        return us.value

    # ---- entity lifecycle (Cosmos::register_entity / condemn_entity) ----
    def create_entities(self, rows):
        """Add the entities of the Scene `rows`; the ones the world holds stay on the device.  Returns how many were new."""
        n = C.c_uint32(0)
        self._ck(self.lib.icx_entities_create(self.handle, rows.n, abi.ptr(np.ascontiguousarray(rows.ids, np.uint32)), abi.ptr(rows.T), abi.ptr(rows.P),
                                              abi.ptr(rows.hasP), abi.ptr(rows.C), abi.ptr(rows.tstate), abi.ptr(rows.tcoh), C.byref(n)))
        self.n = len(self.entity_ids())
        return n.value

    def remove_entities(self, ids):
        ids = np.ascontiguousarray(ids, np.uint32)
        n = C.c_uint32(0)
        self._ck(self.lib.icx_entities_remove(self.handle, len(ids), abi.ptr(ids), C.byref(n)))
        self.n = len(self.entity_ids())
        return n.value

    def entity_ids(self):
        k = C.c_size_t(0)
        self.lib.icx_entity_ids_fetch(self.handle, None, 0, C.byref(k))
        out = np.zeros(k.value, np.uint32)
        self.lib.icx_entity_ids_fetch(self.handle, abi.ptr(out), len(out), C.byref(k))
        return out

    def _ck_tl(self, tl, rc):
        if rc != abi.OK:
            raise IcxError(rc, (self.lib.icx_timeline_last_error(tl.handle) or self.lib.icx_last_error(self.handle)).decode())


class Timeline:
    """Device snapshot ring behind the TimelineApi record / replay hooks (include/icx.h)."""

    def __init__(self, depth_ticks, n_max, device=0):
        self.lib = abi.load()
        self.handle = C.c_void_p()
        rc = self.lib.icx_timeline_create(device, depth_ticks, n_max, C.byref(self.handle))
        if rc != abi.OK:
            raise IcxError(rc, self.lib.icx_status_string(rc).decode() + " (icx_timeline_create)")

    def set_breakpoint(self):
        self.lib.icx_breakpoint_set(self.handle)

    def clear_breakpoint(self):
        self.lib.icx_breakpoint_clear(self.handle)

    def export_slot(self, coherency, dev_ptr, capacity_bytes, stream=None):
        """Pack the slot stamped `coherency` into the device buffer at dev_ptr (int); returns the entity count."""
        n = C.c_uint32(0)
        rc = self.lib.icx_timeline_slot_export(self.handle, coherency, C.c_void_p(dev_ptr), capacity_bytes, C.byref(n), C.c_void_p(stream or 0))
        if rc != abi.OK:
            raise IcxError(rc, self.lib.icx_timeline_last_error(self.handle).decode())
        return n.value

    def import_slot(self, coherency, n, dev_ptr, nbytes, stream=None):
        rc = self.lib.icx_timeline_slot_import(self.handle, coherency, n, C.c_void_p(dev_ptr), nbytes, C.c_void_p(stream or 0))
        if rc != abi.OK:
            raise IcxError(rc, self.lib.icx_timeline_last_error(self.handle).decode())

    def _ck(self, rc):
        if rc != abi.OK:
            raise IcxError(rc, self.lib.icx_timeline_last_error(self.handle).decode())

    def push_events(self, coherency, events):
        """Event records (abi.event_dt) of the tick stamped `coherency`: they live in that tick's ring slot on the device."""
        events = np.ascontiguousarray(events, abi.event_dt)
        self._ck(self.lib.icx_timeline_event_push(self.handle, coherency, abi.ptr(events), len(events)))

    def pop_events(self, now, mode=abi.REPLAY_SERVER):
        out = np.zeros(4 * abi.TIMELINE_EVENTS_PER_TICK, abi.event_dt)
        n = C.c_size_t(0)
        self._ck(self.lib.icx_timeline_events_pop(self.handle, now, mode, abi.ptr(out), len(out), C.byref(n)))
        if n.value > len(out):
            raise RuntimeError("more event records due than the pop buffer holds")
        return out[:n.value]

    def fetch_entity(self, coherency, entity):
        """The recorded components of one entity of the tick stamped `coherency` as a 1-row Scene (None if the record does not
        hold it); row.have_collider tells whether the record carried its ColliderComponent."""
        from .scenes import Scene, _empty
        row = _empty(1)
        row["ids"] = np.array([entity], np.uint32)
        found = C.c_int(0)
        self._ck(self.lib.icx_timeline_entity_fetch(self.handle, coherency, int(entity), abi.ptr(row.T), abi.ptr(row.P), abi.ptr(row.hasP), abi.ptr(row.C), C.byref(found)))
        if not (found.value & 1):
            return None
        row["have_collider"] = bool(found.value & 2)
        return row

    def id_mismatches(self):
        v = C.c_uint64(0)
        self.lib.icx_timeline_id_mismatches(self.handle, C.byref(v))
        return v.value

    def scan_fetch(self, n_worldlines, reset_ever=False):
        """Results of the last enqueued scan + the per-world-line 'flagged by some scan since the last reset' bytes."""
        flags, ever = np.zeros(n_worldlines, np.uint8), np.zeros(n_worldlines, np.uint8)
        count, score = np.zeros(n_worldlines, np.uint32), np.zeros(n_worldlines, np.float64)
        rc = self.lib.icx_divergence_scan_fetch(self.handle, n_worldlines, abi.ptr(flags), abi.ptr(ever), abi.ptr(count), abi.ptr(score),
                                                1 if reset_ever else 0)
        if rc != abi.OK:
            raise IcxError(rc, self.lib.icx_timeline_last_error(self.handle).decode())
        return flags, ever, count, score

    def profile(self, on=True):
        self.lib.icx_timeline_profile(self.handle, 1 if on else 0)

    def kernel_ms(self, which):
        """Duration of the last k_record (0) / k_replay (1) / k_divergence_scan (2) launch, CUDA events on the world's stream."""
        ms = C.c_float(0)
        rc = self.lib.icx_timeline_kernel_ms(self.handle, which, C.byref(ms))
        if rc != abi.OK:
            raise IcxError(rc, self.lib.icx_timeline_last_error(self.handle).decode())
        return ms.value

    def slot_info(self, coherency):
        v, n = C.c_int(0), C.c_uint32(0)
        self.lib.icx_timeline_slot_info(self.handle, coherency, C.byref(v), C.byref(n))
        return bool(v.value), n.value

    def close(self):
        if self.handle:
            self.lib.icx_timeline_destroy(self.handle)
            self.handle = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()
