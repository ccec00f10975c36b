"""Config 5 end to end (BASELINE.json configs[4]): a chain of timeslices with a time traveller.

What the reference's timeline exists for, on the device pieces of this package:

  a traveller arrives in a past timeslice (chain-link 0, no future timestream)          server_network_dynamo.cpp:748-812
  -> it collides with local entities: TIMESTREAM_INTERCEPTION, recipient -> forking      collision_physics_relay.h:157-228, server_network_dynamo.cpp:813-838
  -> fork ageing: forking -> forked after 13 ticks, resolution needed after 13 more      server_network_dynamo.cpp:681-705
  -> parallel_notify_divergence -> INIT to the past-most timeslice                      parallel_cosmos_context.cpp:425-453
  -> load_and_link (clone, breakpoints on that timeslice's future stream)                server_network_dynamo.cpp:321-337, parallel_cosmos_context.cpp:60-169
  -> the parallel cosmos re-simulates up to the future neighbour's coherency, consuming old history at the breakpoint and
     writing new history in front of it                                                  parallel_network_dynamo.cpp:51-89, :369-385
  -> extract into the future neighbour at chain-link - 1, INIT on that timeslice, ... down to timeslice 0
                                                                                         parallel_cosmos_context.cpp:171-423, server_network_dynamo.cpp:339-364

The reference runs the parallel cosmos on its own thread "as fast as the CPU can" (parallel_cosmos_context.cpp:534-536) and
re-targets it until its coherency meets the future neighbour's (server_network_dynamo.cpp:339-364): WHEN inside a wall tick
that happens is a race there.  The schedule here is the deterministic limit of it -- the cycle runs to completion, down the
whole chain, at the start of the wall tick in which resolution is needed (phase 1 of the networker, before replay / record) --
so that the same schedule can be driven through two back ends and compared tick by tick: DeviceBackend (this file: IcxWorld /
IcxTimeline through the C ABI) and the CPU restatement on the oracle procedures and the reference's real EntityTimestreamMap
(tests/cpu_timetravel.py).  Jump events themselves (JUMP_REQUEST / DEPARTURE / ARRIVAL) are host-side in the reference too and
are scripted: an arrival is a row inserted into the destination timeslice.
"""
import time

import numpy as np

from . import abi, pipeline, world
from .scenes import Scene

DT = 1.0 / 45
KEYS = ("ids", "T", "P", "hasP", "C", "tstate", "tcoh")


def insert_row(scene, row):
    """Scene with `row` (a 1-entity Scene) inserted at its place in ascending id order."""
    at = int(np.searchsorted(scene.ids, row.ids[0]))
    out = Scene({k: v for k, v in scene.items() if k not in KEYS})
    for k in KEYS:
        # not np.concatenate: its dtype promotion drops the trailing padding of the aligned component rows
        v = np.empty(scene.n + 1, scene[k].dtype)
        v[:at], v[at], v[at + 1:] = scene[k][:at], row[k][0], scene[k][at:]
        out[k] = v
    return out, at


def insert_rows(scene, rows):
    """Scene with every entity of the Scene `rows` inserted at its place in ascending id order (one pass per array)."""
    ids = np.concatenate([scene.ids, rows.ids])
    order = np.argsort(ids, kind="stable")
    out = Scene({k: v for k, v in scene.items() if k not in KEYS})
    for k in KEYS:
        v = np.empty(scene.n + rows.n, scene[k].dtype)                      # (not np.concatenate: see insert_row)
        v[:scene.n], v[scene.n:] = scene[k], rows[k]
        out[k] = v[order]
    return out


def arrival_events(ids):
    """What a timeslice pushes into its past when entities `ids` (as the PAST knows them: chain-link already incremented) arrive:
    ENTITY_CREATED then JUMP_ARRIVAL per entity (_entity_created_handler, _jump_arrival_handler: server_network_dynamo.cpp:748-777,
    :940-976).  The arrival's serialised components are not in the event: the tick's ENTITY_UPDATE record carries them."""
    ids = np.asarray(ids, np.uint32)
    ev = np.zeros(2 * len(ids), abi.event_dt)
    ev["type"][0::2], ev["type"][1::2] = abi.EVENT_ENTITY_CREATED, abi.EVENT_JUMP_ARRIVAL
    ev["entity"][0::2] = ids; ev["entity"][1::2] = ids
    ev["source"][:] = abi.NULL_ENTITY
    ev["payload"][0::2, 0] = 0xC1                                            # signature: Transform | Physics | Collider
    ev["payload"][1::2, 0] = (ids >> 4) & 0xFFFFFF                           # tripId
    ev["payload"][1::2, 1] = 1                                               # timesliceDelta
    ev["payload"][1::2, 3] = 0xC1
    return ev


def removal_events(ids):
    """ENTITY_REMOVED for entities `ids` (as the past knows them), _entity_removed_handler: server_network_dynamo.cpp:779-811."""
    ids = np.asarray(ids, np.uint32)
    ev = np.zeros(len(ids), abi.event_dt)
    ev["type"] = abi.EVENT_ENTITY_REMOVED
    ev["entity"] = ids
    ev["source"] = abi.NULL_ENTITY                                           # forced: nothing inside the cosmos caused it
    return ev


def traveller_row(scene, index, genesis, timeslice):
    """A copy of entity `index` arriving as a time traveller: chain-link 0, its own genesis id, merged state."""
    row = Scene({k: scene[k][index:index + 1].copy() for k in KEYS})
    row["ids"] = np.array([abi.compose_entity(timeslice, genesis, 0)], np.uint32)
    row["tstate"][:] = abi.TS_MERGED
    row["tcoh"][:] = 0
    return row


class DeviceBackend:
    """Timeslice k = one IcxWorld (+ the ring of records its future neighbour pushes) on CUDA device devices[k] (all in this
    process).  The parallel cosmos is one more world on the device of the timeslice it is shadowing: load_and_link and the
    replay / record at the breakpoint then stay on that GPU (the ring is never read over NVLink), and only the extract into the
    future neighbour crosses to devices[k - 1] as one peer copy."""

    def __init__(self, scenes_per_slice, delay, base=100, devices=None, capacity_slack=8, host_rows=True):
        """host_rows=False: the host keeps only the ids of every timeslice (what a server needs for entity_has_timestream);
        state() and the parallel cosmos' end state are then not available."""
        self.n = len(scenes_per_slice)
        self.delay, self.base = delay, base
        self.devices = list(devices) if devices else [0] * self.n
        self.scene = list(scenes_per_slice)
        self.host_rows = host_rows
        cap = max(s.n for s in scenes_per_slice) + capacity_slack
        self.cap = cap
        self.w = [world.World(cap, host_timeslice=k, device=self.devices[k]) for k in range(self.n)]
        for k in range(self.n):
            self.w[k].upload(self.scene[k])
            self.w[k].set_coherency(pipeline.start_coherency(k, self.n, delay, base))
        self.tl_in = [None] + [world.Timeline(2 * delay + 4, cap, device=self.devices[k]) for k in range(1, self.n)]
        self.par = {}                                                       # device -> the parallel cosmos' world there
        self._pool = None
        self.lifecycle_ticks = set()                                        # wall ticks in which entities were created / removed (bench)
        self.keep_parallel_end = host_rows
        self.link = {}
        if not host_rows:
            self.scene = [Scene(ids=s.ids.copy()) for s in self.scene]
        if any(self.devices[k] != self.devices[k + 1] for k in range(self.n - 1)):
            import torch
            for k in range(self.n - 1):
                nb = abi.snapshot_capacity(cap, cap)
                self.link[k] = (torch.empty(nb, dtype=torch.uint8, device=torch.device("cuda", self.devices[k])),
                                torch.empty(nb, dtype=torch.uint8, device=torch.device("cuda", self.devices[k + 1])),
                                world.Timeline(2, cap, device=self.devices[k]))

    def close(self):
        if self._pool is not None:
            self._pool.shutdown()
        for o in self.w + list(self.par.values()) + [t for t in self.tl_in if t is not None] + [v[2] for v in self.link.values()]:
            o.close()

    def coherency(self, k):
        return self.w[k].coherency()

    def _has(self, k, entity):
        """entity_exists: the ids of a timeslice are kept ascending (np.isin on a scalar builds a 2^28-entry table)."""
        ids = self.scene[k].ids
        j = int(np.searchsorted(ids, entity))
        return j < len(ids) and ids[j] == entity

    # ---- networker phases ----
    def replay(self, k):
        """Phase 2: the events of the future stream that have come due, then the ENTITY_UPDATE records.  ENTITY_CREATED: the entity
        is created from the components the same tick's record holds, and the creation travels on into this timeslice's own past."""
        if k == 0:
            return False
        c = self.coherency(k)
        evs = self.tl_in[k].pop_events(c)
        new = [int(e["entity"]) for e in evs if e["type"] == abi.EVENT_ENTITY_CREATED and not self._has(k, e["entity"])]
        if new:
            rows = []
            for ent in new:
                row = self.tl_in[k].fetch_entity(c, ent)
                assert row is not None and row.pop("have_collider"), "the record of the tick an entity is created in carries all its components"
                rows.append(row)
            self._create(k, rows)
        gone = [int(e["entity"]) for e in evs if e["type"] == abi.EVENT_ENTITY_REMOVED and self._has(k, e["entity"])]
        if gone:
            self.depart(k, gone)
        return self.w[k].replay(self.tl_in[k], c, abi.REPLAY_SERVER)

    def depart(self, k, ids):
        """Cosmos::condemn_entity for `ids` (a departure, or ENTITY_REMOVED from the future stream), and the removal pushed on."""
        self.lifecycle_ticks.add(getattr(self, "_tick", -1))
        ids = np.asarray(ids, np.uint32)
        keep = ~np.isin(self.scene[k].ids, ids, kind="sort")
        self.scene[k] = Scene({key: (v[keep] if isinstance(v, np.ndarray) and key in KEYS else v) for key, v in self.scene[k].items()})
        assert self.w[k].remove_entities(ids) == len(ids)
        ring = self._past_ring(k)
        if ring is not None:
            ring.push_events(self.coherency(k), removal_events(pipeline.past_ids(ids)))

    def _past_ring(self, k):
        """The ring this timeslice's pushes into the past go to (events ride with the tick's record)."""
        if k + 1 >= self.n:
            return None
        return self.link[k][2] if self.devices[k] != self.devices[k + 1] else self.tl_in[k + 1]

    def _create(self, k, rows):
        """Cosmos::register_entity + components for the entities `rows` (1-row Scenes), and -- _entity_created_handler -- the
        creation pushed on into the past."""
        self.lifecycle_ticks.add(getattr(self, "_tick", -1))
        all_rows = rows[0]
        for row in rows[1:]:
            all_rows, _ = insert_row(all_rows, row)
        self.scene[k] = insert_rows(self.scene[k], all_rows) if self.host_rows else Scene(ids=np.sort(np.concatenate([self.scene[k].ids, all_rows.ids])))
        assert self.w[k].create_entities(all_rows) == len(rows)
        assert np.array_equal(self.w[k].entity_ids(), self.scene[k].ids)
        ring = self._past_ring(k)
        if ring is not None:
            ring.push_events(self.coherency(k), arrival_events(pipeline.past_ids(all_rows.ids)))

    def record_to_past(self, k):
        if k + 1 >= self.n:
            return
        c = self.coherency(k)
        if self.devices[k] == self.devices[k + 1]:
            self.w[k].record(self.tl_in[k + 1], c)                           # push_past_timestream straight into the neighbour's ring
            return
        import torch
        src, dst, tl_out = self.link[k]
        self.w[k].record(tl_out, c)
        n = tl_out.export_slot(c, src.data_ptr(), src.numel(), self.w[k].stream())
        self.w[k].sync()
        dst.copy_(src)                                                      # peer copy (NVLink)
        torch.cuda.synchronize(self.devices[k + 1])
        self.tl_in[k + 1].import_slot(c, n, dst.data_ptr(), dst.numel(), self.w[k + 1].stream())

    def fork_age(self, k):
        return self.w[k].fork_age()

    def step(self, k):
        self.w[k].step(DT)
        return self.w[k].interceptions()

    def state(self, k):
        return self.w[k].download(self.scene[k])

    def arrive(self, k, rows):
        """Travellers arrive: icx_entities_create -- the timeslice's own entities stay on the device (the host keeps only the
        row template: ids, hasP) -- and the arrival is pushed into the past."""
        self._create(k, rows)

    def has_future(self, k):
        """entity_has_timestream for every entity of timeslice k: the future neighbour pushes its entities at chain-link + 1."""
        return np.isin(self.scene[k].ids, pipeline.past_ids(self.scene[k - 1].ids), kind="sort").astype(np.uint8)

    def pmap(self, fn, ks):
        """One phase of a wall tick over the timeslices: concurrently (one host thread per timeslice; the C ABI calls release
        the GIL and every world has its own device / stream) when they live on different GPUs."""
        ks = list(ks)
        if len(set(self.devices)) == 1:
            return [fn(k) for k in ks]
        if self._pool is None:
            from concurrent.futures import ThreadPoolExecutor
            self._pool = ThreadPoolExecutor(max_workers=self.n)
        return list(self._pool.map(fn, ks))

    # ---- one parallel cycle step: shadow timeslice k, re-simulate up to its future neighbour, extract into it ----
    def cycle_step(self, k):
        tl, src, dst = self.tl_in[k], self.w[k], self.w[k - 1]
        if self.devices[k] not in self.par:
            self.par[self.devices[k]] = world.World(self.cap, host_timeslice=abi.NULL_TIMESLICE, device=self.devices[k])
        par = self.par[self.devices[k]]
        hf = self.has_future(k)
        t0 = time.perf_counter()
        tl.set_breakpoint()                                                 # link_timestreams(future stream of k)
        n = src.clone_into(par, hf)
        par_scene = None
        if self.keep_parallel_end:                                          # (tests) the parallel cosmos' entities, for the download below
            keep = hf.astype(bool) | ((self.scene[k].ids & 15) != 0)
            par_scene = Scene({key: (v[keep] if isinstance(v, np.ndarray) and key in KEYS else v) for key, v in self.scene[k].items()})
            assert par_scene.n == n
        ints = []
        c, target = self.coherency(k), self.coherency(k - 1)
        t1 = time.perf_counter()
        while c != target:
            tl.pop_events(c, abi.REPLAY_PARALLEL)                           # (the parallel cosmos consumes the old history's events too; not acted on here)
            par.replay(tl, c, abi.REPLAY_PARALLEL)                          # pop_timestream_at_breakpoint: consumes old history
            par.record(tl, c, at_breakpoint=True)                           # push_timestream_at_breakpoint: new history
            par.step(DT)
            ints.append(par.interceptions())
            c = (c + 1) & 0xFFFF
        t2 = time.perf_counter()
        end = par.download(par_scene) if par_scene is not None else None
        n_ex, n_miss = par.extract_into(dst)
        tl.clear_breakpoint()
        t3 = time.perf_counter()
        self.cycle_split = getattr(self, "cycle_split", []) + [(t1 - t0, t2 - t1, t3 - t2)]
        return {"cloned": n, "extracted": n_ex, "missing": n_miss, "interceptions": ints, "parallel_end": end}


def run(backend, wall_ticks, arrivals=(), observe=None, with_states=True, departures=()):
    """Drive the chain for `wall_ticks` wall ticks.  arrivals: [(wall tick, timeslice, row)]; departures: [(wall tick, timeslice, ids)].  observe(tick, phase, data) is called
    with what both back ends must agree on.  Returns the number of parallel cycles that ran."""
    cycles = 0
    n = backend.n
    pmap = getattr(backend, "pmap", None) or (lambda fn, ks: [fn(k) for k in ks])
    for t in range(wall_ticks):
        backend._tick = t
        for k in range(n):
            rows = [row for (ta, ka, row) in arrivals if ta == t and ka == k]
            if rows:
                backend.arrive(k, rows)
            for (td, kd, ids) in departures:
                if td == t and kd == k:
                    backend.depart(k, ids)
        # fork ageing of the PREVIOUS tick decided whether a resolution is needed; the cycle runs before this tick's networker
        need = getattr(backend, "_need", False)
        if need:
            t0 = time.perf_counter()
            results = [backend.cycle_step(k) for k in range(n - 1, 0, -1)]  # INIT to the past-most timeslice, then down the chain
            backend.cycle_seconds = getattr(backend, "cycle_seconds", []) + [time.perf_counter() - t0]
            cycles += 1
            backend._need = False
            if observe:
                observe(t, "cycle", results)
        pmap(backend.replay, range(n))                                      # phase 2
        pmap(backend.record_to_past, range(n))                              # phase 4
        ages = [tuple(a) for a in pmap(backend.fork_age, range(n))]         # phase 5
        backend._need = any(nd for _, nd in ages)
        ints = pmap(backend.step, range(n))                                 # physicser
        if observe:
            observe(t, "tick", {"ages": ages, "interceptions": ints, "states": [backend.state(k) for k in range(n)] if with_states else None})
    return cycles
