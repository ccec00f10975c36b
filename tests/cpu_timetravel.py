"""TEST INFRASTRUCTURE: the CPU side of intercession_b200/timetravel.py -- the same schedule driven over the oracle procedures
(tick, interception predicate + handlers, fork ageing), the reference's REAL EntityTimestreamMap / TsBreakpointQueue when
oracle/_ref is built (else their restatement), and oracle/clone_oracle.py (pinned against the reference's entity loops)."""
import os
import sys

import numpy as np

from intercession_b200 import abi, pipeline, timetravel
from intercession_b200.scenes import Scene
import refs

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "oracle"))
import clone_oracle  # noqa: E402


class CpuBackend:
    def __init__(self, scenes_per_slice, delay, base=100):
        self.n, self.delay = len(scenes_per_slice), delay
        self.scene = [s.copy() for s in scenes_per_slice]
        self.coh = [pipeline.start_coherency(k, self.n, delay, base) for k in range(self.n)]
        # the reference's Entity is 16 bits wide (genesis < 255): its real containers serve the small scenarios, their restatement
        # in the oracle (pinned to them by tests/test_timeline_oracle.py) the larger ones
        small = max(s.n for s in scenes_per_slice) + 8 < 250 and all(((s.ids >> 4) & 0xFFFFFF).max() < 250 for s in scenes_per_slice)
        lib, prefix = (refs.ref(), "icxref") if (refs.ref() is not None and small) else (refs.oracle(), "icxo")
        self.containers = prefix
        self.tl_in = [None] + [refs.CpuTimeline(lib, prefix) for _ in range(1, self.n)]
        self.stash = {}

    def close(self):
        for t in self.tl_in:
            if t is not None:
                t.close()

    def coherency(self, k):
        return self.coh[k]

    def _has(self, k, entity):
        ids = self.scene[k].ids
        j = int(np.searchsorted(ids, entity))
        return j < len(ids) and ids[j] == entity

    def replay(self, k):
        if k == 0:
            return False
        evs = self.tl_in[k].pop_events(self.coh[k])
        new = [int(e["entity"]) for e in evs if e["type"] == abi.EVENT_ENTITY_CREATED and not self._has(k, e["entity"])]
        if new:
            self._create(k, [self.stash.pop((k, ent)) for ent in new])
        gone = [int(e["entity"]) for e in evs if e["type"] == abi.EVENT_ENTITY_REMOVED and self._has(k, e["entity"])]
        if gone:
            self.depart(k, gone)
        self.scene[k], n = self.tl_in[k].replay(self.scene[k], self.coh[k])
        return n > 0

    def depart(self, k, ids):
        ids = np.asarray(ids, np.uint32)
        keep = ~np.isin(self.scene[k].ids, ids, kind="sort")
        self.scene[k] = Scene({key: (v[keep] if isinstance(v, np.ndarray) and key in timetravel.KEYS else v) for key, v in self.scene[k].items()})
        if k + 1 < self.n:
            self.tl_in[k + 1].push_events(self.coh[k], timetravel.removal_events(pipeline.past_ids(ids)))

    def _create(self, k, rows):
        """register_entity + the serialised components the arrival message carries (here: stashed beside the event), then the
        creation pushed on into the past (_entity_created_handler / _jump_arrival_handler)."""
        for row in rows:
            self.scene[k], _ = timetravel.insert_row(self.scene[k], row)
        if k + 1 < self.n:
            ids = np.array([int(r.ids[0]) for r in rows], np.uint32)
            past = pipeline.past_ids(ids)
            self.tl_in[k + 1].push_events(self.coh[k], timetravel.arrival_events(past))
            for row, pid in zip(rows, past):
                r = row.copy(); r["ids"] = np.array([pid], np.uint32); r["tstate"][:] = abi.TS_MERGED; r["tcoh"][:] = 0
                self.stash[(k + 1, int(pid))] = r

    def record_to_past(self, k):
        if k + 1 < self.n:
            self.tl_in[k + 1].record(self.scene[k], self.coh[k], link_inc=True)

    def fork_age(self, k):
        self.scene[k], promoted, need = refs.oracle_fork_age(self.scene[k], self.coh[k])
        return promoted, need

    def step(self, k):
        out = refs.procedures_tick(refs.oracle(), "icxo", self.scene[k], host=k, coherency=self.coh[k])
        ints = out["ints"]
        self.scene[k] = Scene({key: v for key, v in out.items() if key not in ("hits", "ints", "seconds")})
        self.coh[k] = (self.coh[k] + 1) & 0xFFFF
        return ints

    def state(self, k):
        return self.scene[k].copy()                                         # extract mutates the future neighbour's scene in place

    def arrive(self, k, rows):
        self._create(k, rows)

    def has_future(self, k):
        return np.isin(self.scene[k].ids, pipeline.past_ids(self.scene[k - 1].ids), kind="sort").astype(np.uint8)

    def cycle_step(self, k):
        tl = self.tl_in[k]
        hf = self.has_future(k)
        tl.breakpoint(True)
        par, c = clone_oracle.load_and_link(self.scene[k], self.coh[k], hf)
        n = par.n
        ints = []
        target = self.coh[k - 1]
        while c != target:
            tl.pop_events(c, mode=abi.REPLAY_PARALLEL)
            tl.replay(par, c, mode=abi.REPLAY_PARALLEL)
            tl.record(par, c, at_breakpoint=True, link_inc=False)
            out = refs.procedures_tick(refs.oracle(), "icxo", par, host=abi.NULL_TIMESLICE, coherency=c)
            ints.append(out["ints"])
            par = Scene({key: v for key, v in out.items() if key not in ("hits", "ints", "seconds")})
            c = (c + 1) & 0xFFFF
        n_ex, n_miss = clone_oracle.extract_entity_updates(par, self.scene[k - 1], self.coh[k - 1], k - 1)
        tl.breakpoint(False)
        return {"cloned": n, "extracted": n_ex, "missing": n_miss, "interceptions": ints, "parallel_end": par}
