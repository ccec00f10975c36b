"""Config 5 with time travellers (bench.py time_travel section), phase by phase: wall time of every networker phase per wall tick
and the resolve statistics of every timeslice, to see where a wall tick goes.
    python tools/run_timetravel.py [bodies] [timeslices] [travellers] [delay] [wall ticks]"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from intercession_b200 import abi, pipeline, scenes, timetravel  # noqa: E402

bodies = int(sys.argv[1]) if len(sys.argv) > 1 else 100000
n_ts = int(sys.argv[2]) if len(sys.argv) > 2 else 7
n_trav = int(sys.argv[3]) if len(sys.argv) > 3 else 64
delay = int(sys.argv[4]) if len(sys.argv) > 4 else 20
ticks = int(sys.argv[5]) if len(sys.argv) > 5 else 40
base = scenes.arena(bodies, seed=20261107, kinds=("sphere", "box"), temporal=True)
chain = []
for k in range(n_ts):
    sc = base.copy(); sc["ids"] = pipeline.past_ids(base.ids, k); chain.append(sc)
kinds = base.C["colliders"]["colliderType"][:, 0]
boxes = np.nonzero((kinds == abi.COLLIDER_BOX) & ((base.ids & 15) != 15))[0]
twins = boxes[np.linspace(0, len(boxes) - 1, n_trav).astype(np.int64)] if n_trav else []
arrivals = []
for i, twin in enumerate(twins):
    k = 1 + i % (n_ts - 1)
    row = timetravel.traveller_row(chain[k], int(twin), genesis=2000000 + i, timeslice=k)
    row.T["origin"][0, 0] += np.float32(0.3)
    arrivals.append((3 + (i // (n_ts - 1)) % 8, k, row))
be = timetravel.DeviceBackend(chain, delay, capacity_slack=n_trav + 8, host_rows=False)
acc = {}


def timed(label, f):
    def g(*a, **kw):
        t0 = time.perf_counter()
        r = f(*a, **kw)
        acc[label] = acc.get(label, 0.0) + time.perf_counter() - t0
        return r
    return g


for wk in be.w:
    wk.create_entities = timed("arrive.create_entities", wk.create_entities)
    wk.replay = timed("replay.icx_replay", wk.replay)
for tk in be.tl_in:
    if tk is not None:
        tk.fetch_entity = timed("replay.fetch_entity", tk.fetch_entity)
        tk.pop_events = timed("replay.pop_events", tk.pop_events)
        tk.push_events = timed("push_events", tk.push_events)
for name in ("replay", "record_to_past", "fork_age", "step", "arrive", "cycle_step"):
    f = getattr(be, name)

    def wrap(*a, _f=f, _n=name):
        t0 = time.perf_counter()
        r = _f(*a)
        acc[_n] = acc.get(_n, 0.0) + time.perf_counter() - t0
        return r
    setattr(be, name, wrap)


def observe(t, ph, d):
    if ph == "tick":
        st = [be.w[k].stats() for k in range(n_ts)]
        print(t, {k: round(v * 1e3, 2) for k, v in acc.items()}, "ints", [len(i) for i in d["interceptions"]],
              "ms", [round(be.w[k].last_step_ms(), 2) for k in range(n_ts)], "att", [s.n_attempts for s in st], "rounds", [s.n_rounds for s in st],
              "retries", [s.n_retries for s in st], "dirty", [s.n_dirty for s in st], "contacts", [s.n_contacts for s in st], flush=True)
    else:
        print(t, "CYCLE", [(r["cloned"], r["extracted"], sum(len(i) for i in r["interceptions"])) for r in d], round(acc.get("cycle_step", 0) * 1e3, 1), "ms",
              "per step (has_future+clone, re-simulation, extract) ms:", [tuple(round(x * 1e3, 2) for x in c) for c in be.cycle_split[-(n_ts - 1):]])
    acc.clear()


timetravel.run(be, ticks, arrivals, observe, with_states=False)
be.close()
