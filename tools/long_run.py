"""Long run of a config-2/3 scene (100 warm-up + 1000 timed ticks): per-window tick time, phase split and
resolve statistics, so the regime after the bodies have reached the floor is measured, not guessed.
    python tools/long_run.py [bodies] [ticks] [window] [config3|config2]
Prints one JSON line per window and a final summary line."""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from intercession_b200 import scenes, world

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1048576
ticks = int(sys.argv[2]) if len(sys.argv) > 2 else 1100
window = int(sys.argv[3]) if len(sys.argv) > 3 else 25
name = sys.argv[4] if len(sys.argv) > 4 else "config3"
budget_s = float(os.environ.get("ICX_LONG_BUDGET_S", "600"))
s = scenes.config3(n) if name == "config3" else scenes.config2(n)
w = world.World(s.n)
w.upload(s)
w.profile(True)
t_start = time.perf_counter()
done = 0
prev = {k: 0.0 for k in ("integrate_prep", "grid_sort", "pairs", "narrow", "resolve")}
rows = []
while done < ticks:
    k = min(window, ticks - done)
    t0 = time.perf_counter()
    w.step(1 / 45, k)
    wall = time.perf_counter() - t0
    ms = w.last_step_ms()
    done += k
    st = w.stats()
    ph, _, _ = w.phases()
    row = {"tick": done, "ms_per_tick": round(ms / k, 4), "wall_ms_per_tick": round(1e3 * wall / k, 4),
           "phases_ms": {p: round((ph[p] - prev[p]) / k, 4) for p in ph},
           "candidates": st.n_candidates, "dirty": st.n_dirty, "contacts": st.n_contacts, "rounds": st.n_rounds,
           "attempts": st.n_attempts, "retries": st.n_retries, "margin": round(st.margin, 4), "cell": round(st.cell_size, 4),
           "max_neighbors": st.max_neighbors, "movers": st.n_movers}
    import ctypes
    import numpy as np
    cn = np.zeros(48, np.uint32)
    w.lib.icx_debug_counters(w.handle, cn.ctypes.data_as(ctypes.c_void_p))
    row["restarts_grew"], row["restarts_budget"], row["xpairs"], row["xmax"], row["pool"] = int(cn[30]), int(cn[31]), int(cn[26]), round(float(cn[27:28].view(np.float32)[0]), 3), int(cn[29])
    prev = dict(ph)
    rows.append(row)
    print(json.dumps(row), flush=True)
    if time.perf_counter() - t_start > budget_s:
        print(json.dumps({"stopped": "time budget", "at_tick": done}), flush=True)
        break
print(json.dumps({"summary": name, "bodies": n, "ticks": done,
                  "mean_ms_after_100": round(sum(r["ms_per_tick"] for r in rows if r["tick"] > 100) / max(1, sum(1 for r in rows if r["tick"] > 100)), 4)}))
