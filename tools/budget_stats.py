"""How good is the predicted push?  Runs a scene to a tick, then prints, for the bodies static resolution moved, percentiles of
actual displacement / predicted push / budget.   python tools/budget_stats.py [bodies] [tick] [config3|config2]"""
import ctypes, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from intercession_b200 import scenes, world
n = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
ticks = [int(x) for x in (sys.argv[2] if len(sys.argv) > 2 else "60").split(",")]
name = sys.argv[3] if len(sys.argv) > 3 else "config2"
s = scenes.config3(n) if name == "config3" else scenes.config2(n)
w = world.World(s.n); w.upload(s)
done = 0
for t in ticks:
    w.step(1 / 45, t - done); done = t
    arr = [np.zeros(s.n, np.float32) for _ in range(3)]
    for k in range(3):
        w.lib.icx_debug_array(w.handle, k, arr[k].ctypes.data_as(ctypes.c_void_p), s.n)
    push, bm, disp = arr
    st = w.stats()
    moved = disp > 1e-6
    print(f"tick {t}: dirty {st.n_dirty} movers {st.n_movers} attempts {st.n_attempts} rounds {st.n_rounds} moved {int(moved.sum())} ms {w.last_step_ms():.3f}")
    q = [50, 90, 99, 99.9, 100]
    if moved.any():
        print("  disp pct", np.round(np.percentile(disp[moved], q), 4), " push pct", np.round(np.percentile(push[moved], q), 4), " bm pct", np.round(np.percentile(bm[moved], q), 4))
        ratio = disp[moved] / np.maximum(push[moved], 1e-6)
        print("  disp/push pct", np.round(np.percentile(ratio, q), 3), " frac disp > 0.05:", round(float((disp[moved] > 0.05).mean()), 4),
              " frac push > 0.025:", round(float((push[moved] > 0.025).mean()), 4), " used/budget p99:", round(float(np.percentile(disp[moved] / bm[moved], 99)), 3))
