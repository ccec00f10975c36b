"""Config-4 world-line batch: phase split of the tick and device trace of the resolve.  python tools/run_batch.py [worldlines] [ticks]"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from intercession_b200 import abi, scenes, world
n_wl = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
ticks = int(sys.argv[2]) if len(sys.argv) > 2 else 8
s = scenes.worldline_batch(n_wl, 1024, seed=20261106)
w = world.World(s.n, worldline_stride=1024, host_timeslice=abi.NULL_TIMESLICE)
w.upload(s)
w.step(1 / 45, 2)
w.profile(True)
w.step(1 / 45, ticks)
ph, nt, _ = w.phases()
st = w.stats(); tot = w.totals()
print(f"worldlines {n_wl} bodies {s.n} ticks {nt} ms/tick {w.last_step_ms() / ticks:.4f} phases " + " ".join(f"{k} {v / nt:.4f}" for k, v in ph.items()))
print(f"  dirty {st.n_dirty} contacts {st.n_contacts} rounds {st.n_rounds} attempts {st.n_attempts} movers {st.n_movers} mean rounds {tot['rounds'] / nt:.1f} attempts {tot['attempts'] / nt:.2f}")
w.trace(True)
w.step(1 / 45)
prev = 0
for t, tag in w.trace(True):
    print(f"  {t / 1e3:9.1f} us (+{(t - prev) / 1e3:7.1f})  tag {tag >> 16:x} arg {tag & 0xffff}")
    prev = t
