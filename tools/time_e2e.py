"""Where an end-to-end step (host AoS in, tick, host AoS out) spends its time: python tools/time_e2e.py [bodies]"""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from intercession_b200 import scenes, world
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1048576
s = scenes.config3(n)
w = world.World(s.n); w.upload(s); w.step(1 / 45, 3)
import bench
keep = []
if len(sys.argv) > 2 and sys.argv[2] == "pinned":
    for k in ("ids", "T", "P", "C", "hasP", "tstate", "tcoh"):
        s[k], t = bench.pinned_like(s[k]); keep.append(t)
T, tT = bench.pinned_like(s.T); P, tP = bench.pinned_like(s.P)
acc = {"upload": 0.0, "step": 0.0, "download": 0.0, "contacts": 0.0}
for it in range(6):
    t0 = time.perf_counter(); w.upload(s)
    t1 = time.perf_counter(); w.step(1 / 45)
    t2 = time.perf_counter(); w.download_into(T=T, P=P)
    t3 = time.perf_counter(); c = w.contacts()
    t4 = time.perf_counter()
    if it:
        for k, v in zip(acc, (t1 - t0, t2 - t1, t3 - t2, t4 - t3)):
            acc[k] += v / 5
print({k: round(v * 1e3, 2) for k, v in acc.items()}, "ms;", "pinned" if keep else "pageable", "host arrays")
