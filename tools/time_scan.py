"""Kernel time of k_divergence_scan (CUDA events, icx_timeline_kernel_ms): one 1 M-entity world against a full record, and the
config-4 batch (4096 world-lines x 1024 bodies against one shared 1024-body record).  ICX_LIB selects a tuning variant."""
import sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from intercession_b200 import abi, scenes, world
reps = 20
s = scenes.config3(1048576)
w = world.World(s.n); w.upload(s)
tl = world.Timeline(6, s.n); tl.profile(True)
for k in range(6):
    w.record(tl, 1000 + k)
ms = []
for k in range(reps):
    f, c, sc = w.divergence_scan(tl, 1000 + k % 6, 1); ms.append(tl.kernel_ms(2))
print("check", int(c.sum()), float(sc.sum()), end=" | ")
print(os.environ.get("ICX_LIB", "product"), "single 1M: us", round(float(np.mean(ms[2:])) * 1e3, 2), "min", round(float(np.min(ms)) * 1e3, 2))
tl.close(); w.close()
per, nwl = 1024, 4096
b = scenes.worldline_batch(nwl, per, seed=1, perturb=False)
wb = world.World(b.base.n, host_timeslice=abi.NULL_TIMESLICE); wb.upload(b.base)
tb = world.Timeline(3, per); tb.profile(True)
wb.step(1 / 45, 3); wb.record(tb, 1)                # the candidates have not moved: every body is 3 ticks off the record
wc = world.World(b.n, worldline_stride=per, host_timeslice=abi.NULL_TIMESLICE); wc.upload(b)
flush = world.Timeline(2, b.n)
ms = []
for k in range(reps):
    wc.record(flush, k)                       # 400 MB through L2 between scans
    f, c, sc = wc.divergence_scan(tb, 1, nwl); ms.append(tb.kernel_ms(2))
print("check", int(c.sum()), float(sc.sum()), int(f.sum()), end=" | ")
print(os.environ.get("ICX_LIB", "product"), "batch 4096x1024: us", round(float(np.mean(ms[2:])) * 1e3, 2), "min", round(float(np.min(ms)) * 1e3, 2))
