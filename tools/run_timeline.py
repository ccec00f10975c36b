"""Record / replay / divergence-scan launches over a resident world (for ncu captures):
    python tools/run_timeline.py [bodies] [reps]
The ring holds 6 slots of 100 B per entity (larger than L2 at 1 M entities), cycled like bench.py's timeline_kernels."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from intercession_b200 import abi, scenes, world
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1048576
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 6
depth, c0 = 6, 1000
s = scenes.config3(n)
w = world.World(s.n); w.upload(s)
w.step(1 / 45, 1)
tl = world.Timeline(depth, s.n)
for k in range(reps):
    w.record(tl, (c0 + k % depth) & 0xFFFF)
for k in range(reps):
    c = (c0 + k % depth) & 0xFFFF
    w.record(tl, c)
    assert w.replay(tl, c, abi.REPLAY_SERVER)
for k in range(depth):
    w.record(tl, (c0 + k) & 0xFFFF)
for k in range(reps):
    f, cnt, sc = w.divergence_scan(tl, (c0 + k % depth) & 0xFFFF, 1)
w.sync()
print("timeline launches done; entities", s.n, "divergent", int(f[0]), "count", int(cnt[0]))
tl.close(); w.close()
