"""A few k_projectile_update launches over a resident world (for ncu captures): python tools/run_projectiles.py [bodies] [reps]"""
import sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from intercession_b200 import abi, scenes, world
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1048576
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 4
s = scenes.config3(n)
w = world.World(s.n); w.upload(s)
rng = np.random.default_rng(5)
p = np.zeros(s.n, abi.projectile_dt)
p["currLifetime"] = rng.uniform(0, 200, s.n); p["maxLifetime"] = 300.0; p["hz"] = rng.uniform(0.5, 4.0, s.n)
p["scaleMin"] = rng.uniform(0.3, 0.5, s.n); p["scaleMax"] = p["scaleMin"] + 0.1
w.projectiles_upload(np.arange(s.n, dtype=np.uint32), p)
for k in range(reps):
    w.projectiles_fixed_update(1.0 / 60.0)
w.sync()
print("projectile updates done;", s.n, "projectiles, last kernel", w.projectiles_kernel_ms() * 1e3, "us, condemned", len(w.projectiles_condemned()))
w.close()
