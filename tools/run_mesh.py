"""The static-mesh pass on the headline scene (for ncu): python tools/run_mesh.py [bodies] [ticks]"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from intercession_b200 import scenes, world  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1048576
ticks = int(sys.argv[2]) if len(sys.argv) > 2 else 4
s = scenes.config3(n)
side = float(s["side"])
tv, tt = scenes.torus_mesh(centre=(0.0, 0.0, 0.0), major=0.25 * side, minor=0.05 * side, n_major=192, n_minor=48)
gv, gt = scenes.height_mesh(side, 256, y=-0.45 * side, bump=0.2, seed=3)
verts = np.concatenate([tv, gv]); tris = np.concatenate([tt, gt + len(tv)]).astype(np.uint32)
with world.World(s.n) as w:
    w.upload(s)
    w.upload_mesh(verts, tris)
    for t in range(ticks):
        w.step(1 / 45)
        print("tick", t, "mesh us", round(w.mesh_kernel_us(), 1), "contacts", len(w.mesh_contacts()), flush=True)
