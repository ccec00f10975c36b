import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__)))); sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import numpy as np
from intercession_b200 import scenes, world, abi
import refs
np.set_printoptions(linewidth=200)
s = scenes.arena(40, seed=1, fill_side=8.0, spacing=1.05, jitter=0.25, kinds=("box",))
with world.World(s.n) as w:
    w.upload(s); w.collision_engage(); hits = w.contacts(); st = w.stats()
exp = refs.procedures_tick(refs.oracle(), "icxo", s, euler=False)
print(len(hits), len(exp.hits), st.n_rounds, st.n_candidates)
for g, e in zip(hits, exp.hits):
    same = g.tobytes() == e.tobytes()
    if not same:
        print("GPU", g); print("EXP", e)
