"""Tick the config-3 scene a few times (for ncu captures): python tools/run_ticks.py [bodies] [ticks]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from intercession_b200 import scenes, world
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1048576
s = scenes.config3(n)
w = world.World(s.n); w.upload(s)
w.step(1 / 45, int(sys.argv[2]) if len(sys.argv) > 2 else 6)
st = w.stats()
print("ticks done; candidates", st.n_candidates, "dirty", st.n_dirty, "contacts", st.n_contacts, "ms", w.last_step_ms())
