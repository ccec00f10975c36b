"""Device-side timeline of the ordered resolve (icx_debug_trace) on the 1 M scene: prints per-round times."""
import sys, os, ctypes
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))  # repo root
from intercession_b200 import scenes, world
s = scenes.config3(int(sys.argv[1]) if len(sys.argv) > 1 else 1048576)
w = world.World(s.n); w.upload(s); w.step(1/45, int(sys.argv[2]) if len(sys.argv) > 2 else 4); w.trace(True)
for k in range(2):
    w.step(1/45); st = w.stats()
    print("tick", k, "dirty", st.n_dirty, "rounds", st.n_rounds, "attempts", st.n_attempts, "contacts", st.n_contacts, "ms", w.last_step_ms())
    cn = np.zeros(48, np.uint32); w.lib.icx_debug_counters(w.handle, cn.ctypes.data_as(ctypes.c_void_p)); pr = cn[32:]
    if pr[0]:
        print("  items", pr[0], "cycles/first pair", 16*pr[1]/pr[0], "cycles/chain", 16*pr[2]/pr[0], "| filter calls", pr[3], "follow-up pairs", pr[4],
              "| load+SAT cycles per exact pair", 16*pr[5]/max(pr[0]+pr[4],1), "not separated", pr[6],
              "| hits", pr[8], "cycles to hit decided", 16*pr[7]/max(pr[8],1), "(box-box:", pr[12], 16*pr[11]/max(pr[12],1), ") response cycles", 16*pr[9]/max(pr[8],1),
              "| rr response: first run", 16*pr[13]/max(pr[14],1), "second run", 16*pr[10]/max(pr[14],1), "n", pr[14])
    prev = 0
    for t, tag in w.trace(True):
        print(f"  {t/1e3:9.1f} us (+{(t-prev)/1e3:7.1f})  tag {tag>>16:x} arg {tag&0xffff}")
        prev = t
