"""Free run of a scene on the device and on the CPU oracle side by side: first tick at which the states differ, with the
contact lists of both for that tick (the states before it are identical, so the lists are comparable).
    python tools/freerun_diff.py [bodies] [ticks] [seed]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import refs  # noqa: E402
from test_gpu_timetravel import make_chain  # noqa: E402
from intercession_b200 import world  # noqa: E402
from util import fields_diff  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 500
ticks = int(sys.argv[2]) if len(sys.argv) > 2 else 60
seed = int(sys.argv[3]) if len(sys.argv) > 3 else 31
s = make_chain(n, 1, seed=seed)[0]
w = world.World(s.n + 8)
w.upload(s)
cpu = s
for t in range(ticks):
    w.step(1 / 45)
    d = w.download(s)
    hd = w.contacts()
    out = refs.procedures_tick(refs.oracle(), "icxo", cpu, cap=64 * cpu.n)
    hc = out["hits"]
    diffs = {k: fields_diff(d[k], out[k]) for k in ("T", "P")}
    same = not any(diffs.values())
    kd = [(int(h["packetA"]), int(h["packetB"])) for h in hd]
    kc = [(int(h["packetA"]), int(h["packetB"])) for h in hc]
    st = w.stats()
    print(t, "same" if same else "DIFF", "contacts dev/cpu", len(hd), len(hc), "order same", kd == kc, "rounds", st.n_rounds, "attempts", st.n_attempts,
          "movers", st.n_movers, flush=True)
    if not same or kd != kc:
        sd, sc = set(kd), set(kc)
        print("  only on device:", sorted(sd - sc)[:20])
        print("  only on cpu   :", sorted(sc - sd)[:20])
        print("  fields that differ:", diffs)
        bad = np.nonzero((d.T["origin"] != out.T["origin"]).any(axis=1) | (d.P["linearVelocity"] != out.P["linearVelocity"]).any(axis=1))[0]
        print("  entities that differ:", bad[:20].tolist())
        pk_ent = np.repeat(np.arange(s.n), s.C["colliders"]["isActive"].sum(axis=1))
        for e in bad[:4]:
            pks = np.nonzero(pk_ent == e)[0].tolist()
            print("  entity", e, "packets", pks, "dev", d.T["origin"][e], "cpu", out.T["origin"][e])
            print("    device contacts:", [(i, k) for i, k in enumerate(kd) if k[0] in pks or k[1] in pks])
            print("    cpu contacts   :", [(i, k) for i, k in enumerate(kc) if k[0] in pks or k[1] in pks])
        for i, (a, b) in enumerate(zip(kd, kc)):
            if a != b:
                print("  first order difference at contact", i, "dev", kd[i:i + 6], "cpu", kc[i:i + 6])
                break
        for i in range(min(len(hd), len(hc))):
            if kd[i] == kc[i] and hd[i].tobytes() != hc[i].tobytes():
                print("  first contact with different values", i, kd[i], hd[i], hc[i])
                break
        np.savez(os.path.join(ROOT, "gpurun_out", "freerun_diff.npz"), T=cpu.T, P=cpu.P, C=cpu.C, ids=cpu.ids, hd=hd, hc=hc, tick=t)
        break
    cpu = type(s)({k: v for k, v in out.items() if k not in ("hits", "ints", "seconds")})
