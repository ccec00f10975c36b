#!/usr/bin/env python
"""Generate the golden fixtures under tests/golden/ by RUNNING THE REFERENCE ITSELF.

The reference ships no tests or golden vectors for this path, so the known
answers are frozen here from oracle/_ref/libicx_ref.so -- the reference's own unmodified sources
(source/physics/*, core/cosmos.cpp, events/event_broker.cpp) compiled headlessly by oracle/Makefile.
Run in the build container (needs /root/reference to build oracle/_ref):

    python tests/golden/make_golden.py

Fixtures (all outputs are the reference's; inputs are regenerated from seeds by the tests and
cross-checked against the stored input hash):
  kat_procedures.npz      per-pair intersectProcedures / responseProcedures answers (random +
                          degenerate cases of tests/katgen.py)
  verbatim_arena.npz      Cosmos + PhysicsSynchro/ColliderSynchro + PhysicsDynamo (the real
                          objects, uint16 ids) on a dense 200-body arena: state, hit list and
                          interception list after tick 1 and tick 12, server and parallel host ids
  verbatim_case_study.npz config-1 scene (moon cosmos floors + point + focal biped), state
                          after 1 / 10 / 100 / 1000 ticks
  procedures_arena.npz    reference procedures over plain arrays (wide ids) on a 1500-body dense
                          arena: one collision pass and 5 free-running ticks
"""
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from intercession_b200 import abi, scenes  # noqa: E402
import katgen  # noqa: E402
import refs  # noqa: E402


def digest(*arrays):
    h = hashlib.sha256()
    for a in arrays:
        h.update(np.ascontiguousarray(a).tobytes())
    return np.frombuffer(h.digest(), np.uint8).copy()


def scene_digest(s):
    return digest(s.ids, s.T, s.P, s.hasP, s.C, s.tstate, s.tcoh)


# ---- scene recipes shared with the tests (tests/test_golden.py imports these) ----
def verbatim_arena_scene():
    s = scenes.arena(200, seed=4242, fill_side=8.3, spacing=1.0, jitter=0.22, kinds=("sphere", "box", "aabb"), temporal=True)
    s.ids[6:90] |= 1                            # chain-link 1
    s.ids[120:150] = (s.ids[120:150] & ~np.uint32(15)) | 2
    s.ids[190:206] |= 15                        # non-temporal
    s.tstate[60:130] = abi.TS_FORKED
    s.tstate[130:140] = abi.TS_FORKING
    s.tcoh[60:140] = 3
    s.hasP[50:54] = 0
    s.C["colliders"]["useBehaviorsResponse"][20:40, 0] = 1
    s.C["colliders"]["collisionType"][40:44, 0] = abi.COLLISION_NOOP
    return s


def procedures_arena_scene():
    s = scenes.arena(1500, seed=777, fill_side=16.0, spacing=1.05, jitter=0.25, kinds=("sphere", "box", "aabb"), temporal=True)
    s.ids[10:700] |= 1
    s.tstate[300:900] = abi.TS_FORKED
    return s


def state_arrays(prefix, s):
    return {prefix + "T": s.T, prefix + "P": s.P, prefix + "C": s.C, prefix + "tstate": s.tstate, prefix + "tcoh": s.tcoh,
            prefix + "hits": s.hits, prefix + "ints": s.ints}


def main():
    lib = refs.ref()
    assert lib is not None, "oracle/_ref/libicx_ref.so missing: run `make -C oracle` where /root/reference exists"
    out = {}

    # 1. per-procedure KATs
    cases = katgen.all_cases()
    kat = katgen.run_kat_isolated(lib, "icxref", cases)
    np.savez_compressed(os.path.join(HERE, "kat_procedures.npz"), input_sha256=digest(cases), out=kat)
    print("kat_procedures:", len(cases), "cases,", int((kat["hit"] == 1).sum()), "hits,",
          int((kat["hit"] == katgen.REF_CRASHED).sum()), "on which the reference itself crashes")

    # 2. verbatim reference objects on a dense arena
    s = verbatim_arena_scene()
    d = {"input_sha256": scene_digest(s)}
    for host in (0, abi.NULL_TIMESLICE):
        for ticks in (1, 12):
            r = refs.verbatim_run(s, n_ticks=ticks, host=host, coherency=40)
            d.update(state_arrays(f"h{host}_t{ticks}_", r))
            print("verbatim_arena host", host, "ticks", ticks, "hits", len(r.hits), "ints", len(r.ints))
    np.savez_compressed(os.path.join(HERE, "verbatim_arena.npz"), **d)

    # 3. case-study scene, verbatim objects
    d = {}
    for push in (False, True):
        s = scenes.case_study(push=push)
        d[f"push{int(push)}_input_sha256"] = scene_digest(s)
        for ticks in (1, 10, 100, 1000):
            r = refs.verbatim_run(s, n_ticks=ticks)
            d[f"push{int(push)}_t{ticks}_T"] = r.T
            d[f"push{int(push)}_t{ticks}_P"] = r.P
            d[f"push{int(push)}_t{ticks}_hits"] = r.hits
        print("case_study push", push, "origin after 1000:", r.T["origin"][2:].tolist())
    np.savez_compressed(os.path.join(HERE, "verbatim_case_study.npz"), **d)

    # 4. reference procedures over plain arrays, wide ids
    s = procedures_arena_scene()
    d = {"input_sha256": scene_digest(s)}
    r = refs.procedures_tick(lib, "icxref", s, euler=False, coherency=9)
    d.update({"pass_T_origin": r.T["origin"], "pass_P_velocity": r.P["velocity"], "pass_P_angularVelocity": r.P["angularVelocity"],
              "pass_tstate": r.tstate, "pass_tcoh": r.tcoh, "pass_hits": r.hits, "pass_ints": r.ints})
    cur = s
    for t in range(5):
        cur = refs.procedures_tick(lib, "icxref", cur, coherency=t)
    d.update({"run5_T": cur.T, "run5_P_velocity": cur.P["velocity"], "run5_P_angularVelocity": cur.P["angularVelocity"],
              "run5_hits": cur.hits, "run5_tstate": cur.tstate})
    print("procedures_arena pass hits", len(r.hits), "ints", len(r.ints), "run5 hits", len(cur.hits))
    np.savez_compressed(os.path.join(HERE, "procedures_arena.npz"), **d)

    for f in sorted(os.listdir(HERE)):
        if f.endswith(".npz"):
            print(f, os.path.getsize(os.path.join(HERE, f)), "bytes")


if __name__ == "__main__":
    main()
