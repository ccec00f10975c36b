"""CPU suite, part 1: the oracle (oracle/icx_oracle.cpp, the CPU restatement) against the golden
fixtures frozen from the reference's own code by tests/golden/make_golden.py.  Needs neither a GPU
nor /root/reference, so it also pins the oracle on the GPU box.  Everything is bit-exact."""
import os

import numpy as np
import pytest

from intercession_b200 import abi, scenes
import katgen
import refs
from util import assert_hits_equal, fields_diff

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load(name):
    return np.load(os.path.join(GOLD, name))


def gm():
    import importlib.util
    spec = importlib.util.spec_from_file_location("make_golden", os.path.join(GOLD, "make_golden.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def oracle_run(s, ticks, host=0, coherency=0, euler=True):
    cur = s
    for t in range(ticks):
        cur = refs.procedures_tick(refs.oracle(), "icxo", cur, host=host, coherency=(coherency + t) & 0xFFFF, euler=euler)
    return cur


def test_kat_procedures(built):
    g = load("kat_procedures.npz")
    cases = katgen.all_cases()
    assert np.array_equal(gm().digest(cases), g["input_sha256"]), "KAT inputs drifted from the frozen fixture"
    got = katgen.run_kat(refs.oracle(), "icxo", cases)
    exp = g["out"]
    assert (exp["hit"] == 1).sum() > 150
    crashed = exp["hit"] == katgen.REF_CRASHED       # the reference divides by zero there (no known answer)
    assert crashed.sum() < 10
    d = fields_diff(got[~crashed], exp[~crashed])
    assert not d, d[:5]


@pytest.mark.parametrize("host", [0, abi.NULL_TIMESLICE])
@pytest.mark.parametrize("ticks", [1, 12])
def test_verbatim_arena(built, host, ticks):
    """The real reference objects (Cosmos + synchros + PhysicsDynamo) vs the oracle's loop."""
    g = load("verbatim_arena.npz")
    m = gm()
    s = m.verbatim_arena_scene()
    assert np.array_equal(m.scene_digest(s), g["input_sha256"])
    r = oracle_run(s, ticks, host=host, coherency=40)
    p = f"h{host}_t{ticks}_"
    for k in ("T", "P", "C"):
        d = fields_diff(r[k], g[p + k])
        assert not d, (k, d[:4])
    assert np.array_equal(r.tstate, g[p + "tstate"]) and np.array_equal(r.tcoh, g[p + "tcoh"])
    assert_hits_equal(r.hits, g[p + "hits"], ignore_responded=True)      # the instrumented relay cannot see `responded`
    assert not fields_diff(r.ints, g[p + "ints"])
    assert len(g[p + "hits"]) > 10


@pytest.mark.parametrize("push", [False, True])
def test_verbatim_case_study(built, push):
    """BASELINE.json configs[0]: the case-study scene for 1000 ticks on the reference CPU tick."""
    g = load("verbatim_case_study.npz")
    s = scenes.case_study(push=push)
    assert np.array_equal(gm().scene_digest(s), g[f"push{int(push)}_input_sha256"])
    cur, done = s, 0
    for ticks in (1, 10, 100, 1000):
        cur = oracle_run(cur, ticks - done, coherency=done)
        done = ticks
        assert not fields_diff(cur.T, g[f"push{int(push)}_t{ticks}_T"]), ticks
        assert not fields_diff(cur.P, g[f"push{int(push)}_t{ticks}_P"]), ticks
        assert_hits_equal(cur.hits, g[f"push{int(push)}_t{ticks}_hits"], ignore_responded=True)


def test_procedures_arena(built):
    g = load("procedures_arena.npz")
    m = gm()
    s = m.procedures_arena_scene()
    assert np.array_equal(m.scene_digest(s), g["input_sha256"])
    r = refs.procedures_tick(refs.oracle(), "icxo", s, euler=False, coherency=9)
    assert np.array_equal(r.T["origin"].view(np.uint32), g["pass_T_origin"].view(np.uint32))
    assert np.array_equal(r.P["velocity"].view(np.uint32), g["pass_P_velocity"].view(np.uint32))
    assert np.array_equal(r.P["angularVelocity"].view(np.uint32), g["pass_P_angularVelocity"].view(np.uint32))
    assert np.array_equal(r.tstate, g["pass_tstate"]) and np.array_equal(r.tcoh, g["pass_tcoh"])
    assert_hits_equal(r.hits, g["pass_hits"])
    assert not fields_diff(r.ints, g["pass_ints"])
    cur = oracle_run(s, 5)
    assert not fields_diff(cur.T, g["run5_T"])
    assert np.array_equal(cur.P["velocity"].view(np.uint32), g["run5_P_velocity"].view(np.uint32))
    assert_hits_equal(cur.hits, g["run5_hits"])
    assert np.array_equal(cur.tstate, g["run5_tstate"])
