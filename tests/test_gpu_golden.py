"""GPU parity against the COMMITTED golden fixtures (tests/golden/*.npz, frozen from the reference's
own code): no oracle in the loop, nothing read from /root/reference."""
import os

import numpy as np
import pytest

from intercession_b200 import abi, scenes, world
from util import assert_hits_equal, fields_diff

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
DT = 1.0 / 45


def gm():
    import importlib.util
    spec = importlib.util.spec_from_file_location("make_golden", os.path.join(GOLD, "make_golden.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def test_collision_pass_bit_exact_vs_reference_procedures(built):
    g = np.load(os.path.join(GOLD, "procedures_arena.npz"))
    s = gm().procedures_arena_scene()
    with world.World(s.n) as w:
        w.upload(s)
        w.set_coherency(9)
        w.collision_engage()
        got, hits, ints = w.download(s), w.contacts(), w.interceptions()
    assert np.array_equal(got.T["origin"].view(np.uint32), g["pass_T_origin"].view(np.uint32))
    assert np.array_equal(got.P["velocity"].view(np.uint32), g["pass_P_velocity"].view(np.uint32))
    assert np.array_equal(got.P["angularVelocity"].view(np.uint32), g["pass_P_angularVelocity"].view(np.uint32))
    assert np.array_equal(got.tstate, g["pass_tstate"]) and np.array_equal(got.tcoh, g["pass_tcoh"])
    assert_hits_equal(hits, g["pass_hits"])
    assert not fields_diff(ints, g["pass_ints"])


@pytest.mark.parametrize("host", [0, abi.NULL_TIMESLICE])
def test_full_tick_vs_reference_objects(built, host):
    """One whole tick against Cosmos + PhysicsDynamo: identical hit / interception sequences, linear
    state within 2e-6 (orientation went through device sin/cos, <= 2 ulp, before the collision pass)."""
    g = np.load(os.path.join(GOLD, "verbatim_arena.npz"))
    s = gm().verbatim_arena_scene()
    with world.World(s.n, host_timeslice=host) as w:
        w.upload(s)
        w.set_coherency(40)
        w.step(DT)
        got, hits, ints = w.download(s), w.contacts(), w.interceptions()
    p = f"h{host}_t1_"
    eh = g[p + "hits"]
    assert len(hits) == len(eh)
    for k in ("entityA", "entityB", "slotA", "slotB", "packetA", "packetB"):
        assert np.array_equal(hits[k], eh[k]), k
    assert np.abs(hits["depth"] - eh["depth"]).max() < 2e-6
    assert np.abs(hits["normal"] - eh["normal"]).max() < 2e-5
    assert not fields_diff(ints, g[p + "ints"])
    assert np.array_equal(got.tstate, g[p + "tstate"]) and np.array_equal(got.tcoh, g[p + "tcoh"])
    assert np.abs(got.T["origin"] - g[p + "T"]["origin"]).max() < 2e-6
    assert np.abs(got.P["velocity"] - g[p + "P"]["velocity"]).max() < 2e-5


@pytest.mark.parametrize("push", [False, True])
def test_case_study_1000_ticks(built, push):
    """BASELINE.json configs[0]: the case-study scene, 1000 ticks, against the reference CPU tick."""
    g = np.load(os.path.join(GOLD, "verbatim_case_study.npz"))
    s = scenes.case_study(push=push)
    with world.World(s.n) as w:
        w.upload(s)
        done = 0
        for ticks in (1, 10, 100, 1000):
            w.step(DT, ticks - done)
            done = ticks
            got = w.download(s)
            eT, eP = g[f"push{int(push)}_t{ticks}_T"], g[f"push{int(push)}_t{ticks}_P"]
            assert np.abs(got.T["origin"] - eT["origin"]).max() < 1e-4, ticks
            assert np.abs(got.P["velocity"] - eP["velocity"]).max() < 1e-4, ticks
            assert np.abs(got.T["orientation"] - eT["orientation"]).max() < 1e-5, ticks
        hits = w.contacts()
    eh = g[f"push{int(push)}_t1000_hits"]
    assert len(hits) == len(eh) and np.array_equal(hits["packetA"], eh["packetA"]) and np.array_equal(hits["packetB"], eh["packetB"])
