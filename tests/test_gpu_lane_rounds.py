"""GPU parity of the LANE-PARALLEL form of the ordered resolve (one work item per lane: serial SAT, thread-local manifold
scratch, scalar cursor walk -- csrc/icx_resolve.cu, process_pair<true> / round_execute_lanes).  Production uses it only for
rounds with >= 4 items per warp, which small scenes never reach, so ICX_F_LANE_ROUNDS forces it for every round and the
scenes of test_gpu_physics.py (dense arenas, rays + springs, interceptions on both hosts, world-line batch, missing physics,
the sphere-sphere flag) are checked bit for bit against the oracle again, plus free-running ticks against the warp form."""
import numpy as np
import pytest

from intercession_b200 import abi, scenes, world
from test_gpu_physics import DT, dense, lockstep_collision

pytestmark = pytest.mark.gpu
LANES = abi.F_LANE_ROUNDS


@pytest.mark.parametrize("seed", [1, 2])
def test_lane_rounds_dense(built, seed):
    exp, st = lockstep_collision(dense(4000, seed), flags=LANES)
    assert len(exp.hits) > 500 and st.n_rounds >= 2


def test_lane_rounds_sphere_sphere_flag(built):
    lockstep_collision(dense(3000, 5), flags=abi.F_SPHERE_SPHERE | LANES)


@pytest.mark.parametrize("host", [0, abi.NULL_TIMESLICE])
def test_lane_rounds_interceptions(built, host):
    s = dense(3000, 3, temporal=True)
    s.ids[10:1500] |= 1
    s.ids[10:1500] &= ~np.uint32(14)            # chain-link 1
    s.ids[2000:2500] = (s.ids[2000:2500] & ~np.uint32(15)) | 2
    s.tstate[1000:2200] = abi.TS_FORKED
    s.tstate[2200:2400] = abi.TS_FORKING
    exp, _ = lockstep_collision(s, host=host, coherency=77, flags=LANES)
    assert len(exp.ints) > 50


def test_lane_rounds_case_study_ray_spring(built):
    n_hits = 0
    for y in (-3.5, -3.0, -1.0):            # box+legs on the floor / legs (ray spring) only / box on the ball
        s = scenes.case_study(push=True)
        s.T["origin"][3] = (12, y, 0)
        s.P["velocity"][3] = (0.3, -1.0, 0.2)
        exp, _ = lockstep_collision(s, flags=LANES)
        n_hits += len(exp.hits)
    assert n_hits >= 4


def test_lane_rounds_worldline_batch(built):
    s = scenes.worldline_batch(8, 256, seed=9)
    lockstep_collision(s, host=abi.NULL_TIMESLICE, stride=256, flags=LANES)


def test_lane_rounds_missing_physics_and_noop(built):
    s = dense(2000, 8)
    s.hasP[100:400] = 0
    s.C["colliders"]["collisionType"][400:700, 0] = abi.COLLISION_NOOP
    s.C["colliders"]["useBehaviorsResponse"][600:900, 0] = 1
    s.C["colliders"]["isActive"][900:950, 0] = 0
    s.C["colliders"]["colliderType"][950:1000, 0] = abi.COLLIDER_NONE
    lockstep_collision(s, flags=LANES)


def test_lane_rounds_equal_warp_rounds_over_ticks(built):
    """40 free-running ticks of a dense arena: both forms of the resolve give the same bits (state, contacts, counters)."""
    s = scenes.arena(6000, seed=33, spacing=1.0, jitter=0.2, fill_side=28.0)
    out = []
    for flags in (0, LANES):
        with world.World(s.n, flags=flags) as w:
            w.upload(s)
            w.step(DT, 40)
            out.append((w.download(s), w.contacts(), w.stats().n_contacts))
    (a, ha, na), (b, hb, nb) = out
    assert na == nb and na > 100
    for f in a.T.dtype.names:
        assert np.array_equal(a.T[f].view(np.uint32), b.T[f].view(np.uint32)), f
    for f in ("velocity", "angularVelocity", "acceleration", "angularAcceleration"):
        assert np.array_equal(a.P[f].view(np.uint32), b.P[f].view(np.uint32)), f
    assert np.array_equal(ha.view(np.uint8), hb.view(np.uint8))
