"""GPU parity tests: the CUDA path through the C ABI against the oracle (CPU restatement, itself
pinned bit-exact to the compiled reference by test_oracle_vs_ref.py) on identical inputs."""
import numpy as np
import pytest

from intercession_b200 import abi, scenes, world
import refs
from util import assert_hits_equal, assert_state_equal, fields_diff, ulp_diff

pytestmark = pytest.mark.gpu
DT = 1.0 / 45


def oracle_tick(s, **kw):
    return refs.procedures_tick(refs.oracle(), "icxo", s, **kw)


def dense(n, seed, **kw):
    side = 2.0 + 1.1 * round(n ** (1 / 3) + 1)
    kw.setdefault("kinds", ("sphere", "box", "aabb"))
    return scenes.arena(n, seed=seed, fill_side=side, spacing=1.05, jitter=0.25, **kw)


def lockstep_collision(s, host=0, flags=0, stride=0, coherency=0, margin=0.0, check_pairs=True):
    """Collision pass only, GPU vs oracle from the same post-integration state: everything bit-exact."""
    with world.World(s.n, host_timeslice=host, flags=flags, worldline_stride=stride, margin=margin) as w:
        w.upload(s)
        w.set_coherency(coherency)
        w.collision_engage()
        got = w.download(s)
        hits, ints, pairs, st = w.contacts(), w.interceptions(), w.pairs(), w.stats()
    exp = oracle_tick(s, euler=False, host=host, flags=flags, stride=stride, coherency=coherency)
    assert_hits_equal(hits, exp.hits)
    assert not fields_diff(ints, exp.ints), "interceptions differ"
    assert_state_equal(got, exp)
    if check_pairs:
        opairs = refs.oracle_pairs(s, st.margin, flags=flags, stride=stride, method=1)
        assert len(pairs) == len(opairs) and not fields_diff(pairs, opairs), (len(pairs), len(opairs))
        assert st.n_candidates == len(opairs)
    return exp, st


def test_euler_parity(built):
    s = dense(3000, 11)
    s.P["acceleration"][10:500] = np.random.default_rng(0).uniform(-3, 3, (490, 3)).astype(np.float32)
    s.P["angularAcceleration"][10:500] = np.random.default_rng(1).uniform(-3, 3, (490, 3)).astype(np.float32)
    s.P["isAsleep"][500:520] = 1
    s.hasP[520:530] = 0
    with world.World(s.n) as w:
        w.upload(s)
        w.euler_engage(DT)
        got = w.download(s)
    exp = oracle_tick(s, collide=False)
    # linear state bit-exact; orientation goes through sin/cos (libm vs device): <= 2 ulp of the larger component
    assert not fields_diff(got.T, exp.T, skip=("orientation",))
    assert not fields_diff(got.P, exp.P)
    q, qe = got.T["orientation"], exp.T["orientation"]
    assert np.abs(q - qe).max() <= 2.4e-7
    frac_exact = (q.view(np.uint32) == qe.view(np.uint32)).all(axis=1).mean()
    assert frac_exact > 0.98, frac_exact


@pytest.mark.parametrize("seed", [1, 2])
def test_collision_lockstep_dense(built, seed):
    exp, st = lockstep_collision(dense(4000, seed))
    assert len(exp.hits) > 500 and st.n_rounds >= 2


def test_collision_lockstep_sphere_sphere_flag(built):
    lockstep_collision(dense(3000, 5), flags=abi.F_SPHERE_SPHERE)


@pytest.mark.parametrize("host", [0, abi.NULL_TIMESLICE])
def test_collision_lockstep_interceptions(built, host):
    s = dense(3000, 3, temporal=True)
    s.ids[10:1500] |= 1
    s.ids[10:1500] &= ~np.uint32(14)            # chain-link 1
    s.ids[2000:2500] = (s.ids[2000:2500] & ~np.uint32(15)) | 2
    s.tstate[1000:2200] = abi.TS_FORKED
    s.tstate[2200:2400] = abi.TS_FORKING
    exp, _ = lockstep_collision(s, host=host, coherency=77)
    assert len(exp.ints) > 50


def test_collision_case_study_ray_spring(built):
    n_hits = 0
    for y in (-3.5, -3.0, -1.0):            # box+legs on the floor / legs (ray spring) only / box on the ball
        s = scenes.case_study(push=True)
        s.T["origin"][3] = (12, y, 0)
        s.P["velocity"][3] = (0.3, -1.0, 0.2)
        exp, _ = lockstep_collision(s)
        n_hits += len(exp.hits)
    assert n_hits >= 4


def test_collision_worldline_batch(built):
    s = scenes.worldline_batch(8, 256, seed=9)
    lockstep_collision(s, host=abi.NULL_TIMESLICE, stride=256)


def test_missing_physics_and_noop(built):
    s = dense(2000, 8)
    s.hasP[100:400] = 0
    s.C["colliders"]["collisionType"][400:700, 0] = abi.COLLISION_NOOP
    s.C["colliders"]["useBehaviorsResponse"][600:900, 0] = 1
    s.C["colliders"]["isActive"][900:950, 0] = 0
    s.C["colliders"]["colliderType"][950:1000, 0] = abi.COLLIDER_NONE
    lockstep_collision(s)


def test_free_run_ticks(built):
    """Whole ticks (Euler + collision) free-running: positions track the oracle within FP32 noise."""
    s = scenes.arena(3000, seed=21)
    with world.World(s.n) as w:
        w.upload(s)
        cur = s
        for t in range(20):
            w.step(DT)
            cur = oracle_tick(cur, coherency=t)
        got = w.download(s)
        assert w.coherency() == 20
    assert np.abs(got.T["origin"] - cur.T["origin"]).max() < 1e-4
    assert np.abs(got.P["velocity"] - cur.P["velocity"]).max() < 1e-4


def test_update_and_retry_paths(built):
    s = dense(1500, 4)
    with world.World(s.n, margin=1e-4, max_neighbors=2) as w:       # force margin + capacity retries
        w.upload(s)
        w.collision_engage()
        got, hits, st = w.download(s), w.contacts(), w.stats()
    exp = oracle_tick(s, euler=False)
    assert st.n_retries > 0
    assert_hits_equal(hits, exp.hits)
    assert_state_equal(got, exp)
    # partial update of a few bodies behaves like a fresh upload
    s2 = s.copy()
    idx = np.array([10, 500, 900], np.uint32)
    s2.T["origin"][idx] += np.float32(0.25)
    s2.P["velocity"][idx] = (1, 2, 3)
    with world.World(s.n) as w:
        w.upload(s)
        w.update(idx, T=np.ascontiguousarray(s2.T[idx]), P=np.ascontiguousarray(s2.P[idx]))
        w.collision_engage()
        got, hits = w.download(s2), w.contacts()
    exp = oracle_tick(s2, euler=False)
    assert_hits_equal(hits, exp.hits)
    assert_state_equal(got, exp)
