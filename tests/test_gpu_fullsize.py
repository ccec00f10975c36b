"""GPU parity at BASELINE.json's full sizes (configs[1] 64 K, configs[2] 1 M colliders).

64 K: the collision pass in lock step against the oracle procedures behind the oracle-side cull
 -- hit list, interceptions, end state and broadphase pair list bit-exact.
1 M: the oracle's own cull takes minutes at this size, so the oracle procedures replay the
DEVICE's candidate list (its definition is pinned at <= 64 K) and everything downstream must be
bit-exact; the list itself is checked through size-independent properties: canonical order, no
duplicates, every listed pair overlaps, and completeness on a random sample of colliders against
a brute-force scan of all bounds.  Plus determinism of free-running ticks (atomics and scheduling
must not leak into results)."""
import numpy as np
import pytest

from intercession_b200 import scenes, world
import refs
from util import assert_hits_equal, assert_state_equal, fields_diff

pytestmark = pytest.mark.gpu
DT = 1.0 / 45


def collide_on_device(s):
    with world.World(s.n) as w:
        w.upload(s)
        w.collision_engage()
        return w.download(s), w.contacts(), w.interceptions(), w.pairs(), w.stats(), w.all_pairs()


def test_config2_64k_collision_lockstep(built):
    s = scenes.config2(65536)
    got, hits, ints, pairs, st, _ = collide_on_device(s)
    opairs = refs.oracle_pairs(s, st.margin, method=1)
    assert len(pairs) == len(opairs) and not fields_diff(pairs, opairs)
    exp = refs.procedures_tick(refs.oracle(), "icxo", s, euler=False, cand=opairs)
    assert len(exp.hits) > 500
    assert_hits_equal(hits, exp.hits)
    assert not fields_diff(ints, exp.ints)
    assert_state_equal(got, exp)


def test_config3_1m_collision_lockstep_and_pair_properties(built):
    s = scenes.config3(1048576)
    got, hits, ints, pairs, st, cand = collide_on_device(s)
    # ---- pair list properties ----
    key = pairs["a"].astype(np.uint64) << np.uint64(32) | pairs["b"].astype(np.uint64)
    assert np.all(pairs["a"] < pairs["b"]) and np.all(np.diff(key.astype(np.int64)) > 0)      # canonical order, unique
    box = refs.oracle_aabbs(s, st.margin)                                                     # per packet: lo xyz, hi xyz (inflated)
    assert len(box) > int(pairs["b"].max())
    lo, hi = box[:, :3], box[:, 3:]
    a, b = pairs["a"], pairs["b"]
    assert np.all((lo[a] <= hi[b]).all(1) & (lo[b] <= hi[a]).all(1))                          # every listed pair overlaps
    active = s.C["colliders"]["isActive"].reshape(-1) != 0
    ent_of = np.repeat(np.arange(s.n), 2)[active]                                             # packet -> entity (submission order)
    assert len(ent_of) == len(box)
    rng = np.random.default_rng(1)
    listed = set(zip(a.tolist(), b.tolist()))
    found = 0
    for p in rng.choice(len(box), size=96, replace=False):
        ov = np.nonzero((lo[p] <= hi).all(1) & (lo <= hi[p]).all(1) & (ent_of != ent_of[p]))[0]
        for q in ov:                                                                           # sphere-sphere pairs are not candidates (G3)
            ta, tb = (s.C["colliders"]["colliderType"].reshape(-1)[active][[p, q]])
            if ta == 2 and tb == 2:
                continue
            assert (min(p, q), max(p, q)) in listed, (p, q)
            found += 1
    assert found > 100
    # ---- everything downstream of the list, bit-exact ----
    exp = refs.procedures_tick(refs.oracle(), "icxo", s, euler=False, cand=cand)      # uniform-margin pairs + the movers' extra pairs
    assert len(exp.hits) > 5000
    assert_hits_equal(hits, exp.hits)
    assert not fields_diff(ints, exp.ints)
    assert_state_equal(got, exp)


def test_config3_1m_free_run_is_deterministic(built):
    s = scenes.config3(1048576)
    outs = []
    for _ in range(2):
        with world.World(s.n) as w:
            w.upload(s)
            w.step(DT, 6)
            outs.append((w.download(s), w.contacts(), w.stats().n_candidates))
    (g0, h0, c0), (g1, h1, c1) = outs
    assert c0 == c1 and len(h0) == len(h1) and not fields_diff(h0, h1)
    assert not fields_diff(g0.T, g1.T) and not fields_diff(g0.P, g1.P)
