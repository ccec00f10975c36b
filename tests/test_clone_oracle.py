"""Hand-derived cases for the numpy restatement of load_and_link / extract_entity_updates
(oracle/clone_oracle.py; reference: source/spacetime/parallel_cosmos_context.cpp:60-169, :171-423).
The reference class cannot be compiled here, so these expectations were worked out
from the source by hand: parity for this row is restated, not executed."""
import os
import sys

import numpy as np
import pytest

from intercession_b200 import abi, scenes
from util import fields_diff

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "oracle"))
import clone_oracle  # noqa: E402


def ids_of(links, timeslice=2):
    return np.array([abi.compose_entity(timeslice, g + 1, l) for g, l in enumerate(links)], np.uint32)


def scene(links, timeslice=2):
    s = scenes.arena(len(links), seed=11, walls=False, temporal=True, fill_side=12.0)
    s["ids"] = ids_of(links, timeslice)
    return s


def test_load_and_link_filters_and_states():
    s = scene([0, 0, 1, 2, 15, 3])
    s.tstate[:] = [abi.TS_MERGED, abi.TS_FORKED, abi.TS_FORKED, abi.TS_FORKING, abi.TS_MERGED, abi.TS_GHOST]
    s.tcoh[:] = [7, 8, 9, 10, 11, 12]
    has_future = np.array([0, 1, 0, 0, 0, 1], np.uint8)
    before = s.copy()
    par, coh = clone_oracle.load_and_link(s, 40, has_future)
    # entity 0: link 0 and no future -> omitted (:110); entity 2 has no future but link 1 -> kept
    assert list(par.ids) == list(before.ids[1:])
    assert coh == 40
    # forked entities stay forked (stamped 40) and are re-merged in the source (:145-150); all others {merged, 0}
    assert list(par.tstate) == [abi.TS_FORKED, abi.TS_FORKED, abi.TS_MERGED, abi.TS_MERGED, abi.TS_MERGED]
    assert list(par.tcoh) == [40, 40, 0, 0, 0]
    assert list(s.tstate) == [abi.TS_MERGED, abi.TS_MERGED, abi.TS_MERGED, abi.TS_FORKING, abi.TS_MERGED, abi.TS_GHOST]
    assert list(s.tcoh) == [7, 40, 40, 10, 11, 12]
    assert not fields_diff(par.T, before.T[1:]) and not fields_diff(par.C, before.C[1:]) and not fields_diff(par.P, before.P[1:])


def test_load_and_link_without_mask_keeps_everything():
    s = scene([0, 1, 15])
    par, _ = clone_oracle.load_and_link(s, 3, None)
    assert par.n == 3


def test_extract_rules():
    dst = scene([1, 0, 15, 2])                       # future-side ids
    par = dst.copy()
    par["ids"] = np.array([dst.ids[0] + 1, dst.ids[1] + 1, dst.ids[2], dst.ids[3] + 1], np.uint32)   # links 2, 1, 15, 3
    par.T["origin"] += np.float32(5.0)
    par.tstate[:] = abi.TS_FORKED
    before = dst.copy()
    n, miss = clone_oracle.extract_entity_updates(par, dst, 77, dst_host_id=3)
    assert (n, miss) == (3, 0)
    assert np.array_equal(dst.T["origin"][[0, 1, 3]], par.T["origin"][[0, 1, 3]])
    assert np.array_equal(dst.T["origin"][2], before.T["origin"][2])              # non-temporal: ignored (:213)
    # state follows only where the decremented link is > 0 (:255)
    assert list(dst.tstate) == [abi.TS_FORKED, before.tstate[1], before.tstate[2], abi.TS_FORKED]
    assert list(dst.tcoh) == [77, before.tcoh[1], before.tcoh[2], 77]


def test_extract_timeslice0_remerges_and_counts_missing():
    dst = scene([1, 2])
    par = scene([2, 3, 4])
    par["ids"] = np.array([dst.ids[0] + 1, dst.ids[1] + 1, abi.compose_entity(2, 99, 4)], np.uint32)
    par.tstate[:] = abi.TS_FORKED
    n, miss = clone_oracle.extract_entity_updates(par, dst, 5, dst_host_id=0)
    assert (n, miss) == (2, 1)
    assert list(dst.tstate) == [abi.TS_MERGED, abi.TS_MERGED] and list(dst.tcoh) == [5, 5]


def test_extract_ignores_parallel_created_and_rejects_link0():
    dst = scene([1])
    par = scene([2])
    par["ids"] = np.array([abi.compose_entity(abi.NULL_TIMESLICE, 1, 2)], np.uint32)
    assert clone_oracle.extract_entity_updates(par, dst, 1, 1) == (0, 0)
    par["ids"] = np.array([abi.compose_entity(2, 1, 0)], np.uint32)
    with pytest.raises(ValueError):
        clone_oracle.extract_entity_updates(par, dst, 1, 1)


# ---- pinned: the numpy restatement against the reference's own entity loops run on the real Cosmos (oracle/ref/ref_clone.cpp) ----
import refs  # noqa: E402

needs_ref = pytest.mark.skipif(refs.ref() is None, reason="oracle/_ref not built (no /root/reference here)")


def random_past_scene(n, seed):
    s = scenes.arena(n, seed=seed, walls=False, temporal=True, timeslice=2, fill_side=14.0)
    rng = np.random.default_rng(seed)
    g = np.arange(n, dtype=np.uint32) + 1                                       # genesis < 255: the real Entity is 16 bits wide
    link = rng.choice([0, 1, 2, 3, 14, 15], n, p=[0.25, 0.25, 0.2, 0.15, 0.05, 0.1]).astype(np.uint32)
    s["ids"] = (np.uint32(2) << np.uint32(28)) | (g << np.uint32(4)) | link
    s.tstate[:] = rng.choice([abi.TS_MERGED, abi.TS_FORKED, abi.TS_FORKING, abi.TS_SUPERPOSITION, abi.TS_GHOST], n, p=[0.5, 0.2, 0.15, 0.1, 0.05])
    s.tcoh[:] = rng.integers(0, 60, n).astype(np.uint16)
    s.tcoh[s.tstate == abi.TS_MERGED] = np.where(rng.random((s.tstate == abi.TS_MERGED).sum()) < 0.5, 0, s.tcoh[s.tstate == abi.TS_MERGED])
    s.hasP[rng.random(n) < 0.1] = 0
    has_future = (rng.random(n) < 0.6).astype(np.uint8)
    return s, has_future


def same_scene(a, b):
    assert a.n == b.n and np.array_equal(a.ids, b.ids)
    for k in ("T", "C"):
        assert not fields_diff(a[k], b[k]), k
    hp = a.hasP.astype(bool)
    assert np.array_equal(a.hasP, b.hasP) and not fields_diff(a.P[hp], b.P[hp])
    assert np.array_equal(a.tstate, b.tstate) and np.array_equal(a.tcoh, b.tcoh)


@needs_ref
@pytest.mark.parametrize("seed", [1, 2, 3])
def test_load_and_link_matches_the_reference_loop(seed):
    s, hf = random_past_scene(200, seed)
    a, b = s.copy(), s.copy()
    par_o, coh_o = clone_oracle.load_and_link(a, 41, hf)
    par_r, coh_r = refs.ref_load_and_link(b, 41, hf)
    assert coh_o == coh_r == 41
    same_scene(par_o, par_r)
    assert np.array_equal(a.tstate, b.tstate) and np.array_equal(a.tcoh, b.tcoh)        # forked entities re-merged in the source, stamped
    assert (a.tstate != s.tstate).any() and par_o.n < s.n


@needs_ref
@pytest.mark.parametrize("seed,host", [(4, 3), (5, 0), (6, 1)])
def test_extract_matches_the_reference_loop(seed, host):
    dst, _ = random_past_scene(200, seed)
    dst["ids"] = (dst.ids & ~np.uint32(0xF0000000)) | (np.uint32(host) << np.uint32(28))
    link = dst.ids & 15
    keep = link != 14                                                           # a past copy of a link-14 id does not exist (increment refused)
    rng = np.random.default_rng(seed)
    par = dst.copy()
    for k in ("ids", "T", "P", "hasP", "C", "tstate", "tcoh"):
        par[k] = dst[k][keep].copy()
    plink = par.ids & 15
    par["ids"] = np.where(plink == 15, par.ids, par.ids + 1).astype(np.uint32)  # the parallel cosmos holds the past's ids (link + 1)
    created = rng.random(par.n) < 0.1                                           # entities created during the parallel run: timeslice 14 (:223-229)
    par["ids"] = np.where(created, (par.ids & ~np.uint32(0xF0000000)) | (np.uint32(abi.NULL_TIMESLICE) << np.uint32(28)), par.ids).astype(np.uint32)
    order = np.argsort(par.ids, kind="stable")
    for k in ("ids", "T", "P", "hasP", "C", "tstate", "tcoh"):
        par[k] = par[k][order].copy()
    par.T["origin"] += np.float32(3.0)
    par.P["velocity"] *= np.float32(-2.0)
    par.C["colliders"]["restLength"] += np.float32(0.5)
    par.tstate[:] = rng.choice([abi.TS_MERGED, abi.TS_FORKED, abi.TS_FORKING], par.n)
    a, b = dst.copy(), dst.copy()
    n_o, miss_o = clone_oracle.extract_entity_updates(par, a, 88, host)
    n_r, reg_r = refs.ref_extract(par, b, 88, host)
    assert miss_o == reg_r == 0 and n_o == n_r > 50
    same_scene(a, b)
    assert (a.T["origin"] != dst.T["origin"]).any()
    # a destination that lacks some of them: the reference registers exactly the ones the restatement registers (:241-245)
    lack = rng.random(dst.n) < 0.15
    small = dst.copy()
    for k in ("ids", "T", "P", "hasP", "C", "tstate", "tcoh"):
        small[k] = dst[k][~lack].copy()
    grown, n_o2, reg_o = clone_oracle.extract_entity_updates(par, small.copy(), 88, host, register=True)
    _, reg_r2 = refs.ref_extract(par, small.copy(), 88, host)
    assert reg_o == reg_r2 > 0 and n_o2 == n_o and grown.n == small.n + reg_o
    sub = np.isin(a.ids, grown.ids)                                             # (what nobody extracts -- non-temporal ids, ids created during the run -- stays lacking)
    part = a.copy()
    for k in ("ids", "T", "P", "hasP", "C", "tstate", "tcoh"):
        part[k] = a[k][sub].copy()
    assert np.array_equal(grown.ids, part.ids)                                  # registered + extracted = what the full destination became
    for k in ("T", "P", "C"):                                                   # (a registered link-0 entity starts {merged, 0}; the full destination's kept its old state)
        rows = grown.hasP.astype(bool) if k == "P" else np.ones(grown.n, bool)     # (no PhysicsComponent: the row's P bytes mean nothing)
        assert not fields_diff(grown[k][rows], part[k][rows]), k
    fresh = np.isin(grown.ids, small.ids, invert=True)
    assert ((grown.tstate[fresh] == abi.TS_MERGED) | ((grown.ids[fresh] & 15) > 0) | (host == 0)).all()


@needs_ref
def test_extract_chain_link_zero_throws_like_the_reference():
    dst, _ = random_past_scene(20, 9)
    par = dst.copy()
    par["ids"] = (par.ids & ~np.uint32(15))                                     # chain-link 0 everywhere
    with pytest.raises(ValueError):
        refs.ref_extract(par, dst.copy(), 5, 1)
    with pytest.raises(ValueError):
        clone_oracle.extract_entity_updates(par, dst.copy(), 5, 1)
