"""GPU parity of icx_world_clone / icx_world_extract (ParallelCosmosContext::load_and_link /
extract_entity_updates, parallel_cosmos_context.cpp:60-169, :171-423) against the numpy
restatement in oracle/clone_oracle.py.  Byte copies and integer state: bit-exact."""
import os
import sys

import numpy as np
import pytest

from intercession_b200 import abi, scenes, world
from util import fields_diff

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "oracle"))
import clone_oracle  # noqa: E402

pytestmark = pytest.mark.gpu
DT = 1.0 / 45


def past_scene(n=400, seed=5):
    """A past timeslice: temporal bodies (links 1..3), some link-0 without a future, walls non-temporal."""
    s = scenes.arena(n, seed=seed, walls=True, temporal=True, timeslice=2, spacing=1.0, jitter=0.2, fill_side=11.0)
    rng = np.random.default_rng(seed)
    nb = n
    link = rng.integers(0, 4, nb).astype(np.uint32)
    s.ids[:nb] = (s.ids[:nb] & ~np.uint32(15)) | link
    s.tstate[:nb] = rng.choice([abi.TS_MERGED, abi.TS_FORKED, abi.TS_FORKING], nb, p=[0.6, 0.25, 0.15]).astype(np.uint8)
    s.tcoh[:nb] = rng.integers(0, 30, nb).astype(np.uint16)
    has_future = (rng.random(s.n) < 0.7).astype(np.uint8)
    return s, has_future


def same(got, exp):
    assert got.n == exp.n
    assert np.array_equal(got.ids, exp.ids)
    for k in ("T", "P", "C"):
        d = fields_diff(got[k], exp[k])
        assert not d, (k, d)
    assert np.array_equal(got.tstate, exp.tstate) and np.array_equal(got.tcoh, exp.tcoh)


def download_all(w, like):
    out = like.copy()
    out["ids"] = like.ids.copy()
    return w.download(out)


def test_clone_matches_load_and_link(built):
    s, hf = past_scene()
    with world.World(s.n) as src, world.World(s.n, host_timeslice=abi.NULL_TIMESLICE) as par:
        src.upload(s)
        src.set_coherency(37)
        src.step(DT, 3)                                   # the copy must carry the CURRENT state, not the uploaded one
        cur = src.download(s)
        exp_src = cur.copy()
        exp_par, coh = clone_oracle.load_and_link(exp_src, src.coherency(), hf)
        n = src.clone_into(par, hf)
        assert n == exp_par.n and par.coherency() == coh == 40
        same(download_all(par, exp_par), exp_par)
        same(src.download(s), exp_src)                    # forked entities re-merged in the source
        # the clone is a working world: ticking both from the same state stays bit-identical where nothing was dropped
        n2 = src.clone_into(par, None)
        assert n2 == s.n
        src.step(DT, 2); par.step(DT, 2)
        a, b = src.download(s), download_all(par, s)
        assert not fields_diff(a.T, b.T) and not fields_diff(a.P, b.P)


def test_extract_matches_reference_rules(built):
    s, hf = past_scene(n=300, seed=9)
    s.ids[:300] = (s.ids[:300] & ~np.uint32(15)) | (1 + (np.arange(300) % 3)).astype(np.uint32)      # links 1..3: extractable
    s.ids[300:] |= 15                                                                 # the walls are non-temporal
    fut = s.copy()
    fut["ids"] = np.where((s.ids & 15) < 15, s.ids - 1, s.ids).astype(np.uint32)     # the future side holds link - 1
    fut.T["origin"][:300] += np.float32(0.25)
    for host in (3, 0):
        with world.World(s.n, host_timeslice=abi.NULL_TIMESLICE) as par, world.World(s.n, host_timeslice=host) as dst:
            par.upload(s); par.step(DT, 2)
            dst.upload(fut); dst.set_coherency(par.coherency())
            exp = fut.copy()
            n_exp = clone_oracle.extract_entity_updates(par.download(s), exp, dst.coherency(), host)
            assert par.extract_into(dst) == n_exp and n_exp[0] == 300
            same(dst.download(fut), exp)


def test_extract_counts_missing_and_rejects_link0(built):
    s, _ = past_scene(n=50, seed=2)
    s.ids[:50] = (s.ids[:50] & ~np.uint32(15)) | 2
    s.ids[50:] |= 15
    fut = s.copy()
    fut["ids"] = np.where((s.ids & 15) < 15, s.ids - 1, s.ids).astype(np.uint32)
    fut.ids[:5] = (np.arange(5, dtype=np.uint32) + 4000) << 4 | 1                 # five entities unknown to the destination
    with world.World(s.n, host_timeslice=abi.NULL_TIMESLICE) as par, world.World(s.n, host_timeslice=1) as dst:
        par.upload(s); dst.upload(fut)
        assert par.extract_into(dst) == (45, 5)
        bad = s.copy(); bad.ids[7] &= ~np.uint32(15)
        par.upload(bad)
        with pytest.raises(world.IcxError) as ei:
            par.extract_into(dst)
        assert ei.value.code == abi.E_INVALID


def test_extract_registers_what_the_destination_lacks(built):
    """parallel_cosmos_context.cpp:241-245: an entity of the past that does not exist in the future is registered there, then
    extracted like the others.  On the device the destination's own rows stay where they are (icx_entities_create)."""
    s, _ = past_scene(n=50, seed=2)
    s.ids[:50] = (s.ids[:50] & ~np.uint32(15)) | 2
    s.ids[50:] |= 15
    assert np.all(np.diff(s.ids.astype(np.int64)) > 0)
    full = s.copy()
    full["ids"] = np.where((s.ids & 15) < 15, s.ids - 1, s.ids).astype(np.uint32)
    full.T["origin"][:50] += np.float32(0.25)
    keep = np.ones(s.n, bool); keep[[3, 10, 20, 30, 44]] = False
    fut = scenes.Scene({k: (v[keep] if isinstance(v, np.ndarray) and k in ("ids", "T", "P", "hasP", "C", "tstate", "tcoh") else v) for k, v in full.items()})
    for host in (1, 0):
        with world.World(s.n, host_timeslice=abi.NULL_TIMESLICE) as par, world.World(s.n, host_timeslice=host) as dst:
            par.upload(s); par.step(DT, 2)
            dst.upload(fut); dst.set_coherency(par.coherency())
            exp, n_ex, n_reg = clone_oracle.extract_entity_updates(par.download(s), fut.copy(), dst.coherency(), host, register=True)
            assert (n_ex, n_reg) == (50, 5) and exp.n == s.n
            assert par.extract_into(dst, register=True) == (50, 0)
            assert np.array_equal(dst.entity_ids(), exp.ids)
            same(dst.download(exp), exp)
            # the registered entities are live: both worlds tick on identically from here
            with world.World(s.n, host_timeslice=host) as fresh:
                fresh.upload(exp); fresh.set_coherency(dst.coherency())
                dst.step(DT, 3); fresh.step(DT, 3)
                same(dst.download(exp), fresh.download(exp))
