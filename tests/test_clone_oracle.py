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
