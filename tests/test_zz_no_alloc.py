"""GPU: icx_divergence_scan_static -- the divergence scan with its per-world-line accumulators in a file-scope static device
array (no cudaMalloc in the call; no counterpart in the reference).  Called
through the C ABI and compared with the existing entry point, bit for bit (the sums are fixed point), on two calls in a row:
the second scans a world with nothing incongruent, so sums left over from the first call would show.  Collected last on
purpose."""
import ctypes as C

import numpy as np
import pytest

from intercession_b200 import abi, scenes, world

pytestmark = pytest.mark.gpu


def scan_static(w, tl, coherency, n_wl, want=(True, True, True)):
    flags = np.full(n_wl, 0xEE, np.uint8)
    count = np.full(n_wl, 0xEEEEEEEE, np.uint32)
    score = np.full(n_wl, -1.0, np.float64)
    rc = w.lib.icx_divergence_scan_static(w.handle, tl.handle, coherency, n_wl, abi.ptr(flags) if want[0] else None,
                                          abi.ptr(count) if want[1] else None, abi.ptr(score) if want[2] else None)
    assert rc == abi.OK, (rc, w.lib.icx_timeline_last_error(tl.handle).decode())
    return flags, count, score


def test_static_scratch_scan_equals_the_allocating_one_twice_in_a_row(built):
    n_wl, per = 6, 1024
    live = scenes.worldline_batch(n_wl, per, seed=11)
    rng = np.random.default_rng(2)
    k = rng.choice(live.n, size=live.n // 7, replace=False)
    live.T["origin"][k] += rng.normal(0, 0.8, (len(k), 3)).astype(np.float32)
    live.T["origin"][k[0]] = np.nan
    live.T["origin"][k[1]] = 1e30
    calm = scenes.worldline_batch(n_wl, per, seed=11, perturb=False)
    with world.World(live.n, worldline_stride=per, host_timeslice=abi.NULL_TIMESLICE) as w, \
            world.World(calm.n, worldline_stride=per, host_timeslice=abi.NULL_TIMESLICE) as w2, \
            world.World(per) as wb, world.Timeline(4, per) as tl:
        wb.upload(live.base)
        wb.record(tl, 77)
        w.upload(live)
        w2.upload(calm)
        # call 1: many incongruent bodies
        f1, c1, s1 = scan_static(w, tl, 77, n_wl)
        ef, ec, es = w.divergence_scan(tl, 77, n_wl)
        assert np.array_equal(f1, ef) and np.array_equal(c1, ec) and np.array_equal(s1, es)
        assert c1.sum() > 10 and s1.max() > 0
        # call 2, straight after: nothing incongruent -- the array still holds call 1's sums unless the entry point clears them
        f2, c2, s2 = scan_static(w2, tl, 77, n_wl)
        ef2, ec2, es2 = w2.divergence_scan(tl, 77, n_wl)
        assert np.array_equal(f2, ef2) and np.array_equal(c2, ec2) and np.array_equal(s2, es2)
        assert c2.sum() == 0 and not f2.any()
        # and back again; flags alone (counts pass through the bounded stack buffer), scores alone
        f3, _, _ = scan_static(w, tl, 77, n_wl, want=(True, False, False))
        _, _, s3 = scan_static(w, tl, 77, n_wl, want=(False, False, True))
        assert np.array_equal(f3, ef) and np.array_equal(s3, es)
        # beyond the static array: a clear error, nothing launched
        rc = w.lib.icx_divergence_scan_static(w.handle, tl.handle, 77, 8193, None, None, None)
        assert rc == abi.E_CAPACITY and b"ICX_SCAN_STATIC_MAX_WORLDLINES" in w.lib.icx_timeline_last_error(tl.handle)


def test_static_scratch_scan_at_the_maximum(built):
    """8192 world-lines x 1024 bodies (the static array's limit) against one shared base record: exactly the moved ones flag."""
    n_wl, per = 8192, 1024
    live = scenes.worldline_batch(n_wl, per, seed=13, perturb=False)
    moved = np.array([0, 1, 999, 4096, 8191])
    live.T["origin"][moved * per + (moved % per), 0] += np.float32(2.0)
    with world.World(live.n, worldline_stride=per, host_timeslice=abi.NULL_TIMESLICE) as w, \
            world.World(per) as wb, world.Timeline(2, per) as tl:
        wb.upload(live.base)
        wb.record(tl, 5)
        w.upload(live)
        for _ in range(2):
            flags, cnt, score = scan_static(w, tl, 5, n_wl)
            assert np.array_equal(np.nonzero(flags)[0], moved) and cnt.sum() == len(moved)
            assert np.array_equal(score[moved], np.full(len(moved), 4.0)) and score.sum() == 4.0 * len(moved)
        ef, ec, es = w.divergence_scan(tl, 5, n_wl)
        assert np.array_equal(flags, ef) and np.array_equal(cnt, ec) and np.array_equal(score, es)
