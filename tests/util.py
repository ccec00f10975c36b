"""Shared comparison helpers for the parity tests."""
import numpy as np

from intercession_b200 import abi


def bits(a):
    a = np.ascontiguousarray(a)
    return a.view(np.uint32) if a.dtype == np.float32 else a


def fields_diff(a, b, prefix="", skip=()):
    """Names of (nested) fields whose bit patterns differ, with a sample."""
    bad = []
    if a.shape != b.shape:
        return [("shape", a.shape, b.shape)]
    for name in a.dtype.names:
        if prefix + name in skip:
            continue
        x, y = a[name], b[name]
        if x.dtype.names:
            bad += fields_diff(x, y, prefix + name + ".", skip)
        elif not np.array_equal(bits(x), bits(y)):
            idx = np.argwhere(bits(x) != bits(y))
            bad.append((prefix + name, len(idx), idx[:3].tolist(), x[tuple(idx[0])], y[tuple(idx[0])]))
    return bad


def ulp_diff(a, b):
    """Max distance in units-in-the-last-place between two float32 arrays (same-sign finite values)."""
    ia = np.ascontiguousarray(a, np.float32).view(np.int32).astype(np.int64)
    ib = np.ascontiguousarray(b, np.float32).view(np.int32).astype(np.int64)
    ia = np.where(ia < 0, -(ia & 0x7FFFFFFF), ia)
    ib = np.where(ib < 0, -(ib & 0x7FFFFFFF), ib)
    return int(np.abs(ia - ib).max()) if ia.size else 0


def assert_state_equal(got, exp, what="state", skip=()):
    for k in ("T", "P", "C"):
        d = fields_diff(got[k], exp[k], skip=skip)
        assert not d, f"{what}: {k} differs: {d[:4]}"
    assert np.array_equal(got.tstate, exp.tstate), f"{what}: tstate differs"
    assert np.array_equal(got.tcoh, exp.tcoh), f"{what}: tcoh differs"


def assert_hits_equal(got, exp, what="hits", ignore_responded=False):
    assert len(got) == len(exp), f"{what}: {len(got)} hits vs {len(exp)} expected"
    g, e = got.copy(), exp.copy()
    if ignore_responded:
        g["flags"] &= ~np.uint16(abi.CONTACT_RESPONDED)
        e["flags"] &= ~np.uint16(abi.CONTACT_RESPONDED)
    d = fields_diff(g, e)
    assert not d, f"{what}: {d[:4]}"
