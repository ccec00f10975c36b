"""CPU suite, part 5: the world-line record / replay / divergence restatement (oracle) against the
reference's REAL timestream containers and predicates (EntityTimestreamMap, TsBreakpointQueue,
Message, coherency_greater_or_equal, is_divergent, TimejumpConditions::operator==) compiled into
oracle/_ref.  The dynamo loops around them are restated on both sides (asio is not vendored)."""
import numpy as np
import pytest

from intercession_b200 import abi, scenes
import refs
from util import fields_diff

needs_ref = pytest.mark.skipif(refs.ref() is None, reason="oracle/_ref not built (no /root/reference here)")


def small_scene(n=60, seed=3):
    s = scenes.arena(n, seed=seed, walls=False, temporal=True)
    s.ids[:] = (np.arange(n, dtype=np.uint32) << 4) | 1          # genesis < 255, chain-link 1
    s.ids[n - 5:] |= 15                                          # a few non-temporal ids (link not incremented)
    s.ids[n - 8:n - 5] = (s.ids[n - 8:n - 5] & ~np.uint32(15)) | 14     # link 14: increment refused
    s.tstate[5:15] = abi.TS_FORKED
    s.tstate[15:20] = abi.TS_FORKING
    s.tstate[20:25] = abi.TS_SUPERPOSITION
    s.tstate[25:28] = abi.TS_GHOST
    return s


def changed_colliders(n, k):
    """Entities whose ColliderComponent evolve(s, k) touches."""
    return (np.arange(n) + k) % 5 == 0


def evolve(s, k, prev=None):
    """State of the future timeslice at tick k: EVERY field of Transform / Physics / ColliderComponent that a tick, a
    behaviour (biped_behaviors.h:96-100 restLength / damping, projectile_behaviors.h:38 scale) or a host edit may change
    differs from the past's copy, so a replay that restores less than the reference's (all components,
    server_network_dynamo.cpp:408-412) cannot pass.  prev: the state of tick k - 1, whose collider edits persist (a world keeps
    what it was told)."""
    out = s.copy()
    if prev is not None:
        out["C"] = prev.C.copy()
    f = np.float32
    out.T["origin"] += f(0.125 * k)
    out.T["scale"] *= f(1.0 + 0.02 * (k % 7))
    out.P["velocity"] *= f(1.0 + 0.01 * k)
    out.P["acceleration"] += f(0.5 * k)
    out.P["mass"] += f(0.25 * (k % 5))
    out.P["linearDrag"] = f(0.01 + 0.001 * k)
    out.P["angularDrag"] = f(0.02 + 0.001 * k)
    out.P["collisionLinearDrag"] = f(0.005 * k)
    out.P["collisionAngularDrag"] = f(0.03 + 0.002 * k)
    i = np.arange(s.n)
    out.P["lockOrigin"] = ((i + k) % 3 == 0).astype(np.uint8)
    out.P["lockedOrigin"] = out.T["origin"] + f(1.0)
    out.P["lockOrientation"] = ((i + k) % 4 == 0).astype(np.uint8)
    out.P["lockedOrientation"][:, 1] = f(0.001 * k)
    out.P["isAsleep"] = ((i + k) % 11 == 0).astype(np.uint8)
    # colliders: only SOME entities change per tick (the sparse path), by different amounts
    ch = changed_colliders(s.n, k)
    c = out.C["colliders"]
    c["restLength"][ch, 0] = f(0.1 + 0.01 * k)
    c["damping"][ch, 0] = f(10.0 + k)
    c["stiffness"][ch, 1] = f(100.0 + k)
    c["staticFriction"][ch, 0] = f(0.45 + 0.001 * k)
    c["localTransform"]["scale"][ch, 0] = f(1.0 + 0.01 * (k % 3))
    c["influenceOrientation"][ch, 0] = k % 2
    return out


def past_ids(ids):
    link = ids & 15
    return np.where(link + 1 >= 15, ids, ids + 1).astype(np.uint32)


def both():
    return [refs.CpuTimeline(refs.ref(), "icxref"), refs.CpuTimeline(refs.oracle(), "icxo")]


@needs_ref
def test_coherency_compare_matches_reference():
    rng = np.random.default_rng(0)
    vals = np.concatenate([rng.integers(0, 65536, 4000), [0, 1, 32767, 32768, 32769, 65535] * 6]).astype(np.uint16)
    for a, b in zip(vals[::2], vals[1::2]):
        assert refs.ref().icxref_coherency_ge(int(a), int(b)) == refs.oracle().icxo_coherency_ge(int(a), int(b))
    assert refs.oracle().icxo_coherency_ge(3, 65530) == 1 and refs.oracle().icxo_coherency_ge(65530, 3) == 0


@needs_ref
def test_congruence_matches_reference():
    rng = np.random.default_rng(1)
    a, b = abi.default_transforms(5000), abi.default_transforms(5000)
    a["origin"] = rng.uniform(-50, 50, (5000, 3)).astype(np.float32)
    b["origin"] = a["origin"] + rng.normal(0, 0.9, (5000, 3)).astype(np.float32)
    b["origin"][:10] = a["origin"][:10] + np.float32(np.sqrt(2.0 / 3.0))      # right at the threshold
    b["origin"][10] = np.nan
    r, o = np.zeros(5000, np.uint8), np.zeros(5000, np.uint8)
    refs.ref().icxref_congruent(5000, abi.ptr(a), abi.ptr(b), abi.ptr(r))
    refs.oracle().icxo_congruent(5000, abi.ptr(a), abi.ptr(b), abi.ptr(o))
    assert np.array_equal(r, o) and 0.2 < r.mean() < 0.9 and o[10] == 0


@needs_ref
def test_record_replay_server_matches_reference_containers():
    """Future timeslice records ticks 100..110 (chain-link incremented on the way into the past); the
    past timeslice replays at various `now`: exact hit, nothing yet, stale records dropped, wrap-around."""
    fut = small_scene()
    past = fut.copy()
    past["ids"] = past_ids(fut.ids)
    tls = both()
    for t in range(100, 111):
        for tl in tls:
            tl.record(evolve(fut, t), t)
    for now in (99, 100, 103, 103, 107, 120):        # 103 twice: second time nothing is left for it
        res = [tl.replay(past, now) for tl in tls]
        assert res[0][1] == res[1][1], now
        assert not fields_diff(res[0][0].T, res[1][0].T) and not fields_diff(res[0][0].P, res[1][0].P), now
        assert not fields_diff(res[0][0].C, res[1][0].C), now
        if res[1][1]:
            keep = ~np.isin(past.tstate, (abi.TS_FORKED, abi.TS_FORKING))
            rec = evolve(fut, now if now <= 110 else 110)
            for name in ("scale",):
                assert np.array_equal(res[1][0].T[name][keep], rec.T[name][keep])
            for name in ("mass", "linearDrag", "lockOrigin", "isAsleep", "collisionLinearDrag"):
                assert np.array_equal(res[1][0].P[name][keep], rec.P[name][keep]), name
            assert np.array_equal(res[1][0].C["colliders"]["restLength"][keep], rec.C["colliders"]["restLength"][keep])
    # divergent entities (forked / forking) kept their state, everyone else took the record stamped `now`
    got, n = tls[1].replay(past, 110)
    assert n == 0                                     # 120 already drained the streams
    for tl in tls:
        tl.close()


@needs_ref
def test_replay_divergent_entities_keep_state():
    fut = small_scene()
    past = fut.copy()
    past["ids"] = past_ids(fut.ids)
    tls = both()
    rec = evolve(fut, 7)
    for tl in tls:
        tl.record(rec, 500)
    (r, nr), (o, no) = [tl.replay(past, 500) for tl in tls]
    assert nr == no == past.n - 15
    div = np.isin(past.tstate, (abi.TS_FORKED, abi.TS_FORKING))
    assert np.array_equal(o.T["origin"][div], past.T["origin"][div])
    assert np.array_equal(o.T["origin"][~div], rec.T["origin"][~div])
    assert not fields_diff(r.T, o.T) and not fields_diff(r.P, o.P) and not fields_diff(r.C, o.C)
    assert np.array_equal(o.P["mass"][div], past.P["mass"][div]) and np.array_equal(o.P["mass"][~div], rec.P["mass"][~div])
    for tl in tls:
        tl.close()


@needs_ref
def test_replay_across_the_coherency_wrap():
    """pop_future_timestream pops every record that is cyclically <= now; the relay then drops `coherency < now` with a PLAIN
    compare (server_network_dynamo.cpp:391), so with now = 2 the older records stamped 65534 / 65535 ARE deserialised, oldest
    first, before 0, 1, 2 are dropped or applied: the end state is the record stamped `now`, or -- when that one has not
    arrived -- the newest of the pre-wrap records."""
    fut = small_scene()
    past = fut.copy()
    past["ids"] = past_ids(fut.ids)
    for last, now in ((2, 2), (65535, 2), (1, 3)):
        tls = both()
        t = 65533
        while True:
            for tl in tls:
                tl.record(evolve(fut, t % 97), t)
            if t == last:
                break
            t = (t + 1) & 0xFFFF
        (r, nr), (o, no) = [tl.replay(past, now) for tl in tls]
        assert nr == no and nr > 0, (last, now)
        assert not fields_diff(r.T, o.T) and not fields_diff(r.P, o.P) and not fields_diff(r.C, o.C), (last, now)
        keep = ~np.isin(past.tstate, (abi.TS_FORKED, abi.TS_FORKING))
        exp = evolve(fut, (now if last == now else 65535) % 97)
        assert np.array_equal(o.T["origin"][keep], exp.T["origin"][keep]), (last, now)
        for tl in tls:
            tl.close()


@needs_ref
def test_breakpoint_parallel_rewrites_history():
    """link_timestreams sets breakpoints: the past timeslice can no longer pop old history, the parallel
    cosmos pops it at the breakpoint and inserts re-simulated history in front of it."""
    fut = small_scene()
    past = fut.copy()
    past["ids"] = past_ids(fut.ids)
    tls = both()
    for t in range(10, 16):
        for tl in tls:
            tl.record(evolve(fut, t), t)
    for tl in tls:
        tl.breakpoint(True)
    res = [tl.replay(past, 10) for tl in tls]                       # blocked at the breakpoint
    assert res[0][1] == res[1][1] == 0
    par = past.copy()                                               # the parallel cosmos holds the past's ids
    for t in range(10, 13):
        res = [tl.replay(par, t, mode=abi.REPLAY_PARALLEL) for tl in tls]      # consumes, applies nothing
        assert res[0][1] == res[1][1] == 0 and not fields_diff(res[0][0].T, par.T)
        for tl in tls:
            tl.record(evolve(par, 100 + t), t, at_breakpoint=True, link_inc=False)
    for now in (10, 11, 12, 13):                                    # new history is poppable, old (13) still is not
        res = [tl.replay(past, now) for tl in tls]
        assert res[0][1] == res[1][1], now
        assert not fields_diff(res[0][0].T, res[1][0].T), now
        assert (res[1][1] > 0) == (now < 13)
    for tl in tls:
        tl.breakpoint(False)
    res = [tl.replay(past, 13) for tl in tls]
    assert res[0][1] == res[1][1] > 0 and not fields_diff(res[0][0].T, res[1][0].T)
    for tl in tls:
        tl.close()


def test_divergence_scan_and_fork_age_oracle(built):
    live = scenes.worldline_batch(6, 128, seed=5)
    rec = live.base.copy()
    live.T["origin"][130] += np.float32(1.0)         # world-line 1: length2 = 3 > 2
    live.T["origin"][300] += np.float32(0.5)         # world-line 2: length2 = 0.75, congruent
    flags, cnt, score = refs.oracle_scan(live, rec, stride=128)
    assert flags.tolist() == [0, 1, 0, 0, 0, 0] and cnt.tolist() == [0, 1, 0, 0, 0, 0]
    assert abs(score[1] - 3.0) < 1e-4 and abs(score[2] - 0.75) < 1e-4 and score[0] == 0
    s = small_scene()
    s.tcoh[5:15] = 100
    s.tcoh[15:20] = 105
    out, n, need = refs.oracle_fork_age(s, 112)
    assert n == 0 and not need                       # 100 + 13 > 112
    out, n, need = refs.oracle_fork_age(s, 113)
    assert n == 0 and need                           # forked since 100: due for resolution
    out, n, need = refs.oracle_fork_age(s, 118)
    assert n == 5 and (out.tstate[15:20] == abi.TS_FORKED).all() and (out.tcoh[15:20] == 118).all()
