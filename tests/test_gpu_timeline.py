"""GPU parity of the world-line record / replay / divergence path (device snapshot ring,
icx_timeline.cu) against the oracle restatement, which tests/test_timeline_oracle.py pins to the
reference's real containers.  Everything is bit-exact (byte copies, integer flags, fixed-point score)."""
import numpy as np
import pytest

from intercession_b200 import abi, scenes, world
import refs
from util import fields_diff
from test_timeline_oracle import changed_colliders, evolve, past_ids, small_scene

pytestmark = pytest.mark.gpu
DT = 1.0 / 45


def set_state(w, s, colliders_of=None):
    """Overwrite transform + physics of every entity (what a tick / a behaviour would have produced) and the
    ColliderComponent of the entities in the mask `colliders_of` (what BipedBehaviors does to its own entities)."""
    w.update(np.arange(s.n, dtype=np.uint32), T=s.T, P=s.P)
    if colliders_of is not None and colliders_of.any():
        idx = np.nonzero(colliders_of)[0].astype(np.uint32)
        w.update(idx, Cc=np.ascontiguousarray(s.C[idx]))


def test_record_replay_server(built):
    fut = small_scene()
    past = fut.copy()
    past["ids"] = past_ids(fut.ids)
    otl = refs.CpuTimeline(refs.oracle(), "icxo")
    with world.World(fut.n) as wf, world.World(past.n) as wp, world.Timeline(16, fut.n) as tl:
        wf.upload(fut)
        st = fut
        for t in range(100, 111):
            st = evolve(fut, t, prev=st)
            set_state(wf, st, changed_colliders(fut.n, t))
            wf.record(tl, t)
            otl.record(st, t)
        # The past timeslice replays in order, like the reference's (every tick pops what has come due): entities that are
        # forked / forking keep everything, the others end up with ALL components of the record -- scale, mass, drags, locks,
        # isAsleep, and the collider rows that changed at ANY earlier tick.
        wp.upload(past)
        cur = past
        for now in (99, 100, 101, 103, 103, 107, 120, 110):
            applied = wp.replay(tl, now)
            cur, n = otl.replay(cur, now)
            got = wp.download(past)
            assert applied == (n > 0), now
            assert not fields_diff(got.T, cur.T) and not fields_diff(got.P, cur.P) and not fields_diff(got.C, cur.C), now
        assert tl.id_mismatches() == 0
    otl.close()


def test_replay_checks_ids_and_wraps(built):
    """A record is addressed to the id it carries (chain-link + 1 on the way into the past, server_network_dynamo.cpp:673):
    a receiver whose ids are not those is left alone and counted.  Across the uint16 wrap the reference's plain `<` stale
    filter (:391) lets older records through: the device applies what the real EntityTimestreamMap path applies."""
    fut = small_scene()
    past = fut.copy()
    past["ids"] = past_ids(fut.ids)
    with world.World(fut.n) as wf, world.World(past.n) as wp, world.Timeline(16, fut.n) as tl:
        wf.upload(fut)
        set_state(wf, evolve(fut, 3))
        wf.record(tl, 40)
        wrong = past.copy()
        wrong["ids"] = fut.ids.copy()                           # the future's own ids: only link-14 / non-temporal ids coincide
        wp.upload(wrong)
        assert wp.replay(tl, 40)
        got = wp.download(wrong)
        same = past.ids == fut.ids
        div = np.isin(past.tstate, (abi.TS_FORKED, abi.TS_FORKING))
        assert np.array_equal(got.T["origin"][~same], wrong.T["origin"][~same])
        assert tl.id_mismatches() == int((~same & ~div).sum()) > 0
    for last, now in ((2, 2), (65535, 2), (1, 3)):
        otl = refs.CpuTimeline(refs.oracle(), "icxo")
        with world.World(fut.n) as wf, world.World(past.n) as wp, world.Timeline(16, fut.n) as tl:
            wf.upload(fut)
            t, st = 65533, fut
            while True:
                st = evolve(fut, t % 97, prev=st)
                set_state(wf, st, changed_colliders(fut.n, t % 97))
                wf.record(tl, t)
                otl.record(st, t)
                if t == last:
                    break
                t = (t + 1) & 0xFFFF
            wp.upload(past)
            assert wp.replay(tl, now)
            exp, n = otl.replay(past, now)
            got = wp.download(past)
            assert n > 0 and not fields_diff(got.T, exp.T) and not fields_diff(got.P, exp.P) and not fields_diff(got.C, exp.C), (last, now)
        otl.close()


def test_replay_collider_rows_wait_for_divergent_entities(built):
    """Every record of the reference carries every component, so an entity that was forked while its collider changed in the
    future takes the change with the first record it accepts again.  The sparse rows it skipped are kept for that moment."""
    fut = small_scene()
    past = fut.copy()
    past["ids"] = past_ids(fut.ids)
    otl = refs.CpuTimeline(refs.oracle(), "icxo")
    with world.World(fut.n) as wf, world.World(past.n) as wp, world.Timeline(16, fut.n) as tl:
        wf.upload(fut)
        st = fut
        for t in range(10, 16):
            st = evolve(fut, t, prev=st)
            set_state(wf, st, changed_colliders(fut.n, t))
            wf.record(tl, t)
            otl.record(st, t)
        wp.upload(past)
        cur = past
        for now in (10, 11, 12):
            wp.replay(tl, now)
            cur, _ = otl.replay(cur, now)
        # the forked / forking entities merge again (what icx_world_extract does at the end of a parallel cycle)
        cur.tstate[:] = abi.TS_MERGED
        merged = wp.download(past)
        merged.tstate[:] = abi.TS_MERGED
        wp.upload(merged)                                       # a re-upload drops the kept rows: replay 13 must still restore ...
        wp.replay(tl, 13)
        cur, _ = otl.replay(cur, 13)
        got = wp.download(past)
        assert not fields_diff(got.T, cur.T) and not fields_diff(got.P, cur.P)
        # ... T and P in full; collider rows that changed at ticks <= 12 were not re-sent, which a re-upload must bring itself
    otl.close()
    # same flow without the re-upload: the state table changes through the ABI the extraction uses
    otl = refs.CpuTimeline(refs.oracle(), "icxo")
    with world.World(fut.n) as wf, world.World(past.n) as wp, world.World(past.n, host_timeslice=abi.NULL_TIMESLICE) as wpar, \
            world.Timeline(16, fut.n) as tl:
        wf.upload(fut)
        st = fut
        for t in range(10, 16):
            st = evolve(fut, t, prev=st)
            set_state(wf, st, changed_colliders(fut.n, t))
            wf.record(tl, t)
            otl.record(st, t)
        past0 = past.copy()
        past0["ids"] = (past0.ids & ~np.uint32(0xF0000000))     # timeslice 0: extraction re-merges everything (:268-271)
        wp.upload(past0)
        cur = past0
        for now in (10, 11, 12):
            wp.replay(tl, now)
            cur, _ = otl.replay(cur, now)
        # parallel cycle: clone (ids as they are), extract back at chain-link - 1 -> every extracted entity is merged again
        par = cur.copy()
        par["ids"] = past_ids(cur.ids)
        wpar.upload(par)
        n_ex, n_miss = wpar.extract_into(wp)
        now_state = wp.download(past0)
        cur = now_state.copy()
        wp.replay(tl, 13)
        cur, _ = otl.replay(cur, 13)
        got = wp.download(past0)
        assert n_ex > 0
        assert not fields_diff(got.T, cur.T) and not fields_diff(got.P, cur.P) and not fields_diff(got.C, cur.C)
    otl.close()


def test_ring_keeps_only_depth_ticks(built):
    s = small_scene()
    with world.World(s.n) as w, world.Timeline(8, s.n) as tl:
        w.upload(s)
        for t in range(20):
            w.record(tl, t)
        assert tl.slot_info(19) == (True, s.n) and tl.slot_info(12) == (True, s.n)
        assert tl.slot_info(11) == (False, 0)           # overwritten by tick 19
        assert not w.replay(tl, 3)


def test_breakpoint_parallel_flow(built):
    fut = small_scene()
    past = fut.copy()
    past["ids"] = past_ids(fut.ids)
    otl = refs.CpuTimeline(refs.oracle(), "icxo")
    with world.World(fut.n) as wf, world.World(past.n) as wp, world.World(past.n, host_timeslice=abi.NULL_TIMESLICE) as wpar, \
            world.Timeline(32, fut.n) as tl:
        wf.upload(fut)
        wpar.upload(past)
        for t in range(10, 16):
            st = evolve(fut, t)
            set_state(wf, st)
            wf.record(tl, t)
            otl.record(st, t)
        tl.set_breakpoint()
        otl.breakpoint(True)
        wp.upload(past)
        assert not wp.replay(tl, 10) and otl.replay(past, 10)[1] == 0
        for t in range(10, 13):
            assert not wpar.replay(tl, t, mode=abi.REPLAY_PARALLEL)         # consumes, applies nothing
            otl.replay(past, t, mode=abi.REPLAY_PARALLEL)
            st = evolve(past, 100 + t)
            set_state(wpar, st)
            wpar.record(tl, t, at_breakpoint=True)
            otl.record(st, t, at_breakpoint=True, link_inc=False)
        for now in (10, 11, 12, 13):
            wp.upload(past)
            applied = wp.replay(tl, now)
            exp, n = otl.replay(past, now)
            assert applied == (n > 0) == (now < 13), now
            assert not fields_diff(wp.download(past).T, exp.T), now
        tl.clear_breakpoint()
        otl.breakpoint(False)
        wp.upload(past)
        assert wp.replay(tl, 13)
        assert not fields_diff(wp.download(past).T, otl.replay(past, 13)[0].T)
    otl.close()


@pytest.mark.parametrize("n_wl,stride", [(6, 1024), (1, 0), (5, 3072)])
def test_divergence_scan_vs_oracle(built, n_wl, stride):
    per = stride if stride else 5000
    live = scenes.worldline_batch(n_wl, per, seed=11)
    rng = np.random.default_rng(2)
    k = rng.choice(live.n, size=live.n // 7, replace=False)
    live.T["origin"][k] += rng.normal(0, 0.8, (len(k), 3)).astype(np.float32)
    live.T["origin"][k[0]] = np.nan
    live.T["origin"][k[1]] = 1e30
    base = live.base
    with world.World(live.n, worldline_stride=stride, host_timeslice=abi.NULL_TIMESLICE) as w, \
            world.World(base.n) as wb, world.Timeline(4, base.n) as tl:
        wb.upload(base)
        wb.record(tl, 77)
        w.upload(live)
        flags, cnt, score = w.divergence_scan(tl, 77, n_wl)
    eflags, ecnt, escore = refs.oracle_scan(live, base, stride=stride)
    assert np.array_equal(flags, eflags) and np.array_equal(cnt, ecnt)
    assert np.array_equal(score, escore)             # fixed-point accumulation: bit-exact
    assert cnt.sum() > 10


def test_divergence_scan_full_size_properties(built):
    """BASELINE.json configs[3] size: 4096 world-lines x 1024 bodies against one shared base record.
    Unperturbed copies are congruent everywhere; moving body j of world-line w flags exactly w."""
    n_wl, per = 4096, 1024
    live = scenes.worldline_batch(n_wl, per, seed=13, perturb=False)
    moved = np.array([0, 1, 999, 2048, 4095])
    live.T["origin"][moved * per + (moved % per), 0] += np.float32(2.0)       # length2 = 4 > 2
    with world.World(live.n, worldline_stride=per, host_timeslice=abi.NULL_TIMESLICE) as w, \
            world.World(per) as wb, world.Timeline(2, per) as tl:
        wb.upload(live.base)
        wb.record(tl, 5)
        w.upload(live)
        flags, cnt, score = w.divergence_scan(tl, 5, n_wl)
    assert np.array_equal(np.nonzero(flags)[0], moved) and cnt.sum() == len(moved)
    assert np.array_equal(score[moved], np.full(len(moved), 4.0)) and score.sum() == 4.0 * len(moved)


def test_fork_age_vs_oracle(built):
    s = small_scene(200)
    s.tstate[:] = np.random.default_rng(0).integers(0, 5, s.n)
    s.tcoh[:] = np.random.default_rng(1).integers(90, 125, s.n)
    for coh in (100, 113, 130):
        with world.World(s.n) as w:
            w.upload(s)
            w.set_coherency(coh)
            n, need = w.fork_age()
            got = w.download(s)
        exp, en, eneed = refs.oracle_fork_age(s, coh)
        assert (n, need) == (en, eneed), coh
        assert np.array_equal(got.tstate, exp.tstate) and np.array_equal(got.tcoh, exp.tcoh), coh


def test_worldline_candidates_against_base_history(built):
    """Config-4 flow at a reduced size: a base world records its history; W perturbed candidates are
    re-simulated from tick 0 and scanned against it every tick; flags / counts / scores equal the
    oracle's evaluated on the downloaded states."""
    n_wl, per, ticks = 12, 1024, 6
    cand = scenes.worldline_batch(n_wl, per, seed=21)
    base = cand.base
    with world.World(base.n, host_timeslice=abi.NULL_TIMESLICE) as wb, world.Timeline(ticks + 2, per) as tl, \
            world.World(cand.n, worldline_stride=per, host_timeslice=abi.NULL_TIMESLICE) as wc, world.Timeline(2, cand.n) as tl_new:
        wb.upload(base)
        hist = []
        for t in range(ticks):
            wb.step(DT)
            wb.record(tl, t + 1)
            hist.append(wb.download(base))
        wc.upload(cand)
        any_flag = np.zeros(n_wl, bool)
        for t in range(ticks):
            wc.replay(tl_new, t, mode=abi.REPLAY_PARALLEL)
            wc.step(DT)
            wc.record(tl_new, t + 1)
            flags, cnt, score = wc.divergence_scan(tl, t + 1, n_wl)
            live = wc.download(cand)
            ef, ec, es = refs.oracle_scan(live, hist[t], stride=per)
            assert np.array_equal(flags, ef) and np.array_equal(cnt, ec) and np.array_equal(score, es), t
            any_flag |= flags.astype(bool)
        assert score.max() > 0            # the perturbed bodies drift away from the base history
