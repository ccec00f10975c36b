"""GPU: entity lifecycle on the device (icx_entities_create / icx_entities_remove; Cosmos::register_entity / condemn_entity as
the timeline's ENTITY_CREATED / ENTITY_REMOVED handlers and a traveller's arrival call them, server_network_dynamo.cpp:422-463,
:748-812).  The entities a world already holds never leave the device; the result must be the world a fresh upload of the merged
arrays gives -- same rows, same state words, same ticks afterwards, and those ticks are the oracle's."""
import numpy as np
import pytest

from intercession_b200 import abi, scenes, timetravel, world
from intercession_b200.scenes import Scene
import refs
from util import fields_diff

pytestmark = pytest.mark.gpu
KEYS = timetravel.KEYS


def scene_rows(s, keep):
    return Scene({k: (v[keep] if k in KEYS else v) for k, v in s.items()})


def assert_same(a, b, what):
    assert a.n == b.n and np.array_equal(a.ids, b.ids), what
    for key in ("T", "P", "C"):
        d = fields_diff(a[key], b[key])
        assert not d, (what, key, d[:3])
    assert np.array_equal(a.tstate, b.tstate) and np.array_equal(a.tcoh, b.tcoh) and np.array_equal(a.hasP, b.hasP), what


def travellers(s, twins, genesis0, timeslice):
    rows = None
    for i, twin in enumerate(twins):
        r = timetravel.traveller_row(s, int(twin), genesis=genesis0 + i, timeslice=timeslice)
        r.T["origin"][0, 0] += np.float32(0.3)
        r["tstate"][:] = (abi.TS_MERGED, abi.TS_FORKING, abi.TS_FORKED)[i % 3]
        r["tcoh"][:] = 40 + i
        rows = r if rows is None else timetravel.insert_row(rows, r)[0]
    return rows


def test_create_and_remove_equal_a_fresh_upload(built):
    s = scenes.arena(3000, seed=5, walls=True, temporal=True, kinds=("sphere", "box", "aabb"))
    with world.World(s.n + 16) as wa, world.World(s.n + 16) as wb:
        wa.upload(s); wa.set_coherency(77)
        wa.step(1 / 45, 6)
        cur = wa.download(s)
        # three arrivals: one sorts into the middle of the live ids (timeslice 0), two behind them; plus an id that exists already
        mid = travellers(cur, [40], 100, 0)
        mid["ids"] = np.array([abi.compose_entity(0, 100, 5)], np.uint32)           # between the live genesis 100 and 101
        end = travellers(cur, [900, 2100], 5000000, 3)
        dup = scene_rows(cur, np.array([17]))
        rows = timetravel.insert_row(timetravel.insert_row(end, mid.copy())[0], dup)[0]
        assert rows.n == 4
        merged = cur
        for r in (mid, scene_rows(end, np.array([0])), scene_rows(end, np.array([1]))):
            merged, _ = timetravel.insert_row(merged, r)
        assert np.all(np.diff(merged.ids.astype(np.int64)) > 0) and 0 < int(np.searchsorted(merged.ids, mid.ids[0])) < s.n
        assert wa.create_entities(rows) == 3
        assert np.array_equal(wa.entity_ids(), merged.ids) and wa.coherency() == 77 + 6
        assert_same(wa.download(merged), merged, "after create")
        wb.upload(merged); wb.set_coherency(77 + 6)
        cpu = merged
        for t in range(8):
            wa.step(1 / 45); wb.step(1 / 45)
            ga, gb = wa.download(merged), wb.download(merged)
            assert_same(ga, gb, ("tick after create", t))
            ha, hb = wa.contacts(), wb.contacts()
            assert len(ha) == len(hb) and not fields_diff(ha, hb), t
            ia, ib = wa.interceptions(), wb.interceptions()
            assert len(ia) == len(ib) and not fields_diff(ia, ib), t
            if t < 3:                                                               # ... and they are the oracle's ticks (pair list from the device)
                out = refs.procedures_tick(refs.oracle(), "icxo", cpu, coherency=77 + 6 + t, cand=wa.all_pairs())
                assert len(out["hits"]) == len(ha) and not fields_diff(out["hits"], ha, skip=("packetA", "packetB")), t
                assert len(out["ints"]) == len(ia) and not fields_diff(out["ints"], ia), t
                cpu = Scene({k: v for k, v in out.items() if k not in ("hits", "ints", "seconds")})
                assert_same(ga, cpu, ("oracle tick", t))
        assert sum(len(x) for x in (ia,)) >= 0
        # remove: the two travellers of timeslice 3, one wall-side body, and an id that is not there
        cur = wa.download(merged)
        gone = np.array([end.ids[0], end.ids[1], merged.ids[1234], 0x7FFFFFF0], np.uint32)
        keep = ~np.isin(merged.ids, gone)
        left = scene_rows(cur, keep)
        assert wa.remove_entities(gone) == 3
        assert np.array_equal(wa.entity_ids(), left.ids)
        assert_same(wa.download(left), left, "after remove")
        wb.upload(left); wb.set_coherency(wa.coherency())
        for t in range(5):
            wa.step(1 / 45); wb.step(1 / 45)
            assert_same(wa.download(left), wb.download(left), ("tick after remove", t))
        assert wa.remove_entities(gone) == 0 and wa.create_entities(dup) == 0


def test_records_find_their_entities_across_a_create(built):
    """The ring is addressed by id: rows that shift when an entity is created in the receiving cosmos still take their records."""
    from intercession_b200.pipeline import past_ids
    fut = scenes.arena(400, seed=9, walls=True, temporal=True, kinds=("sphere", "box"))
    past = fut.copy(); past["ids"] = past_ids(fut.ids)
    otl = refs.CpuTimeline(refs.oracle(), "icxo")
    with world.World(fut.n) as wf, world.World(past.n + 4) as wp, world.Timeline(8, fut.n) as tl:
        wf.upload(fut); wp.upload(past)
        st = fut
        for c in (50, 51):
            wf.step(1 / 45)
            st = wf.download(fut)
            wf.record(tl, c); otl.record(st, c)
        assert wp.replay(tl, 50)
        cur, _ = otl.replay(past, 50)
        row = travellers(cur, [60], 77, 1)
        row["ids"] = np.array([abi.compose_entity(0, 77, 9)], np.uint32)            # sorts in front of most rows
        wp.create_entities(row)
        merged, at = timetravel.insert_row(cur, row)
        assert at < past.n // 2
        assert wp.replay(tl, 51)
        exp, n = otl.replay(merged, 51)
        assert n == fut.n - 6 or n > 0
        assert_same(wp.download(merged), exp, "replay after create")
        assert tl.id_mismatches() == 1                                              # the traveller has no future stream
    otl.close()
