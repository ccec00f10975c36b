"""CPU: the event-record rules (due: cyclic <=; stale: plain <; old history behind the breakpoint) of the list that stands in for
event streams on the oracle side (tests/refs.py CpuTimeline with wide ids) against the reference's REAL EntityTimestreamMap /
TsBreakpointQueue holding REAL EventMessages (oracle/ref/ref_timeline.cpp) -- the same script tests/test_gpu_events.py runs
against the device."""
import numpy as np
import pytest

from intercession_b200 import abi, timetravel
import refs

needs_ref = pytest.mark.skipif(refs.ref() is None, reason="oracle/_ref not built")


def ev(kind, entity, source=abi.NULL_ENTITY, payload=(0, 0, 0, 0)):
    e = np.zeros(1, abi.event_dt)
    e["type"], e["entity"], e["source"], e["payload"] = kind, entity, source, payload
    return e


def canon(events):
    rows = [(int(e["entity"]), i, int(e["type"]), int(e["source"]), int(e["coherency"]), tuple(int(x) for x in e["payload"])) for i, e in enumerate(events)]
    rows.sort(key=lambda r: (r[0], r[1]))
    return [(r[0], r[2], r[3], r[4], r[5]) for r in rows]


def ids(genesis, link=1):
    return abi.compose_entity(0, genesis, link)


@needs_ref
def test_event_list_follows_the_real_streams():
    real, mine = refs.CpuTimeline(refs.ref(), "icxref"), refs.CpuTimeline(refs.oracle(), "icxo")
    script = [(100, ev(abi.EVENT_JUMP_REQUEST, ids(7), payload=(0xABCD, 1, 1, 0xC1))),
              (100, ev(abi.EVENT_ENTITY_CREATED, ids(200), ids(7), payload=(0xC1, 0, 0, 0))),
              (101, ev(abi.EVENT_JUMP_DEPARTURE, ids(7), payload=(0xABCD, 1, 1, 0xC1))),
              (101, ev(abi.EVENT_ENTITY_REMOVED, ids(7), ids(7))),
              (103, ev(abi.EVENT_JUMP_ARRIVAL, ids(200), payload=(0xABCD, 0xFFFFFFFF, 0, 0xC1))),
              (104, ev(abi.EVENT_ENTITY_REMOVED, ids(9))),
              (106, ev(abi.EVENT_ENTITY_CREATED, ids(201), abi.NULL_ENTITY, payload=(0x41, 0, 0, 0)))]
    for c, r in script:
        real.push_events(c, r); mine.push_events(c, r)
    for now in (99, 100, 100, 102, 105, 105, 110):
        assert canon(mine.pop_events(now)) == canon(real.pop_events(now)), now
    for tl in (real, mine):
        tl.close()
    # breakpoint
    real, mine = refs.CpuTimeline(refs.ref(), "icxref"), refs.CpuTimeline(refs.oracle(), "icxo")
    for c in (40, 41, 42):
        r = ev(abi.EVENT_ENTITY_CREATED, ids(50 + c), ids(3), payload=(0xC1, 0, 0, 0))
        real.push_events(c, r); mine.push_events(c, r)
    real.breakpoint(True); mine.breakpoint(True)
    assert len(mine.pop_events(40)) == len(real.pop_events(40)) == 0
    for now in (40, 41):
        assert canon(mine.pop_events(now, mode=abi.REPLAY_PARALLEL)) == canon(real.pop_events(now, mode=abi.REPLAY_PARALLEL)) != []
    real.breakpoint(False); mine.breakpoint(False)
    assert canon(mine.pop_events(42)) == canon(real.pop_events(42)) != []
    for tl in (real, mine):
        tl.close()
    # coherency wrap
    real, mine = refs.CpuTimeline(refs.ref(), "icxref"), refs.CpuTimeline(refs.oracle(), "icxo")
    for c in (65534, 65535, 0, 1):
        r = ev(abi.EVENT_ENTITY_REMOVED, ids(60 + (c & 7)), ids(1))
        real.push_events(c, r); mine.push_events(c, r)
    got, exp = mine.pop_events(1), real.pop_events(1)
    assert canon(got) == canon(exp) and len(got) == 3
    for tl in (real, mine):
        tl.close()


def test_arrival_and_removal_records():
    there = np.array([abi.compose_entity(2, 77, 1), abi.compose_entity(2, 78, 3)], np.uint32)
    a = timetravel.arrival_events(there)
    assert a["type"].tolist() == [abi.EVENT_ENTITY_CREATED, abi.EVENT_JUMP_ARRIVAL] * 2
    assert a["entity"].tolist() == [there[0], there[0], there[1], there[1]] and (a["source"] == abi.NULL_ENTITY).all()
    assert (a["payload"][0::2, 0] == 0xC1).all() and a["payload"][1::2, 0].tolist() == [77, 78] and (a["payload"][1::2, 1] == 1).all()
    r = timetravel.removal_events(there)
    assert (r["type"] == abi.EVENT_ENTITY_REMOVED).all() and r["entity"].tolist() == there.tolist()


def test_insert_rows_equals_repeated_insert_row():
    from intercession_b200 import scenes
    from util import fields_diff
    s = scenes.arena(60, seed=3, walls=True, temporal=True, kinds=("sphere", "box"))
    r1 = timetravel.traveller_row(s, 10, genesis=5000, timeslice=0)
    r2 = timetravel.traveller_row(s, 20, genesis=3, timeslice=0)
    r2["ids"] = np.array([int(s.ids[30]) + 1], np.uint32)                   # sorts into the middle
    rows, _ = timetravel.insert_row(r1, r2)
    a = timetravel.insert_rows(s, rows)
    b, _ = timetravel.insert_row(s, r1)
    b, _ = timetravel.insert_row(b, r2)
    assert np.array_equal(a.ids, b.ids) and np.all(np.diff(a.ids.astype(np.int64)) > 0)
    for k in ("T", "P", "C"):
        assert not fields_diff(a[k], b[k]), k
    assert np.array_equal(a.tstate, b.tstate) and np.array_equal(a.hasP, b.hasP)
