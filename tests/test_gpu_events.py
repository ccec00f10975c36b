"""GPU: event records in the device timeline (ENTITY_CREATED / ENTITY_REMOVED / JUMP_*) against the
reference's REAL EntityTimestreamMap / TsBreakpointQueue holding REAL EventMessages with the reference's own params structs
(oracle/ref/ref_timeline.cpp): which records come out at which coherency, in which per-entity order, what is dropped as stale,
what the breakpoint hides -- and that the records travel inside the snapshot transport buffer."""
import numpy as np
import pytest

from intercession_b200 import abi, scenes, world
from intercession_b200.pipeline import past_ids
import refs

pytestmark = pytest.mark.gpu


def ev(kind, entity, source=abi.NULL_ENTITY, payload=(0, 0, 0, 0)):
    e = np.zeros(1, abi.event_dt)
    e["type"], e["entity"], e["source"], e["payload"] = kind, entity, source, payload
    return e


def canon(events):
    """Per-entity order is the container's; the order streams are visited in is not defined by the reference."""
    rows = [(int(e["entity"]), i, int(e["type"]), int(e["source"]), int(e["coherency"]), tuple(int(x) for x in e["payload"])) for i, e in enumerate(events)]
    rows.sort(key=lambda r: (r[0], r[1]))
    return [(r[0], r[2], r[3], r[4], r[5]) for r in rows]


def ids(genesis, link=1, host=0):
    return abi.compose_entity(host, genesis, link)


@pytest.fixture()
def real_containers(built):
    if refs.ref() is None:
        pytest.skip("oracle/_ref not built")
    return refs.CpuTimeline(refs.ref(), "icxref")


def test_event_records_follow_the_real_entity_timestream_map(real_containers):
    otl = real_containers
    script = [  # (push coherency, records)
        (100, [ev(abi.EVENT_JUMP_REQUEST, ids(7), payload=(0xABCD, 1, 1, 0xC1))]),
        (100, [ev(abi.EVENT_ENTITY_CREATED, ids(200), ids(7), payload=(0xC1, 0, 0, 0))]),
        (101, [ev(abi.EVENT_JUMP_DEPARTURE, ids(7), payload=(0xABCD, 1, 1, 0xC1)), ev(abi.EVENT_ENTITY_REMOVED, ids(7), ids(7))]),
        (103, [ev(abi.EVENT_JUMP_ARRIVAL, ids(200), payload=(0xABCD, -1 & 0xFFFFFFFF, 0, 0xC1))]),
        (104, [ev(abi.EVENT_ENTITY_REMOVED, ids(9))]),
        (106, [ev(abi.EVENT_ENTITY_CREATED, ids(201), abi.NULL_ENTITY, payload=(0x41, 0, 0, 0))]),
    ]
    with world.Timeline(16, 8) as tl:
        for c, recs in script:
            for r in recs:
                tl.push_events(c, r); otl.push_events(c, r)
        total = 0
        for now in (99, 100, 100, 102, 105, 105, 110):          # nothing due; due; already popped; 101 is stale at 102; 103 / 104 stale at 105; 106 stale at 110
            got, exp = tl.pop_events(now), otl.pop_events(now)
            assert canon(got) == canon(exp), now
            total += len(got)
        assert total == 2                                          # only the two records of tick 100 were popped on time
    # breakpoint: what was pushed before it is old history, reachable only at the breakpoint
    otl2 = refs.CpuTimeline(refs.ref(), "icxref")
    with world.Timeline(16, 8) as tl:
        for c in (40, 41, 42):
            r = ev(abi.EVENT_ENTITY_CREATED, ids(50 + c), ids(3), payload=(0xC1, 0, 0, 0))
            tl.push_events(c, r); otl2.push_events(c, r)
        tl.set_breakpoint(); otl2.breakpoint(True)
        assert len(tl.pop_events(40)) == 0 and len(otl2.pop_events(40)) == 0
        for now in (40, 41):
            assert canon(tl.pop_events(now, abi.REPLAY_PARALLEL)) == canon(otl2.pop_events(now, abi.REPLAY_PARALLEL)) != []
        tl.clear_breakpoint(); otl2.breakpoint(False)
        assert canon(tl.pop_events(42)) == canon(otl2.pop_events(42)) != []
    # across the coherency wrap: pop_from_timestream compares cyclically, the stale filter plainly (server_network_dynamo.cpp:391)
    otl3 = refs.CpuTimeline(refs.ref(), "icxref")
    with world.Timeline(16, 8) as tl:
        for c in (65534, 65535, 0, 1):
            r = ev(abi.EVENT_ENTITY_REMOVED, ids(60 + (c & 7)), ids(1))
            tl.push_events(c, r); otl3.push_events(c, r)
        got, exp = tl.pop_events(1), otl3.pop_events(1)
        assert canon(got) == canon(exp) and len(got) == 3        # 65534 / 65535 are not "< 1": they come out, 0 is dropped
    for t in (otl2, otl3):
        t.close()
    with world.Timeline(4, 8) as tl:
        with pytest.raises(world.IcxError):
            tl.push_events(5, np.concatenate([ev(abi.EVENT_ENTITY_REMOVED, ids(i)) for i in range(abi.TIMELINE_EVENTS_PER_TICK + 1)]))


def test_event_records_travel_with_the_snapshot(built):
    import torch
    fut = scenes.arena(300, seed=3, walls=True, temporal=True, kinds=("sphere", "box"))
    past = fut.copy(); past["ids"] = past_ids(fut.ids)
    buf = torch.empty(abi.snapshot_capacity(fut.n, fut.n), dtype=torch.uint8, device="cuda")
    with world.World(fut.n) as wf, world.World(past.n) as wp, world.Timeline(4, fut.n) as tl_out, world.Timeline(8, fut.n) as tl_in:
        wf.upload(fut); wp.upload(past)
        recs = np.concatenate([ev(abi.EVENT_ENTITY_CREATED, ids(900), ids(12), payload=(0xC1, 0, 0, 0)), ev(abi.EVENT_JUMP_ARRIVAL, ids(900), payload=(77, 1, 0, 0xC1))])
        for c in (20, 21, 22):
            wf.step(1 / 45)
            if c == 21:
                tl_out.push_events(c, recs)                       # before the record of the same tick
            wf.record(tl_out, c)
            n = tl_out.export_slot(c, buf.data_ptr(), buf.numel(), wf.stream())
            wf.sync()
            tl_in.import_slot(c, n, buf.data_ptr(), buf.numel(), wp.stream())
        assert len(tl_in.pop_events(20)) == 0
        assert wp.replay(tl_in, 20)
        got = tl_in.pop_events(21)
        assert len(got) == 2 and np.array_equal(got["type"], recs["type"]) and np.array_equal(got["entity"], recs["entity"])
        assert np.array_equal(got["payload"], recs["payload"]) and (got["coherency"] == 21).all()
        assert wp.replay(tl_in, 21)                               # the tick's snapshot is still there: events and states are consumed separately
        assert len(tl_in.pop_events(22)) == 0 and wp.replay(tl_in, 22)
        exp = wf.download(fut)
        gotp = wp.download(past)
        assert np.array_equal(gotp.T["origin"], exp.T["origin"])
