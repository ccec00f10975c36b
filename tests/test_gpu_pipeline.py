"""GPU: the timeslice pipeline (config 5) with three timeslices in one process.  Without
interceptions every past timeslice is overwritten each tick by the record its future neighbour
pushed `delay` ticks earlier (server_network_dynamo.cpp:377-421, :642-679), so after the replay at
coherency c its state must equal, bit for bit, the state the neighbour held when its own coherency
was c -- through record, export, device transport, import and replay."""
import numpy as np
import pytest

from intercession_b200 import abi, pipeline, scenes, world
from util import fields_diff

pytestmark = pytest.mark.gpu


def test_three_timeslices_replay_their_future(built):
    delay, n_ts, ticks = 3, 3, 12
    base = scenes.arena(300, seed=21, walls=True, temporal=True, spacing=1.0, jitter=0.2, fill_side=10.0)
    slices, links = [], []
    for k in range(n_ts):
        s = base.copy()
        s["ids"] = pipeline.past_ids(base.ids, k)
        if k:
            s.T["origin"][:300] += np.float32(0.01 * k)          # the past starts somewhere else: the replay must overwrite it
        slices.append(pipeline.Timeslice(s, k, n_ts, delay))
        if k < n_ts - 1:
            links.append(pipeline.LoopbackLink(s.n))
    pushed = {}                                                   # (timeslice, coherency) -> state it recorded at that coherency
    checked = 0

    def on_tick(t, phase):
        nonlocal checked
        if phase == "replayed":                                   # phase 2 done, phase 4 (record) comes next: what is about to be pushed
            for k, ts in enumerate(slices):
                c = ts.w.coherency()
                cur = ts.w.download(ts.scene)
                pushed[(k, c)] = cur
                if k > 0 and t >= delay:
                    # right after the replay at coherency c, timeslice k holds what k - 1 pushed at c, bit for bit
                    exp = pushed[(k - 1, c)]
                    assert not fields_diff(cur.T, exp.T) and not fields_diff(cur.P, exp.P) and not fields_diff(cur.C, exp.C), (k, c)
                    checked += 1
    try:
        pipeline.run_local(slices, ticks, links, on_tick=on_tick)
        for k in range(1, n_ts):
            assert slices[k].replayed == ticks - delay            # one record came due per tick once the pipeline is full
            assert slices[k].tl_in.id_mismatches() == 0
        assert checked == (ticks - delay) * (n_ts - 1)
    finally:
        for ts in slices:
            ts.close()


def test_middle_timeslice_forwards_what_it_replayed(built):
    """Reference order inside the networker: replay from the future (:377-421), THEN push to the past (:642-679).  A change
    injected into timeslice 0 at coherency c (a behaviour, an input: anything the tick itself does not derive) must therefore
    reach timeslice 2 through timeslice 1's record stamped c -- with push before replay the middle timeslice would forward
    its own stale state instead."""
    delay, n_ts = 2, 3
    base = scenes.arena(120, seed=5, walls=True, temporal=True, spacing=1.0, jitter=0.2, fill_side=8.0)
    slices, links = [], []
    for k in range(n_ts):
        s = base.copy()
        s["ids"] = pipeline.past_ids(base.ids, k)
        slices.append(pipeline.Timeslice(s, k, n_ts, delay))
        if k < n_ts - 1:
            links.append(pipeline.LoopbackLink(s.n))
    inject_at = slices[0].w.coherency() + 3
    idx = np.array([10, 50, 90], np.uint32)
    marker = np.float32(123.5)
    seen, before = {}, {}

    def on_tick(t, phase):
        if phase == "replayed":
            w0 = slices[0].w
            if w0.coherency() == inject_at:                       # external edit on the future-most timeslice, before it records
                cur = w0.download(slices[0].scene)
                before["T"] = cur.T.copy()
                P = np.ascontiguousarray(cur.P[idx]); P["linearDrag"] = marker
                T = np.ascontiguousarray(cur.T[idx]); T["scale"] *= np.float32(1.25)
                w0.update(idx, T=T, P=P)
            for k, ts in enumerate(slices):
                seen[(k, ts.w.coherency())] = ts.w.download(ts.scene)
    try:
        pipeline.run_local(slices, 3 + 2 * delay + 3, links, on_tick=on_tick)
        got = seen[(2, inject_at)]                                # timeslice 2 right after ITS replay at that coherency
        src = seen[(0, inject_at)]                                # timeslice 0 right after the injection
        assert np.all(got.P["linearDrag"][idx] == marker)
        mid = seen[(1, inject_at)]
        assert np.all(mid.P["linearDrag"][idx] == marker)
        assert np.array_equal(got.T["scale"][idx], mid.T["scale"][idx]) and np.array_equal(got.T["scale"][idx], src.T["scale"][idx])
        assert not np.array_equal(got.T["scale"][idx], before["T"]["scale"][idx])
    finally:
        for ts in slices:
            ts.close()


def test_slot_export_import_roundtrip(built):
    import torch
    s = scenes.arena(200, seed=3, walls=False, temporal=True, fill_side=12.0)
    with world.World(s.n) as a, world.World(s.n) as b, world.Timeline(4, s.n) as ta, world.Timeline(6, s.n + 10) as tb:
        a.upload(s); a.step(1 / 45, 2)
        a.record(ta, 9); a.sync()
        buf = torch.empty(abi.snapshot_capacity(s.n, s.n), dtype=torch.uint8, device="cuda")      # the first record of a ring carries every collider row
        assert ta.export_slot(9, buf.data_ptr(), buf.numel()) == s.n
        with pytest.raises(world.IcxError):
            ta.export_slot(8, buf.data_ptr(), buf.numel())        # nothing recorded with that stamp
        with pytest.raises(world.IcxError):
            ta.export_slot(9, buf.data_ptr(), 10)                 # buffer too small
        tb.import_slot(9, s.n, buf.data_ptr(), buf.numel())
        assert tb.slot_info(9) == (True, s.n)
        sb = s.copy(); sb["ids"] = pipeline.past_ids(s.ids)      # the record is addressed to the past's ids (chain-link + 1)
        b.upload(sb); b.set_coherency(9)
        assert b.replay(tb, 9, abi.REPLAY_SERVER)
        ga, gb = a.download(s), b.download(s)
        assert not fields_diff(ga.T, gb.T) and not fields_diff(ga.P, gb.P) and tb.id_mismatches() == 0
