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
    seen = {}                                                     # (timeslice, coherency) -> state pushed at that coherency
    checked = 0

    def on_tick(t, phase):
        nonlocal checked
        if phase == "pushed":
            for k, ts in enumerate(slices):
                seen[(k, ts.w.coherency())] = ts.w.download(ts.scene)
    try:
        for t in range(ticks):
            pipeline.run_local(slices, 1, links, on_tick=on_tick)
            for k in range(1, n_ts):
                ts = slices[k]
                if t >= delay:
                    assert ts.replayed == t - delay + 1           # one record came due per tick once the pipeline is full
                    checked += 1
        # one more wall tick by hand: right after the replay at coherency c, timeslice k holds what k-1 pushed at c
        counts = [ts.push(links[k].ptr(), links[k].nbytes()) if k < n_ts - 1 else 0 for k, ts in enumerate(slices)]
        on_tick(ticks, "pushed")
        for k in range(1, n_ts):
            slices[k].pull(links[k - 1].ptr(), counts[k - 1])
        for k in range(1, n_ts):
            ts = slices[k]
            c = ts.w.coherency()
            assert ts.w.replay(ts.tl_in, c, abi.REPLAY_SERVER)
            got, exp = ts.w.download(ts.scene), seen[(k - 1, c)]
            assert not fields_diff(got.T, exp.T) and not fields_diff(got.P, exp.P), k
            # (the tick is deterministic: once synchronised, the past re-derives the same state on its own)
            assert not fields_diff(seen[(k, c)].T, exp.T)
        assert checked == (ticks - delay) * (n_ts - 1)
    finally:
        for ts in slices:
            ts.close()


def test_slot_export_import_roundtrip(built):
    import torch
    s = scenes.arena(200, seed=3, walls=False, temporal=True, fill_side=12.0)
    with world.World(s.n) as a, world.World(s.n) as b, world.Timeline(4, s.n) as ta, world.Timeline(6, s.n + 10) as tb:
        a.upload(s); a.step(1 / 45, 2)
        a.record(ta, 9); a.sync()
        buf = torch.empty(s.n * abi.SNAPSHOT_BYTES, dtype=torch.uint8, device="cuda")
        assert ta.export_slot(9, buf.data_ptr(), buf.numel()) == s.n
        with pytest.raises(world.IcxError):
            ta.export_slot(8, buf.data_ptr(), buf.numel())        # nothing recorded with that stamp
        with pytest.raises(world.IcxError):
            ta.export_slot(9, buf.data_ptr(), 10)                 # buffer too small
        tb.import_slot(9, s.n, buf.data_ptr())
        assert tb.slot_info(9) == (True, s.n)
        b.upload(s); b.set_coherency(9)
        assert b.replay(tb, 9, abi.REPLAY_SERVER)
        ga, gb = a.download(s), b.download(s)
        assert not fields_diff(ga.T, gb.T) and not fields_diff(ga.P, gb.P)
