"""CPU suite: host logic of the timeslice pipeline (intercession_b200/pipeline.py) at world_size 2 and 3
over gloo.  Every rank plays one timeslice and moves a small tensor stamped with the coherency it
pushes; the neighbour must receive exactly the stamp schedule() predicts, and the stamp it is told
to replay must be one it received `delay` wall ticks earlier (the snapshot bytes themselves are
device buffers and are covered by tests/test_gpu_pipeline.py)."""
import os
import socket
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world_size, port, delay, ticks, q):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    from intercession_b200 import pipeline
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world_size)
    ok = True
    ring = {}
    for t in range(ticks):
        s = pipeline.schedule(rank, world_size, delay, t)
        out = torch.tensor([s["send"] if s["send"] is not None else -1], dtype=torch.int64)
        inp = torch.full((1,), -2, dtype=torch.int64)
        reqs = []
        if rank < world_size - 1:
            reqs.append(dist.isend(out, rank + 1))
        if rank > 0:
            reqs.append(dist.irecv(inp, rank - 1))
        for r in reqs:
            r.wait()
        if rank > 0:
            ok &= int(inp[0]) == s["recv"]
            ring[int(inp[0]) % (delay + 2)] = int(inp[0])
            if s["replay"] is not None:
                ok &= ring.get(s["replay"] % (delay + 2)) == s["replay"]        # still in a ring of depth delay + 2
            else:
                ok &= t < delay
        else:
            ok &= s["recv"] is None and s["replay"] is None
        ok &= s["coherency"] == (pipeline.start_coherency(rank, world_size, delay) + t) & 0xFFFF
    q.put((rank, bool(ok)))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world_size,delay", [(2, 3), (3, 2)])
def test_pipeline_schedule_gloo(world_size, delay):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world_size, port, delay, 12, q)) for r in range(world_size)]
    [p.start() for p in procs]
    res = [q.get(timeout=120) for _ in procs]
    [p.join(timeout=60) for p in procs]
    assert sorted(res) == [(r, True) for r in range(world_size)]


def test_schedule_wraps_and_past_ids():
    import numpy as np
    sys.path.insert(0, ROOT)
    from intercession_b200 import pipeline
    s = pipeline.schedule(1, 2, 5, 10, base=0xFFFA)
    assert s["coherency"] == (0xFFFA + 10) & 0xFFFF and s["recv"] == (0xFFFA + 15) & 0xFFFF and s["replay"] == s["coherency"]
    ids = np.array([0x10, 0x1D, 0x1E, 0x1F], np.uint32)
    assert list(pipeline.past_ids(ids)) == [0x11, 0x1E, 0x1E, 0x1F]            # refused at link >= 14 (ecs_types.h:127-133)
    assert list(pipeline.past_ids(ids, 2)) == [0x12, 0x1E, 0x1E, 0x1F]
