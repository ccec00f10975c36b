"""CPU suite, part 6: the N > 1 path.  World-line candidates are sharded round-robin over ranks with
no data-path collective; the per-world-line divergence results are gathered (NCCL on the GPU box).
Here: world_size 2 and 3 over gloo, each rank computing its shard's results with the ORACLE scan
(test infrastructure standing in for the CUDA scan), gathered results equal to the unsharded ones."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world_size, port, n_wl, per, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch.distributed as dist
    from intercession_b200 import scenes, shard
    import refs
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world_size)
    full = scenes.worldline_batch(n_wl, per, seed=31)
    rng = np.random.default_rng(4)
    k = rng.choice(full.n, size=40, replace=False)
    full.T["origin"][k] += np.float32(1.5)
    mine = shard.worldlines_of_rank(n_wl, rank, world_size)
    rows = (mine[:, None] * per + np.arange(per)[None, :]).reshape(-1)
    local = scenes.Scene({kk: (v[rows] if isinstance(v, np.ndarray) and len(v) == full.n else v) for kk, v in full.items()})
    f, c, s = refs.oracle_scan(local, full.base, stride=per)
    gf, gc, gs = shard.gather_results(f, c, s, n_wl, rank, world_size)
    ef, ec, es = refs.oracle_scan(full, full.base, stride=per)
    ok = np.array_equal(gf, ef) and np.array_equal(gc, ec) and np.array_equal(gs, es) and int(gc.sum()) > 0
    q.put((rank, bool(ok), shard.most_congruent(gf, gs)))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world_size,n_wl", [(2, 8), (3, 7)])
def test_sharded_scan_gather_gloo(built, world_size, n_wl):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world_size, port, n_wl, 64, q)) for r in range(world_size)]
    [p.start() for p in procs]
    res = [q.get(timeout=180) for _ in procs]
    [p.join(timeout=60) for p in procs]
    assert all(ok for _, ok, _ in res), res
    assert len({best for _, _, best in res}) == 1          # every rank agrees on the most congruent world-line


def test_round_robin_assignment():
    from intercession_b200 import shard
    for ws in (1, 2, 4, 8):
        got = np.sort(np.concatenate([shard.worldlines_of_rank(4096, r, ws) for r in range(ws)]))
        assert np.array_equal(got, np.arange(4096))
    assert shard.most_congruent([1, 0, 0], [0.1, 3.0, 2.0]) == 2
    assert shard.most_congruent([1, 1], [5.0, 4.0]) == 1
