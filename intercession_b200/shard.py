"""Sharding of independent world-line candidates over the GPUs of one box.

A single scene does not shard: the reference's pair loop is sequential (replicas only).  A batch
of independent world-line candidates does, with zero coupling during the run: world-line w lives
on rank w % world_size.  The ONLY collective is the gather of the per-world-line divergence
results {flag u8, incongruent u32, score f64} after a scan -- NCCL over NVLink on the GPU box,
gloo in the CPU tests.  torch.distributed is plumbing here; the data path never crosses ranks.
"""
import numpy as np


def worldlines_of_rank(n_worldlines, rank, world_size):
    """Round-robin assignment: the global indices of the world-lines rank `rank` simulates."""
    return np.arange(rank, n_worldlines, world_size, dtype=np.int64)


def gather_results(flags, count, score, n_worldlines, rank, world_size, device=None):
    """All-gather the per-rank scan results into global per-world-line arrays (every rank gets them).

    flags/count/score: local arrays of this rank's world-lines, in the order of
    worldlines_of_rank().  Ranks may own different numbers of world-lines; they are padded to the
    maximum for the collective.  Returns (flags u8[n], count u32[n], score f64[n]).
    """
    mine = worldlines_of_rank(n_worldlines, rank, world_size)
    assert len(flags) == len(count) == len(score) == len(mine)
    if world_size == 1:
        return np.asarray(flags, np.uint8), np.asarray(count, np.uint32), np.asarray(score, np.float64)
    import torch
    import torch.distributed as dist
    per = (n_worldlines + world_size - 1) // world_size
    # one packed f64 buffer per rank: [flag, count, score] x per  (counts < 2^53 are exact in f64)
    local = torch.zeros(3 * per, dtype=torch.float64)
    local[0:len(mine)] = torch.from_numpy(np.asarray(flags, np.float64))
    local[per:per + len(mine)] = torch.from_numpy(np.asarray(count, np.float64))
    local[2 * per:2 * per + len(mine)] = torch.from_numpy(np.asarray(score, np.float64))
    if device is not None:
        local = local.to(device)
    out = torch.empty(world_size * 3 * per, dtype=torch.float64, device=local.device)
    dist.all_gather_into_tensor(out, local)
    out = out.cpu().numpy().reshape(world_size, 3, per)
    gflags = np.zeros(n_worldlines, np.uint8)
    gcount = np.zeros(n_worldlines, np.uint32)
    gscore = np.zeros(n_worldlines, np.float64)
    for r in range(world_size):
        idx = worldlines_of_rank(n_worldlines, r, world_size)
        gflags[idx] = out[r, 0, :len(idx)].astype(np.uint8)
        gcount[idx] = out[r, 1, :len(idx)].astype(np.uint32)
        gscore[idx] = out[r, 2, :len(idx)]
    return gflags, gcount, gscore


def most_congruent(flags, score):
    """The candidate the reference would keep: 'most congruent' world-line (README.md:48) = a
    non-divergent one with the lowest score, else the lowest score overall."""
    flags = np.asarray(flags, bool)
    score = np.asarray(score)
    ok = np.nonzero(~flags)[0]
    pool = ok if len(ok) else np.arange(len(score))
    return int(pool[np.argmin(score[pool])])
