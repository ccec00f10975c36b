"""Timeslice pipeline (BASELINE.json configs[4]): one timeslice per GPU / rank,
coupled only by the future -> past snapshot stream.

Reference: every timeslice is a ServerCosmosContext whose fixed step runs networker, behaver,
physicser in that order (source/server/server_cosmos_context.cpp:59-66).  Inside the networker
(source/server/server_network_dynamo.cpp) phase 2 pops the records of the FUTURE neighbour whose
coherency has come due (:377-421), phase 4 pushes this tick's state, stamped with the current
coherency, to the PAST neighbour (:642-679), phase 5 ages forks (:681-705).  Timeslice k runs
`delay` ticks behind timeslice k-1 (timeline_config.h:34), so the record it replays at coherency c
is the one its neighbour pushed `delay` wall ticks earlier.

Here the per-entity message lists are dense device snapshots (icx_record / icx_replay); between
ranks a snapshot travels as ONE buffer of 100 B per entity (icx_timeline_slot_export / _import):
ncclSend / ncclRecv over NVLink when the ranks are processes, a device copy when several
timeslices share a process.  No other data-path collective exists.
"""
import numpy as np

from . import abi, world

DT = 1.0 / 45


def start_coherency(rank, n_ranks, delay, base=100):
    """Timeslice k starts k * delay ticks in the past of timeslice 0."""
    return (base + (n_ranks - 1 - rank) * delay) & 0xFFFF


def schedule(rank, n_ranks, delay, tick, base=100):
    """Stamps moved at wall tick `tick` (0-based) by timeslice `rank`: what it pushes to the past,
    what it receives from the future, and which record it can replay (None while the pipeline fills)."""
    c = (start_coherency(rank, n_ranks, delay, base) + tick) & 0xFFFF
    send = c if rank < n_ranks - 1 else None
    recv = (c + delay) & 0xFFFF if rank > 0 else None            # the future neighbour's current coherency
    replay = c if (rank > 0 and tick >= delay) else None         # pushed `delay` wall ticks ago
    return {"coherency": c, "send": send, "recv": recv, "replay": replay}


def past_ids(ids, steps=1):
    """increment_causal_chain_link (ecs/ecs_types.h:118-136) applied `steps` times: refused at link >= 14."""
    ids = np.asarray(ids, np.uint32).copy()
    for _ in range(steps):
        link = ids & 15
        ids = np.where(link + 1 >= 15, ids, ids + 1).astype(np.uint32)
    return ids


class LoopbackLink:
    """Snapshot transport between two timeslices of ONE process: a device buffer both sides see."""

    def __init__(self, n_entities, device=0):
        import torch
        self.buf = torch.empty(max(n_entities, 1) * abi.SNAPSHOT_BYTES, dtype=torch.uint8, device=torch.device("cuda", device))
        self.pending = None

    def ptr(self):
        return self.buf.data_ptr()

    def nbytes(self):
        return self.buf.numel()


class Timeslice:
    """One ServerCosmosContext worth of device state: the world, the ring of records received from
    the future neighbour, and a two-slot ring for the record it pushes to the past."""

    def __init__(self, scene, rank, n_ranks, delay, device=0, base=100):
        self.rank, self.n_ranks, self.delay, self.base = rank, n_ranks, delay, base
        self.scene = scene
        self.w = world.World(scene.n, host_timeslice=rank, device=device)
        self.w.upload(scene)
        self.w.set_coherency(start_coherency(rank, n_ranks, delay, base))
        self.tl_in = world.Timeline(delay + 2, scene.n, device=device) if rank > 0 else None
        self.tl_out = world.Timeline(2, scene.n, device=device) if rank < n_ranks - 1 else None
        self.tick_no = 0
        self.replayed = 0
        self.forks_promoted = 0
        self.resolution_needed = False

    def close(self):
        for o in (self.tl_in, self.tl_out, self.w):
            if o is not None:
                o.close()

    # -- the three networker phases and the physics step; transport happens between push and pull --
    def push(self, out_ptr, out_bytes, stream=None):
        """Phase 4: record the current state (stamped with the current coherency) and pack it for the past neighbour."""
        if self.tl_out is None:
            return 0
        c = self.w.coherency()
        self.w.record(self.tl_out, c)
        self.w.sync()
        return self.tl_out.export_slot(c, out_ptr, out_bytes, stream)

    def pull(self, in_ptr, n, stream=None):
        """Store the snapshot the future neighbour pushed this wall tick (stamped with ITS coherency)."""
        if self.tl_in is None:
            return
        self.tl_in.import_slot((self.w.coherency() + self.delay) & 0xFFFF, n, in_ptr, stream)

    def advance(self, dt=DT):
        """Phase 2 (replay what has come due), phase 5 (fork ageing), then PhysicsDynamo::run_relays."""
        s = schedule(self.rank, self.n_ranks, self.delay, self.tick_no, self.base)
        assert s["coherency"] == self.w.coherency()
        if s["replay"] is not None and self.w.replay(self.tl_in, s["replay"], abi.REPLAY_SERVER):
            self.replayed += 1
        promoted, need = self.w.fork_age()
        self.forks_promoted += promoted
        self.resolution_needed |= need
        self.w.step(dt)
        self.tick_no += 1


def run_local(slices, ticks, links, dt=DT, on_tick=None):
    """Drive several timeslices of one process in lock step (the reference runs one thread per
    timeslice; the order inside a wall tick is irrelevant because a record is consumed `delay`
    ticks after it was pushed).  links[k] connects timeslice k to k + 1."""
    for t in range(ticks):
        counts = [ts.push(links[k].ptr(), links[k].nbytes()) if k < len(slices) - 1 else 0 for k, ts in enumerate(slices)]
        for k, ts in enumerate(slices):
            if k > 0:
                ts.pull(links[k - 1].ptr(), counts[k - 1])
        if on_tick:
            on_tick(t, "pushed")
        for ts in slices:
            ts.advance(dt)
        if on_tick:
            on_tick(t, "advanced")


def run_distributed(ts, ticks, rank, n_ranks, device, dt=DT):
    """One timeslice per rank: per wall tick ONE ncclSend to the past neighbour and ONE ncclRecv from the
    future neighbour (batched so neighbouring ranks cannot deadlock), each of 100 B per entity."""
    import torch
    import torch.distributed as dist
    dev = torch.device("cuda", device)
    nbytes = max(ts.scene.n, 1) * abi.SNAPSHOT_BYTES
    out = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    inp = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    sent = 0
    for _ in range(ticks):
        n = ts.push(out.data_ptr(), nbytes)
        ops = []
        if rank < n_ranks - 1:
            ops.append(dist.P2POp(dist.isend, out, rank + 1))
        if rank > 0:
            ops.append(dist.P2POp(dist.irecv, inp, rank - 1))
        if ops:
            for r in dist.batch_isend_irecv(ops):
                r.wait()
            torch.cuda.synchronize(dev)
        if rank < n_ranks - 1:
            sent += n * abi.SNAPSHOT_BYTES
        ts.pull(inp.data_ptr(), ts.scene.n)
        ts.advance(dt)
    return sent
