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
ranks a snapshot travels as ONE buffer of 164 B per entity + the changed collider rows
(icx_timeline_slot_export / _import): ncclSend / ncclRecv over NVLink when the ranks are processes, a
device copy when several timeslices share a process.  No other data-path collective exists, and a wall
tick needs no host synchronisation: record -> export -> transport -> import -> replay are ordered by
CUDA events / stream order.
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


DELTA_ROWS = None       # changed ColliderComponents a transport buffer has room for per tick; None = every entity (the first record of a
                        # ring, and the one after a parallel cycle wrote into the timeslice, carry all of them)


class LoopbackLink:
    """Snapshot transport between two timeslices of ONE process: a device buffer both sides see."""

    def __init__(self, n_entities, device=0, rows=DELTA_ROWS):
        import torch
        self.buf = torch.empty(abi.snapshot_capacity(max(n_entities, 1), n_entities if rows is None else rows), dtype=torch.uint8, device=torch.device("cuda", device))
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

    def close(self):
        for o in (self.tl_in, self.tl_out, self.w):
            if o is not None:
                o.close()

    # -- the networker phases in the reference's order (server_network_dynamo.cpp: replay :377-421, record :642-679,
    #    fork ageing :681-705), then the physics step; the transport sits between record and the next tick's replay --
    def replay(self):
        """Phase 2: take over the record of the future neighbour that has come due."""
        s = schedule(self.rank, self.n_ranks, self.delay, self.tick_no, self.base)
        assert s["coherency"] == self.w.coherency()
        if s["replay"] is not None and self.w.replay(self.tl_in, s["replay"], abi.REPLAY_SERVER):
            self.replayed += 1

    def push(self, out_ptr, out_bytes, stream=None):
        """Phase 4: record the current (just replayed) state stamped with the current coherency and pack it for the past
        neighbour.  Asynchronous: the export waits for the record through the slot's event."""
        if self.tl_out is None:
            return 0
        c = self.w.coherency()
        self.w.record(self.tl_out, c)
        return self.tl_out.export_slot(c, out_ptr, out_bytes, stream if stream is not None else self.w.stream())

    def pull(self, in_ptr, n, nbytes, stream=None):
        """Store the snapshot the future neighbour pushed this wall tick (stamped with ITS coherency)."""
        if self.tl_in is None:
            return
        self.tl_in.import_slot((self.w.coherency() + self.delay) & 0xFFFF, n, in_ptr, nbytes, stream if stream is not None else self.w.stream())

    def step(self, dt=DT):
        """Phase 5 (fork ageing), then PhysicsDynamo::run_relays."""
        self.w.fork_age_async()
        self.w.step(dt)
        self.tick_no += 1


def run_local(slices, ticks, links, dt=DT, on_tick=None):
    """Drive several timeslices of one process in lock step (the reference runs one thread per
    timeslice; a record is consumed `delay` ticks after it was pushed).  links[k] connects timeslice k
    to k + 1.  Per wall tick and timeslice: replay, record + export, import, fork ageing, tick."""
    for t in range(ticks):
        for ts in slices:
            ts.replay()
        if on_tick:
            on_tick(t, "replayed")
        for k, ts in enumerate(slices):
            if k < len(slices) - 1:
                st = ts.w.stream()                                  # export and import on ONE stream: ordered without a host sync
                n = ts.push(links[k].ptr(), links[k].nbytes(), st)
                slices[k + 1].pull(links[k].ptr(), n, links[k].nbytes(), st)
        if on_tick:
            on_tick(t, "pushed")
        for ts in slices:
            ts.step(dt)
        if on_tick:
            on_tick(t, "advanced")


def run_distributed(ts, ticks, rank, n_ranks, device, dt=DT, rows=DELTA_ROWS):
    """One timeslice per rank: per wall tick ONE ncclSend to the past neighbour and ONE ncclRecv from the
    future neighbour (batched so neighbouring ranks cannot deadlock), each of 164 B per entity + the changed
    collider rows.  Export, send / recv and import are all enqueued on torch's current stream (NCCL orders
    itself against it), the world's stream is tied in by the slot events: no host synchronisation per tick."""
    import torch
    import torch.distributed as dist
    dev = torch.device("cuda", device)
    nbytes = abi.snapshot_capacity(max(ts.scene.n, 1), ts.scene.n if rows is None else rows)
    out = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    inp = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    st = torch.cuda.current_stream(dev).cuda_stream
    sent = 0
    for _ in range(ticks):
        ts.replay()
        n = ts.push(out.data_ptr(), nbytes, st)
        ops = []
        if rank < n_ranks - 1:
            ops.append(dist.P2POp(dist.isend, out, rank + 1))
        if rank > 0:
            ops.append(dist.P2POp(dist.irecv, inp, rank - 1))
        if ops:
            for r in dist.batch_isend_irecv(ops):
                r.wait()                                            # stream-level wait, the host does not block
        if rank < n_ranks - 1:
            sent += n * abi.SNAPSHOT_BYTES + abi.SNAPSHOT_HEADER_BYTES
        ts.pull(inp.data_ptr(), ts.scene.n, nbytes, st)
        ts.step(dt)
    torch.cuda.synchronize(dev)
    return sent
