#!/usr/bin/env python
"""bench.py -- collider-steps/s of the physics tick on the 1 M mixed-collider scene (BASELINE.json).

One step = one tick (PhysicsDynamo::run_relays: Euler + collision pass) over one scene.
  value      whole-job collider-steps/s with the scene resident in HBM (CUDA events on the world's
             stream around exactly K ticks, max over ranks)
  e2e        the same through the C ABI with HOST buffers every step: icx_bodies_upload (pinned AoS
             components in) + icx_step + icx_bodies_download (T, P out) + icx_contacts_fetch
  roofline   algorithmic bytes (308 B/collider + 8 B/pair + 48 B/contact) over the
             measured tick, against the measured HBM copy bandwidth; per-kernel shares alongside
  cpu_baseline  the reference's own procedures (oracle/_ref) ticking a bounded sample on the host

--gpus N > 1: a single scene does not shard (the reference's pair loop is sequential), so each rank
ticks its own replica of the scene (weak scaling, no data-path collective); the only NCCL traffic
is the gather of per-rank timings / stats.
--impl reference: the reference CPU code on the host cores (see cpu_reference()).
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

DT = 1.0 / 45
METRIC = "collider-steps/sec at 1M bodies"
UNIT = "collider-steps/s"
B_PER_COLLIDER, B_PER_PAIR, B_PER_CONTACT = 308, 8, 48


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def build_scene(name, n, seed_offset=0):
    from intercession_b200 import scenes
    if name == "config3":
        return scenes.config3(n, seed=20261105 + seed_offset)
    if name == "config2":
        return scenes.config2(n, seed=20261104 + seed_offset)
    raise SystemExit(f"unknown workload {name}")


class ClockSampler:
    """SM clock and throttle reasons sampled DURING the timed region (B200_PROFILING.md's clocks line) through
    NVML in-process (pynvml): spawning nvidia-smi every few ms from every rank stalls the driver and the
    measurement with it.  Falls back to one nvidia-smi query per 0.25 s if NVML cannot be loaded."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index, poll=True):
        # poll=False (ranks > 0 of a multi-GPU run): one sample on entry and one on exit instead of a polling
        # thread -- eight processes polling NVML at once serialise on the driver and slow every rank's launches
        self.index, self.stop, self.t, self.poll = index, threading.Event(), None, poll
        self.sm, self.mx, self.reasons = [], [], set()
        self.nvml = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.nvml = pynvml
        except Exception:
            self.nvml = None

    def _sample_nvml(self):
        n = self.nvml
        self.sm.append(float(n.nvmlDeviceGetClockInfo(self.h, n.NVML_CLOCK_SM)))
        self.mx.append(float(n.nvmlDeviceGetMaxClockInfo(self.h, n.NVML_CLOCK_SM)))
        get = getattr(n, "nvmlDeviceGetCurrentClocksEventReasons", None) or n.nvmlDeviceGetCurrentClocksThrottleReasons
        r = int(get(self.h))
        for name, bit in (("hw_slowdown", 0x8), ("sw_thermal_slowdown", 0x20), ("hw_thermal_slowdown", 0x40), ("sw_power_cap", 0x4)):
            if r & bit:
                self.reasons.add(name)

    def _sample_smi(self):
        out = subprocess.run(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits"],
                             capture_output=True, text=True, timeout=5).stdout
        r = [x.strip() for x in out.strip().split(",")]
        if len(r) >= 6 and r[0].replace(".", "").isdigit():
            self.sm.append(float(r[0])); self.mx.append(float(r[1]))
            for i, name in enumerate(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]):
                if r[2 + i].lower().startswith("active"):
                    self.reasons.add(name)

    def _run(self):
        while not self.stop.is_set():
            try:
                self._sample_nvml() if self.nvml else self._sample_smi()
            except Exception:
                pass
            self.stop.wait(0.02 if self.nvml else 0.25)

    def _once(self):
        try:
            self._sample_nvml() if self.nvml else self._sample_smi()
        except Exception:
            pass

    def __enter__(self):
        if self.poll:
            self.t = threading.Thread(target=self._run, daemon=True)
            self.t.start()
        else:
            self._once()
        return self

    def __exit__(self, *a):
        if self.poll:
            self.stop.set()
            self.t.join(timeout=6)
        else:
            self._once()

    def summary(self):
        return {"sm_mhz": float(np.median(self.sm)) if self.sm else None, "sm_max_mhz": max(self.mx) if self.mx else None,
                "reasons": sorted(self.reasons), "samples": len(self.sm), "source": "nvml" if self.nvml else "nvidia-smi"}


def pinned_like(arr):
    """Copy a numpy (structured) array into page-locked host memory (torch owns the allocation)."""
    import torch
    t = torch.empty(arr.nbytes, dtype=torch.uint8, pin_memory=True)
    v = t.numpy().view(arr.dtype).reshape(arr.shape)
    v[...] = arr
    return v, t


def cpu_reference(scene_name, n_sample, ticks, threads, seed_offset=0):
    """Reference CPU tick on a bounded sample: the reference's own EulerPhysicsRelay + intersect/response
    procedures (oracle/_ref, unmodified sources) in the reference's pair order behind the oracle-side
    cull -- the verbatim all-pairs loop is O(N^2) (~47 h per tick at 1 M).  The reference
    tick is single-threaded by construction; `threads` independent sample scenes run in parallel."""
    import refs
    lib = refs.ref()
    kind = "reference"
    if lib is None:
        lib, kind = refs.oracle(), "port"
    prefix = "icxref" if kind == "reference" else "icxo"
    scenes_ = [build_scene(scene_name, n_sample, seed_offset + 100 + k) for k in range(threads)]
    done = [0] * threads
    busy = [0.0] * threads          # seconds inside the reference's own code (its chrono around the relay loops) + the cull it runs
                                    # behind; the harness' per-call array copies (tests/refs.py) are not the reference's time

    def work(k):
        s = scenes_[k]
        for _ in range(ticks):
            e = refs.procedures_tick(lib, prefix, s, collide=False)
            c0 = time.perf_counter()
            cand = refs.oracle_pairs(e, 0.25)
            c1 = time.perf_counter()
            s = refs.procedures_tick(lib, prefix, e, euler=False, cand=cand, cap=4 * s.n)
            busy[k] += e["seconds"] + s["seconds"] + (c1 - c0)
            done[k] += s.n - s["n_walls"]
    t0 = time.perf_counter()
    ths = [threading.Thread(target=work, args=(k,)) for k in range(threads)]
    [t.start() for t in ths]
    [t.join() for t in ths]
    sec = time.perf_counter() - t0
    return sum(done) / max(max(busy), 1e-9), max(busy), kind


def worldline_batch_bench(n_worldlines, per, ticks, rank, world_size, local_rank, barrier, parts=0):
    """BASELINE.json configs[3]: independent world-line candidates x `per` bodies, sharded round-robin over
    the ranks (strong scaling, no data-path collective).  Per tick and rank: replay (parallel mode:
    consume) + physics tick + record + divergence scan against the shared base history; then ONE
    gather of {flag, incongruent, score} per world-line (NCCL over NVLink when world_size > 1).
    A rank's share runs as `parts` worlds ticked concurrently by as many host threads on their own streams, each with 1 / parts
    of the device for its persistent resolve kernels (icx_device_share_set): a small share is latency-bound on its dependency
    rounds, and independent candidates need not wait for each other's.  parts = 0: one world per 2 048 world-lines or less, at most 4,
    and no more than the host has spare cores for (measured on one B200: 512 world-lines 0.151 s as one world, 0.140 s as four;
    with 8 ranks on 16 cores the four threads per rank cost what they gain)."""
    import threading

    import torch
    import torch.distributed as dist
    from intercession_b200 import abi, scenes, shard, world as icxworld
    mine = shard.worldlines_of_rank(n_worldlines, rank, world_size)
    if parts <= 0:                                # (a host thread per world spins in its stream sync: no more worlds than spare cores)
        parts = max(1, min(4, 2048 // max(len(mine), 1), (os.cpu_count() or 1) // (2 * world_size)))
    parts = max(1, min(parts, len(mine)))
    full_local = scenes.worldline_batch(len(mine), per, seed=20261106, perturb=False)
    for k, wl in enumerate(mine):                 # candidate w perturbs body (w mod per), seed + 1 + w  (same as scenes.worldline_batch)
        r = np.random.default_rng(20261106 + 1 + int(wl))
        full_local.P["velocity"][k * per + (int(wl) % per)] += r.uniform(-2, 2, 3).astype(np.float32)
    base = full_local.base
    hist_depth = ticks + 2
    keys = ("ids", "T", "P", "hasP", "C", "tstate", "tcoh")
    bounds = [len(mine) * q // parts for q in range(parts + 1)]          # contiguous runs of this rank's candidates
    part = []
    for q in range(parts):
        lo, hi = bounds[q] * per, bounds[q + 1] * per
        sc = scenes.Scene({k: (v[lo:hi].copy() if k in keys else v) for k, v in full_local.items()})
        wb = icxworld.World(base.n, host_timeslice=abi.NULL_TIMESLICE, device=local_rank)
        tl_base = icxworld.Timeline(hist_depth, per, device=local_rank)
        wb.upload(base)
        for t in range(ticks):                    # the base history every candidate is compared with (replicated, read-only; one copy per
            wb.step(DT)                           # part because the scan's accumulators live in the ring object)
            wb.record(tl_base, t + 1, keep_link=True)
        wb.close()
        wc = icxworld.World(sc.n, worldline_stride=per, host_timeslice=abi.NULL_TIMESLICE, device=local_rank)
        wc.share_device(parts)
        tl_new = icxworld.Timeline(2, sc.n, device=local_rank)
        wc.upload(sc)
        part.append({"wc": wc, "tl_new": tl_new, "tl_base": tl_base, "n_wl": bounds[q + 1] - bounds[q], "dev_ms": 0.0, "scan_ms": []})
    dev = torch.device("cuda", local_rank) if world_size > 1 else None
    z = np.zeros(len(mine))
    shard.gather_results(z, z, z, n_worldlines, rank, world_size, device=dev)      # warm the communicator up (untimed)

    def run_part(pt):
        wc, tl_new, tl_base = pt["wc"], pt["tl_new"], pt["tl_base"]
        tl_base.profile(True)
        for t in range(ticks):
            wc.replay(tl_new, t, mode=abi.REPLAY_PARALLEL)
            wc.step(DT)
            pt["dev_ms"] += wc.last_step_ms()
            wc.record(tl_new, t + 1, at_breakpoint=False, keep_link=True)
            wc.divergence_scan_enqueue(tl_base, t + 1, pt["n_wl"])     # verdicts stay on the device: no host round trip per tick
            if t % 16 == 15 or t == ticks - 1:
                pt["scan_ms"].append(tl_base.kernel_ms(2))
        pt["out"] = tl_base.scan_fetch(pt["n_wl"])                       # the last tick's counts / scores + "ever flagged"

    barrier()
    t0 = time.perf_counter()
    if parts == 1:
        run_part(part[0])
    else:
        ths = [threading.Thread(target=run_part, args=(pt,)) for pt in part]
        [t.start() for t in ths]
        [t.join() for t in ths]
    any_flag = np.concatenate([pt["out"][1] for pt in part]); cnt = np.concatenate([pt["out"][2] for pt in part]); score = np.concatenate([pt["out"][3] for pt in part])
    gflags, gcnt, gscore = shard.gather_results(any_flag, cnt, score, n_worldlines, rank, world_size, device=dev)
    barrier()
    sec = time.perf_counter() - t0
    st = part[0]["wc"].stats()
    n_dirty, n_contacts = sum(pt["wc"].stats().n_dirty for pt in part), sum(pt["wc"].stats().n_contacts for pt in part)
    dev_ms = max(pt["dev_ms"] for pt in part)
    scan_ms = [m for pt in part for m in pt["scan_ms"]]
    scan_bodies = part[0]["n_wl"] * per
    for pt in part:
        pt["wc"].close(); pt["tl_new"].close(); pt["tl_base"].close()
    sec_t = torch.tensor([sec], dtype=torch.float64, device="cuda")
    if world_size > 1:
        dist.all_reduce(sec_t, op=dist.ReduceOp.MAX)
    sec = float(sec_t.item())
    bodies = n_worldlines * per
    return {"workload": f"config4: {n_worldlines} world-line candidates x {per} bodies, {ticks} ticks of replay + tick + record + divergence scan, "
                        f"sharded round-robin over {world_size} GPU(s), base history replicated",
            "scaling": "strong", "n_gpus": world_size, "ticks": ticks, "seconds": round(sec, 4),
            "worlds_per_gpu": parts, "worldlines_per_world": part[0]["n_wl"],
            "worldline_ticks_per_s": n_worldlines * ticks / sec, "collider_steps_per_s": bodies * ticks / sec,
            "physics_tick_ms_rank0": round(dev_ms / ticks, 4), "dirty_bodies_rank0": n_dirty, "contacts_rank0": n_contacts,
            "divergent_worldlines": int(gflags.sum()), "most_congruent": shard.most_congruent(gflags, gscore),
            "gather": {"collective": "all_gather_into_tensor (NCCL)" if world_size > 1 else "none (1 rank)",
                       "bytes_per_rank": 24 * ((n_worldlines + world_size - 1) // world_size)},
            "algorithmic_bytes_per_body_tick": {"record": 296, "scan": 26},
            # k_divergence_scan over one world's candidates: 16 B of live position per body from HBM (the tick and the record
            # ran in between, > L2); the 16 KB base record is shared by every world-line and stays in L2
            "scan_kernel_rank0": {"us": round(float(np.mean(scan_ms)) * 1e3, 2), "bodies": int(scan_bodies),
                                  "alg_GBps": round(16.0 * scan_bodies / (float(np.mean(scan_ms)) * 1e-3) / 1e9, 1),
                                  "timing": "CUDA events per launch (icx_timeline_kernel_ms), mean over the ticks"}}


def timeline_kernels_bench(w, n, peak, device, reps=24, depth=6):
    """The world-line kernels on their own over the resident 1 M-entity world: record (k_record), replay (k_replay),
    divergence scan (k_divergence_scan).  `us` = mean kernel duration from CUDA events on the world's stream
    (icx_timeline_kernel_ms) over `reps` launches cycling through `depth` ring slots (depth x 100 B x n > L2);
    `call_us` = host clock per ABI call over the same batch (record / replay are asynchronous launches; the scan returns its
    reduction to the host, so its call includes one read-back and stream sync).
    Algorithmic bytes per body: record / replay 148 B read + 148 B written (the whole TransformComponent +
    PhysicsComponent; the kernels move 164 B each way: vec3 fields padded to float4, plus the id), scan 26 B."""
    from intercession_b200 import abi, world as icxworld
    tl = icxworld.Timeline(depth, n, device=device)
    tl.profile(True)
    c0 = 1000

    def run(fn, which, reps):
        fn(0); w.sync()
        kern = []
        t0 = time.perf_counter()
        for k in range(reps):
            fn(k)
        w.sync()
        call = (time.perf_counter() - t0) / reps
        for k in range(reps):                      # events hold the LAST launch only: time each launch on its own
            fn(k)
            kern.append(tl.kernel_ms(which))
        return float(np.mean(kern)) * 1e-3, call

    def entry(sec, call, nbytes, extra=None):
        e = {"us": round(sec * 1e6, 2), "call_us": round(call * 1e6, 2), "alg_bytes": nbytes,
             "alg_GBps": round(nbytes / sec / 1e9, 1), "frac": round(nbytes / sec / 1e9 / peak, 4)}
        if extra:
            e["note"] = extra
        return e

    out = {}
    sec, call = run(lambda k: w.record(tl, (c0 + k % depth) & 0xFFFF, keep_link=True), 0, reps)
    out["record"] = entry(sec, call, 296 * n)
    rec_call = call

    def pair(k):                                   # a replay consumes its record (pop): re-record before each one
        c = (c0 + k % depth) & 0xFFFF
        w.record(tl, c, keep_link=True)
        assert w.replay(tl, c, abi.REPLAY_SERVER)
    sec, call = run(pair, 1, reps)
    out["replay"] = entry(sec, max(call - rec_call, 1e-9), 296 * n, "call_us = record + replay pair minus the record call")
    for k in range(depth):
        w.record(tl, (c0 + k) & 0xFFFF, keep_link=True)
    sec, call = run(lambda k: w.divergence_scan(tl, (c0 + k % depth) & 0xFFFF, 1), 2, reps)
    out["divergence_scan"] = entry(sec, call, 26 * n, "call_us includes the result read-back and stream sync")
    out["entities"] = n
    out["timing"] = f"CUDA events per launch, mean of {reps}; {depth} ring slots of {164 * n / 1e6:.0f} MB (larger than L2)"
    tl.close()
    return out


def projectile_behaviors_bench(w, n, peak, reps=24):
    """ProjectileBehaviors::on_fixed_update on the device for one projectile per entity of the resident
    world: mean kernel duration from CUDA events on the world's stream (icx_projectiles_kernel_ms).  Algorithmic bytes per
    projectile: 40 B component row read + 8 B lifetime written + 4 B entity index + 16 B + 16 B scale/mass word read and
    written = 84 B.  88 MB per launch at 1 M: below L2 capacity, so the rows are written
    flushed out of L2 by a 256 MB read pass between launches."""
    import torch
    from intercession_b200 import abi
    rng = np.random.default_rng(5)
    p = np.zeros(n, abi.projectile_dt)
    p["currLifetime"] = rng.uniform(0, 200, n); p["maxLifetime"] = 300.0; p["hz"] = rng.uniform(0.5, 4.0, n)
    p["scaleMin"] = rng.uniform(0.3, 0.5, n); p["scaleMax"] = p["scaleMin"] + 0.1
    w.projectiles_upload(np.arange(n, dtype=np.uint32), p)
    flush = torch.zeros(256 << 18, dtype=torch.float32, device="cuda")
    us = []
    for k in range(reps + 3):
        flush.sum(); torch.cuda.synchronize()              # a READ pass: the L2 is left full of clean lines (a fill would leave dirty ones to drain)
        w.projectiles_fixed_update(1.0 / 60.0)
        if k >= 3:
            us.append(w.projectiles_kernel_ms() * 1e3)
    sec = float(np.mean(us)) * 1e-6
    nbytes = 84 * n
    w.projectiles_upload(np.zeros(0, np.uint32), np.zeros(0, abi.projectile_dt))
    return {"kernel": "k_projectile_update", "projectiles": n, "us": round(sec * 1e6, 2), "alg_bytes": nbytes,
            "alg_GBps": round(nbytes / sec / 1e9, 1), "frac": round(nbytes / sec / 1e9 / peak, 4),
            "timing": f"CUDA events per launch, mean of {reps}, L2 flushed (256 MB read pass) before each launch"}


def mesh_pass_bench(w, s, peak, ticks=6):
    """The static triangle-mesh pass (extension) on the headline scene: a torus of 18 432 triangles in the middle
    of the arena and a ground sheet of 131 072 under it; mean device time of k_mesh_pass over a few ticks."""
    from intercession_b200 import scenes
    side = float(s["side"]) if "side" in s else 200.0
    tv, tt = scenes.torus_mesh(centre=(0.0, 0.0, 0.0), major=0.25 * side, minor=0.05 * side, n_major=192, n_minor=48)
    gv, gt = scenes.height_mesh(side, 256, y=-0.45 * side, bump=0.2, seed=3)
    verts = np.concatenate([tv, gv]); tris = np.concatenate([tt, gt + len(tv)]).astype(np.uint32)
    w.upload_mesh(verts, tris)
    us, contacts = [], 0
    for _ in range(ticks):
        w.step(DT)
        us.append(w.mesh_kernel_us())
        contacts = len(w.mesh_contacts())
    w.remove_mesh()
    sec = float(np.mean(us[1:])) * 1e-6
    nbytes = s.n * (16 + 16 + 16 + 16 + 4 + 2 * 16)     # what the selection kernel streams per entity: prm1, prm0, origin, origin after integration, first_packet, packet bounds
    return {"kernel": "k_mesh_select + k_mesh_pass", "entities": s.n, "triangles": int(len(tris)), "us": round(sec * 1e6, 2), "mesh_contacts_last_tick": contacts,
            "alg_bytes": int(nbytes), "alg_GBps": round(nbytes / sec / 1e9, 1), "frac": round(nbytes / sec / 1e9 / peak, 4),
            "bound": "latency of the ordered walk over the selected entities (k_mesh_select streams alg_bytes; k_mesh_pass runs one warp per selected entity)",
            "timing": "CUDA events around the two launches (icx_mesh_kernel_us), mean of %d ticks" % (ticks - 1),
            "parity": "self-oracle (no mesh collider in the reference): tests/test_gpu_mesh.py, tests/test_mesh_oracle.py"}


def timeslice_pipeline_bench(bodies, ticks, delay, rank, world_size, local_rank, barrier):
    """BASELINE.json configs[4] (reduced: no time travellers, no parallel cosmos): one timeslice per rank, each
    `delay` ticks behind its future neighbour; per wall tick and link ONE snapshot of 164 B per entity travels
    future -> past (ncclSend / ncclRecv over NVLink between ranks; a device copy when world_size == 1, where two
    timeslices share the GPU so that the path is exercised).  Per tick and timeslice, in the reference's order: replay of the
    record that came due, record + export, import, fork ageing, physics tick -- ordered by CUDA events, no host sync per tick."""
    import torch
    import torch.distributed as dist
    from intercession_b200 import abi, pipeline, scenes
    n_ts = world_size if world_size > 1 else 2
    base = scenes.arena(bodies, seed=20261107, kinds=("sphere", "box"), temporal=True)
    mine = [rank] if world_size > 1 else list(range(n_ts))
    slices = []
    for k in mine:
        sc = base.copy()
        sc["ids"] = pipeline.past_ids(base.ids, k)
        slices.append(pipeline.Timeslice(sc, k, n_ts, delay, device=local_rank))
    warm = delay + 2
    if world_size > 1:
        pipeline.run_distributed(slices[0], warm, rank, n_ts, local_rank)
    else:
        links = [pipeline.LoopbackLink(base.n, local_rank) for _ in range(n_ts - 1)]
        pipeline.run_local(slices, warm, links)
    barrier()
    t0 = time.perf_counter()
    if world_size > 1:
        sent = pipeline.run_distributed(slices[0], ticks, rank, n_ts, local_rank)
    else:
        pipeline.run_local(slices, ticks, links)
        sent = (n_ts - 1) * ticks * (base.n * abi.SNAPSHOT_BYTES + abi.SNAPSHOT_HEADER_BYTES)
    barrier()
    sec = time.perf_counter() - t0
    sec_t = torch.tensor([sec], dtype=torch.float64, device="cuda")
    sent_t = torch.tensor([float(sent)], dtype=torch.float64, device="cuda")
    rep_t = torch.tensor([float(min(ts.replayed for ts in slices if ts.rank > 0) if any(ts.rank > 0 for ts in slices) else 1e9)],
                         dtype=torch.float64, device="cuda")
    if world_size > 1:
        dist.all_reduce(sec_t, op=dist.ReduceOp.MAX)
        dist.all_reduce(sent_t, op=dist.ReduceOp.SUM)
        dist.all_reduce(rep_t, op=dist.ReduceOp.MIN)
    sec = float(sec_t.item())

    # ---- one parallel-cosmos cycle: load_and_link from a past timeslice, re-simulate `delay` ticks up to its
    # future neighbour's coherency, extract at chain-link - 1.  Rank 0 only, parallel cosmos on the same GPU (the peer-GPU
    # copies are covered by tests/test_gpu_clone.py; the idle ranks' GPUs sit in the NCCL barrier meanwhile). ----
    cycle = None
    if rank == 0:
        from intercession_b200 import world as icxworld
        dst = slices[0]
        if world_size == 1:
            src_w, own_src = slices[1].w, False
        else:
            sc = base.copy(); sc["ids"] = pipeline.past_ids(base.ids, 1)
            src_w, own_src = icxworld.World(sc.n, host_timeslice=1, device=local_rank), True
            src_w.upload(sc); src_w.set_coherency(dst.w.coherency())
        par = icxworld.World(base.n, host_timeslice=abi.NULL_TIMESLICE, device=local_rank)
        best = None
        for rep in range(3):
            torch.cuda.synchronize()
            c0 = time.perf_counter()
            n_cl = src_w.clone_into(par); par.sync()
            c1 = time.perf_counter()
            par.step(pipeline.DT, delay); par.sync()
            c2 = time.perf_counter()
            n_ex, n_miss = par.extract_into(dst.w); dst.w.sync()
            c3 = time.perf_counter()
            cur = (c3 - c0, c1 - c0, c2 - c1, c3 - c2)
            best = cur if best is None or cur[0] < best[0] else best
        cycle = {"workload": f"load_and_link of {n_cl} entities -> {delay} re-simulated ticks -> extract at chain-link - 1 ({n_ex} written, {n_miss} missing); best of 3",
                 "cycle_ms": round(best[0] * 1e3, 3), "clone_ms": round(best[1] * 1e3, 3), "run_ms": round(best[2] * 1e3, 3),
                 "extract_ms": round(best[3] * 1e3, 3), "state_bytes_each_way": int(n_cl) * 316}
        par.close()
        if own_src:
            src_w.close()
    barrier()
    for ts in slices:
        ts.close()
    return {"workload": f"config5 (reduced): {n_ts} timeslices x {bodies} temporal bodies + 6 walls, delay {delay} ticks, {ticks} wall ticks; "
                        f"no time travellers; one parallel-cosmos cycle timed separately",
            "parallel_cycle": cycle,
            "n_timeslices": n_ts, "timeslices_per_gpu": len(mine), "ticks": ticks, "seconds": round(sec, 4),
            "wall_ticks_per_s": ticks / sec, "collider_steps_per_s": n_ts * base.n * ticks / sec,
            "snapshot_bytes_per_link_tick": base.n * abi.SNAPSHOT_BYTES, "snapshot_bytes_moved": int(sent_t.item()),
            "transport": "ncclSend/ncclRecv (batch_isend_irecv)" if world_size > 1 else "device copy (one process, one GPU)",
            "records_replayed_min": int(rep_t.item())}


def time_travel_bench(bodies, n_ts, n_travellers, delay, wall_ticks, rank, world_size, local_rank, barrier):
    """BASELINE.json configs[4] end to end (intercession_b200/timetravel.py): `n_ts` timeslices of `bodies` temporal bodies,
    `n_travellers` time travellers arriving at scripted ticks one timeslice into their past (JUMP_REQUEST{timesliceDelta=+1}),
    interceptions -> fork ageing -> parallel_notify_divergence -> load_and_link -> re-simulation with record / replay at the
    breakpoint -> extract, down the whole chain.  ONE host process (rank 0) drives every timeslice: timeslice k on GPU
    k mod N, one host thread per timeslice inside a phase, the parallel cosmos on the GPU of the timeslice it shadows.  The
    other ranks wait on the rendezvous store (host side: a pending NCCL kernel on their GPUs would keep the cooperative
    launches of rank 0 from ever becoming resident there)."""
    import datetime

    import numpy as np
    import torch
    import torch.distributed as dist
    from intercession_b200 import abi, pipeline, scenes, timetravel
    barrier()
    store = dist.distributed_c10d._get_default_store() if world_size > 1 else None
    if rank != 0:
        store.wait(["icx_time_travel_done"], datetime.timedelta(seconds=900))
        return None
    out = None
    try:
        n_dev = max(1, min(world_size, torch.cuda.device_count()))
        devices = [k % n_dev for k in range(n_ts)]
        base = scenes.arena(bodies, seed=20261107, kinds=("sphere", "box"), temporal=True)
        chain = []
        for k in range(n_ts):
            sc = base.copy()
            sc["ids"] = pipeline.past_ids(base.ids, k)
            chain.append(sc)
        kinds = base.C["colliders"]["colliderType"][:, 0]
        boxes = np.nonzero((kinds == abi.COLLIDER_BOX) & ((base.ids & 15) != 15))[0]
        twins = boxes[np.linspace(0, len(boxes) - 1, n_travellers).astype(np.int64)]
        arrivals = []
        for i, twin in enumerate(twins):                            # traveller i: a copy of `twin`, 0.3 units inside it, chain-link 0
            k = 1 + i % (n_ts - 1)
            row = timetravel.traveller_row(chain[k], int(twin), genesis=2000000 + i, timeslice=k)
            row.T["origin"][0, 0] += np.float32(0.3)
            arrivals.append((3 + (i // (n_ts - 1)) % 8, k, row))
        be = timetravel.DeviceBackend(chain, delay, devices=devices, capacity_slack=n_travellers + 8, host_rows=False)
        seen = {"ints": 0, "promoted": 0, "cycle_ints": 0, "extracted": 0, "missing": 0}

        stamps, cycle_ticks = [], []

        def observe(t, phase, d):
            if phase == "cycle":
                cycle_ticks.append(t)
            if phase == "tick":
                stamps.append(time.perf_counter())
                seen["ints"] += sum(len(i) for i in d["interceptions"])
                seen["promoted"] += sum(p for p, _ in d["ages"])
            else:
                for r in d:
                    seen["cycle_ints"] += sum(len(i) for i in r["interceptions"])
                    seen["extracted"] += r["extracted"]; seen["missing"] += r["missing"]
        for dv in set(devices):
            torch.cuda.synchronize(dv)
        t0 = time.perf_counter()
        cycles = timetravel.run(be, wall_ticks, arrivals, observe, with_states=False)
        for k in range(n_ts):
            be.w[k].sync()
        sec = time.perf_counter() - t0
        cyc = getattr(be, "cycle_seconds", [])
        mism = [int(t.id_mismatches()) for t in be.tl_in if t is not None]
        be.close()
        tick_sec = sec - sum(cyc)
        # wall time of the ticks with neither an arrival nor a cycle in them, before and after the first cycle (after it the
        # future-most timeslices hold the re-simulated history: delay x (n_ts - 1) more ticks of the drop, bodies piled on the floor)
        per_tick = np.diff(np.array([t0] + stamps))
        arrival_ticks = {ta for ta, _, _ in arrivals}
        first_cycle = cycle_ticks[0] if cycle_ticks else wall_ticks
        busy = set(be.lifecycle_ticks)                          # arrivals and the ENTITY_CREATED records that come due further down the chain
        quiet_before = [per_tick[t] for t in range(wall_ticks) if t not in busy and t < first_cycle and t >= 2]
        quiet_after = [per_tick[t] for t in range(wall_ticks) if t not in cycle_ticks and t not in busy and t > first_cycle]
        out = {"workload": f"config5: {n_ts} timeslices x {bodies} temporal bodies + 6 walls, {n_travellers} time travellers arriving one timeslice "
                           f"into their past at wall ticks 3-10, timeslice delay {delay} ticks"
                           + ("" if delay >= 225 else " (config 5 says 5 s = 225 ticks: reduced so that the default run stays within minutes; --tt-delay 225 runs it as written)")
                           + f", {wall_ticks} wall ticks; interceptions, fork ageing and the parallel cycle down the whole chain inside the run",
               "n_timeslices": n_ts, "gpus_used": n_dev, "devices": devices, "host": "one process, one thread per timeslice and phase",
               "wall_ticks": wall_ticks, "seconds": round(sec, 4),
               "wall_ticks_per_s_outside_cycles": round(wall_ticks / tick_sec, 2),
               "ticks_per_s_per_timeslice_before_first_cycle": round(len(quiet_before) / max(sum(quiet_before), 1e-9), 1) if quiet_before else None,
               "ticks_per_s_per_timeslice_after_first_cycle": round(len(quiet_after) / max(sum(quiet_after), 1e-9), 1) if quiet_after else None,
               "ms_per_tick_with_entity_lifecycle": round(1e3 * float(np.mean([per_tick[t] for t in sorted(busy) if 0 <= t < wall_ticks and t not in cycle_ticks])), 2) if busy else None,
               "ticks_with_entity_lifecycle": len(busy),
               "collider_steps_per_s_outside_cycles": round(n_ts * base.n * wall_ticks / tick_sec, 1),
               "parallel_cycles": cycles, "cycle_latency_ms": [round(c * 1e3, 2) for c in cyc],
               "cycle_work": f"{n_ts - 1} x (load_and_link + {delay} re-simulated ticks with replay / record at the breakpoint + extract)",
               "interceptions": seen["ints"], "forks_promoted": seen["promoted"], "interceptions_in_cycles": seen["cycle_ints"],
               "entities_extracted": seen["extracted"], "entities_missing_on_extract": seen["missing"],
               "replays_that_left_an_entity_alone": mism,      # per ring, summed over the ticks: chain-link-0 travellers have no future stream (icx_timeline_id_mismatches)
               "parity": "tests/test_gpu_timetravel.py: the same schedule at 150 / 500 bodies, every tick bit-exact against the CPU restatement"}
    finally:
        if store is not None:
            store.set("icx_time_travel_done", "1")
    return out


def long_run_bench(a, device):
    """SURVEY 8(d) protocol (100 warm-up + 1000 timed ticks) as far as a wall-clock budget allows: the scene of the headline
    (config 3, full size) and the 64 K scene of config 2, ticked from the drop; per window of 25 ticks the mean tick time, the
    phase split and what drives the ordered resolve (dirty bodies, contacts, dependency rounds, replay attempts, movers).
    Bodies fall up to 200 units: after ~tick 60 they reach the floor at up to 60 units/s and the pile is violent until it
    settles -- the regime in which the reference's sequential single-pass resolution is hardest to reproduce in parallel."""
    from intercession_b200 import world as icxworld
    out = {}
    for name, scene_name, bodies, share in (("headline_scene", a.workload, a.bodies, 0.6), ("config2_64k", "config2", 65536, 0.4)):
        sc = build_scene(scene_name, bodies)
        w = icxworld.World(sc.n, device=device)
        w.upload(sc)
        w.profile(True)
        rows, done, t_start, window = [], 0, time.perf_counter(), 25
        prev = {k: 0.0 for k in ("integrate_prep", "grid_sort", "pairs", "narrow", "resolve")}
        prev_tot = w.totals()
        status = "complete (100 + 1000 ticks)"
        while done < 1100:
            w.step(DT, window)
            ms = w.last_step_ms()
            done += window
            st = w.stats()
            ph, _, _ = w.phases()
            tot = w.totals()
            dt_ = {k: tot[k] - prev_tot[k] for k in tot}
            if done in (25, 50, 100, 150, 200, 300, 400, 500, 700, 1000, 1100) or time.perf_counter() - t_start > share * a.long_seconds:
                rows.append({"tick": done, "ms_per_tick": round(ms / window, 4),
                             "phases_ms": {p_: round((ph[p_] - prev[p_]) / window, 4) for p_ in ph},
                             "candidate_pairs": int(dt_["candidates"] / window), "contacts": int(dt_["contacts"] / window),
                             "dirty_bodies": int(dt_["dirty"] / window), "rounds": round(dt_["rounds"] / window, 1),
                             "attempts": round(dt_["attempts"] / window, 2), "movers": st.n_movers, "retries": st.n_retries})
            prev, prev_tot = dict(ph), tot
            if time.perf_counter() - t_start > share * a.long_seconds:
                status = f"stopped by the wall-clock budget ({share * a.long_seconds:.0f} s) at tick {done}"
                break
        timed = [r for r in rows if r["tick"] > 100]
        out[name] = {"workload": f"{scene_name}: {bodies} bodies, ticked from the drop", "status": status, "ticks": done,
                     "colliders": int(sc.C["colliders"]["isActive"].sum()), "windows": rows,
                     "slowest_window_ms_per_tick": max(r["ms_per_tick"] for r in rows)}
        w.close()
    return out


def run_reference_arm(a, rank, world_size):
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    threads = max(1, min(cores, 32))
    n_sample = min(a.bodies, 32768)
    # warm-up
    for _ in range(min(a.warmup, 1)):
        cpu_reference(a.workload, n_sample, 1, threads)
    t0 = time.perf_counter()
    v, sec, kind = cpu_reference(a.workload, n_sample, a.steps, threads)
    sample = (f"{threads} independent {n_sample}-body samples of {a.workload} (same recipe), {a.steps} ticks each: reference "
              f"EulerPhysicsRelay + intersect/response procedures in reference pair order behind the oracle-side cull; "
              f"the verbatim all-pairs tick is O(N^2)")
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps,
            "warmup": a.warmup, "ms_per_step": 1e3 * sec / max(a.steps, 1), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"{a.workload}: {a.bodies} mixed sphere/AABB/OBB colliders in a closed arena (CPU sample {n_sample} x {threads})"},
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": threads, "kind": kind, "sample": sample},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="config3")
    ap.add_argument("--bodies", type=int, default=1048576)
    ap.add_argument("--e2e-steps", type=int, default=5)
    ap.add_argument("--cpu-seconds", type=float, default=15.0)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--long-seconds", type=float, default=75.0, help="wall-clock budget of the long-run section (0 = skip it)")
    ap.add_argument("--no-timeline", action="store_true", help="skip the record / replay / scan kernel timings")
    ap.add_argument("--worldlines", type=int, default=4096, help="config-4 world-line candidates (0 = skip that section)")
    ap.add_argument("--worldline-bodies", type=int, default=1024)
    ap.add_argument("--worldline-parts", type=int, default=0, help="worlds per GPU the rank's candidates are ticked as, concurrently (0 = one per 2 048 world-lines or less, at most 4)")
    ap.add_argument("--worldline-ticks", type=int, default=225, help="config 4 as written: 225 ticks against a 225-deep base history")
    ap.add_argument("--pipeline-bodies", type=int, default=100000, help="config-5 bodies per timeslice (0 = skip that section)")
    ap.add_argument("--pipeline-ticks", type=int, default=12)
    ap.add_argument("--pipeline-delay", type=int, default=5)
    ap.add_argument("--tt-timeslices", type=int, default=7, help="config 5 with time travellers: timeslices (0 = skip that section)")
    ap.add_argument("--tt-travellers", type=int, default=64)
    ap.add_argument("--tt-delay", type=int, default=20, help="timeslice delay in ticks (config 5 as written: 225)")
    ap.add_argument("--tt-ticks", type=int, default=50)
    a = ap.parse_args()
    a.warmup = max(a.warmup, 3) if a.impl == "ours" else a.warmup

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world_size = int(os.environ.get("WORLD_SIZE", "1"))

    if a.impl == "reference":
        run_reference_arm(a, rank, world_size)
        return

    import torch
    import torch.distributed as dist
    from intercession_b200 import world as icxworld
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- there is no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    if world_size > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        if world_size > 1:
            dist.barrier()
        torch.cuda.synchronize()

    s = build_scene(a.workload, a.bodies)             # every rank runs the SAME scene (BASELINE config 3 fixes the seed): true replicas
    n_colliders = int(s.C["colliders"]["isActive"].sum())
    w = icxworld.World(s.n, device=local_rank)
    w.upload(s)
    w.step(DT, a.warmup)
    w.profile(True)
    launches0 = w.phases()[2]
    barrier()
    with ClockSampler(local_rank, poll=(rank == 0)) as clk:
        w.step(DT, a.steps)
        ms = w.last_step_ms()
        barrier()
    phases, pticks, launches1 = w.phases()
    st = w.stats()
    tot = w.totals()
    w.profile(False)
    ms_t = torch.tensor([ms], dtype=torch.float64, device="cuda")
    if world_size > 1:
        dist.all_reduce(ms_t, op=dist.ReduceOp.MAX)
    ms_max = float(ms_t.item())
    # per-rank view of the replicas
    mine = torch.tensor([ms / a.steps, float(st.n_retries), float(st.n_attempts), float(min(clk.sm) if clk.sm else 0.0),
                         float(len(clk.reasons - {"sw_power_cap"}))], dtype=torch.float64, device="cuda")
    per_rank = [mine]
    if world_size > 1:
        per_rank = [torch.zeros_like(mine) for _ in range(world_size)]
        dist.all_gather(per_rank, mine)
    per_rank = [[round(float(x), 4) for x in t.tolist()] for t in per_rank]
    value = world_size * n_colliders * a.steps / (ms_max * 1e-3)

    # ---- roofline: algorithmic bytes of the tick over its measured duration; pairs and contacts are the MEAN over the timed
    # ticks (summed on the device side by the library), not the last tick's ----
    peak, peak_src = measured_peak()
    mean_pairs, mean_contacts = tot["candidates"] / max(tot["ticks"], 1), tot["contacts"] / max(tot["ticks"], 1)
    alg_bytes = B_PER_COLLIDER * n_colliders + B_PER_PAIR * mean_pairs + B_PER_CONTACT * mean_contacts
    tick_ms = ms / a.steps
    achieved = alg_bytes / (tick_ms * 1e-3) / 1e9
    per_kernel = {}
    kernel_bytes = {"integrate_prep": B_PER_COLLIDER * n_colliders, "grid_sort": 0, "pairs": B_PER_PAIR * mean_pairs,
                    "narrow": 0, "resolve": B_PER_CONTACT * mean_contacts}
    for k, v in phases.items():
        kms = v / max(pticks, 1)
        per_kernel[k] = {"ms": round(kms, 4), "share": round(v / max(sum(phases.values()), 1e-9), 3),
                         "alg_GBps": round(kernel_bytes[k] / max(kms * 1e-3, 1e-12) / 1e9, 1)}
    dom = max(phases, key=phases.get)
    traffic, traffic_src = None, None
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tp) and a.workload == "config3" and a.bodies == 1048576:
        tj = json.load(open(tp))
        traffic, traffic_src = tj["tick_dram_bytes"], tj["source"]
    roofline = {"bound": "hbm", "kernel": "tick (all kernels; dominant phase: %s)" % dom, "achieved": round(achieved, 2),
                "peak": peak, "peak_source": peak_src, "unit": "GB/s", "frac": round(achieved / peak, 5), "traffic": traffic,
                "traffic_source": traffic_src,
                "algorithmic_bytes_per_tick": alg_bytes, "phases": per_kernel}

    # ---- e2e: host buffers in / out through the C ABI every step.  The colliders are resident after the first upload (the
    # reference's ColliderSynchro resubmits references, not copies: collider_synchro.cpp:28-42); every step the host hands in
    # the whole TransformComponent + PhysicsComponent arrays (what its behaviours may have edited: 148 B per entity, pinned) and
    # takes them back in place plus the ordered contact records.  (icx_motion_download returns only the 76 B per entity a tick
    # writes; scattering those rows into AoS components on the host costs more than the 72 B per entity it saves on PCIe.) ----
    T, _k1 = pinned_like(s.T)
    P, _k2 = pinned_like(s.P)
    w.upload(s)                                        # untimed: scene load
    n_contacts = 0
    for it in range(2 + a.e2e_steps):
        if it == 2:
            barrier()
            t0 = time.perf_counter()
        w.update_all(T=T, P=P)
        w.step(DT)
        w.download_into(T=T, P=P)                      # straight into the caller's component arrays (the device packs the AoS rows)
        n_contacts = len(w.contacts())
    barrier()
    e2e_sec = time.perf_counter() - t0
    e2e_t = torch.tensor([e2e_sec], dtype=torch.float64, device="cuda")
    if world_size > 1:
        dist.all_reduce(e2e_t, op=dist.ReduceOp.MAX)
    e2e_val = world_size * n_colliders * a.e2e_steps / float(e2e_t.item())
    h2d = T.nbytes + P.nbytes
    d2h = T.nbytes + P.nbytes + n_contacts * 48
    # ---- the same loop with device-shared component pools (host/component_pools.hpp): the device copy is
    # authoritative between ticks, the host sends only the rows it edited (here 1 % of the entities per step, a behaviours-sized
    # share) and takes Transform + Physics back in place.  Reported beside e2e, not instead of it. ----
    rng = np.random.default_rng(5)
    k_edit = max(1, s.n // 100)
    Te, _k3 = pinned_like(s.T[:k_edit])
    Pe, _k4 = pinned_like(s.P[:k_edit])
    Tb, Pb = T.view(np.uint8).reshape(s.n, -1), P.view(np.uint8).reshape(s.n, -1)
    Teb, Peb = Te.view(np.uint8).reshape(k_edit, -1), Pe.view(np.uint8).reshape(k_edit, -1)
    edit_rows = [np.sort(rng.choice(s.n, size=k_edit, replace=False)).astype(np.uint32) for _ in range(2 + a.e2e_steps)]     # which rows: not the host's per-step work
    for it in range(2 + a.e2e_steps):
        if it == 2:
            barrier()
            t0 = time.perf_counter()
        rows = edit_rows[it]
        np.take(Tb, rows, axis=0, out=Teb); np.take(Pb, rows, axis=0, out=Peb)      # the host's edit: the rows as it last saw them (gathered like ComponentPools::to_device; as bytes: numpy's structured take is slow)
        w.update(rows, T=Te, P=Pe)
        w.step(DT)
        w.download_into(T=T, P=P)
        n_contacts2 = len(w.contacts())
    barrier()
    pools_sec = time.perf_counter() - t0
    pools_t = torch.tensor([pools_sec], dtype=torch.float64, device="cuda")
    if world_size > 1:
        dist.all_reduce(pools_t, op=dist.ReduceOp.MAX)
    e2e_pools = {"value": world_size * n_colliders * a.e2e_steps / float(pools_t.item()), "unit": UNIT,
                 "h2d_bytes_per_step": int(k_edit * (4 + 148)), "d2h_bytes_per_step": int(T.nbytes + P.nbytes + n_contacts2 * 48),
                 "what": "device-shared component pools: 1 % of the rows edited by the host and resubmitted per step (icx_bodies_update with a row list), "
                         "Transform + Physics of every entity read back in place; C++ form: icx_host::ComponentPools bound to PhysicsDynamo::run_relays "
                         "(tests/host/test_host_pools.cpp)"}
    tlk = timeline_kernels_bench(w, s.n, peak, local_rank) if (rank == 0 and not a.no_timeline) else None
    beh = projectile_behaviors_bench(w, s.n, peak) if (rank == 0 and not a.no_timeline) else None
    mesh = mesh_pass_bench(w, s, peak) if (rank == 0 and not a.no_timeline) else None
    barrier()
    w.close()

    long_run = None
    if rank == 0 and world_size == 1 and a.long_seconds > 0:          # (N = 1 only: the other ranks would sit in the barrier)
        long_run = long_run_bench(a, local_rank)
    barrier()

    wlb = None
    if a.worldlines > 0:
        wlb = worldline_batch_bench(a.worldlines, a.worldline_bodies, a.worldline_ticks, rank, world_size, local_rank, barrier, a.worldline_parts)

    tsp = None
    if a.pipeline_bodies > 0:
        tsp = timeslice_pipeline_bench(a.pipeline_bodies, a.pipeline_ticks, a.pipeline_delay, rank, world_size, local_rank, barrier)

    tt = None
    if a.pipeline_bodies > 0 and a.tt_timeslices > 1:
        tt = time_travel_bench(a.pipeline_bodies, a.tt_timeslices, a.tt_travellers, a.tt_delay, a.tt_ticks, rank, world_size, local_rank, barrier)
        barrier()

    cpu = None
    if rank == 0 and world_size == 1 and not a.no_cpu:          # reported at N = 1 only (the reference arm covers every N)
        cores = os.cpu_count() or 1
        threads = max(1, min(cores, 32))
        n_sample = min(a.bodies, 32768)
        v1, sec1, kind = cpu_reference(a.workload, n_sample, 1, threads)
        ticks = int(max(1, min(20, a.cpu_seconds / max(sec1, 1e-3))))
        v, sec, kind = cpu_reference(a.workload, n_sample, ticks, threads)
        cpu = {"value": v, "unit": UNIT, "cores": threads, "kind": kind,
               "sample": f"{threads} independent {n_sample}-body samples of {a.workload}, {ticks} ticks each ({sec:.1f} s): reference "
                         f"procedures (oracle/_ref) in reference pair order behind the oracle-side cull; verbatim all-pairs tick is O(N^2)"}

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world_size, "steps": a.steps, "warmup": a.warmup,
                "ms_per_step": ms_max / a.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "f32", "data": "synthetic",
                "config": {"workload": f"{a.workload}: {a.bodies} mixed sphere/AABB/OBB colliders + 6 walls in a closed arena, "
                                       f"one scene replica per GPU, gravity on, dt=1/45",
                           "l2": "inputs larger than L2 (per-tick state and derived arrays > 126 MB)" if a.bodies >= 400000 else "inputs fit L2 (reduced size)",
                           "colliders_per_gpu": n_colliders, "candidate_pairs": st.n_candidates, "contacts": st.n_contacts,
                           "dirty_bodies": st.n_dirty, "window": f"ticks {a.warmup + 1}-{a.warmup + a.steps} after the drop (the scene is still in free fall; see long_run for later ticks)",
                           "mean_candidate_pairs": round(mean_pairs, 1), "mean_contacts": round(mean_contacts, 1),
                           "mean_rounds": round(tot["rounds"] / max(tot["ticks"], 1), 2), "mean_attempts": round(tot["attempts"] / max(tot["ticks"], 1), 2),
                           "resolve_rounds": st.n_rounds, "resolve_attempts": st.n_attempts, "retries": st.n_retries, "margin": st.margin,
                           "per_rank": {"columns": ["ms_per_step", "retries", "resolve_attempts_last_tick", "sm_mhz_min", "slowdown_reasons"],
                                        "rows": per_rank}},
                "clocks": clk.summary(), "roofline": roofline,
                "e2e": {"value": e2e_val, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h)},
                "e2e_device_shared_pools": e2e_pools,
                "gpu_launches": int(launches1 - launches0),
                "long_run": long_run,
                "timeline_kernels": tlk,
                "behaviors": beh,
                "mesh_pass": mesh,
                "worldline_batch": wlb,
                "timeslice_pipeline": tsp,
                "time_travel": tt,
                "cpu_baseline": cpu}
        print(json.dumps(line), flush=True)
    if world_size > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
