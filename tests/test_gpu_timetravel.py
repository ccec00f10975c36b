"""GPU: config 5 end to end at a reduced size ("interception / flag parity vs the CPU restatement at <= 4 K
bodies"): three timeslices, a time traveller arrives in the past-most one, collides, its victims fork, the forks age, the
parallel cosmos is loaded from the past-most timeslice, re-simulates with the breakpoint record / replay, extracts up the chain.
Device (IcxWorld / IcxTimeline through the C ABI, intercession_b200/timetravel.py) and CPU restatement (oracle procedures + the
reference's real EntityTimestreamMap + clone oracle, tests/cpu_timetravel.py) run the SAME schedule; every tick the interception
lists, the timestream states and stamps, and the whole component state of every timeslice must be identical, and so must the
parallel cosmos at the end of every re-simulation."""
import numpy as np
import pytest

from intercession_b200 import abi, pipeline, scenes, timetravel
from util import fields_diff
from cpu_timetravel import CpuBackend

pytestmark = pytest.mark.gpu


def make_chain(n_bodies, n_ts, seed=31):
    side = 2.0 + 1.1 * round(n_bodies ** (1 / 3) + 1)
    base = scenes.arena(n_bodies, seed=seed, walls=True, temporal=True, fill_side=side, spacing=1.05, jitter=0.25, kinds=("sphere", "box", "aabb"))
    base.T["orientation"][:] = (1, 0, 0, 0)                       # no rotation anywhere: the tick is bit-exact end to end (no sin / cos ulp),
    base.P["lockOrientation"][:] = 1                              # so device and CPU can free-run side by side for the whole scenario
    base.P["lockedOrientation"][:] = (1, 0, 0, 0)
    base.P["angularVelocity"][:] = 0
    chain = []
    for k in range(n_ts):
        s = base.copy()
        s["ids"] = pipeline.past_ids(base.ids, k)
        chain.append(s)
    return chain


def same_scene(a, b, what):
    assert a.n == b.n and np.array_equal(a.ids, b.ids), what
    for key in ("T", "P", "C"):
        d = fields_diff(a[key], b[key])
        assert not d, (what, key, d[:3])
    assert np.array_equal(a.tstate, b.tstate) and np.array_equal(a.tcoh, b.tcoh), what


def _device_layouts():
    import torch
    out = [pytest.param([0, 0, 0], id="one-gpu")]
    if torch.cuda.is_available() and torch.cuda.device_count() >= 2:
        out.append(pytest.param([0, 1, 0], id="two-gpus"))          # every link and every extract crosses NVLink
    return out


# 150 bodies, genesis ids < 255: the CPU side keeps its records in the reference's REAL EntityTimestreamMap / TsBreakpointQueue
# (16-bit Entity); 500 bodies, traveller genesis 2000: in their restatement (pinned to them by tests/test_timeline_oracle.py)
@pytest.mark.parametrize("n_bodies,genesis", [(150, 240), (500, 2000)], ids=["reference-containers", "wide-ids"])
@pytest.mark.parametrize("devices", _device_layouts())
def test_traveller_forks_its_victims_and_the_parallel_cosmos_resolves_them(built, devices, n_bodies, genesis):
    import refs
    n_ts, delay, ticks = 3, 4, 38
    chain = make_chain(n_bodies, n_ts)
    # the traveller: a copy of a box of the past-most timeslice, pushed 0.3 units into it, arriving at wall tick 3
    kinds = chain[0].C["colliders"]["colliderType"][:, 0]
    twin = next(i for i in range(6 + n_bodies // 5, chain[0].n) if kinds[i] == abi.COLLIDER_BOX)
    row = timetravel.traveller_row(chain[n_ts - 1], twin, genesis=genesis, timeslice=n_ts - 1)
    row.T["origin"][0, 0] += np.float32(0.3)
    arrivals = [(3, n_ts - 1, row)]
    dev = timetravel.DeviceBackend(chain, delay, devices=devices)
    cpu = CpuBackend(chain, delay)
    if genesis < 255 and refs.ref() is not None:
        assert cpu.containers == "icxref"
    log = {"dev": [], "cpu": []}
    try:
        cd = timetravel.run(dev, ticks, arrivals, lambda t, ph, d: log["dev"].append((t, ph, d)))
        cc = timetravel.run(cpu, ticks, arrivals, lambda t, ph, d: log["cpu"].append((t, ph, d)))
    finally:
        dev.close(); cpu.close()
    assert cd == cc >= 1                                           # at least one parallel cycle ran, on the same wall tick(s)
    assert len(log["dev"]) == len(log["cpu"])
    n_ints = n_forked = 0
    for (t, ph, d), (t2, ph2, c) in zip(log["dev"], log["cpu"]):
        assert (t, ph) == (t2, ph2)
        if ph == "tick":
            assert d["ages"] == c["ages"], t
            for k in range(n_ts):
                assert len(d["interceptions"][k]) == len(c["interceptions"][k]) and not fields_diff(d["interceptions"][k], c["interceptions"][k]), (t, k)
                same_scene(d["states"][k], c["states"][k], (t, k))
                n_ints += len(d["interceptions"][k])
                n_forked += int((d["states"][k].tstate == abi.TS_FORKED).sum())
        else:
            for rd, rc in zip(d, c):
                assert (rd["cloned"], rd["extracted"], rd["missing"]) == (rc["cloned"], rc["extracted"], rc["missing"]), t
                assert len(rd["interceptions"]) == len(rc["interceptions"])
                for a, b in zip(rd["interceptions"], rc["interceptions"]):
                    assert len(a) == len(b) and not fields_diff(a, b), t
                same_scene(rd["parallel_end"], rc["parallel_end"], (t, "parallel"))
    assert n_ints > 0 and n_forked > 0
    # after the cycle every timeslice but the traveller's own victims' copies in timeslice 0 ... : timeslice 0 is fully merged again
    last = [e for e in log["dev"] if e[1] == "tick"][-1][2]["states"]
    assert (last[0].tstate == abi.TS_MERGED).all()


@pytest.mark.parametrize("devices", _device_layouts())
def test_an_arrival_in_a_middle_timeslice_is_created_down_the_chain(built, devices):
    """The event records in the device timeline at work: a traveller arrives in timeslice 1 of 3; ENTITY_CREATED + JUMP_ARRIVAL ride
    in the tick's snapshot slot into timeslice 2's ring; `delay` wall ticks later they come due there, the entity is created on
    the device from the components the same tick's record holds (icx_timeline_entity_fetch -> icx_entities_create) and follows
    its future self from then on.  Device and CPU restatement (real EntityTimestreamMap for records AND events) identical on
    every tick."""
    import refs
    n_ts, delay, ticks = 3, 4, 22
    chain = make_chain(150, n_ts)
    kinds = chain[0].C["colliders"]["colliderType"][:, 0]
    twin = next(i for i in range(36, chain[0].n) if kinds[i] == abi.COLLIDER_BOX)
    row = timetravel.traveller_row(chain[1], twin, genesis=240, timeslice=1)
    row.T["origin"][0, 0] += np.float32(0.3)
    arrivals = [(3, 1, row)]
    dev = timetravel.DeviceBackend(chain, delay, devices=devices)
    cpu = CpuBackend(chain, delay)
    if refs.ref() is not None:
        assert cpu.containers == "icxref"
    log = {"dev": [], "cpu": []}
    try:
        leaves = [(12, 1, row.ids.copy())]                         # ... and leaves again at wall tick 12: ENTITY_REMOVED follows it down
        timetravel.run(dev, ticks, arrivals, lambda t, ph, d: log["dev"].append((t, ph, d)), departures=leaves)
        timetravel.run(cpu, ticks, arrivals, lambda t, ph, d: log["cpu"].append((t, ph, d)), departures=leaves)
        mism = [int(t.id_mismatches()) for t in dev.tl_in if t is not None]
    finally:
        dev.close(); cpu.close()
    assert len(log["dev"]) == len(log["cpu"])
    n_ints = 0
    for (t, ph, d), (t2, ph2, c) in zip(log["dev"], log["cpu"]):
        assert (t, ph) == (t2, ph2) and ph == "tick"
        assert d["ages"] == c["ages"], t
        for k in range(n_ts):
            assert len(d["interceptions"][k]) == len(c["interceptions"][k]) and not fields_diff(d["interceptions"][k], c["interceptions"][k]), (t, k)
            same_scene(d["states"][k], c["states"][k], (t, k))
            n_ints += len(d["interceptions"][k])
    sizes = [[s.n for s in d["states"]] for _, _, d in log["dev"]]
    assert sizes[2] == [156, 156, 156] and sizes[3] == [156, 157, 156] and sizes[3 + delay] == [156, 157, 157]
    assert sizes[11] == [156, 157, 157] and sizes[12] == [156, 156, 157] and sizes[12 + delay] == [156, 156, 156]
    assert n_ints > 0
    # while the traveller was unknown in timeslice 2 (delay ticks: its records were on their way, not due) nothing was missed:
    # once it exists every record finds its entity
    assert mism[1] == 0
