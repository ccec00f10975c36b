"""CPU: the time-travel schedule (intercession_b200/timetravel.py run()) over the CPU back end alone -- the oracle procedures,
the reference's real EntityTimestreamMap / TsBreakpointQueue (oracle/_ref) and the pinned clone oracle.  What the reference's
chain of events must produce is asserted on its own here; tests/test_gpu_timetravel.py then holds the device to it
tick by tick."""
import numpy as np

from intercession_b200 import abi, timetravel
from cpu_timetravel import CpuBackend
from test_gpu_timetravel import make_chain


def test_one_traveller_one_cycle_on_the_reference_containers():
    import refs
    n_ts, delay = 3, 4
    chain = make_chain(150, n_ts)
    kinds = chain[0].C["colliders"]["colliderType"][:, 0]
    twin = next(i for i in range(36, chain[0].n) if kinds[i] == abi.COLLIDER_BOX)
    row = timetravel.traveller_row(chain[n_ts - 1], twin, genesis=240, timeslice=n_ts - 1)
    row.T["origin"][0, 0] += np.float32(0.3)
    cpu = CpuBackend(chain, delay)
    if refs.ref() is not None:
        assert cpu.containers == "icxref"
    log = []
    try:
        cycles = timetravel.run(cpu, 34, [(3, n_ts - 1, row)], lambda t, ph, d: log.append((t, ph, d)))
    finally:
        cpu.close()
    ticks = [(t, d) for t, ph, d in log if ph == "tick"]
    # the traveller exists in the past-most timeslice only, and only there anything is intercepted
    assert all(len(d["interceptions"][0]) == 0 and len(d["interceptions"][1]) == 0 for _, d in ticks)
    first = next(t for t, d in ticks if len(d["interceptions"][2]))
    assert first >= 3
    # forking -> forked 13 ticks after the first interception, resolution needed 13 ticks after that (server_network_dynamo.cpp:681-705)
    promoted = next(t for t, d in ticks if d["ages"][2][0] > 0)
    needed = next(t for t, d in ticks if d["ages"][2][1])
    assert promoted == first + 13 and needed == promoted + 13
    # ... and the cycle runs at the start of the next wall tick, down the whole chain
    assert cycles == 1
    (tc, _, res), = [e for e in log if e[1] == "cycle"]
    assert tc == needed + 1 and len(res) == n_ts - 1
    for r in res:
        assert r["cloned"] == chain[0].n                       # the traveller (chain-link 0, no future timestream) is not cloned
        assert r["extracted"] == 150 and r["missing"] == 0     # every temporal body goes up the chain at chain-link - 1
        assert len(r["interceptions"]) == delay                # re-simulated up to the future neighbour's coherency
    # load_and_link reset the source timeslice: nothing is forked right after the cycle
    after = next(d for t, d in ticks if t == tc)
    assert (after["states"][2].tstate != abi.TS_FORKED).all()


def test_an_arrival_in_a_middle_timeslice_travels_on_into_the_past():
    """_entity_created_handler / _jump_arrival_handler push the arrival into the past stream (server_network_dynamo.cpp:748-777,
    :940-976): `delay` wall ticks later the traveller exists in the next timeslice too, one causal chain-link further."""
    n_ts, delay = 3, 4
    chain = make_chain(150, n_ts)
    kinds = chain[0].C["colliders"]["colliderType"][:, 0]
    twin = next(i for i in range(36, chain[0].n) if kinds[i] == abi.COLLIDER_BOX)
    row = timetravel.traveller_row(chain[1], twin, genesis=240, timeslice=1)
    row.T["origin"][0, 0] += np.float32(0.3)
    cpu = CpuBackend(chain, delay)
    sizes = []
    try:
        timetravel.run(cpu, 18, [(3, 1, row)], lambda t, ph, d: sizes.append([s.n for s in d["states"]]) if ph == "tick" else None,
                       departures=[(12, 1, row.ids.copy())])
        there = int(row.ids[0]) + 1
        assert sizes[2] == [156, 156, 156] and sizes[3] == [156, 157, 156] and sizes[3 + delay - 1] == [156, 157, 156] and sizes[3 + delay] == [156, 157, 157]
        # ... and when it leaves again (wall tick 12), ENTITY_REMOVED follows it into the past
        assert sizes[11] == [156, 157, 157] and sizes[12] == [156, 156, 157] and sizes[12 + delay] == [156, 156, 156]
        assert there not in cpu.scene[2].ids.tolist() and int(row.ids[0]) not in cpu.scene[1].ids.tolist()
    finally:
        cpu.close()
