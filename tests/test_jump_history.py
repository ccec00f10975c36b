"""a21, second half: the jump bookkeeping of ParallelNetworkDynamo (cached departures, per-trip arrival history
with the paradox search, resolved trips; parallel_network_dynamo.cpp:155-161, :303-355, :488-593) in the host
mirror (intercession_b200/host/jump_history.hpp) against the reference-side harness, which keeps the reference's
three containers and compares with the REAL TimejumpConditions::operator== (oracle/_ref)."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

import refs

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tests", "host", "test_jump_history.cpp")
EXE = os.path.join(ROOT, "build", "test_jump_history")


@pytest.fixture(scope="module")
def exe():
    os.makedirs(os.path.dirname(EXE), exist_ok=True)
    subprocess.run(["g++", "-std=c++17", "-O2", "-Wall", "-Wextra", "-Werror", SRC, "-o", EXE], check=True)
    return EXE


@pytest.fixture(scope="module")
def ref():
    if not os.path.exists(refs.REF_PATH):
        pytest.skip("oracle/_ref not built")
    lib = C.CDLL(refs.REF_PATH)
    pf, pu = C.POINTER(C.c_float), C.POINTER(C.c_uint32)
    lib.icxref_jumps_create.restype = C.c_void_p
    lib.icxref_jumps_destroy.argtypes = [C.c_void_p]
    lib.icxref_jumps_clear_departures.argtypes = [C.c_void_p]
    lib.icxref_jumps_departure.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, pf]
    lib.icxref_jumps_arrival.argtypes = [C.c_void_p, C.c_uint32, pf]
    lib.icxref_jumps_request.argtypes = [C.c_void_p, C.c_uint32, pf, pu]
    return lib


def run_ref(lib, script):
    h = lib.icxref_jumps_create()
    out = []
    for op in script:
        if op[0] == "C":
            lib.icxref_jumps_clear_departures(h); out.append("C")
            continue
        o = (C.c_float * 3)(*[float(x) for x in op[-1]])
        if op[0] == "D":
            lib.icxref_jumps_departure(h, op[1], op[2], o); out.append("D")
        elif op[0] == "A":
            out.append("A %d" % lib.icxref_jumps_arrival(h, op[1], o))
        elif op[0] == "R":
            t = C.c_uint32(0)
            k = lib.icxref_jumps_request(h, op[1], o, C.byref(t))
            out.append("R %d %u" % (k, t.value))
    lib.icxref_jumps_destroy(h)
    return out


def run_host(exe, script):
    lines = []
    for op in script:
        if op[0] == "C":
            lines.append("C")
            continue
        bits = " ".join("%x" % int(b) for b in np.asarray(op[-1], np.float32).view(np.uint32))
        lines.append(" ".join([op[0]] + [str(x) for x in op[1:-1]] + [bits]))
    r = subprocess.run([exe], input="\n".join(lines) + "\n", capture_output=True, text=True, check=True)
    return r.stdout.split("\n")[:-1]


def test_two_phase_paradox_by_hand(exe):
    """A -> B -> A on one trip: the third arrival differs from the second and matches the first."""
    a, b = np.float32([0, 0, 0]), np.float32([5, 0, 0])
    near_a = np.float32([1, 1, 0])                   # length2 = 2.0 exactly: still congruent (<=)
    far_a = np.float32([1, 1, 0.01])                 # just outside
    script = [("A", 7, a), ("A", 7, a), ("A", 7, b), ("A", 7, near_a),          # 0 0 0 1 (paradox: != b, == a)
              ("A", 7, near_a), ("A", 7, b), ("A", 7, far_a),                   # history restarted with near_a: 0, 0, then far_a != b ...
              ("D", 0x10000011, 7, a), ("D", 0x10000011, 9, b),                 # second insert for the same entity is ignored
              ("R", 0x10000011, far_a), ("R", 0x10000011, a), ("C",)]
    got = run_host(exe, script)
    # far_a vs history [near_a, near_a, b]: != b; vs near_a: d = (0,0,0.01) -> congruent -> paradox again
    assert got[:7] == ["A 0", "A 0", "A 0", "A 1", "A 0", "A 0", "A 1"]
    # request 1: cached trip 7 with origin a; far_a is not congruent; trip 7 is resolved -> kind 2, entry erased
    # request 2: nothing cached any more -> missing, tripId 91339
    assert got[7:] == ["D", "D", "R 2 7", "R 0 91339", "C"]


@pytest.mark.parametrize("seed", [1, 2, 3, 4])
def test_random_scripts_match_reference_harness(exe, ref, seed):
    rng = np.random.default_rng(seed)
    # a few anchor points ~1.4 apart so that congruence flips often; jitter puts many cases near the 2.0 threshold
    anchors = rng.uniform(-2, 2, (6, 3)).astype(np.float32)
    script = []
    for _ in range(4000):
        o = (anchors[rng.integers(6)] + rng.normal(0, 0.45, 3)).astype(np.float32)
        k = rng.random()
        if k < 0.5:
            script.append(("A", int(rng.integers(5)), o))
        elif k < 0.72:
            script.append(("D", int(0x10000010 + rng.integers(8)), int(rng.integers(5)), o))
        elif k < 0.98:
            script.append(("R", int(0x10000010 + rng.integers(8)), o))
        else:
            script.append(("C",))
    want = run_ref(ref, script)
    got = run_host(exe, script)
    assert len(got) == len(want)
    bad = [i for i, (g, w) in enumerate(zip(got, want)) if g != w]
    assert not bad, (bad[:5], [(script[i], got[i], want[i]) for i in bad[:3]])
    assert sum(1 for w in want if w == "A 1") > 20 and sum(1 for w in want if w.startswith("R 2")) > 3   # the interesting branches ran


def test_threshold_is_inclusive_and_unfused(exe, ref):
    """length2 <= 2.0f at the boundary, on values where an FMA-contracted sum would round differently."""
    rng = np.random.default_rng(11)
    script = []
    for i in range(3000):
        d = rng.normal(0, 1, 3); d = d / np.linalg.norm(d) * np.sqrt(2.0) * (1 + rng.uniform(-3e-7, 3e-7))
        base = rng.uniform(-50, 50, 3).astype(np.float32)
        script.append(("A", i, base))
        script.append(("A", i, (base + d.astype(np.float32)).astype(np.float32)))      # differs from back()? then history scan
        script.append(("D", i, i, base))
        script.append(("R", i, (base + d.astype(np.float32)).astype(np.float32)))
    assert run_host(exe, script) == run_ref(ref, script)
