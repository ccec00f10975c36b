"""CPU suite, part 2: live cross-checks between the three CPU implementations, where the compiled
reference is present (oracle/_ref/libicx_ref.so; built in the container that has /root/reference and
shipped prebuilt to the GPU box):

  verbatim   the reference OBJECTS: Cosmos + PhysicsSynchro/ColliderSynchro + PhysicsDynamo
  icxref_*   the reference PROCEDURES (EulerPhysicsRelay, intersect/response tables) over plain
             arrays, pair loop restated in oracle/ref/ref_procedures.cpp (needed beyond 65 535 ids)
  icxo_*     the oracle: plain C++ restatement without glm (oracle/icx_oracle.cpp)

plus the oracle's own broadphase definition (grid cull == brute force, and conservative)."""
import numpy as np
import pytest

from intercession_b200 import abi, scenes
import katgen
import refs
from util import assert_hits_equal, assert_state_equal, fields_diff

needs_ref = pytest.mark.skipif(refs.ref() is None, reason="oracle/_ref not built (no /root/reference here)")


def dense(n, seed, **kw):
    side = 2.0 + 1.1 * round(n ** (1 / 3) + 1)
    kw.setdefault("kinds", ("sphere", "box", "aabb"))
    return scenes.arena(n, seed=seed, fill_side=side, spacing=1.05, jitter=0.25, **kw)


def temporal_mix(s):
    n = s.n
    s.ids[n // 20: n // 2] |= 1
    s.ids[n // 20: n // 2] &= ~np.uint32(14)
    s.tstate[n // 3: 2 * n // 3] = abi.TS_FORKED
    s.tstate[2 * n // 3: 3 * n // 4] = abi.TS_FORKING
    return s


@needs_ref
@pytest.mark.parametrize("host", [0, abi.NULL_TIMESLICE])
def test_verbatim_objects_vs_procedures_harness(built, host):
    """The restated pair loop / interception predicate of the procedures harness reproduces the
    reference objects exactly (this is what licenses using it beyond the uint16 id limit)."""
    s = temporal_mix(scenes.arena(220, seed=31, fill_side=9.0, spacing=1.0, jitter=0.22,
                                  kinds=("sphere", "box", "aabb"), temporal=True))
    v = refs.verbatim_run(s, n_ticks=1, host=host, coherency=5)
    p = refs.procedures_tick(refs.ref(), "icxref", s, host=host, coherency=5)
    assert len(v.hits) > 30 and len(v.ints) > 5
    assert_state_equal(v, p, "verbatim vs procedures")
    assert_hits_equal(v.hits, p.hits, ignore_responded=True)
    assert not fields_diff(v.ints, p.ints)


@needs_ref
@pytest.mark.parametrize("seed", [0, 1, 2])
@pytest.mark.parametrize("flags", [0, abi.F_SPHERE_SPHERE])
def test_oracle_vs_reference_procedures(built, seed, flags):
    s = temporal_mix(dense(1200, seed, temporal=True))
    for host in (0, abi.NULL_TIMESLICE):
        r = refs.procedures_tick(refs.ref(), "icxref", s, host=host, coherency=7, flags=flags)
        o = refs.procedures_tick(refs.oracle(), "icxo", s, host=host, coherency=7, flags=flags)
        assert len(r.hits) > 100
        assert_state_equal(o, r, "oracle vs reference procedures")
        assert_hits_equal(o.hits, r.hits)
        assert not fields_diff(o.ints, r.ints)


@needs_ref
def test_oracle_vs_reference_free_run(built):
    r = o = dense(600, 5)
    for t in range(40):
        r = refs.procedures_tick(refs.ref(), "icxref", r, coherency=t)
        o = refs.procedures_tick(refs.oracle(), "icxo", o, coherency=t)
    assert_state_equal(o, r, "40 free-running ticks")


@needs_ref
def test_oracle_vs_reference_kats_fresh_seed(built):
    cases = katgen.random_cases(4711, per_combo=24)
    exp = katgen.run_kat_isolated(refs.ref(), "icxref", cases)
    got = katgen.run_kat(refs.oracle(), "icxo", cases)
    ok = exp["hit"] != katgen.REF_CRASHED
    assert not fields_diff(got[ok], exp[ok])


@needs_ref
def test_oracle_cull_equals_all_pairs_relay(built):
    """the oracle-side cull is validated against the true all-pairs loop -- identical
    hit sequences and end state when the reference procedures visit only the culled candidates."""
    s = temporal_mix(dense(1500, 9, temporal=True))
    full = refs.procedures_tick(refs.ref(), "icxref", s, euler=False)
    cand = refs.oracle_pairs(s, 0.75)      # dense scene: static resolution moves bodies by up to ~0.5
    culled = refs.procedures_tick(refs.ref(), "icxref", s, euler=False, cand=cand)
    assert len(cand) < 0.05 * s.n * s.n
    assert_state_equal(culled, full, "culled vs all-pairs")
    assert_hits_equal(culled.hits, full.hits)
    assert not fields_diff(culled.ints, full.ints)


@pytest.mark.parametrize("margin", [0.0, 0.05, 0.3])
def test_broadphase_grid_equals_brute_force(built, margin):
    s = dense(900, 12)
    brute = refs.oracle_pairs(s, margin, method=0)
    grid = refs.oracle_pairs(s, margin, method=1)
    assert len(brute) > 500
    assert not fields_diff(grid, brute)
    # every hit of the sequential pass is a listed candidate (conservativeness) once the margin
    # covers the static-resolution displacement
    if margin >= 0.3:
        o = refs.procedures_tick(refs.oracle(), "icxo", s, euler=False)
        keys = set((int(p["a"]), int(p["b"])) for p in brute)
        assert all((int(h["packetA"]), int(h["packetB"])) in keys for h in o.hits)


def test_broadphase_bounds_contain_shapes(built):
    """The bound definition (oracle packet_aabb) really bounds each collider: corners of boxes,
    sphere centre +- radius, both ray ends lie inside the (margin 0) box."""
    s = dense(400, 3)
    s.C["colliders"]["isActive"][10:60, 1] = 1
    s.C["colliders"]["colliderType"][10:60, 1] = abi.COLLIDER_RAY
    s.C["colliders"]["localTransform"]["scale"][10:60, 1] = (1, 1, 2.5)
    bb = refs.oracle_aabbs(s, 0.0)
    assert len(bb) == int(s.C["colliders"]["isActive"].sum())
    assert (bb[:, 3:] >= bb[:, :3]).all()
    # boxes with identity local transform: corners = origin + R * (+-0.5 * scale)
    k = 0
    for i in range(s.n):
        for slot in range(2):
            c = s.C["colliders"][i, slot]
            if not c["isActive"]:
                continue
            if c["colliderType"] == abi.COLLIDER_BOX and slot == 0:
                w, x, y, z = s.T["orientation"][i].astype(np.float64)
                R = np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)],
                              [2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)],
                              [2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)]])
                corners = np.array([[sx, sy, sz] for sx in (-.5, .5) for sy in (-.5, .5) for sz in (-.5, .5)]) * s.T["scale"][i]
                wc = corners @ R.T + s.T["origin"][i]
                assert (wc >= bb[k, :3] - 1e-4).all() and (wc <= bb[k, 3:] + 1e-4).all()
            k += 1
