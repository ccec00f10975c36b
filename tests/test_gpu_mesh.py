"""GPU: the static triangle-mesh pass (extension without a counterpart in the reference: ColliderType::mesh is commented out
in source/physics/collider.h:18-29).  Checked against the SELF-ORACLE icxo_mesh_pass -- an independent CPU restatement of
the definition in include/icx.h, brute force over all triangles: parity "unpinned" by construction -- bit for bit: the
contacts (entity, slot, triangle, point, normal, depth) in the pass's order and the whole component state, after the pair
loop of the same collision pass, which itself is checked against the pinned oracle."""
import numpy as np
import pytest

from intercession_b200 import abi, scenes, world
from intercession_b200.scenes import Scene
import refs
from util import assert_hits_equal, assert_state_equal, fields_diff

pytestmark = pytest.mark.gpu
DT = 1.0 / 45


def mesh_scene(n, seed, rotate=True):
    s = scenes.arena(n, seed=seed, walls=False, kinds=("sphere", "box", "aabb"))
    rng = np.random.default_rng(seed)
    if rotate:
        q = rng.standard_normal((s.n, 4)).astype(np.float32)
        q /= np.linalg.norm(q, axis=1, keepdims=True)
        s.T["orientation"][:] = q.astype(np.float32)
    s.P["velocity"][:] = rng.uniform(-2, 2, (s.n, 3)).astype(np.float32)
    s.P["velocity"][:, 1] -= 3.0
    lo, hi = s.T["origin"].min(0), s.T["origin"].max(0)
    fv, ft = scenes.height_mesh(float((hi - lo).max()) + 6.0, 24, y=float(lo[1]) + 0.4, bump=0.25, seed=seed)       # cuts through the lowest bodies
    tv, tt = scenes.torus_mesh(centre=tuple(0.5 * (lo + hi)), major=0.3 * float((hi - lo).max()), minor=0.9, n_major=40, n_minor=16)
    verts = np.concatenate([fv, tv])
    tris = np.concatenate([ft, tt + len(fv)]).astype(np.uint32)
    return s, verts, tris


def lockstep(s, verts, tris, material=None):
    """Collision pass (pair loop + mesh pass) from the same post-integration state: everything bit-exact."""
    with world.World(s.n) as w:
        w.upload(s)
        w.upload_mesh(verts, tris, material)
        w.collision_engage()
        got = w.download(s)
        hits, mhits, us = w.contacts(), w.mesh_contacts(), w.mesh_kernel_us()
    pairs_done = refs.procedures_tick(refs.oracle(), "icxo", s, euler=False)
    assert_hits_equal(hits, pairs_done.hits)
    exp, ehits = refs.oracle_mesh_pass(Scene({k: v for k, v in pairs_done.items() if k not in ("hits", "ints", "seconds")}), verts, tris, material)
    assert len(mhits) == len(ehits), (len(mhits), len(ehits))
    d = fields_diff(mhits, ehits)
    assert not d, d[:4]
    assert_state_equal(got, exp)
    assert us > 0
    return exp, ehits


@pytest.mark.parametrize("seed", [1, 2])
def test_mesh_pass_lockstep(built, seed):
    s, verts, tris = mesh_scene(3000, seed)
    exp, hits = lockstep(s, verts, tris)
    kinds = s.C["colliders"]["colliderType"][hits["row"], 0]
    assert (kinds == abi.COLLIDER_SPHERE).sum() > 20 and (kinds == abi.COLLIDER_BOX).sum() > 20
    assert len(np.unique(hits["triangle"])) > 50
    # some body met several triangles in one pass (the later ones saw it moved)
    assert np.unique(hits["row"], return_counts=True)[1].max() >= 2
    # every contact pushes its body away from the triangle's surface: after the pass nobody is deeper than it was reported
    assert (hits["depth"] > 0).all() and np.isfinite(hits["point"]).all()


def test_mesh_pass_skips_what_the_definition_skips(built):
    s, verts, tris = mesh_scene(1500, 5)
    s.hasP[100:400] = 0                                                             # no PhysicsComponent: not visited
    s.C["colliders"]["collisionType"][400:700, 0] = abi.COLLISION_NOOP
    s.C["colliders"]["isActive"][700:800, 0] = 0
    s.P["mass"][800:900] = 0                                                        # infinite mass: not visited
    exp, hits = lockstep(s, verts, tris, material=(0.9, 0.2, 0.1))
    rows = set(hits["row"].tolist())
    assert not rows & set(range(100, 900)) and len(rows) > 50


def test_mesh_body_over_many_triangles(built):
    """A body whose bounds cover more triangles than one staging window holds (and triangles registered in many cells): the
    windows partition one ascending sequence."""
    s, verts, tris = mesh_scene(300, 7)
    gv, gt = scenes.height_mesh(40.0, 64, y=float(s.T["origin"][:, 1].min()) + 0.3, bump=0.02, seed=1)       # 8 192 small triangles
    big = np.array([[-30, -9, -30], [30, -9, -30], [0, -9, 40]], np.float32)                                 # + one huge triangle under everything
    verts = np.concatenate([gv, big]); tris = np.concatenate([gt, np.array([[len(gv), len(gv) + 1, len(gv) + 2]])]).astype(np.uint32)
    s.T["scale"][5] = (14.0, 1.0, 14.0)                                             # a slab over ~1 000 triangles
    s.T["origin"][5, 1] = float(gv[:, 1].mean()) + 0.45
    s.T["orientation"][5] = (1, 0, 0, 0)
    s.P["mass"][5] = 1.0e6                                                          # the pair loop barely moves it
    s.C["colliders"]["colliderType"][5, 0] = abi.COLLIDER_BOX
    exp, hits = lockstep(s, verts, tris)
    assert (hits["row"] == 5).sum() >= 3                                            # pushed clear by the first few; ~1 200 candidates were walked


def test_mesh_pass_ray_spring_legs(built):
    """Bipeds over mesh ground: a second collider slot with ray-spring legs (client_focal_entity.h:30-91) on 400 of the bodies.
    The ray keeps its nearest hit (minParametricValue carries over from the pair loop) and answers it as a spring."""
    s, verts, tris = mesh_scene(1500, 6)
    h = np.float32(np.sqrt(0.5))
    legs = np.arange(50, 450)
    c = s.C["colliders"]
    c["isActive"][legs, 1] = 1
    c["colliderType"][legs, 1] = abi.COLLIDER_RAY
    c["collisionType"][legs, 1] = abi.COLLISION_SPRING
    c["localTransform"]["orientation"][legs, 1] = (h, h, 0, 0)
    c["localTransform"]["scale"][legs, 1] = (1, 1, 2.0)
    c["stiffness"][legs, 1], c["damping"][legs, 1], c["restLength"][legs, 1] = 2500, 400, 0.3
    c["staticFriction"][legs, 1], c["dynamicFriction"][legs, 1] = 0, 0
    c["inheritOrientation"][legs, 1], c["influenceOrientation"][legs, 1] = 0, 0
    c["minParametricValue"][legs, 1] = 1.0
    exp, hits = lockstep(s, verts, tris)
    ray = hits[hits["slot"] == 1]
    assert len(ray) > 10 and (ray["depth"] > 0).all()
    assert (exp.C["colliders"]["minParametricValue"][legs, 1] < 1.0).sum() >= len(np.unique(ray["row"]))


def test_mesh_contact_list_grows_after_an_overflow(built):
    """More contacts in one pass than the list holds: the states are complete, the fetch says the list was not and enlarges it."""
    s, _, _ = mesh_scene(3000, 4, rotate=False)
    lo, hi = s.T["origin"].min(0), s.T["origin"].max(0)
    span = float((hi - lo)[[0, 2]].max()) + 4
    vs, ts_, nv = [], [], 0
    for y in np.arange(float(lo[1]), float(hi[1]) + 0.5, 1.0):                      # sheets of small triangles through the whole arena:
        v, tr = scenes.height_mesh(span, 64, y=float(y), bump=0.0,                  # every body sits in one
                                   centre=(float(0.5 * (lo[0] + hi[0])), float(0.5 * (lo[2] + hi[2]))))
        vs.append(v); ts_.append(tr + nv); nv += len(v)
    verts, tris = np.concatenate(vs), np.concatenate(ts_).astype(np.uint32)
    exp_pairs = refs.procedures_tick(refs.oracle(), "icxo", s, euler=False)
    exp, ehits = refs.oracle_mesh_pass(Scene({k: v for k, v in exp_pairs.items() if k not in ("hits", "ints", "seconds")}), verts, tris)
    assert len(ehits) > 4096
    with world.World(s.n) as w:
        w.upload(s); w.upload_mesh(verts, tris)
        w.collision_engage()
        assert_state_equal(w.download(s), exp)                                      # the pass itself was complete
        with pytest.raises(world.IcxError) as ei:
            w.mesh_contacts()
        assert ei.value.code == abi.E_CAPACITY
        w.upload(s)
        w.collision_engage()
        got = w.mesh_contacts()
        assert len(got) == len(ehits) and not fields_diff(got, ehits)


def test_mesh_free_run_against_oracle(built):
    """Whole ticks with the mesh, device and oracle side by side from the drop: locked orientations keep the tick bit-exact
    (no sin / cos), so the two stay identical while bodies land on the ground mesh and roll off the torus."""
    s, verts, tris = mesh_scene(600, 9, rotate=False)
    s.P["lockOrientation"][:] = 1
    s.P["lockedOrientation"][:] = (1, 0, 0, 0)
    s.P["angularVelocity"][:] = 0
    cur, n_hits = s, 0
    with world.World(s.n) as w:
        w.upload(s)
        w.upload_mesh(verts, tris)
        for t in range(40):
            w.step(DT)
            got = w.download(s)
            mh = w.mesh_contacts()
            out = refs.procedures_tick(refs.oracle(), "icxo", cur, coherency=t)
            cur, eh = refs.oracle_mesh_pass(Scene({k: v for k, v in out.items() if k not in ("hits", "ints", "seconds")}), verts, tris)
            assert len(mh) == len(eh) and not fields_diff(mh, eh), t
            assert_state_equal(got, cur, what=f"tick {t}")
            n_hits += len(eh)
        # the ground is a two-sided sheet: a body whose centre starts below it is pushed out downwards; everyone who started
        # clearly above it is still above it
        n_ground = 2 * 24 * 24
        above = s.T["origin"][:, 1] > verts[:3 * n_ground, 1][: (24 + 1) ** 2].max() + 1.0
        assert n_hits > 200 and above.sum() > 300 and (got.T["origin"][above, 1] > verts[:, 1].min() - 1.0).all()
        w.remove_mesh()
        w.step(DT)
        assert len(w.mesh_contacts()) == 0
