"""Synthetic scenes of the shapes BASELINE.json names, as AoS numpy arrays.

All generators are deterministic (numpy PCG64, fixed draw order) and vectorised so
the 1 M-collider scene builds in about a second.  Entity order == submission order
== ascending id (the reference iterates a std::set<Entity>, ecs/i_synchro.h:48).
"""
import numpy as np

from . import abi


class Scene(dict):
    """ids, T, P, hasP, C, tstate, tcoh (+ meta) -- attribute access for brevity."""
    __getattr__ = dict.__getitem__

    @property
    def n(self):
        return len(self["ids"])

    def copy(self):
        return Scene({k: (v.copy() if isinstance(v, np.ndarray) else v) for k, v in self.items()})


def _empty(n):
    return Scene(ids=np.zeros(n, np.uint32), T=abi.default_transforms(n), P=abi.default_physics(n),
                 hasP=np.ones(n, np.uint8), C=abi.default_colliders(n),
                 tstate=np.zeros(n, np.uint8), tcoh=np.zeros(n, np.uint16))


def _unit_quats(rng, n):
    q = rng.standard_normal((n, 4)).astype(np.float32)
    q /= np.linalg.norm(q, axis=1, keepdims=True).astype(np.float32)
    return q.astype(np.float32)


def _lattice_positions(rng, n, L, spacing=1.25, jitter=0.1):
    """n distinct jittered lattice sites inside (-L/2+1, L/2-1)^3; any two are > spacing-2*jitter apart."""
    m = int(np.floor((L - 2.0) / spacing))
    assert m ** 3 >= n, "arena too small for the requested body count"
    sites = rng.choice(m ** 3, size=n, replace=False)
    ijk = np.stack([sites // (m * m), (sites // m) % m, sites % m], axis=1).astype(np.float32)
    lo = -0.5 * (m - 1) * spacing
    pos = lo + ijk * np.float32(spacing) + rng.uniform(-jitter, jitter, (n, 3)).astype(np.float32)
    return pos.astype(np.float32)


def _walls(s, first, L, thickness=2.0):
    """Six infinite-mass, fully locked wall boxes closing the cube of side L."""
    k = first
    for axis in range(3):
        for sign in (-1.0, 1.0):
            o = np.zeros(3, np.float32)
            o[axis] = sign * (L / 2 + thickness / 2)
            sc = np.full(3, L + 2 * thickness, np.float32)
            sc[axis] = thickness
            s.T["origin"][k] = o
            s.T["scale"][k] = sc
            s.P["mass"][k] = 0.0
            s.P["lockOrigin"][k] = 1
            s.P["lockedOrigin"][k] = o
            s.P["lockOrientation"][k] = 1
            c = s.C["colliders"][k, 0]
            c["isActive"] = 1
            s.C["colliders"][k, 0] = c
            k += 1
    return k


def arena(n_bodies, seed=20261104, kinds=("sphere", "box"), mix=None, walls=True, temporal=False,
          timeslice=0, fill_side=None, spacing=1.25, jitter=0.1):
    """Config-2/3 recipe: closed cubic arena of side 2*cbrt(N) holding N dynamic bodies.

    kinds/mix: body i draws its kind from `kinds` (round-robin when mix is None, else
    by the given probabilities): "sphere", "box" (random orientation = OBB) or "aabb"
    (box, identity orientation, lockOrientation) -- collider.h:18-29 has no AABB type,
    an AABB is a degenerate box.
    """
    rng = np.random.default_rng(seed)
    nw = 6 if walls else 0
    n = n_bodies + nw
    s = _empty(n)
    L = float(fill_side) if fill_side else 2.0 * n_bodies ** (1.0 / 3.0)
    link = 0 if temporal else abi.NULL_CHAINLINK
    g = np.arange(n, dtype=np.uint32)
    s["ids"] = ((np.uint32(timeslice) << np.uint32(28)) | (g << np.uint32(4))
                | np.where(g < nw, abi.NULL_CHAINLINK, link).astype(np.uint32))
    if walls:
        _walls(s, 0, L)
    b = slice(nw, n)
    if mix is None:
        kind = np.arange(n_bodies) % len(kinds)
    else:
        kind = rng.choice(len(kinds), size=n_bodies, p=mix)
    s.T["origin"][b] = _lattice_positions(rng, n_bodies, L, spacing, jitter)
    scale = rng.uniform(0.5, 1.0, n_bodies).astype(np.float32)
    s.T["scale"][b] = scale[:, None]
    q = _unit_quats(rng, n_bodies)
    is_aabb = np.array([kinds[k] == "aabb" for k in kind])
    q[is_aabb] = np.array([1, 0, 0, 0], np.float32)
    s.T["orientation"][b] = q
    s.P["velocity"][b] = rng.uniform(-1, 1, (n_bodies, 3)).astype(np.float32)
    w = rng.uniform(-0.5, 0.5, (n_bodies, 3)).astype(np.float32)
    w[is_aabb] = 0
    s.P["angularVelocity"][b] = w
    s.P["mass"][b] = scale ** 3 * np.float32(10.0)
    s.P["lockOrientation"][b] = is_aabb.astype(np.uint8)
    cc = s.C["colliders"]
    cc["isActive"][b, 0] = 1
    is_sphere = np.array([kinds[k] == "sphere" for k in kind])
    cc["colliderType"][b, 0] = np.where(is_sphere, abi.COLLIDER_SPHERE, abi.COLLIDER_BOX)
    s["side"] = L
    s["n_walls"] = nw
    return s


def config2(n=65536, seed=20261104):
    """64 K rigid spheres + boxes in a closed arena (BASELINE.json configs[1])."""
    return arena(n, seed, kinds=("sphere", "box"))


def config3(n=1048576, seed=20261105):
    """1 M mixed: 40 % sphere, 30 % AABB, 30 % OBB (BASELINE.json configs[2]; mesh colliders
    do not exist in the reference)."""
    return arena(n, seed, kinds=("sphere", "aabb", "box"), mix=(0.4, 0.3, 0.3))


def case_study(push=False):
    """Config 1: physics entities of build_moon_cosmos (source/staging/moon_cosmos.cpp:25-72,
    :190-210) + create_client_focal_entity (source/staging/client_focal_entity.h:30-91), as
    restated in SURVEY.md 8d.  Ids follow the reference allocator for timeslice 0."""
    s = _empty(4)
    half = np.float32(np.sqrt(0.5))
    # floor_1, floor_2 : non-temporal infinite-mass locked boxes, rotated -90deg about x
    for k, (o, sc) in enumerate([((12, -4, 0), (24, 24, 2)), ((-12, -6, 0), (24, 24, 0.05))]):
        s.ids[k] = abi.compose_entity(0, k, abi.NULL_CHAINLINK)
        s.T["origin"][k] = o
        s.T["orientation"][k] = (half, -half, 0, 0)      # angleAxis(90deg, (-1,0,0))
        s.T["scale"][k] = sc
        s.P["mass"][k] = 0
        s.P["lockOrigin"][k] = 1
        s.P["lockedOrigin"][k] = o
        s.P["lockOrientation"][k] = 1
        s.P["lockedOrientation"][k] = (half, -half, 0, 0)
        s.C["colliders"]["isActive"][k, 0] = 1
        s.C["colliders"]["localTransform"]["origin"][k, 0] = (0, 0, -0.499)
    # point: temporal sphere
    s.ids[2] = abi.compose_entity(0, 2, 0)
    s.T["origin"][2] = (12, -1, 0)
    s.P["mass"][2] = 100
    c = s.C["colliders"]
    c["isActive"][2, 0] = 1
    c["colliderType"][2, 0] = abi.COLLIDER_SPHERE
    c["dynamicFriction"][2, 0] = 0.9
    c["restitution"][2, 0] = 0.1
    # focal biped: box rigid + ray spring "legs"
    s.ids[3] = abi.compose_entity(0, 3, 0)
    s.T["origin"][3] = (12 if push else 0, 3, 0)
    s.P["mass"][3] = 40
    s.P["lockOrientation"][3] = 1
    c["isActive"][3, 0] = 1
    c["isActive"][3, 1] = 1
    c["colliderType"][3, 1] = abi.COLLIDER_RAY
    c["collisionType"][3, 1] = abi.COLLISION_SPRING
    c["localTransform"]["orientation"][3, 1] = (half, half, 0, 0)    # angleAxis(90deg, (1,0,0))
    c["localTransform"]["scale"][3, 1] = (1, 1, 1.5)
    c["stiffness"][3, 1] = 2500
    c["damping"][3, 1] = 400
    c["restLength"][3, 1] = 0.3
    c["staticFriction"][3, 1] = 0
    c["dynamicFriction"][3, 1] = 0
    c["inheritOrientation"][3, 1] = 0
    c["influenceOrientation"][3, 1] = 0
    c["useBehaviorsResponse"][3, 1] = 1
    return s


def worldline_batch(n_worldlines, bodies_per_worldline=1024, seed=20261106, perturb=True):
    """Config 4: W copies of one base scene (config-2 recipe, no walls, side ~ 20), candidate w
    perturbs the velocity of body (w mod n) by U(-2,2)^3 (seed + w).  Entities are temporal with
    chain-link 1 (a parallel cosmos never holds link-0 entities, collision_physics_relay.h:182)."""
    base = arena(bodies_per_worldline, seed, kinds=("sphere", "box"), walls=False)
    n = bodies_per_worldline
    s = _empty(n * n_worldlines)
    for k in ("T", "P", "hasP", "C", "tstate", "tcoh"):
        s[k] = np.tile(base[k], n_worldlines)
    g = np.arange(n * n_worldlines, dtype=np.uint32)
    s["ids"] = (np.uint32(abi.NULL_TIMESLICE) << np.uint32(28)) | (g << np.uint32(4)) | np.uint32(1)
    if perturb:
        for w in range(n_worldlines):
            r = np.random.default_rng(seed + 1 + w)
            s.P["velocity"][w * n + (w % n)] += r.uniform(-2, 2, 3).astype(np.float32)
    s["base"] = base
    s["stride"] = n
    return s


def torus_mesh(centre=(0.0, 0.0, 0.0), major=6.0, minor=2.0, n_major=48, n_minor=24, axis=1):
    """A torus as a triangle soup (what resources/torus.obj is in the reference's asset folder; generated, not loaded:
    BASELINE config 3 "plus torus.obj triangle-mesh colliders").  Returns (vertices [n, 3] float32, triangles [m, 3] uint32);
    the ring lies in the plane normal to `axis` (1 = y: it lies flat)."""
    u = np.arange(n_major, dtype=np.float64) * (2 * np.pi / n_major)
    v = np.arange(n_minor, dtype=np.float64) * (2 * np.pi / n_minor)
    uu, vv = np.meshgrid(u, v, indexing="ij")
    ring = major + minor * np.cos(vv)
    p = np.stack([ring * np.cos(uu), minor * np.sin(vv), ring * np.sin(uu)], axis=-1)          # y up
    if axis != 1:
        p = np.roll(p, axis - 1, axis=-1)
    verts = (p.reshape(-1, 3) + np.asarray(centre, np.float64)).astype(np.float32)
    i, j = np.meshgrid(np.arange(n_major), np.arange(n_minor), indexing="ij")
    a = (i * n_minor + j).ravel(); b = (((i + 1) % n_major) * n_minor + j).ravel()
    c = (((i + 1) % n_major) * n_minor + (j + 1) % n_minor).ravel(); d = (i * n_minor + (j + 1) % n_minor).ravel()
    tris = np.concatenate([np.stack([a, b, c], axis=1), np.stack([a, c, d], axis=1)]).astype(np.uint32)
    return verts, tris


def height_mesh(side, cells, y=0.0, bump=0.0, seed=0, centre=(0.0, 0.0)):
    """A (cells x cells) grid of triangles over a square of edge `side`: a floor (bump = 0) or a rolling ground."""
    rng = np.random.default_rng(seed)
    g = np.linspace(-0.5 * side, 0.5 * side, cells + 1)
    xx, zz = np.meshgrid(g + centre[0], g + centre[1], indexing="ij")
    yy = y + bump * rng.standard_normal(xx.shape)
    verts = np.stack([xx, yy, zz], axis=-1).reshape(-1, 3).astype(np.float32)
    i, j = np.meshgrid(np.arange(cells), np.arange(cells), indexing="ij")
    a = (i * (cells + 1) + j).ravel(); b = a + 1; c = a + cells + 1; d = c + 1
    tris = np.concatenate([np.stack([a, b, d], axis=1), np.stack([a, d, c], axis=1)]).astype(np.uint32)
    return verts, tris
