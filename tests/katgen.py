"""Deterministic inputs for the per-procedure known-answer tests (KATs): pairs of colliders placed
so that roughly half of them intersect, plus the hand-made degenerate cases SURVEY.md 8c lists
(axis-aligned boxes -> NaN cross-product axes, exactly coincident shapes, ray origin inside a box).
Shared by tests/golden/make_golden.py (which freezes the reference's answers) and the live tests."""
import ctypes as C

import numpy as np

from intercession_b200 import abi

BOX, SPHERE, RAY = abi.COLLIDER_BOX, abi.COLLIDER_SPHERE, abi.COLLIDER_RAY
RIGID, SPRING = abi.COLLISION_RIGID, abi.COLLISION_SPRING

kat_in_dt = np.dtype([("tA", abi.transform_dt), ("cA", abi.collider_dt), ("pA", abi.physics_dt),
                      ("tB", abi.transform_dt), ("cB", abi.collider_dt), ("pB", abi.physics_dt), ("flags", "<u4")])
kat_out_dt = np.dtype([("hit", "<i4"), ("point", "<f4", 3), ("normal", "<f4", 3), ("depth", "<f4"),
                       ("minParamA", "<f4"), ("minParamB", "<f4"),
                       # after the response procedure (only when hit and a response exists)
                       ("r_point", "<f4", 3), ("r_normal", "<f4", 3), ("r_depth", "<f4"),
                       ("r_tA", abi.transform_dt), ("r_pA", abi.physics_dt), ("r_tB", abi.transform_dt), ("r_pB", abi.physics_dt)])

TYPE_COMBOS = [(BOX, BOX, 0), (SPHERE, BOX, 0), (BOX, SPHERE, 0), (RAY, BOX, 0), (BOX, RAY, 0),
               (SPHERE, SPHERE, abi.F_SPHERE_SPHERE), (SPHERE, SPHERE, 0), (RAY, SPHERE, 0), (RAY, RAY, 0)]
RESPONSE_COMBOS = [(RIGID, RIGID), (SPRING, RIGID), (RIGID, SPRING), (SPRING, SPRING), (abi.COLLISION_FORCE, RIGID)]


def _quat(rng, n):
    q = rng.standard_normal((n, 4)).astype(np.float32)
    return (q / np.linalg.norm(q, axis=1, keepdims=True)).astype(np.float32)


def random_cases(seed, per_combo=48):
    rng = np.random.default_rng(seed)
    n = per_combo * len(TYPE_COMBOS)
    k = np.zeros(n, kat_in_dt)
    for side in ("A", "B"):
        k["t" + side] = abi.default_transforms(n)
        k["p" + side] = abi.default_physics(n)
        k["c" + side] = abi.default_colliders(n)["colliders"][:, 0]
        k["c" + side]["isActive"] = 1
    combo = np.arange(n) % len(TYPE_COMBOS)
    k["cA"]["colliderType"] = np.array([TYPE_COMBOS[c][0] for c in combo])
    k["cB"]["colliderType"] = np.array([TYPE_COMBOS[c][1] for c in combo])
    k["flags"] = np.array([TYPE_COMBOS[c][2] for c in combo])
    rc = (np.arange(n) // len(TYPE_COMBOS)) % len(RESPONSE_COMBOS)
    k["cA"]["collisionType"] = np.array([RESPONSE_COMBOS[c][0] for c in rc])
    k["cB"]["collisionType"] = np.array([RESPONSE_COMBOS[c][1] for c in rc])
    k["tA"]["origin"] = rng.uniform(-4, 4, (n, 3)).astype(np.float32)
    k["tA"]["orientation"] = _quat(rng, n)
    k["tB"]["orientation"] = _quat(rng, n)
    k["tA"]["scale"] = rng.uniform(0.4, 2.0, (n, 3)).astype(np.float32)
    k["tB"]["scale"] = rng.uniform(0.4, 2.0, (n, 3)).astype(np.float32)
    d = rng.standard_normal((n, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True).astype(np.float32)
    k["tB"]["origin"] = k["tA"]["origin"] + d * rng.uniform(0.1, 1.9, (n, 1)).astype(np.float32)
    # every fourth case: non-trivial local transforms, orientation flags, materials
    m = np.arange(n) % 4 == 3
    for side in ("A", "B"):
        c = k["c" + side]
        c["localTransform"]["origin"][m] = rng.uniform(-0.3, 0.3, (m.sum(), 3)).astype(np.float32)
        c["localTransform"]["orientation"][m] = _quat(rng, m.sum())
        c["localTransform"]["scale"][m] = rng.uniform(0.5, 1.5, (m.sum(), 3)).astype(np.float32)
        c["inheritOrientation"][m] = rng.integers(0, 2, m.sum())
        c["influenceOrientation"][m] = rng.integers(0, 2, m.sum())
        c["staticFriction"] = rng.uniform(0, 1, n).astype(np.float32)
        c["dynamicFriction"] = rng.uniform(0, 1, n).astype(np.float32)
        c["restitution"] = rng.uniform(0, 1, n).astype(np.float32)
        c["stiffness"] = rng.uniform(50, 3000, n).astype(np.float32)
        c["damping"] = rng.uniform(1, 400, n).astype(np.float32)
        c["restLength"] = rng.uniform(0.0, 0.5, n).astype(np.float32)
        p = k["p" + side]
        p["velocity"] = rng.uniform(-3, 3, (n, 3)).astype(np.float32)
        p["angularVelocity"] = rng.uniform(-2, 2, (n, 3)).astype(np.float32)
        p["acceleration"] = rng.uniform(-1, 1, (n, 3)).astype(np.float32)
        p["angularAcceleration"] = rng.uniform(-1, 1, (n, 3)).astype(np.float32)
        p["mass"] = rng.uniform(0.5, 50, n).astype(np.float32)
    # infinite masses (mass 0): one side, other side, both
    k["pA"]["mass"][np.arange(n) % 11 == 5] = 0
    k["pB"]["mass"][np.arange(n) % 13 == 7] = 0
    # long rays so that ray/box cases do hit
    for side in ("A", "B"):
        r = k["c" + side]["colliderType"] == RAY
        k["c" + side]["localTransform"]["scale"][r, 2] = 3.0
    return k


def degenerate_cases():
    out = []

    def case(tyA, tyB, oA, oB, sA=(1, 1, 1), sB=(1, 1, 1), qA=(1, 0, 0, 0), qB=(1, 0, 0, 0), flags=0, cA=RIGID, cB=RIGID):
        k = np.zeros(1, kat_in_dt)
        for side, ty, o, s, q, ct in (("A", tyA, oA, sA, qA, cA), ("B", tyB, oB, sB, qB, cB)):
            k["t" + side] = abi.default_transforms(1)
            k["c" + side] = abi.default_colliders(1)["colliders"][:, 0]
            k["p" + side] = abi.default_physics(1)
            k["c" + side]["isActive"] = 1
            k["c" + side]["colliderType"] = ty
            k["c" + side]["collisionType"] = ct
            k["t" + side]["origin"] = o
            k["t" + side]["scale"] = s
            k["t" + side]["orientation"] = q
        k["pA"]["velocity"] = (0.5, -1.0, 0.25)
        k["pB"]["velocity"] = (-0.5, 1.0, 0.0)
        k["flags"] = flags
        out.append(k)

    h = np.float32(np.sqrt(0.5))
    case(BOX, BOX, (0, 0, 0), (0.9, 0, 0))                              # aligned boxes, face contact: NaN cross axes
    case(BOX, BOX, (0, 0, 0), (0.9, 0.9, 0.9))                          # aligned, corner overlap
    case(BOX, BOX, (0, 0, 0), (0, 0, 0))                                # exactly coincident
    case(BOX, BOX, (0, 0, 0), (0, 0.95, 0), sB=(4, 1, 4))               # small box on a slab
    case(BOX, BOX, (-5, 0, 0), (0, -5, 0), sA=(2, 12, 12), sB=(12, 2, 12))   # two walls meeting in an edge
    case(BOX, BOX, (0, 0, 0), (0, 1.2, 0), qB=(h, 0, 0, h))             # rotated 90 deg about z: edge-aligned
    case(BOX, BOX, (0, 0, 0), (3, 0, 0))                                # clearly apart
    case(SPHERE, BOX, (0, 0, 0), (0, 0, 0))                             # sphere centre == box centre
    case(SPHERE, BOX, (0, 0.9, 0), (0, 0, 0), sB=(4, 1, 4))             # sphere resting on a slab
    case(BOX, SPHERE, (0, 0, 0), (0, 0.9, 0), sA=(4, 1, 4))
    case(SPHERE, SPHERE, (0, 0, 0), (0, 0, 0), flags=abi.F_SPHERE_SPHERE)     # exactly overlapping spheres
    case(SPHERE, SPHERE, (0, 0, 0), (0.5, 0, 0), flags=abi.F_SPHERE_SPHERE)
    case(SPHERE, SPHERE, (0, 0, 0), (0.5, 0, 0))                        # unwired: no hit
    case(RAY, BOX, (0, 0, 0), (0, 0, 0), cA=SPRING)                     # ray origin inside the box
    case(RAY, BOX, (0, 0, -1), (0, 0, 0.2), cA=SPRING)                  # ray entering a face head-on
    case(BOX, RAY, (0, 0, 0.2), (0, 0, -1), cB=SPRING)
    case(RAY, BOX, (0, 0, -3), (0, 0, 0))                               # too short to reach
    case(RAY, BOX, (0, 2, 0), (0, 0, 0), qA=(h, h, 0, 0), sA=(1, 1, 1.8), cA=SPRING)   # biped-leg-like, pointing down
    return _stack(out)


def _stack(parts):
    """np.concatenate would canonicalise (re-pack) the padded nested struct dtypes; copy by slices."""
    k = np.zeros(sum(len(p) for p in parts), kat_in_dt)
    at = 0
    for p in parts:
        k[at:at + len(p)] = p
        at += len(p)
    return k


def all_cases():
    return _stack([random_cases(20261104), degenerate_cases()])


def run_kat(lib, prefix, cases):
    """Feed every case to `<prefix>_intersect` and, when it hits, `<prefix>_response`."""
    out = np.zeros(len(cases), kat_out_dt)
    f3 = lambda: (C.c_float * 3)()
    for i in range(len(cases)):
        k = np.zeros(1, kat_in_dt)
        k[0] = cases[i]
        pt, nm = np.zeros(3, np.float32), np.zeros(3, np.float32)
        dp, ma, mb = C.c_float(0), C.c_float(0), C.c_float(0)
        hit = getattr(lib, prefix + "_intersect")(abi.ptr(k["tA"]), abi.ptr(k["cA"]), abi.ptr(k["tB"]), abi.ptr(k["cB"]),
                                                  int(k["flags"][0]), abi.ptr(pt), abi.ptr(nm), C.byref(dp), C.byref(ma), C.byref(mb))
        o = out[i]
        o["hit"], o["point"], o["normal"], o["depth"] = hit, pt, nm, dp.value
        o["minParamA"], o["minParamB"] = ma.value, mb.value
        o["r_tA"], o["r_pA"], o["r_tB"], o["r_pB"] = k["tA"][0], k["pA"][0], k["tB"][0], k["pB"][0]
        if hit == 1:
            tA, pA, cA = k["tA"].copy(), k["pA"].copy(), k["cA"].copy()
            tB, pB, cB = k["tB"].copy(), k["pB"].copy(), k["cB"].copy()
            d = C.c_float(dp.value)
            getattr(lib, prefix + "_response")(abi.ptr(tA), abi.ptr(pA), abi.ptr(cA), abi.ptr(tB), abi.ptr(pB), abi.ptr(cB),
                                               abi.ptr(pt), abi.ptr(nm), C.byref(d))
            o["r_point"], o["r_normal"], o["r_depth"] = pt, nm, d.value
            o["r_tA"], o["r_pA"], o["r_tB"], o["r_pB"] = tA[0], pA[0], tB[0], pB[0]
        out[i] = o
    return out


REF_CRASHED = -2


def run_kat_isolated(lib, prefix, cases):
    """run_kat with every case in a forked child: the reference itself dies (SIGFPE, modulo by a
    zero manifold size at collision_procedures.h:243) when pseudo_clip_polyhedra clips the clippee
    away entirely.  Such cases are recorded with hit = REF_CRASHED and carry no known answer."""
    import os
    out = np.zeros(len(cases), kat_out_dt)
    for i in range(len(cases)):
        r, w = os.pipe()
        pid = os.fork()
        if pid == 0:
            try:
                os.close(r)
                os.write(w, run_kat(lib, prefix, cases[i:i + 1]).tobytes())
            finally:
                os._exit(0)
        os.close(w)
        buf = b""
        while True:
            chunk = os.read(r, 1 << 16)
            if not chunk:
                break
            buf += chunk
        os.close(r)
        _, status = os.waitpid(pid, 0)
        if os.WIFSIGNALED(status) or len(buf) != kat_out_dt.itemsize:
            out[i]["hit"] = REF_CRASHED
        else:
            out[i] = np.frombuffer(buf, kat_out_dt)[0]
    return out
