"""CPU suite: ProjectileBehaviors::on_fixed_update (behaviors/projectile_behaviors.h:17-54).  The oracle's
restatement against the reference's own class (compiled unmodified into oracle/_ref and driven through a real
Cosmos) and against the golden fixture frozen from it (tests/golden/make_golden_projectiles.py)."""
import hashlib
import os

import numpy as np
import pytest

from intercession_b200 import abi
import projgen
import refs

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "projectiles.npz")
N, SEED, DT = 4096, 11, 1.0 / 60.0
needs_ref = pytest.mark.skipif(refs.ref() is None, reason="oracle/_ref not built (no /root/reference here)")


def same_bits(a, b):
    return np.array_equal(np.ascontiguousarray(a).view(np.uint8), np.ascontiguousarray(b).view(np.uint8))


def test_oracle_matches_golden(built):
    g = np.load(GOLD)
    p = projgen.projectile_cases(N, SEED)
    assert np.array_equal(np.frombuffer(hashlib.sha256(p.tobytes()).digest(), np.uint8), g["input_sha256"]), "case generator drifted"
    T = abi.default_transforms(N)
    for steps in (1, 6):
        scale, comp, dead = refs.projectile_update(refs.oracle(), "icxo", T, p, DT, steps)
        assert same_bits(scale[:, 0], g[f"scale_{steps}"]) and same_bits(scale[:, 1], g[f"scale_{steps}"]) and same_bits(scale[:, 2], g[f"scale_{steps}"])
        assert same_bits(comp["currLifetime"], g[f"life_{steps}"])
        assert np.array_equal(dead, g[f"condemned_{steps}"])
        assert dead.sum() > 0 and np.isnan(scale[:, 0]).sum() > 0          # the fixture exercises both


@needs_ref
def test_oracle_matches_reference_class(built):
    for seed, n, steps, dt in ((1, 1000, 1, 1.0 / 60.0), (2, 333, 9, 1.0 / 45.0), (3, 64, 40, 0.25), (4, 1, 1, 0.0)):
        p = projgen.projectile_cases(n, seed)
        T = abi.default_transforms(n)
        r = refs.projectile_update(refs.ref(), "icxref", T, p, dt, steps)
        o = refs.projectile_update(refs.oracle(), "icxo", T, p, dt, steps)
        assert same_bits(r[0], o[0]) and same_bits(r[1], o[1]) and np.array_equal(r[2], o[2])


def test_defaults_pulse_around_two(built):
    """The component's defaults (hz 1, scaleMax = scaleMin = 1) give scale = 0 * sin + 2 exactly."""
    p = np.zeros(4, abi.projectile_dt)
    p["maxLifetime"], p["hz"], p["scaleMax"], p["scaleMin"] = 300.0, 1.0, 1.0, 1.0
    scale, comp, dead = refs.projectile_update(refs.oracle(), "icxo", abi.default_transforms(4), p, DT, 3)
    assert (scale == 2.0).all() and not dead.any() and same_bits(comp["currLifetime"], np.full(4, DT + DT + DT))
