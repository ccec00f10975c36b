"""GPU parity: icx_projectiles_* (ProjectileBehaviors::on_fixed_update on the device) through the C ABI against the
oracle and the golden fixture frozen from the reference's own class.  Lifetimes and condemned sets are bit-exact.
The scale goes through sin() in DOUBLE and is rounded to float once: the device's double sin is within 1-2 ulp
(double) of glibc's, so the float can differ only when the double lands within ~2^-29 (relative) of a float
rounding boundary -- tolerance: <= 1 float ulp, and at most 1 in 10 000 values may differ at all."""
import os

import numpy as np
import pytest

from intercession_b200 import abi, scenes, world
import projgen
import refs

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "projectiles.npz")
DT = 1.0 / 60.0


def ulp_diff(a, b):
    """Float32 distance in ulps; NaN == NaN (any payload), otherwise ordered-integer distance."""
    a, b = np.ascontiguousarray(a, np.float32), np.ascontiguousarray(b, np.float32)
    both_nan = np.isnan(a) & np.isnan(b)
    ia, ib = a.view(np.int32).astype(np.int64), b.view(np.int32).astype(np.int64)
    ia = np.where(ia < 0, -(ia & 0x7FFFFFFF), ia); ib = np.where(ib < 0, -(ib & 0x7FFFFFFF), ib)
    d = np.abs(ia - ib)
    d[both_nan] = 0
    d[np.isnan(a) ^ np.isnan(b)] = 1 << 40
    return d


def check_scale(got, exp):
    d = ulp_diff(got, exp)
    assert d.max() <= 1, f"scale differs by {d.max()} ulp"
    assert (d != 0).sum() <= max(1, d.size // 10000), f"{(d != 0).sum()} of {d.size} scales differ"


def test_projectiles_golden(built):
    g = np.load(GOLD)
    n = 4096
    p = projgen.projectile_cases(n, 11)
    s = scenes.arena(n, seed=5, walls=False)
    index = np.arange(n, dtype=np.uint32)[::-1].copy()              # projectile i rides on entity n - 1 - i
    with world.World(max_entities=s.n) as w:
        w.upload(s)
        w.projectiles_upload(index, p)
        for step in range(1, 7):
            w.projectiles_fixed_update(DT)
            if step in (1, 6):
                # the reference destroys a condemned entity at the next Cosmos::update, so its fixture rows stop at the
                # condemning step; entity lifecycle stays with the host ECS here and the device keeps updating the row
                live = g[f"condemned_{step}"] == 0
                comp = w.projectiles_download()
                assert np.array_equal(comp["currLifetime"].view(np.uint64)[live], g[f"life_{step}"].view(np.uint64)[live])
                got = w.download(s)
                sc = got.T["scale"][index]
                for k in range(3):
                    check_scale(sc[live, k], g[f"scale_{step}"][live])
                if step == 1:
                    assert np.array_equal(w.projectiles_condemned(), np.nonzero(g["condemned_1"])[0])
                    assert np.array_equal(comp["currLifetime"].view(np.uint64), g["life_1"].view(np.uint64))
        w.projectiles_fixed_update(0.0)
        dead = w.projectiles_condemned()
        comp = w.projectiles_download()
        assert np.array_equal(dead, np.nonzero(comp["currLifetime"] > comp["maxLifetime"])[0])
        assert set(np.nonzero(g["condemned_6"])[0]) <= set(dead.tolist())


def test_projectiles_first_step_condemned_and_next_tick_sees_scale(built):
    n = 3000
    p = projgen.projectile_cases(n, 21)
    p["hz"] = np.where(np.isfinite(p["hz"]), p["hz"], 1.0); p["currLifetime"] = np.nan_to_num(p["currLifetime"]); p["scaleMin"] = np.clip(np.nan_to_num(p["scaleMin"]), 0.2, 1.0)
    p["scaleMax"] = p["scaleMin"] + 0.3                              # scale = sin * 0.15 + scaleMin + 1 <= 2.15
    s = scenes.arena(n, seed=8, spacing=2.5, jitter=0.2, fill_side=40.0)
    rng = np.random.default_rng(3)
    index = (6 + rng.permutation(n)[:n - 100]).astype(np.uint32)          # not the 6 walls; 100 bodies carry no projectile
    p = p[:n - 100]; n = n - 100
    exp_scale, exp_comp, exp_dead = refs.projectile_update(refs.oracle(), "icxo", abi.default_transforms(n), p, DT, 1)
    with world.World(max_entities=s.n) as w:
        w.upload(s)
        w.projectiles_upload(index, p)
        w.projectiles_fixed_update(DT)
        assert np.array_equal(w.projectiles_condemned(), np.nonzero(exp_dead)[0])
        assert np.array_equal(w.projectiles_download()["currLifetime"].view(np.uint64), exp_comp["currLifetime"].view(np.uint64))
        got = w.download(s)
        check_scale(got.T["scale"][index], exp_scale)
        untouched = np.setdiff1d(np.arange(s.n), index)
        assert np.array_equal(got.T["scale"][untouched], s.T["scale"][untouched])
        # the tick after the update must use the new scales: same result as uploading them from the host
        s2 = s.copy(); s2.T["scale"][:] = got.T["scale"]
        w.step(1 / 45, 1)
        a = w.download(s)
        with world.World(max_entities=s.n) as w2:
            w2.upload(s2)
            w2.step(1 / 45, 1)
            b = w2.download(s2)
        for comp, x, y in (("T", a.T, b.T), ("P", a.P, b.P)):
            for f in x.dtype.names:
                bad = np.nonzero((x[f].reshape(len(x), -1).view(np.uint8) != y[f].reshape(len(y), -1).view(np.uint8)).any(axis=1))[0]
                assert len(bad) == 0, f"{comp}.{f}: {len(bad)} rows differ, first {bad[:5]}: {x[f][bad[:2]]} vs {y[f][bad[:2]]}"


def test_projectiles_errors(built):
    s = scenes.arena(50, seed=1, walls=False)
    with world.World(max_entities=s.n) as w:
        w.upload(s)
        with pytest.raises(world.IcxError):
            w.projectiles_upload(np.array([s.n], np.uint32), projgen.projectile_cases(1, 1))
        w.projectiles_upload(np.arange(10, dtype=np.uint32), projgen.projectile_cases(10, 1))
        with pytest.raises(world.IcxError):
            w.projectiles_fixed_update(-1.0)
        w.projectiles_fixed_update(DT)
        w.upload(s)                                                  # a new upload drops the projectile rows
        w.projectiles_fixed_update(DT)
        assert len(w.projectiles_condemned()) == 0
