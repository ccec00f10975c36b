"""Seeded ProjectileComponent cases (behaviors/projectile_component.h) for the projectile-behaviours checks."""
import numpy as np

from intercession_b200 import abi


def projectile_cases(n, seed=11):
    """n components: mostly plausible projectiles, plus the defaults, expiring lifetimes, huge sine arguments
    (argument reduction), zero / negative frequencies, inverted scale ranges and non-finite values."""
    rng = np.random.default_rng(seed)
    p = np.zeros(n, abi.projectile_dt)
    p["currLifetime"] = rng.uniform(0.0, 300.0, n)
    p["maxLifetime"] = 300.0
    p["hz"] = rng.uniform(0.1, 20.0, n)
    p["scaleMin"] = rng.uniform(0.2, 1.0, n)
    p["scaleMax"] = p["scaleMin"] + rng.uniform(0.0, 2.0, n)
    k = np.arange(n)
    d = k % 16 == 1                                   # the struct's defaults: scale pulses between 2 and 2
    p["currLifetime"][d] = 0.0; p["hz"][d] = 1.0; p["scaleMax"][d] = 1.0; p["scaleMin"][d] = 1.0
    e = k % 16 == 2                                   # about to expire: straddles maxLifetime within a few steps
    p["maxLifetime"][e] = p["currLifetime"][e] + rng.uniform(-0.01, 0.08, int(e.sum()))
    h = k % 16 == 3                                   # huge arguments
    p["hz"][h] = 10.0 ** rng.uniform(3, 18, int(h.sum()))
    z = k % 16 == 4
    p["hz"][z] = np.where(rng.random(int(z.sum())) < 0.5, 0.0, -p["hz"][z])
    i = k % 16 == 5                                   # inverted range (negative amplitude)
    p["scaleMax"][i] = p["scaleMin"][i] - rng.uniform(0.0, 1.0, int(i.sum()))
    if n >= 64:
        p["hz"][6] = np.inf; p["hz"][22] = np.nan; p["currLifetime"][38] = np.nan; p["scaleMin"][54] = -np.inf
        p["maxLifetime"][7] = -1.0; p["maxLifetime"][23] = np.nan
    return p
