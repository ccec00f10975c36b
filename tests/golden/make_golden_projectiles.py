#!/usr/bin/env python
"""Golden fixture of ProjectileBehaviors::on_fixed_update, produced by RUNNING THE REFERENCE'S OWN CLASS
(behaviors/projectile_behaviors.h compiled unmodified into oracle/_ref, driven through a real Cosmos by
oracle/ref/ref_behaviors.cpp).  Run in the build container:  python tests/golden/make_golden_projectiles.py
  projectiles.npz   4096 seeded components (tests/projgen.py): scale, lifetime and condemned flags after 1 and
                    after 6 fixed steps of dt = 1/60."""
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from intercession_b200 import abi  # noqa: E402
import projgen  # noqa: E402
import refs  # noqa: E402

N, SEED, DT = 4096, 11, 1.0 / 60.0


def main():
    lib = refs.ref()
    assert lib is not None, "oracle/_ref not built"
    p = projgen.projectile_cases(N, SEED)
    T = abi.default_transforms(N)
    d = {"input_sha256": np.frombuffer(hashlib.sha256(p.tobytes()).digest(), np.uint8)}
    for steps in (1, 6):
        scale, comp, dead = refs.projectile_update(lib, "icxref", T, p, DT, steps)
        d[f"scale_{steps}"] = scale[:, 0].copy()
        assert (scale[:, 0].view(np.uint32) == scale[:, 1].view(np.uint32)).all() and (scale[:, 0].view(np.uint32) == scale[:, 2].view(np.uint32)).all()
        d[f"life_{steps}"] = comp["currLifetime"].copy()
        d[f"condemned_{steps}"] = dead
        print("steps", steps, "condemned", int(dead.sum()), "nan scales", int(np.isnan(scale[:, 0]).sum()))
    np.savez_compressed(os.path.join(HERE, "projectiles.npz"), **d)
    print("projectiles.npz", os.path.getsize(os.path.join(HERE, "projectiles.npz")), "bytes")


if __name__ == "__main__":
    main()
