"""CPU: the contact half of BipedBehaviors::on_collision (behaviors/biped_behaviors.h:336-366).  The oracle evaluates it inside
its restated pair loop; oracle/_ref evaluates the same three lines INSIDE the reference relay's real on_collision callback
(real CollisionPhysicsRelay, real ColliderPacket with the live transform): which call is the last per entity, which normal /
point each side is handed, and the caller's origin at callback time (after static resolution) are thereby pinned."""
import numpy as np
import pytest

from intercession_b200 import abi, scenes
import refs
from util import fields_diff

needs_ref = pytest.mark.skipif(refs.ref() is None, reason="oracle/_ref not built (no /root/reference here)")


def behaving_scene(n=180, seed=6):
    s = scenes.arena(n, seed=seed, walls=True, temporal=False, fill_side=8.2 if n <= 180 else 2.0 + 1.1 * round(n ** (1 / 3) + 1), spacing=1.0, jitter=0.25, kinds=("sphere", "box", "aabb"))
    c = s.C["colliders"]
    c["useBehaviorsResponse"][6::3, 0] = 1                 # every third body wants its collisions reported
    c["useBehaviorsResponse"][2, 0] = 1                    # and a wall (immutable caller)
    return s


@needs_ref
def test_grounds_oracle_matches_the_reference_callback():
    s = behaving_scene()
    v = refs.verbatim_run(s, n_ticks=1)
    gr = refs.last_grounds(refs.ref(), "icxref")
    o = refs.procedures_tick(refs.oracle(), "icxo", s)
    go = refs.last_grounds(refs.oracle(), "icxo")
    assert len(v.hits) == len(o.hits) > 40
    assert len(gr) == len(go) > 15
    assert not fields_diff(gr, go)
    assert (go["side"] == 1).any() and (go["side"] == 0).any()


@needs_ref
def test_grounds_case_study_legs():
    """The focal entity's ray-spring "legs" (useBehaviorsResponse) on the floor: grounded, normal = +y as handed to the caller."""
    s = scenes.case_study(push=True)
    s.T["origin"][3] = (12, -3.0, 0)
    refs.verbatim_run(s, n_ticks=1)
    gr = refs.last_grounds(refs.ref(), "icxref")
    refs.procedures_tick(refs.oracle(), "icxo", s)
    go = refs.last_grounds(refs.oracle(), "icxo")
    assert len(gr) == 1 and not fields_diff(gr, go)
    assert gr["entity"][0] == s.ids[3] and abs(abs(gr["normal"][0][1]) - 1.0) < 1e-5 and 0.5 < gr["dist"][0] < 2.0
