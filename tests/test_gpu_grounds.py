"""GPU: the contact half of BipedBehaviors::on_collision on the device (icx_grounds_fetch) against the oracle,
which tests/test_grounds_oracle.py pins to the reference relay's real callback: per entity the LAST call in pair order --
normal as the caller is handed it, groundDist from the caller's origin after static resolution -- bit-exact."""
import numpy as np
import pytest

from intercession_b200 import scenes, world
import refs
from util import fields_diff
from test_grounds_oracle import behaving_scene

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n,seed", [(180, 6), (3000, 2)])
def test_grounds_match_oracle(built, n, seed):
    s = behaving_scene(n, seed) if n < 1000 else scenes.arena(n, seed=seed, fill_side=2.0 + 1.1 * round(n ** (1 / 3) + 1), spacing=1.05, jitter=0.25)
    if n >= 1000:
        s.C["colliders"]["useBehaviorsResponse"][6::3, 0] = 1
    with world.World(s.n) as w:
        w.upload(s)
        w.collision_engage()
        got, hits = w.grounds(), w.contacts()
    exp = refs.procedures_tick(refs.oracle(), "icxo", s, euler=False)
    ge = refs.last_grounds(refs.oracle(), "icxo")
    assert len(hits) == len(exp.hits)
    assert len(got) == len(ge) > 10 and not fields_diff(got, ge)


def test_grounds_case_study_lockstep(built):
    """The focal entity's legs over 60 ticks of the case-study scene, device re-seeded from the oracle every tick: grounded
    whenever the oracle says so, same numbers."""
    cur = scenes.case_study(push=True)
    grounded = 0
    with world.World(cur.n) as w:
        for t in range(60):
            w.upload(cur)
            w.set_coherency(t)
            w.step(1 / 45)
            got = w.grounds()
            cur = refs.procedures_tick(refs.oracle(), "icxo", cur, coherency=t)
            ge = refs.last_grounds(refs.oracle(), "icxo")
            assert len(got) == len(ge), (t, got, ge, w.contacts(), cur.hits)
            if len(got):
                assert np.array_equal(got["entity"], ge["entity"]) and np.abs(got["dist"] - ge["dist"]).max() < 1e-5 and np.abs(got["normal"] - ge["normal"]).max() < 1e-6, t
                grounded += 1
    assert grounded > 5
