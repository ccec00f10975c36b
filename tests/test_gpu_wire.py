"""GPU: the device-packed ENTITY_UPDATE messages (icx_entity_update_messages) are byte-identical to the
messages the reference's own Cosmos::serialize_entity_components + EventMessage << ENTITY_UPDATE_params produce
(oracle/_ref: the real classes) for the same entities -- header, component stack in descending type order, params with the
chain-link incremented -- after the device has ticked, i.e. from device state, not from the uploaded arrays."""
import numpy as np
import pytest

from intercession_b200 import abi, scenes, world
import refs

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("hard_config", [True, False])
def test_entity_update_messages_match_the_reference_bytes(built, hard_config):
    if refs.ref() is None:
        pytest.skip("oracle/_ref not built")
    s = scenes.arena(150, seed=4, walls=True, temporal=True, timeslice=3, fill_side=9.0, spacing=1.0, jitter=0.2)
    s.ids[100:120] = (s.ids[100:120] & ~np.uint32(15)) | 14                  # increment refused at link 14
    s.hasP[40:50] = 0                                                        # no PhysicsComponent: shorter message, bit 6 clear
    with world.World(s.n, host_timeslice=3) as w:
        w.upload(s)
        w.step(1 / 45, 3)
        cur = w.download(s)
        got = w.entity_update_messages(77, types=(0, 6, 7) if hard_config else (0, 1, 2))
    exp = refs.ref_entity_update_messages(cur, 77, hard_config_types=hard_config)
    assert len(got) == len(exp) == s.n
    sizes = set()
    for k, (g, e) in enumerate(zip(got, exp)):
        assert g == e, (k, len(g), len(e))
        sizes.add(len(g))
    assert sizes == {12 + 168 + 108 + 40 + 24, 12 + 168 + 40 + 24}
    # the id on the wire is the past's: chain-link + 1
    params = np.frombuffer(got[10][-24:], np.uint16)
    assert params[0] == refs.narrow16(s.ids[10:11])[0] + 1
