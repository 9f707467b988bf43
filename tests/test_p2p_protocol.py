"""The ack / signal handshake of the peer-mapped ghost-plane exchange under random schedules (tests/p2p_model.py)."""
import pytest

import p2p_model as pm

# ghost buffers refreshed by consecutive operators of a V-cycle on one level: ping-pong smoother passes alternate between the
# two v buffers; the residual+restriction and the in-place prolongation+correction refresh the same one twice in a row
SCHEDULES = {
    "pingpong": ["A", "B", "A", "B", "A", "B"],
    "in_place": ["A", "A", "A", "B", "B", "A"],
    "vcycle": ["B", "f1", "B1", "f2", "B", "B", "A", "A"],
}


@pytest.mark.parametrize("world", [2, 3, 4, 8])
@pytest.mark.parametrize("name", sorted(SCHEDULES))
def test_handshake_is_safe_and_live(world, name):
    for seed in range(40):
        pm.run(world, SCHEDULES[name], seed)


def test_model_catches_the_missing_ack():
    """without the ack a fast rank refreshes a buffer its neighbour is still reading whenever two consecutive exchanges
    target the same buffer"""
    with pytest.raises(pm.Hazard, match="overwrites|reads"):
        for seed in range(200):
            pm.run(3, SCHEDULES["in_place"], seed, use_ack=False)
