"""The ring protocols of the tuned kernels under random warp schedules (tests/ring_model.py): no stale or destroyed read,
no deadlock — and the model does catch the two mistakes made while writing them."""
import pytest

import ring_model as rm


@pytest.mark.parametrize("nzo", [1, 2, 3, 5, 6, 7, 13, 40])
@pytest.mark.parametrize("vz", [False, True])
def test_smooth2u_protocol(nzo, vz):
    for seed in range(25):
        rm.smooth2u(nzo, nwarps=4, seed=seed, vz=vz)


@pytest.mark.parametrize("nzc", [1, 2, 3, 4, 9, 32])
@pytest.mark.parametrize("ns", [5, 6])
def test_resid_restrict_protocol(nzc, ns):
    for seed in range(25):
        rm.resid_restrict(nzc, nwarps=4, seed=seed, ns=ns)


@pytest.mark.parametrize("nz", [1, 2, 5, 64])
def test_producer_ring_protocol(nz):
    for seed in range(25):
        rm.producer_ring(nz, nwarps=4, seed=seed)


def test_model_catches_a_ring_that_is_too_short():
    """thread-0 issue with four stages: the first coarse plane needs v plane 4 before anything was released (the
    deadlock hit on the GPU while writing the kernel; RRCfg now static_asserts five stages)"""
    with pytest.raises(rm.Hazard, match="deadlock"):
        rm.resid_restrict(4, ns=4)


def test_model_catches_the_zero_input_parity_slip():
    """smooth2u VZ without the plain arrival for the f-less plane 0: stage 0's full barrier is one phase behind and the
    wait for plane 6 never passes (the hang seen on the GPU before the fix)"""
    with pytest.raises(rm.Hazard, match="deadlock"):
        rm.smooth2u(7, vz=True, plain_arrive=False)


def test_model_catches_a_refill_that_runs_ahead():
    """thread 0 issuing one plane further than the barrier covers destroys a stage a slower warp still reads"""
    hits = 0
    for seed in range(200):
        try:
            rm.resid_restrict(9, seed=seed, ns=5, refill_ahead=1)
        except rm.Hazard:
            hits += 1
    assert hits > 0

@pytest.mark.parametrize("nzo", [1, 2, 5, 6, 7, 13, 40])
def test_prolongation_folded_into_the_smoother(nzo):
    """smooth2u_tma<PC> (kernels_fused.cu): correcting plane q+2 in place at step q needs no new barrier ..."""
    for seed in range(25):
        rm.smooth2_prolong(nzo, seed=seed)


def test_correcting_one_plane_ahead_is_a_race():
    """... whereas correcting plane q+1 at step q lets a fast warp read a neighbour's point before it was corrected"""
    hits = 0
    for seed in range(100):
        try:
            rm.smooth2_prolong(13, seed=seed, correct_ahead=1)
        except rm.Hazard as e:
            assert "corrected" in str(e) or "stage" in str(e)
            hits += 1
    assert hits > 0
