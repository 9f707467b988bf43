"""gs_residual_static (include/gpusolve_cuda.h; no counterpart in the reference):
the residual + norm with its reduction scratch in a static __device__ array, so that nothing on the call path allocates.
Checked through the C ABI against the existing entry point gs_residual on two calls in a row with different iterates (the
array keeps the last call's contents: a stale ticket, result or partial sum would show in the second call), for all three
modes, a ragged grid and a coarse level; and the size limit is refused with a message.  The file name makes pytest collect
it last."""
import numpy as np
import pytest

from cases import seeded_field

pytestmark = pytest.mark.gpu

V, F, R, NEWTONV = 0, 1, 2, 5
NORM_TOL = 1e-12    # the tolerance of tests/test_parity_gpu.py: only the summation tree differs between the two kernels
FIELD_TOL = 1e-12


def _rel(a, b):
    s = float(np.max(np.abs(b)))
    d = float(np.max(np.abs(a - b)))
    return d / s if s > 0 else d


@pytest.mark.parametrize("mode,n,level", [(0, 63, 0), (0, 127, 0), (0, (70, 45, 33), 0), (0, 127, 1), (1, 63, 0), (2, 63, 0), (0, 255, 0)])
def test_static_scratch_residual_equals_gs_residual(gs, mode, n, level):
    p = gs.make_params(n, mode=mode, pre=2, post=2, maxiter=1, tol=1e-30)
    s = gs.GpuSolve(p, device=0)
    try:
        shape = s.shape(level)
        if mode == 2:
            s.set(level, NEWTONV, seeded_field(shape, 7))
        s.set(level, F, seeded_field(shape, 8, scale=1.0))
        for call, seed in enumerate((21, 22)):          # two calls in a row, different inputs
            s.set(level, V, seeded_field(shape, seed))
            want = s.residual(level)                    # existing entry point (tuned kernel, cudaMalloc'ed scratch)
            r_want = s.get(level, R)
            s.set(level, R, np.zeros(shape))            # so that the new entry point must write r itself
            got = s.residual_static(level)
            r_got = s.get(level, R)
            assert want > 0.0
            assert abs(got - want) <= NORM_TOL * want, f"call {call}: norm {got!r} vs {want!r}"
            if mode == 0:
                assert np.array_equal(r_got + 0.0, r_want + 0.0), f"call {call}: r differs (max rel {_rel(r_got, r_want):.3e})"
            else:
                assert _rel(r_got, r_want) <= FIELD_TOL, f"call {call}: r rel err {_rel(r_got, r_want):.3e}"
        # and the existing entry point is untouched by the static one
        assert s.residual(level) == want
    finally:
        s.close()


def test_static_scratch_is_shared_by_handles_used_in_turn(gs):
    """two handles of different sizes calling alternately (serialised by the synchronous ABI): each call must see a clean array"""
    a = gs.GpuSolve(gs.make_params(95, mode=0, pre=2, post=2, maxiter=1, tol=1e-30), device=0)
    b = gs.GpuSolve(gs.make_params(31, mode=0, pre=2, post=2, maxiter=1, tol=1e-30), device=0)
    try:
        for h, seed in ((a, 3), (b, 4)):
            h.set(0, V, seeded_field(h.shape(0), seed))
        wa, wb = a.residual(0), b.residual(0)
        for _ in range(3):
            assert abs(a.residual_static(0) - wa) <= NORM_TOL * wa
            assert abs(b.residual_static(0) - wb) <= NORM_TOL * wb
    finally:
        a.close()
        b.close()


def test_static_scratch_limit_is_refused_with_a_message(gs):
    """300^3 needs 5 * 75 * 300 = 112500 reduction slots, the array holds 65536"""
    s = gs.GpuSolve(gs.make_params(300, mode=0, pre=2, post=2, maxiter=1, tol=1e-30), device=0)
    try:
        with pytest.raises(gs.GpuSolveError, match="static scratch holds 65536"):
            s.residual_static(0)
        want = s.residual(1)                             # the next level (150^3: 3 * 38 * 150 slots) fits
        assert abs(s.residual_static(1) - want) <= NORM_TOL * want
    finally:
        s.close()
