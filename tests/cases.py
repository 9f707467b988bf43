"""Shared definitions for the parity tests: the case matrix, seeded inputs, checksums."""
import hashlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

CANON_VALUES = (6.0, -1.0, -1.0, -1.0, -1.0, -1.0, -1.0)
CANON_OFFSETS = ((0, 0, 0), (1, 0, 0), (-1, 0, 0), (0, 1, 0), (0, -1, 0), (0, 0, 1), (0, 0, -1))
# a permuted, non-symmetric 7-point stencil: exercises config-driven offsets and summation order
SKEW_VALUES = (6.5, -1.1, -0.9, -1.2, -0.8, -1.3, -0.7)
SKEW_OFFSETS = ((0, 0, 0), (0, 0, 1), (1, 0, 0), (0, -1, 0), (-1, 0, 0), (0, 1, 0), (0, 0, -1))
# the same with a zero row sum (a consistent second-order operator plus first-order terms): multigrid still converges
NC_VALUES = (6.0, -1.1, -0.9, -1.2, -0.8, -1.3, -0.7)
# offsets off the axes (still inside the one-cell halo)
DIAG_OFFSETS = ((0, 0, 0), (1, 1, 0), (-1, -1, 0), (0, 1, 1), (0, -1, -1), (1, 0, 1), (-1, 0, -1))

# name -> kwargs of make_params; sizes the CPU oracle finishes in seconds
SOLVE_CASES = {
    "lin31_v22": dict(n=31, mode=0, pre=2, post=2, maxiter=4, tol=1e-30),
    "lin63_v33": dict(n=63, mode=0, pre=3, post=3, maxiter=3, tol=1e-30),
    "lin_ragged": dict(n=(9, 12, 17), mode=0, pre=1, post=3, maxiter=3, tol=1e-30),
    "lin_even18": dict(n=18, mode=0, pre=2, post=1, maxiter=3, tol=1e-30),
    "lin_tiny1": dict(n=1, mode=0, pre=2, post=2, maxiter=2, tol=1e-30),
    "lin_tiny3": dict(n=(3, 5, 4), mode=0, pre=2, post=2, maxiter=2, tol=1e-30),
    "lin_pre0": dict(n=15, mode=0, pre=0, post=2, maxiter=2, tol=1e-30),
    "lin_post0": dict(n=15, mode=0, pre=3, post=0, maxiter=2, tol=1e-30),
    "lin_skew31": dict(n=31, mode=0, pre=2, post=2, maxiter=3, tol=1e-30, values=SKEW_VALUES, offsets=SKEW_OFFSETS),
    "lin_diag15": dict(n=15, mode=0, pre=2, post=2, maxiter=2, tol=1e-30, values=SKEW_VALUES, offsets=DIAG_OFFSETS),
    "lin_tol": dict(n=31, mode=0, pre=2, post=2, maxiter=10, tol=1e-3),
    "nl31_v22": dict(n=31, mode=1, pre=2, post=2, maxiter=4, tol=1e-30),
    "nl_ragged": dict(n=(9, 12, 17), mode=1, pre=1, post=3, maxiter=3, tol=1e-30),
    "nl_even18": dict(n=18, mode=1, pre=2, post=1, maxiter=3, tol=1e-30),
    "nl_gamma": dict(n=15, mode=1, pre=2, post=2, maxiter=3, tol=1e-30, gamma=2.5, omega=0.7),
    "nt31_v22": dict(n=31, mode=2, pre=2, post=2, maxiter=3, tol=1e-30),
    "nt_ragged": dict(n=(9, 12, 17), mode=2, pre=1, post=3, maxiter=3, tol=1e-30),
    "nt_even18": dict(n=18, mode=2, pre=2, post=1, maxiter=3, tol=1e-30),
    "nt_tol": dict(n=31, mode=2, pre=3, post=3, maxiter=10, tol=1e-5),
}

# larger cases: too slow for every CPU run, used by the GPU suite and golden norms only
BIG_CASES = {
    "lin127_v33": dict(n=127, mode=0, pre=3, post=3, maxiter=2, tol=1e-30),
    "nl127_v22": dict(n=127, mode=1, pre=2, post=2, maxiter=1, tol=1e-30),
    "shipped127": dict(n=127, mode=2, pre=3, post=3, maxiter=10, tol=1e-5),  # examples/data-2nd_order.conf
    # two-sweeps-per-pass smoother (kernels_fused.cu) on the finest levels: even sweep counts, partial tiles
    "lin127_v22": dict(n=127, mode=0, pre=2, post=2, maxiter=3, tol=1e-30),
    "lin_wide_v44": dict(n=(131, 97, 63), mode=0, pre=4, post=4, maxiter=2, tol=1e-30),
    "nl_wide_v42": dict(n=(121, 64, 70), mode=1, pre=4, post=2, maxiter=2, tol=1e-30),
    "nt_wide_v24": dict(n=(100, 90, 47), mode=2, pre=2, post=4, maxiter=2, tol=1e-30),
    # config-driven (non-canonical) offsets on grids large enough for the tuned GEN kernels (kernels_fast.cu stencil_tma<.., GEN>)
    "lin_skew127_v22": dict(n=127, mode=0, pre=2, post=2, maxiter=2, tol=1e-30, values=NC_VALUES, offsets=SKEW_OFFSETS),
    "lin_diag_wide_v31": dict(n=(131, 97, 63), mode=0, pre=3, post=1, maxiter=2, tol=1e-30, values=NC_VALUES, offsets=DIAG_OFFSETS),
    "nl_diag100_v22": dict(n=(100, 90, 47), mode=1, pre=2, post=2, maxiter=1, tol=1e-30, values=NC_VALUES, offsets=DIAG_OFFSETS),
    "nt_skew100_v22": dict(n=(100, 90, 47), mode=2, pre=2, post=2, maxiter=1, tol=1e-30, values=NC_VALUES, offsets=SKEW_OFFSETS),
}

# BASELINE.json configs[1]-[3] at their full size (arrays 513^3): one script per mode (backends.full_size_script), run
# by the reference here (tests/golden/make_golden.py -> norms, digest, strided sample of the field) and by the C port
# next to the GPU in the -m gpu suite (full-field comparison)
FULL_CASES = {
    "full511_lin_v22": dict(n=511, mode=0, pre=2, post=2, maxiter=2, tol=1e-30),
    "full511_nl_v22": dict(n=511, mode=1, pre=2, post=2, maxiter=1, tol=1e-30),
    "full511_nt_v33": dict(n=511, mode=2, pre=3, post=3, maxiter=1, tol=1e-30),
}
FULL_SAMPLE_STRIDE = 16  # field[::16, ::16, ::16] of the 513^3 result goes into fields.npz (33^3 values)

OPERATOR_CASES = {
    "op_lin17": dict(n=17, mode=0),
    "op_lin_ragged": dict(n=(10, 7, 13), mode=0),
    "op_lin_skew": dict(n=(12, 9, 8), mode=0, values=SKEW_VALUES, offsets=SKEW_OFFSETS),
    "op_lin_diag": dict(n=9, mode=0, values=SKEW_VALUES, offsets=DIAG_OFFSETS),
    "op_nl17": dict(n=17, mode=1),
    "op_nl_ragged": dict(n=(10, 7, 13), mode=1),
    "op_nt17": dict(n=17, mode=2),
    "op_nt_ragged": dict(n=(10, 7, 13), mode=2),
    # large enough for the tuned kernels (TMA z-march smoother/residual, fused residual+restriction,
    # paired prolongation); odd, even and ragged extents exercise their edge masks
    "op_lin_mid": dict(n=(70, 53, 41), mode=0),
    "op_lin_mid_even": dict(n=(64, 66, 34), mode=0),
    "op_nl_mid": dict(n=(67, 50, 38), mode=1),
    "op_nt_mid": dict(n=(63, 49, 35), mode=2),
    # wide enough for the two-sweeps-per-pass smoother: several 60-column tiles with a partial last one,
    # partial 14-row tiles, more than one z chunk
    "op_lin_wide": dict(n=(130, 61, 40), mode=0),
    "op_lin_wide_even": dict(n=(120, 56, 34), mode=0),
    "op_nl_wide": dict(n=(125, 40, 33), mode=1),
    "op_nt_wide": dict(n=(100, 35, 20), mode=2),
    # config-driven offsets through the tuned GEN kernels
    "op_lin_skew_mid": dict(n=(70, 53, 41), mode=0, values=NC_VALUES, offsets=SKEW_OFFSETS),
    "op_lin_diag_wide": dict(n=(130, 61, 40), mode=0, values=NC_VALUES, offsets=DIAG_OFFSETS),
    "op_nl_diag_mid": dict(n=(67, 50, 38), mode=1, values=NC_VALUES, offsets=DIAG_OFFSETS),
    "op_nt_skew_mid": dict(n=(63, 49, 35), mode=2, values=NC_VALUES, offsets=SKEW_OFFSETS),
}


def digest(a):
    """sha256 of the fp64 bytes of a dense [z,y,x] array, -0.0 folded into +0.0"""
    a = np.ascontiguousarray(a, dtype=np.float64) + 0.0
    return hashlib.sha256(a.tobytes()).hexdigest()


def seeded_field(shape, seed, scale=0.05, interior_only=True):
    """uniform(-scale, scale) field; halo zero like every array the V-cycle touches"""
    rs = np.random.RandomState(seed)
    a = rs.uniform(-scale, scale, size=shape)
    if interior_only:
        a[0, :, :] = a[-1, :, :] = 0.0
        a[:, 0, :] = a[:, -1, :] = 0.0
        a[:, :, 0] = a[:, :, -1] = 0.0
    return a


def rel_err(a, b):
    d = float(np.max(np.abs(a - b)))
    s = float(np.max(np.abs(b)))
    return d / s if s > 0 else d
