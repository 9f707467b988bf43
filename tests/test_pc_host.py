"""The interpolation state machine of the fused prolongation + post-smoothing pass (gs_device.cuh PcInterp) on the CPU:
carried x-interpolants, odd/even planes, the rim rows of the first and last row pair — against prolong_point for every
interior point of odd, even and ragged grids (tests/pc_host_check.cu).  The kernel itself is checked on the GPU
(test_prolongation_folded_into_the_post_smoother_keeps_the_bits)."""
import ctypes as C
import os
import subprocess

import pytest

from cases import ROOT

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def harness(tmp_path_factory):
    out = str(tmp_path_factory.mktemp("pc") / "libpc_host_check.so")
    cmd = ["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-O2", "-std=c++17", "-fmad=false", "-Xcompiler", "-fPIC",
           "-Xcompiler", "-ffp-contract=off", "-shared", os.path.join(HERE, "pc_host_check.cu"), "-o", out, "-cudart", "static"]
    subprocess.run(cmd, check=True, cwd=ROOT)
    lib = C.CDLL(out)
    lib.pc_host_check.restype = C.c_long
    lib.pc_host_check.argtypes = [C.c_int, C.c_int, C.c_int, C.POINTER(C.c_long)]
    return lib


@pytest.mark.parametrize("n", [(15, 15, 15), (16, 18, 14), (9, 12, 17), (31, 7, 10), (2, 3, 4), (33, 20, 21)])
def test_marched_interpolation_equals_the_closed_form(harness, n):
    checked = C.c_long()
    bad = harness.pc_host_check(n[0], n[1], n[2], C.byref(checked))
    assert checked.value > 0 and bad == 0, f"{bad} of {checked.value} interpolated values differ from prolong_point"
