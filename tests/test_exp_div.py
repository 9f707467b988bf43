"""gs_exp / gs_div (gs_device.cuh): the exp() and the per-point divide of the non-linear modes, run on the CPU through the
same explicit-fma sequences as on the device (tests/exp_div_check.cu), against glibc's exp and the IEEE divide.
North_star's tolerance for the exp() modes is 1e-12 relative on the fields; these stay within an ulp (1.1e-16)."""
import ctypes as C
import math
import os
import subprocess

import pytest

from cases import ROOT

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def harness(tmp_path_factory):
    out = str(tmp_path_factory.mktemp("expdiv") / "libexp_div_check.so")
    cmd = ["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-O2", "-std=c++17", "-fmad=false", "-Xcompiler", "-fPIC",
           "-Xcompiler", "-ffp-contract=off", "-shared", os.path.join(HERE, "exp_div_check.cu"), "-o", out, "-cudart", "static"]
    subprocess.run(cmd, check=True, cwd=ROOT)
    lib = C.CDLL(out)
    lib.exp_check.restype = C.c_longlong
    lib.exp_check.argtypes = [C.c_double, C.c_double, C.c_longlong, C.POINTER(C.c_longlong)]
    lib.div_check.restype = C.c_longlong
    lib.div_check.argtypes = [C.c_double, C.c_double, C.c_double, C.c_longlong, C.POINTER(C.c_longlong)]
    lib.exp_one.restype = C.c_double
    lib.exp_one.argtypes = [C.c_double]
    lib.div_one.restype = C.c_double
    lib.div_one.argtypes = [C.c_double, C.c_double]
    return lib


@pytest.mark.parametrize("lo,hi", [(-0.1, 0.1), (-1.0, 1.0), (-30.0, 30.0), (-699.0, 699.0)])
def test_exp_within_one_ulp_of_glibc(harness, lo, hi):
    differing = C.c_longlong()
    worst = harness.exp_check(lo, hi, 2_000_000, C.byref(differing))
    assert worst <= 1, f"gs_exp is {worst} ulps from glibc's exp somewhere in [{lo}, {hi}]"
    assert differing.value < 0.2 * 2_000_000   # and equal to it for most arguments


def test_exp_special_arguments_go_to_the_library(harness):
    assert harness.exp_one(0.0) == 1.0
    assert harness.exp_one(800.0) == math.inf and harness.exp_one(-800.0) == 0.0
    assert math.isnan(harness.exp_one(math.nan))
    assert harness.exp_one(709.0) == math.exp(709.0) and harness.exp_one(-745.0) == math.exp(-745.0)
    for x in (1.0, -1.0, 0.5 * math.log(2.0), 0.34657359027997264, -0.34657359027997264, 699.999):
        assert abs(harness.exp_one(x) - math.exp(x)) <= 2.3e-16 * math.exp(x)


@pytest.mark.parametrize("blo,bhi,amax", [(1.0, 1e7, 1e4), (1e-3, 1e3, 1.0), (1e-250, 1e250, 1e20)])
def test_divide_is_the_ieee_quotient(harness, blo, bhi, amax):
    differing = C.c_longlong()
    worst = harness.div_check(blo, bhi, amax, 4_000_000, C.byref(differing))
    assert worst <= 1, f"gs_div is {worst} ulps from the IEEE quotient"
    assert differing.value <= 4, f"{differing.value} of 4e6 quotients are not correctly rounded"


def test_divide_special_operands_go_to_the_library(harness):
    assert harness.div_one(1.0, 0.0) == math.inf and harness.div_one(-1.0, 0.0) == -math.inf
    assert math.isnan(harness.div_one(0.0, 0.0)) and math.isnan(harness.div_one(math.nan, 3.0))
    assert harness.div_one(1.0, math.inf) == 0.0 and harness.div_one(1e300, 1e-300) == math.inf
    assert harness.div_one(0.0, 5.0) == 0.0 and harness.div_one(7.0, 7.0) == 1.0
