"""The coarse tail of a V-cycle (levels of a few thousand points) can run as ONE single-CTA kernel: gs_api.cu records
the one-thread-per-point operators into a program (kernels_simple.cu ks_tail_*) and a small interpreter executes it.
This test compiles that translation unit into a harness that executes recorded programs over host arrays
(tests/tail_host_check.cu) and demands the oracle's V-cycle: bit for bit wherever the arithmetic is the
reference's (every mode here, since host and oracle share glibc's exp), 1e-12 for the precomputed-coefficient
variant of mode 2, which rounds one product differently."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from cases import ROOT, SKEW_OFFSETS, SKEW_VALUES, seeded_field

HERE = os.path.dirname(os.path.abspath(__file__))
V, F, R, E, RESTV, NEWTONV, NEWTONF = range(7)


@pytest.fixture(scope="module")
def harness(tmp_path_factory):
    out = str(tmp_path_factory.mktemp("tail") / "libtail_host_check.so")
    cmd = ["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-O2", "-std=c++17", "-fmad=false", "-Xcompiler", "-fPIC",
           "-Xcompiler", "-ffp-contract=off", "-shared", os.path.join(HERE, "tail_host_check.cu"), "-o", out, "-cudart", "static"]
    subprocess.run(cmd, check=True, cwd=ROOT)
    lib = C.CDLL(out)
    lib.tail_host_vcycle.restype = C.c_int
    lib.tail_host_vcycle.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    return lib


def run_case(harness, oracle_mod, n, mode, pre, post, tail, cycles=2, km_coef=0, **kw):
    p = oracle_mod.make_params(n, mode=mode, pre=pre, post=post, maxiter=1, tol=1e-30, **kw)
    o = oracle_mod.Oracle(p, "port")
    nlev = o.num_levels
    v0 = seeded_field(o.shape(0), 21)
    f0 = seeded_field(o.shape(0), 22, scale=1.0)
    o.set(0, V, v0)
    o.set(0, F, f0)
    nvs = None
    if mode == 2:
        nvs = [seeded_field(o.shape(l), 30 + l) for l in range(nlev)]
        for l in range(nlev):
            o.set(l, NEWTONV, nvs[l])
    for _ in range(cycles):
        o.vcycle()
    want = o.get(0, V)
    o.close()
    got = np.empty_like(v0)
    stats = (C.c_int * 2)()
    ptrs = None
    if nvs is not None:
        ptrs = (C.c_void_p * nlev)(*[a.ctypes.data for a in nvs])
    rc = harness.tail_host_vcycle(C.byref(p), tail, km_coef, cycles, v0.ctypes.data, f0.ctypes.data, ptrs, got.ctypes.data, stats)
    assert rc == 0
    return got, want, stats[0], stats[1], nlev


@pytest.mark.parametrize("n,mode,pre,post,tail", [
    (15, 0, 2, 2, 0),            # the whole cycle as one program
    (15, 0, 2, 2, 1),
    (15, 0, 2, 2, -1),           # no tail: every operator its own program
    (31, 0, 2, 2, 1),
    (63, 0, 2, 2, 0),            # six levels in one level table
    ((9, 12, 17), 0, 1, 3, 0),   # ragged, h*h not a power of two
    (18, 0, 2, 1, 0),            # even extents, odd sweep count (out-of-place correction, parity copy)
    ((3, 5, 4), 0, 2, 2, 0),
    (15, 0, 0, 2, 0),
    (15, 0, 3, 0, 1),
    (15, 1, 2, 2, 0),            # FAS: two-output restriction, coarse right-hand side, v -= restV
    ((9, 12, 17), 1, 1, 3, 0),
    (18, 1, 2, 1, 1),
    (15, 2, 2, 2, 0),            # Newton inner cycle, exp() in the sweep
    ((9, 12, 17), 2, 1, 3, 1),
])
def test_recorded_tail_reproduces_the_oracle_vcycle(harness, oracle_mod, n, mode, pre, post, tail):
    got, want, programs, ops, nlev = run_case(harness, oracle_mod, n, mode, pre, post, tail)
    assert np.array_equal(got, want), f"max |diff| = {np.abs(got - want).max():.3e}"
    if tail == 0:
        assert programs == 2    # one program per cycle


def test_skewed_stencil_through_the_tail(harness, oracle_mod):
    got, want, *_ = run_case(harness, oracle_mod, 15, 0, 2, 2, 0, values=SKEW_VALUES, offsets=SKEW_OFFSETS)
    assert np.array_equal(got, want)


def test_long_programs_are_split_not_truncated(harness, oracle_mod):
    """V(30,30) on four levels is more than the 44 operators one program holds: the recorder launches what it has and
    starts the next program; the result must not change"""
    got, want, programs, ops, nlev = run_case(harness, oracle_mod, 15, 0, 30, 30, 0, cycles=1)
    assert ops > 44 and programs == (ops + 43) // 44
    assert np.array_equal(got, want)


def test_precomputed_coefficient_variant(harness, oracle_mod):
    got, want, *_ = run_case(harness, oracle_mod, 15, 2, 2, 2, 0, km_coef=1)
    assert np.abs(got - want).max() <= 1e-12 * np.abs(want).max()
