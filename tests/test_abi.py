"""CPU suite: the C-ABI library loads, exports every symbol include/gpusolve_cuda.h declares,
and its host-side logic (config parser, validation, slab partition) behaves like the
reference driver (src/main.cpp).  No compute calls: there is no GPU here and no CPU fallback."""
import ctypes as C
import os
import re
import subprocess

import pytest

from cases import GOLDEN_DIR, ROOT

HEADER = os.path.join(ROOT, "include", "gpusolve_cuda.h")
HEADERS = sorted(os.path.join(ROOT, "include", f) for f in os.listdir(os.path.join(ROOT, "include")) if f.endswith(".h"))
CONF = os.path.join(GOLDEN_DIR, "data-2nd_order.conf")


def declared_symbols():
    src = "\n".join(open(h).read() for h in HEADERS)
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(gs_[A-Za-z_0-9]+)\s*\(", src)))


def test_header_declares_a_reasonable_surface():
    syms = declared_symbols()
    for must in ("gs_create", "gs_solve", "gs_smooth", "gs_residual", "gs_restrict_residual", "gs_prolong_correct",
                 "gs_vcycle", "gs_newton_step", "gs_get_field", "gs_set_field", "gs_destroy", "gs_last_error"):
        assert must in syms


def test_library_exports_every_declared_symbol(gs):
    lib = C.CDLL(gs.LIB)
    for name in declared_symbols():
        assert hasattr(lib, name), f"{name} declared in include/*.h but not exported"
    assert set(gs.SIGNATURES) == set(declared_symbols()), "python binding and header drifted apart"


def test_library_has_sm100a_code_and_no_nccl_link_dependency(gs):
    out = subprocess.run(["cuobjdump", "-lelf", gs.LIB], capture_output=True, text=True).stdout
    assert "sm_100a" in out
    needed = subprocess.run(["readelf", "-d", gs.LIB], capture_output=True, text=True).stdout
    assert "libnccl" not in needed  # bound with dlopen only when world > 1


def test_parse_shipped_config(gs):
    p = gs.parse_config(CONF)
    assert (p.maxiter, p.tol, tuple(p.dim), p.mode, p.pre, p.post) == (10, 1e-5, (127, 127, 127), 2, 3, 3)
    assert (p.omega, p.gamma) == (0.8, 1.0)
    assert p.h == 1.0 / 128  # src/main.cpp:84
    assert list(p.stencil_val) == [6, -1, -1, -1, -1, -1, -1]
    assert [tuple(o) for o in p.stencil_off] == list(gs.CANONICAL_OFFSETS)
    assert p.print_progress == 1


def _write(tmp_path, lines):
    f = tmp_path / "c.conf"
    f.write_text("\n".join(lines) + "\n")
    return f


BASE = ["10", "10e-5", "15 15 15", "0", "2", "2", "0.8", "1.0", "6 -1 -1 -1 -1 -1 -1", "0 1 -1 0 0 0 0", "0 0 0 1 -1 0 0",
        "0 0 0 0 0 1 -1"]


def test_parser_is_whitespace_driven_like_operator_shift(gs, tmp_path):
    p = gs.parse_config(_write(tmp_path, BASE))  # dims on one line, "10e-5" as runExperiments.py writes it
    assert tuple(p.dim) == (15, 15, 15) and p.tol == 10e-5


@pytest.mark.parametrize("mutate,needle", [
    (lambda l: l[:3] + ["7"] + l[4:], "Invalid mode"),
    (lambda l: l[:-1] + ["0 0 0 0 0 2 -1"], "outside {-1,0,1}"),
    (lambda l: l[:-2], "truncated"),
    (lambda l: l[:2] + ["0 15 15"] + l[3:], "out of range"),
])
def test_parser_rejects_bad_configs(gs, tmp_path, mutate, needle):
    with pytest.raises(gs.GpuSolveError, match=re.escape(needle)):
        gs.parse_config(_write(tmp_path, mutate(list(BASE))))


def test_missing_config_file(gs, tmp_path):
    with pytest.raises(gs.GpuSolveError, match="does not exist"):
        gs.parse_config(tmp_path / "nope.conf")


def test_create_fails_loudly_without_a_gpu(gs):
    if gs.load().gs_device_count() > 0:
        pytest.skip("a GPU is visible")
    with pytest.raises(gs.GpuSolveError):
        gs.GpuSolve(gs.make_params(7))


def test_driver_binary_usage(gs):
    assert os.path.exists(gs.EXE)
    r = subprocess.run([gs.EXE], capture_output=True, text=True)
    assert r.returncode == 1 and "Missing config file" in r.stderr
    r = subprocess.run([gs.EXE, "/nonexistent.conf"], capture_output=True, text=True)
    assert r.returncode == 1 and "does not exist or is not a file" in r.stderr


@pytest.mark.parametrize("n,world", [(1023, 8), (1023, 4), (1023, 2), (511, 8), (511, 2), (127, 2), (255, 3), (100, 4), (63, 8)])
def test_partition_is_a_coarse_aligned_cover(gs, n, world):
    p = gs.make_params(n)
    import math
    nlev = int(math.floor(math.log(n) / math.log(2.0))) + 1
    _, _, rep = gs.partition(p, world, 0, 0)
    assert 0 <= rep <= nlev - 2 or rep == 0
    dims = n
    for level in range(nlev):
        Z = dims + 2
        prev_end = 0
        for r in range(world):
            zb, ze, rep2 = gs.partition(p, world, r, level)
            assert rep2 == rep
            if level >= rep:
                assert (zb, ze) == (0, Z)  # replicated: every rank holds the whole level
                continue
            assert zb == prev_end and ze > zb, "slabs must tile [0, Z) without gaps"
            prev_end = ze
            if level + 1 < rep:
                cb, ce, _ = gs.partition(p, world, r, level + 1)
                assert zb == 2 * cb, "fine slab boundary must be twice the coarse one"
            if r < world - 1:
                assert ze % 2 == 0
        if level < rep:
            assert prev_end == Z
        dims //= 2


def test_partition_world1_and_small_grids_are_unpartitioned(gs):
    for n, world in ((31, 1), (7, 8)):
        zb, ze, rep = gs.partition(gs.make_params(n), world, 0, 0)
        assert (zb, ze, rep) == (0, n + 2, 0)


@pytest.mark.parametrize("tiles,planes,occ,overhead,lo,hi,even", [
    (387, 511, 2, 6, 16, 128, False),     # smooth2u at 513^3
    (110, 255, 2, 6, 16, 128, False),     # ... at 257^3
    (256, 255, 2, 3, 4, 64, False),       # resid+restrict at 513^3 (coarse planes)
    (256, 511, 4, 2, 8, 128, True),       # prolongation at 513^3
    (4, 31, 2, 3, 4, 64, False),          # a small level
    (1, 1, 1, 0, 1, 1, False),
])
def test_zchunk_plan_fills_whole_waves(gs, tiles, planes, occ, overhead, lo, hi, even):
    """grid sizing of the z-marching kernels (gs_plan_zchunk): the chunk respects its bounds and parity, and no other
    admissible chunk count gives fewer waves x (chunk + overhead)"""
    sm = 148
    chunk = gs.plan_zchunk(tiles, planes, sm, occ, overhead, lo, hi, even)
    assert 1 <= chunk <= max(min(hi, planes), 1) + (1 if even else 0)
    assert chunk >= min(lo, planes) or chunk == planes
    if even:
        assert chunk % 2 == 0

    def cost(c):
        nch = -(-planes // c)
        waves = -(-tiles * nch // (sm * occ))
        return waves * (c + overhead)

    best = cost(chunk)
    for n in range(1, planes + 1):
        c = -(-planes // n)
        if even:
            c += c & 1
        if c > min(hi, planes) or c < min(lo, min(hi, planes)):
            continue
        assert best <= cost(c), f"chunk {c} would be cheaper than {chunk}"
    # the headline launch: 513^3 smoother grid is within 2 % of a whole number of waves
    if (tiles, planes) == (387, 511):
        ctas = tiles * -(-planes // chunk)
        assert ctas / (sm * occ) > 0.97 * -(-ctas // (sm * occ))


def test_product_sources_never_touch_the_oracle_or_a_host_compute_path():
    """the oracle is test infrastructure: nothing under the package (library sources, driver, ctypes view) or in the
    public header may include, import, link or name it; and the library exports no host-side compute entry (the
    __host__ __device__ point functions of kernels_simple.cu are only ever called on the host by tests/pc_host_check.cu)"""
    import re
    pkg = os.path.join(ROOT, "gpu-solve_b200")
    files = [os.path.join(ROOT, "include", "gpusolve_cuda.h"), os.path.join(pkg, "__init__.py"), os.path.join(pkg, "Makefile")]
    files += [os.path.join(pkg, "csrc", f) for f in os.listdir(os.path.join(pkg, "csrc"))]
    pat = re.compile(r"oracle|gs_oracle|gso_|_ref/", re.I)
    for f in files:
        with open(f) as fh:
            for i, line in enumerate(fh, 1):
                assert not pat.search(line), f"{f}:{i} mentions the oracle: {line.strip()}"
    # host code of the library that loops over lattice points would be a CPU path: the only per-point code is device code
    api = open(os.path.join(pkg, "csrc", "gs_api.cu")).read()
    assert "_point(" not in api


def _check_partition(gs, n, world):
    import math
    p = gs.make_params(n)
    dims = [n] * 3 if isinstance(n, int) else list(n)
    nlev = int(math.floor(math.log(min(dims)) / math.log(2.0))) + 1
    _, _, rep = gs.partition(p, world, 0, 0)
    assert 0 <= rep <= max(nlev - 2, 0)
    nz = dims[2]
    for level in range(nlev):
        Z = nz + 2
        prev = 0
        for r in range(world):
            zb, ze, rep2 = gs.partition(p, world, r, level)
            assert rep2 == rep
            if level >= rep:
                assert (zb, ze) == (0, Z)
                continue
            # every rank holds at least the four planes asked for, the last one at least two (one interior plane and
            # the halo plane: still exchangeable two planes deep, tests/test_slab_gloo.py runs that case)
            assert zb == prev and ze - zb >= (4 if r < world - 1 else 2), f"n={n} world={world} level={level} rank={r}: [{zb},{ze})"
            prev = ze
            if r < world - 1 and level + 1 < rep:
                # coarse aligned wherever the coarse level is partitioned too; the deepest partitioned level may end a slab
                # on an odd plane (its coarse level is held whole, gs_api.cu coarse_range rounds: tests/test_slab_gloo.py n=69)
                _, ce, _ = gs.partition(p, world, r, level + 1)
                assert ze == 2 * ce
        if level < rep:
            assert prev == Z
        nz //= 2


def test_partition_of_the_bench_grids(gs):
    """the grids bench.py runs at N = 2, 4, 8 (weak) and the strong-scaling 1023^3"""
    for world, n in ((2, (511, 511, 1023)), (4, (511, 1023, 1023)), (8, 1023), (2, 1023), (4, 1023), (8, (511, 511, 4000))):
        _check_partition(gs, n, world)
        _, _, rep = gs.partition(gs.make_params(n), world, 0, 0)
        assert rep >= 1, f"{n} on {world} ranks must be partitioned"


def test_partition_invariants_on_random_grids(gs):
    import random
    rnd = random.Random(7)
    for _ in range(300):
        n = tuple(rnd.randint(1, 1200) for _ in range(3))
        _check_partition(gs, n, rnd.randint(2, 8))


