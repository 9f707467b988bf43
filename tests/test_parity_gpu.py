"""GPU suite: the CUDA backend, called through the C ABI, against the CPU oracle on the same
inputs, against the golden vectors generated from the reference, and — at BASELINE.json's
full sizes — through size-independent properties.

Bars (BASELINE.json north_star): linear mode is BIT-EXACT (non-contracted, reference-ordered
arithmetic); modes with exp() agree to <= 1e-12 relative (CUDA exp vs glibc exp, <= 1 ulp
each); every residual norm agrees to <= 1e-12 relative with a compensated sum over the
oracle's residual field."""
import os

import numpy as np
import pytest

from backends import F, NEWTONV, R, RESTV, V, CudaBackend, OracleBackend, full_size_script, operator_script
from cases import BIG_CASES, FULL_CASES, OPERATOR_CASES, SOLVE_CASES, digest, rel_err, seeded_field

pytestmark = pytest.mark.gpu

FIELD_TOL = 1e-12   # relative (max-norm) tolerance for modes 1 and 2; mode 0 must be bitwise
NORM_TOL = 1e-12


def check_field(name, a, b, mode, exact_ok=True):
    if mode == 0 and exact_ok:
        assert np.array_equal(a + 0.0, b + 0.0), f"{name}: linear-mode field not bit-identical (max rel {rel_err(a, b):.3e})"
    else:
        assert rel_err(a, b) <= FIELD_TOL, f"{name}: rel err {rel_err(a, b):.3e}"


def check_norm(name, a, b):
    assert abs(a - b) <= NORM_TOL * max(abs(b), 1e-300) or a == b, f"{name}: {a!r} vs {b!r}"


@pytest.mark.parametrize("kernels", [0, 1])
@pytest.mark.parametrize("name", sorted(OPERATOR_CASES))
def test_operators_match_oracle(name, kernels, gs, oracle_mod, golden, golden_fields):
    kw = OPERATOR_CASES[name]
    mode = kw["mode"]
    dev = CudaBackend(gs, kernels=kernels, maxiter=1, **kw)
    ref = OracleBackend("port", maxiter=1, **kw)
    got, want = operator_script(dev, mode), operator_script(ref, mode)
    assert set(got) == set(want)
    for k in want:
        if isinstance(want[k], float):
            check_norm(k, got[k], want[k])
        else:
            # the tuned Newton smoother reads a precomputed coefficient: one rounding apart
            check_field(k, got[k], want[k], mode)
            if mode == 0:
                assert digest(got[k]) == golden["operators"][name]["outputs"][k], k
    dev.close()
    ref.close()


@pytest.mark.parametrize("kernels,graph", [(0, 0), (1, 0), (1, 1)])
@pytest.mark.parametrize("name", sorted(SOLVE_CASES))
def test_solves_match_oracle_and_golden(name, kernels, graph, gs, oracle_mod, golden):
    kw, g = SOLVE_CASES[name], golden["solve"][name]
    mode = kw["mode"]
    dev = CudaBackend(gs, kernels=kernels, graph=graph, **kw)
    ref = OracleBackend("port", **kw)
    assert dev.num_levels == ref.num_levels == g["num_levels"]
    for l in range(dev.num_levels):
        assert dev.level_info(l) == ref.level_info(l)
    f_dev, f_ref = dev.get(0, F), ref.get(0, F)
    check_field("rhs", f_dev, f_ref, mode)
    info = dev.solve()
    hist = dev.history()
    if mode == 2:
        norms, inner = ref.newton_solve()
        assert info.inner_cycles == inner == g["inner_cycles"], "inner V-cycle counts (early exit) differ"
        sol_dev, sol_ref = dev.get(0, NEWTONV), ref.get(0, NEWTONV)
    else:
        norms = ref.solve()
        sol_dev, sol_ref = dev.get(0, V), ref.get(0, V)
    assert info.iterations == len(norms) - 1 == len(g["norms"]) - 1, "number of cycles run differs"
    assert len(hist) == len(norms)
    r0 = norms[0] if norms[0] > 0 else 1.0
    for i, (a, b) in enumerate(zip(hist, norms)):
        if mode == 0:
            check_norm(f"norm[{i}]", a, b)
        else:
            # with exp() ulps the late-cycle norm is a cancellation; 1e-12 holds while
            # |r|/|r0| >~ 1e-3, beyond that the floor is |d|r||/|r0|
            assert abs(a - b) <= NORM_TOL * abs(b) or abs(a - b) <= 1e-13 * r0, (i, a, b)
    check_field("solution", sol_dev, sol_ref, mode)
    if mode == 0:
        assert digest(sol_dev) == g["solution_digest"], "not bit-identical to the reference's field"
    dev.close()
    ref.close()


def test_final_norm_matches_compensated_sum(gs, oracle_mod):
    kw = dict(n=63, mode=0, pre=2, post=2, maxiter=2, tol=1e-30)
    dev, ref = CudaBackend(gs, **kw), OracleBackend("port", **kw)
    dev.solve()
    ref.solve()
    check_norm("final", dev.history()[-1], ref.norm_exact(0, R))
    n = dev.residual(0)  # stores r
    assert np.array_equal(dev.get(0, R), ref.get(0, R))
    check_norm("residual()", n, ref.norm_exact(0, R))


@pytest.mark.parametrize("name", sorted(BIG_CASES))
def test_big_cases_against_golden(name, gs, oracle_mod, golden):
    """127^3-class configs incl. the shipped example (Newton V(3,3)): residual history and compensated final norm from the
    reference (golden.json), and the WHOLE solution field against the C port run next to the GPU"""
    kw, g = BIG_CASES[name], golden["solve"][name]
    mode = kw["mode"]
    dev = CudaBackend(gs, **kw)
    info = dev.solve()
    hist = dev.history()
    want = [float.fromhex(x) for x in g["norms"]]
    assert info.iterations == len(want) - 1
    for i, (a, b) in enumerate(zip(hist, want)):
        # the reference's own printed norm moves by ~1.3e-12 with its OpenMP reduction order: 4e-12 here,
        # the 1e-12 bar is held against the compensated sum below
        assert abs(a - b) <= 4e-12 * abs(b) or abs(a - b) <= 1e-13 * want[0], (i, a, b)
    exact = float.fromhex(g["norm_exact_final"])
    assert abs(hist[-1] - exact) <= NORM_TOL * exact or (mode != 0 and abs(hist[-1] - exact) <= 1e-13 * want[0]), (hist[-1], exact)
    ref = OracleBackend("port", **kw)
    if mode == 2:
        assert info.inner_cycles == g["inner_cycles"]
        ref.newton_solve()
        sol, rsol = dev.get(0, NEWTONV), ref.get(0, NEWTONV)
    else:
        ref.solve()
        sol, rsol = dev.get(0, V), ref.get(0, V)
    check_field("solution (whole field vs port)", sol, rsol, mode)
    z, y, x = sol.shape
    centre = float.fromhex(g["centre"])
    assert abs(sol[z // 2, y // 2, x // 2] - centre) <= FIELD_TOL * abs(centre)
    if mode == 0:
        assert digest(sol) == g["solution_digest"]
    dev.close()
    ref.close()


@pytest.mark.parametrize("name", sorted(FULL_CASES))
def test_full_size_511_against_port_and_reference_golden(name, gs, oracle_mod, golden, golden_fields):
    """BASELINE.json configs[1]-[3] at their full size (arrays 513^3; linear V(2,2), FAS V(2,2), Newton-linearised
    V(3,3)): the whole level-0 field against the C port run next to the GPU (bitwise in mode 0, <= 1e-12 with exp()),
    every norm against the port's compensated sum at 1e-12, and the reference's own record of the same script
    (tests/golden/make_golden.py: norms, sha256 of the field in mode 0, a 33^3 strided sample of it)."""
    kw, g = FULL_CASES[name], golden["full"][name]
    mode, st = kw["mode"], g["sample_stride"]
    dev = CudaBackend(gs, **kw)
    norms, fid = full_size_script(dev, mode, kw["maxiter"])
    sol = dev.get(0, fid)
    dev.close()
    want = [float.fromhex(x) for x in g["norms"]]
    assert len(norms) == len(want)
    for i, (a, b) in enumerate(zip(norms, want)):
        assert abs(a - b) <= 4e-12 * abs(b), (i, a, b)      # the reference's OpenMP-order norm (see test_big_cases)
    exact = float.fromhex(g["norm_exact_final"])
    check_norm("final norm vs the reference's compensated sum", norms[-1], exact)
    check_field("strided sample vs the reference", sol[::st, ::st, ::st], golden_fields[name + "/sample"], mode)
    if mode == 0:
        assert digest(sol) == g["solution_digest"], "513^3 field is not bit-identical to the reference's"
    ref = OracleBackend("port", **kw)
    ref.set_threads(os.cpu_count())
    ex = []
    rn, _ = full_size_script(ref, mode, kw["maxiter"], exact=ex)
    for i, (a, b) in enumerate(zip(norms, ex)):
        check_norm(f"norm[{i}] vs compensated sum of the port's residual field", a, b)
    check_field("whole 513^3 field vs port", sol, ref.get(0, fid), mode)
    ref.close()


def test_shipped_config_through_the_driver_binary(gs, golden):
    """GpuSolve-cuda <conf> prints the reference's stdout lines (src/cpu/NewtonSolver.cpp:16,27)"""
    import os
    import re
    import subprocess
    from cases import GOLDEN_DIR
    r = subprocess.run([gs.EXE, os.path.join(GOLDEN_DIR, "data-2nd_order.conf")], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    out = r.stdout.splitlines()
    assert out[0].startswith('Using config file "') and out[1] == "Solving newton problem"
    assert out[2] == "Inital newton residual: 281.289"
    its = [re.fullmatch(r"newton iter: (\d+) residual: (\S+) Took (\d+)ms", l) for l in out[3:]]
    assert all(its) and [m.group(2) for m in its] == ["12.3816", "0.527004", "0.0222718", "0.00093763"]


def test_driver_partial_timers_keep_the_experiment_script_regexes(gs, tmp_path):
    """GPUSOLVE_TIMERS=1: Timer::stop's partial-timer tail (src/Timer.cpp:17-24) after "Took Nms"; the regexes of
    runExperiments.py:46 and :81 must still match and the residuals must not change"""
    import os
    import re
    import subprocess
    conf = tmp_path / "lin.conf"
    conf.write_text("3\n1e-30\n63 63 63\n0\n2\n2\n0.8\n1.0\n6 -1 -1 -1 -1 -1 -1\n0 1 -1 0 0 0 0\n0 0 0 1 -1 0 0\n0 0 0 0 0 1 -1\n")
    outs = []
    for timers in ("0", "1"):
        r = subprocess.run([gs.EXE, str(conf)], capture_output=True, text=True, timeout=120, env=dict(os.environ, GPUSOLVE_TIMERS=timers))
        assert r.returncode == 0, r.stderr
        outs.append(r.stdout)
    it = re.compile(r"iter: (\d+) residual: ([\d\.e-]+) Took (\d+)ms")
    plain, timed = it.findall(outs[0]), it.findall(outs[1])
    assert len(plain) == 3 and [m[:2] for m in plain] == [m[:2] for m in timed]
    assert len(re.findall(r"sumBuffer: (\d+)ms", outs[1])) == 3 and "sumBuffer" not in outs[0]
    assert re.search(r"Took \d+ms, smooth: \d+ms \(\d+x\) ", outs[1])


def test_get_set_roundtrip_and_layouts(gs):
    dev = CudaBackend(gs, n=(5, 7, 6), mode=1)
    a = seeded_field(dev.shape(0), 7, interior_only=False)
    dev.set(0, RESTV, a)
    assert np.array_equal(dev.get(0, RESTV), a)
    zf = dev.get_zfastest(0, RESTV)          # [x, y, z] like the reference's Vector3
    assert np.array_equal(zf, a.transpose(2, 1, 0))
    with pytest.raises(gs.GpuSolveError):
        dev.get(0, NEWTONV)                  # not allocated in mode 1
    with pytest.raises(gs.GpuSolveError):
        dev.get(9, V)
    dev.close()


def test_dump_format_matches_vector3_dump(gs, tmp_path):
    dev = CudaBackend(gs, n=3, mode=0)
    dev.vcycle()
    path = tmp_path / "v.txt"
    dev.dump(0, V, path)
    lines = path.read_text().splitlines()
    assert lines[0] == "5 5 5" and len(lines) == 1 + 125
    v = dev.get(0, V)
    x, y, z, val = lines[1 + 2 * 25 + 2 * 5 + 2].split()  # x outer, z inner (src/cpu/Vector3.cpp:67-75)
    assert (x, y, z) == ("2", "2", "2") and float(val) == pytest.approx(v[2, 2, 2], rel=1e-5)
    dev.close()


def test_smoother_is_linear_in_mode0(gs):
    """S(a*v1 + v2; a*f1 + f2) == a*S(v1; f1) + S(v2; f2) up to rounding, any size"""
    outs = []
    shape = None
    for sv, sf in ((11, 12), (13, 14)):
        dev = CudaBackend(gs, n=(40, 33, 21), mode=0)
        shape = dev.shape(0)
        dev.set(0, V, seeded_field(shape, sv))
        dev.set(0, F, seeded_field(shape, sf, scale=1.0))
        dev.smooth(0, 3)
        outs.append(dev.get(0, V))
        dev.close()
    dev = CudaBackend(gs, n=(40, 33, 21), mode=0)
    dev.set(0, V, 2.0 * seeded_field(shape, 11) + seeded_field(shape, 13))
    dev.set(0, F, 2.0 * seeded_field(shape, 12, scale=1.0) + seeded_field(shape, 14, scale=1.0))
    dev.smooth(0, 3)
    assert rel_err(dev.get(0, V), 2.0 * outs[0] + outs[1]) < 1e-13
    dev.close()


def test_full_size_513_properties(gs, golden):
    """n=511 (arrays 513^3), the benchmark grid: tuned == simple kernels bit for bit after a V(2,2)
    cycle, residual history equal to the reference's (golden.json full511_lin_v22; printed: 8328.23 -> 2840.32 ->
    849.574), field digest equal to the reference's, and the CUDA-graph path equals the plain-launch path."""
    kw = dict(n=511, mode=0, pre=2, post=2, maxiter=2, tol=1e-30)
    g = golden["full"]["full511_lin_v22"]
    a = CudaBackend(gs, kernels=1, graph=1, **kw)
    ia = a.solve()
    ha = a.history()
    assert [f"{x:.6g}" for x in ha] == ["8328.23", "2840.32", "849.574"]
    for x, y in zip(ha[1:], [float.fromhex(t) for t in g["norms"]]):
        assert abs(x - y) <= 4e-12 * y
    va = digest(a.get(0, V))
    assert va == g["solution_digest"]
    a.close()
    b = CudaBackend(gs, kernels=0, graph=0, **kw)
    b.solve()
    # fields are bit-identical; the norms differ only by the summation tree of the two residual kernels
    assert ia.iterations == 2 and len(b.history()) == len(ha)
    for x, y in zip(b.history(), ha):
        check_norm("simple vs tuned norm", x, y)
    assert digest(b.get(0, V)) == va
    b.close()


@pytest.mark.parametrize("mode,n", [(0, 255), (0, (200, 150, 300)), (1, 191), (2, 191)])
def test_back_to_back_cycles_are_race_free(gs, mode, n):
    """V-cycles queued back to back (no host synchronisation between operators, plain launches and CUDA graph)
    from a seeded iterate, three times over: every run must give the same bits as the one-thread-per-point
    kernels in linear mode (<= 1e-12 with exp()), and the same bits as every other run.  A shared-memory ring
    refilled while a slow warp still reads it only shows up under this load (tools/debug_ops.py)."""
    kw = dict(n=n, mode=mode, pre=2, post=2, maxiter=1, tol=1e-30)
    ref = CudaBackend(gs, kernels=0, graph=0, **kw)
    v0 = seeded_field(ref.shape(0), 21)
    field = NEWTONV if mode == 2 else V

    def run(b):
        b.set(0, V, v0)
        if mode == 2:
            b.set(0, NEWTONV, v0)
            return [b.newton_step()[0] for _ in range(2)], b.get(0, field)
        return [b.vcycle() for _ in range(3)], b.get(0, field)

    rn, rv = run(ref)
    ref.close()
    for graph in (0, 1):
        b = CudaBackend(gs, kernels=1, graph=graph, **kw)
        seen = set()
        for rep in range(3):
            norms, v = run(b)
            check_field(f"rep {rep} graph {graph}", v, rv, mode)
            for x, y in zip(norms, rn):
                check_norm(f"rep {rep} graph {graph} norm", x, y)
            seen.add(digest(v))
        assert len(seen) == 1, "tuned path is not deterministic"
        b.close()


@pytest.mark.parametrize("name", ["lin_tol", "lin31_v22", "lin63_v33", "lin_ragged", "lin_pre0", "nl31_v22", "nl_even18", "nt_tol", "nt31_v22"])
def test_device_side_stopping_equals_the_host_loop(name, gs, golden):
    """device_stop = 1 (default): a solve that prints nothing — every inner solve of NewtonSolver::findError (<= 10 cycles,
    tol 0.1, src/cpu/NewtonSolver.cpp:94-100), and CpuSolver::solve with print_progress = 0 — is ONE conditional CUDA graph
    whose WHILE body is the V-cycle and whose stopping test `res <= r0/(1/tol)` runs on the device (kernels_ctl.cu).
    Cycle counts, residual history and fields must equal the host loop's (and through it the reference's) exactly."""
    kw, g = SOLVE_CASES[name], golden["solve"][name]
    mode = kw["mode"]
    res = []
    for dstop in (1, 0):
        dev = CudaBackend(gs, kernels=1, graph=1, **kw)
        dev.set_option("device_stop", dstop)
        l0 = dev.launches
        info = dev.solve()
        res.append((info.iterations, info.inner_cycles, info.converged, dev.history(), digest(dev.get(0, NEWTONV if mode == 2 else V)),
                    dev.launches - l0))
        dev.close()
    assert res[0][:5] == res[1][:5], "device-side loop and host loop disagree"
    assert res[0][5] > res[1][5], "the device loop adds its stop_init / stop_check launches to the same cycles"
    assert res[0][0] == len(g["norms"]) - 1
    if mode == 2:
        assert res[0][1] == g["inner_cycles"]


@pytest.mark.parametrize("mode,n,pre,post", [
    (0, 127, 2, 2), (0, 255, 2, 2), (0, (131, 97, 63), 4, 4), (0, (200, 150, 300), 1, 3), (0, 120, 2, 2), (0, 513, 2, 2),
    (1, (121, 64, 70), 4, 2), (1, 191, 2, 2), (2, (100, 90, 47), 2, 4), (2, 191, 3, 3),
])
def test_prolongation_folded_into_the_post_smoother_keeps_the_bits(gs, mode, n, pre, post):
    """fuse_pc: prolongation + correction + the first two post-smoothing sweeps as ONE pass (kernels_fused.cu smooth2u_tma<.., PC>;
    src/cpu/CpuSolver.cpp:128-134) — the corrected iterate only ever exists in shared memory.  Same per-point arithmetic in
    the same order, so the fields must equal the separate passes bit for bit (also with exp(): identical instructions), back to
    back, with and without the graph, and one launch per level fewer."""
    kw = dict(n=n, mode=mode, pre=pre, post=post, maxiter=1, tol=1e-30)
    ref = CudaBackend(gs, kernels=1, graph=0, **kw)
    ref.set_option("fuse_pc", 0)
    v0 = seeded_field(ref.shape(0), 21)
    field = NEWTONV if mode == 2 else V

    def run(b):
        b.set(0, V, v0)
        if mode == 2:
            b.set(0, NEWTONV, v0)
            return [b.newton_step()[0] for _ in range(2)], b.get(0, field)
        return [b.vcycle() for _ in range(3)], b.get(0, field)

    rn, rv = run(ref)
    ref.close()
    for graph in (0, 1):
        # launches of the unfused path under the same graph setting (a device-side solve loop adds its control kernels)
        u = CudaBackend(gs, kernels=1, graph=graph, **kw)
        u.set_option("fuse_pc", 0)
        run(u)
        l0 = u.launches
        run(u)
        per_run = u.launches - l0
        u.close()
        b = CudaBackend(gs, kernels=1, graph=graph, **kw)
        b.set_option("fuse_pc", 2)   # 2: also in mode 1, where the default keeps the separate passes (slower fused)
        for rep in range(2):
            l0 = b.launches
            norms, v = run(b)
            assert digest(v) == digest(rv), f"graph {graph} rep {rep}: fields differ (max rel {rel_err(v, rv):.3e})"
            assert norms == rn
            assert b.launches - l0 < per_run, "no prolongation launch was folded away"
        b.close()


@pytest.mark.parametrize("n,pre,post", [(127, 2, 2), (191, 2, 2), ((121, 64, 70), 4, 2), (120, 3, 3), (511, 2, 2)])
def test_fas_restriction_of_v_folded_into_the_residual_restriction(gs, n, pre, post):
    """fuse_rv (survey K5, src/cpu/CpuSolver.cpp:106-107): in FAS cycles the fused residual+restriction pass also forms
    restV = v_{l+1} = R(v_l) — the same 27 exact products as the reference's running sum, added plane by plane, so equal
    to the separate restriction within rounding (1e-12 is the tolerance of the exp() mode), one fine-grid launch per level
    fewer, and deterministic from run to run"""
    kw = dict(n=n, mode=1, pre=pre, post=post, maxiter=1, tol=1e-30)
    ref = CudaBackend(gs, kernels=1, graph=0, **kw)
    ref.set_option("fuse_rv", 0)
    v0 = seeded_field(ref.shape(0), 21)

    def run(b):
        b.set(0, V, v0)
        norms = [b.vcycle() for _ in range(3)]
        return norms, b.get(0, V), b.get(1, RESTV)

    rn, rv, rrest = run(ref)
    l0 = ref.launches
    run(ref)
    per_run = ref.launches - l0
    ref.close()
    for graph in (0, 1):
        b = CudaBackend(gs, kernels=1, graph=graph, **kw)
        b.set_option("fuse_rv", 1)
        seen = set()
        for rep in range(2):
            l0 = b.launches
            norms, v, rest = run(b)
            assert rel_err(v, rv) <= FIELD_TOL, f"graph {graph} rep {rep}: v rel err {rel_err(v, rv):.3e}"
            assert rel_err(rest, rrest) <= FIELD_TOL, f"graph {graph} rep {rep}: restV rel err {rel_err(rest, rrest):.3e}"
            for x, y in zip(norms, rn):
                check_norm(f"graph {graph} rep {rep} norm", x, y)
            if graph == 0:
                assert b.launches - l0 < per_run, "no restriction launch was folded away"
            seen.add(digest(v))
        assert len(seen) == 1, "fused path is not deterministic"
        b.close()
