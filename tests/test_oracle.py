"""CPU suite: pins the oracle (oracle/gs_oracle.c) against golden vectors generated from the
reference's own translation units, and — where oracle/_ref was built — against those units
directly.  Bit-exact for every field; norms to 1e-13 (summation order only)."""
import numpy as np
import pytest

from backends import F, NEWTONV, OracleBackend, R, V, operator_script
from cases import OPERATOR_CASES, SOLVE_CASES, digest

NORM_TOL = 1e-13


def _close(a, b, tol=NORM_TOL):
    return abs(a - b) <= tol * max(abs(b), 1e-300) or a == b


@pytest.mark.parametrize("name", sorted(SOLVE_CASES))
def test_port_matches_golden_solves(name, oracle_mod, golden, golden_fields):
    kw, g = SOLVE_CASES[name], golden["solve"][name]
    b = OracleBackend("port", **kw)
    assert b.num_levels == g["num_levels"]
    for l in range(b.num_levels):
        dims, h = b.level_info(l)
        assert list(dims) == g["levels"][l][0] and h == float.fromhex(g["levels"][l][1])
    assert digest(b.get(0, F)) == g["f0_digest"], "right-hand side differs from the reference"
    if kw["mode"] == 2:
        norms, inner = b.newton_solve()
        assert inner == g["inner_cycles"], "inner V-cycle count (early exit at tol 0.1) differs"
        sol, exact = b.get(0, NEWTONV), b.norm_exact(0, F)
    else:
        norms = b.solve()
        sol, exact = b.get(0, V), b.norm_exact(0, R)
    ref = [float.fromhex(x) for x in g["norms"]]
    assert len(norms) == len(ref), "number of cycles run (early exit) differs"
    for a, r in zip(norms, ref):
        assert _close(a, r), (a, r)
    assert exact == float.fromhex(g["norm_exact_final"])
    assert digest(sol) == g["solution_digest"], "solution field is not bit-identical to the reference"
    key = name + "/solution"
    if key in golden_fields:
        assert np.array_equal(sol, golden_fields[key])
    b.close()


@pytest.mark.parametrize("name", sorted(OPERATOR_CASES))
def test_port_matches_golden_operators(name, oracle_mod, golden, golden_fields):
    kw, g = OPERATOR_CASES[name], golden["operators"][name]
    b = OracleBackend("port", maxiter=1, **kw)
    out = operator_script(b, kw["mode"])
    assert set(out) == set(g["outputs"])
    for k, v in out.items():
        if isinstance(v, float):
            assert _close(v, float.fromhex(g["outputs"][k])), k
        else:
            if name + "/" + k in golden_fields:
                assert np.array_equal(v, golden_fields[name + "/" + k]), k
            assert digest(v) == g["outputs"][k], k
    b.close()


def _have_ref(oracle_mod):
    return oracle_mod.available("reference")


@pytest.mark.parametrize("name", ["lin31_v22", "lin_ragged", "nl_even18", "nt31_v22", "lin_diag15"])
def test_port_matches_reference_units_directly(name, oracle_mod):
    if not _have_ref(oracle_mod):
        pytest.skip("oracle/_ref not built (no reference tree on this machine)")
    kw = SOLVE_CASES[name]
    a, b = OracleBackend("port", **kw), OracleBackend("reference", **kw)
    if kw["mode"] == 2:
        (na, ia), (nb, ib) = a.newton_solve(), b.newton_solve()
        assert ia == ib
        fields = (V, F, R, NEWTONV)
    else:
        na, nb = a.solve(), b.solve()
        fields = (V, F, R, 3, 4)
    assert len(na) == len(nb) and all(_close(x, y) for x, y in zip(na, nb))
    for l in range(a.num_levels):
        for f in fields:
            if f == 3 and l == a.num_levels - 1:
                continue
            assert np.array_equal(a.get(l, f), b.get(l, f)), (l, f)


@pytest.mark.parametrize("mode", [0, 1, 2])
def test_stepped_loops_equal_reference_entry_points(mode, oracle_mod, capfd):
    """the harness' stepped solve loops reproduce CpuSolver::solve / NewtonSolver::solve exactly"""
    if not _have_ref(oracle_mod):
        pytest.skip("oracle/_ref not built")
    kw = dict(n=15, mode=mode, pre=2, post=2, maxiter=6, tol=1e-3)
    a, b = OracleBackend("reference", **kw), OracleBackend("reference", print_progress=1, **kw)
    if mode == 2:
        norms, _ = a.newton_solve()
    else:
        norms = a.solve()
    b.native_solve()
    printed = capfd.readouterr().out
    sol = NEWTONV if mode == 2 else V
    assert np.array_equal(a.get(0, sol), b.get(0, sol))
    lines = [l for l in printed.splitlines() if "iter:" in l]
    assert len(lines) == len(norms) - 1, printed
    for line, n in zip(lines, norms[1:]):
        assert float(line.split("residual:")[1].split()[0]) == pytest.approx(n, rel=1e-5)


def test_norm_depends_on_threads_only_in_last_digits(oracle_mod):
    """fields are thread-count independent, the norm is a running sum"""
    kw = dict(n=31, mode=0, pre=2, post=2, maxiter=2, tol=1e-30)
    a = OracleBackend("port", **kw)
    a.set_threads(1)
    n1 = a.solve()
    v1 = a.get(0, V)
    b = OracleBackend("port", **kw)
    b.set_threads(4)
    n4 = b.solve()
    assert np.array_equal(v1, b.get(0, V))
    assert n1 == n4  # the port sums plane partials in plane order: deterministic
