"""Uniform adapters over the CPU oracle (either implementation) and the CUDA backend so one
operator script drives both.  Field ids are identical in both APIs (V,F,R,E,RESTV,NEWTONV,NEWTONF)."""
import numpy as np

from cases import ROOT, seeded_field  # noqa: F401  (adds ROOT to sys.path)
from oracle import oracle as O

V, F, R, E, RESTV, NEWTONV, NEWTONF = range(7)


class OracleBackend:
    """composites spelled with the reference's own primitives, in the order CpuSolver::vcycle uses them"""

    def __init__(self, kind="port", **kw):
        self.o = O.Oracle(O.make_params(**kw), kind)
        self.mode = kw.get("mode", 0)

    def close(self):
        self.o.close()

    def __getattr__(self, name):
        return getattr(self.o, name)

    def restrict_residual(self, level):           # src/cpu/CpuSolver.cpp:93-98
        self.o.residual(level)
        self.o.restrict(level, R, F)

    def fas_coarse_rhs(self, level):              # :106-112
        self.o.restrict(level, V, RESTV)
        self.o.restrict(level, V, V)
        self.o.apply_stencil(level + 1)
        self.o.axpy(level + 1, F, R, +1)

    def prolong(self, level):                     # :128
        self.o.interpolate(level)
        return self.o.get(level, E)

    def prolong_correct(self, level):             # :121-132
        if self.mode == 1:
            self.o.axpy(level + 1, V, RESTV, -1)
        self.o.interpolate(level)
        self.o.axpy(level, V, E, +1)

    def newton_step(self):                        # src/cpu/NewtonSolver.cpp:21-26
        self.o.newton_compF()
        self.o.set(0, V, np.zeros(self.o.shape(0)))
        inner = self.o.newton_find_error()
        return self.o.newton_compF(), inner

    def set_newton_f(self, arr):
        self.o.set(0, NEWTONF, arr)


class CudaBackend:
    def __init__(self, gs, kernels=None, graph=None, **kw):
        self.gs = gs
        self.o = gs.GpuSolve(gs.make_params(**kw), device=0)
        self.mode = kw.get("mode", 0)
        if kernels is not None:
            self.o.set_option("kernels", kernels)
        if graph is not None:
            self.o.set_option("graph", graph)

    def close(self):
        self.o.close()

    def __getattr__(self, name):
        return getattr(self.o, name)

    def set_newton_f(self, arr):
        self.o.set(0, NEWTONF, arr)


def operator_script(b, mode):
    """apply every level operator once to seeded inputs; returns {name: array or float}"""
    out = {}
    nlev = b.num_levels
    assert nlev >= 2
    b.set(0, V, seeded_field(b.shape(0), 1))
    b.set(0, F, seeded_field(b.shape(0), 2, scale=1.0))
    if mode == 2:
        b.set(0, NEWTONV, seeded_field(b.shape(0), 3))
        b.set(1, NEWTONV, seeded_field(b.shape(1), 4))
    out["res0_norm"] = b.residual(0)
    out["res0_r"] = b.get(0, R)
    b.smooth(0, 1)
    out["smooth1_v"] = b.get(0, V)
    b.smooth(0, 2)
    out["smooth3_v"] = b.get(0, V)
    # the fused CUDA sweeps never store r (the reference's jacobi leaves the residual of the last-but-one
    # iterate there, which nothing reads: CpuSolver::vcycle recomputes it at :93); define r before using it
    out["res3_norm"] = b.residual(0)
    b.restrict(0, R, F)
    out["restrict_r_f1"] = b.get(1, F)
    b.restrict(0, V, V)
    out["restrict_v_v1"] = b.get(1, V)
    b.restrict_residual(0)
    out["restrict_residual_f1"] = b.get(1, F)
    if mode == 1:
        b.fas_coarse_rhs(0)
        out["fas_f1"] = b.get(1, F)
        out["fas_restv1"] = b.get(1, RESTV)
        out["fas_v1"] = b.get(1, V)
    b.smooth(1, 2)
    out["coarse_smooth_v1"] = b.get(1, V)
    out["prolong_e0"] = b.prolong(0)
    b.prolong_correct(0)
    out["correct_v0"] = b.get(0, V)
    if mode == 1:
        out["correct_v1"] = b.get(1, V)
    if mode == 2:
        b.set_newton_f(seeded_field(b.shape(0), 5, scale=1.0))
        out["compF_norm"] = b.newton_compF()
        out["compF_f0"] = b.get(0, F)
    return out


def full_size_script(b, mode, cycles, exact=None):
    """The V-cycle path of one mode on the analytic right-hand side, as CpuSolver::solve / NewtonSolver::findError drive
    it.  Modes 0/1: `cycles` V-cycles from v = 0.  Mode 2: a seeded Newton iterate on level 0, restricted down the
    hierarchy (src/cpu/NewtonSolver.cpp:88-92), F(newtonV) (:48-81), then `cycles` V-cycles of the linearised
    problem from v = 0 (:94-100).  Returns (norms, solution field id); with `exact` (a list) the oracle's compensated
    norm of the field each returned norm was reduced from is appended to it."""
    norms = []
    if mode == 2:
        b.set_newton_f(b.get(0, F))                       # NewtonSolver.cpp:12, newtonF = f_0
        b.set(0, NEWTONV, seeded_field(b.shape(0), 3))
        for l in range(b.num_levels - 2):                 # :88-92 (not onto the coarsest level)
            b.restrict(l, NEWTONV, NEWTONV)
        norms.append(b.newton_compF())
        if exact is not None:
            exact.append(b.norm_exact(0, F))
    for _ in range(cycles):
        norms.append(b.vcycle())
        if exact is not None:
            exact.append(b.norm_exact(0, R))
    return norms, V
