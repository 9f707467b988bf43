#!/usr/bin/env python
"""Fused residual+restriction against the unfused one-thread-per-point kernels, repeated (debug aid for races).
    python tools/debug_rr.py [--n 511] [--reps 4] [--opt name=value ...]"""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as entry  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, nargs="+", default=[511])
ap.add_argument("--reps", type=int, default=4)
ap.add_argument("--level", type=int, default=0)
ap.add_argument("--opt", action="append", default=[])
a = ap.parse_args()
n = a.n[0] if len(a.n) == 1 else tuple(a.n)
gs = entry.load_package()
V, F = 0, 1
s = gs.GpuSolve(gs.make_params(n, mode=0, pre=2, post=2, maxiter=1, tol=1e-30))
s.set_option("graph", 0)
for o in a.opt:
    k, v = o.split("=")
    s.set_option(k, int(v))
L = a.level
shape = s.shape(L)
rs = np.random.RandomState(7)
v = np.zeros(shape)
v[1:-1, 1:-1, 1:-1] = rs.uniform(-0.05, 0.05, size=tuple(d - 2 for d in shape))
f = np.zeros(shape)
f[1:-1, 1:-1, 1:-1] = rs.uniform(-1, 1, size=tuple(d - 2 for d in shape))
s.set(L, V, v)
s.set(L, F, f)
s.set_option("kernels", 0)
s.restrict_residual(L)
ref = s.get(L + 1, F).copy()
s.set_option("kernels", 1)
for r in range(a.reps):
    s.set(L + 1, F, np.zeros(s.shape(L + 1)))
    s.restrict_residual(L)
    got = s.get(L + 1, F)
    d = got != ref
    print("rep", r, "mismatches", int(d.sum()), "of", d.size)
    if d.any():
        z, y, x = np.nonzero(d)
        print("  z:", np.unique(z)[:48], "\n  y:", np.unique(y)[:48], "\n  x:", np.unique(x)[:48], "\n  max abs", float(np.abs(got - ref).max()))
s.close()
