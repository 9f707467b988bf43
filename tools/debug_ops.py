#!/usr/bin/env python
"""Step a V-cycle operator by operator through the C ABI on two handles (tuned / one-thread-per-point kernels) and
report the first operator whose output differs (debug aid).   python tools/debug_ops.py [--n 255] [--reps 3] [--opt k=v]"""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as entry  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=255)
ap.add_argument("--reps", type=int, default=3)
ap.add_argument("--cmp", type=int, default=1, help="0: only compare at the end of the cycle (no host sync between operators)")
ap.add_argument("--opt", action="append", default=[])
a = ap.parse_args()
gs = entry.load_package()
V, F = 0, 1
hs = []
for kern in (0, 1):
    s = gs.GpuSolve(gs.make_params(a.n, mode=0, pre=2, post=2, maxiter=1, tol=1e-30))
    s.set_option("kernels", kern)
    s.set_option("graph", 0)
    if kern:
        for o in a.opt:
            k, v = o.split("=")
            s.set_option(k, int(v))
    hs.append(s)
nlev = hs[0].num_levels
rs = np.random.RandomState(3)
shape = hs[0].shape(0)
v0 = np.zeros(shape)
v0[1:-1, 1:-1, 1:-1] = rs.uniform(-0.05, 0.05, size=tuple(d - 2 for d in shape))


def check(tag, level, field):
    if not a.cmp and tag != "end":
        return True
    x, y = hs[0].get(level, field), hs[1].get(level, field)
    d = x != y
    if d.any():
        z, yy, xx = np.nonzero(d)
        print(f"  MISMATCH after {tag}: {int(d.sum())} points; z {np.unique(z)[:12]} y {np.unique(yy)[:12]} x {np.unique(xx)[:12]} "
              f"max abs {float(np.abs(x - y).max()):.3e}")
        return False
    return True


for rep in range(a.reps):
    print("rep", rep)
    for s in hs:
        s.set(0, V, v0)
    ok = True
    for l in range(nlev - 1):
        for s in hs:
            s.smooth(l, 2)
        ok = ok and check(f"smooth L{l} pre", l, V)
        for s in hs:
            s.restrict_residual(l)
        ok = ok and check(f"resid+restrict L{l}", l + 1, F)
        for s in hs:
            s.set(l + 1, V, np.zeros(s.shape(l + 1)))
    for s in hs:
        s.smooth(nlev - 1, 4)
    ok = ok and check("coarsest", nlev - 1, V)
    for l in range(nlev - 2, -1, -1):
        for s in hs:
            s.prolong_correct(l)
        ok = ok and check(f"prolong L{l}", l, V)
        for s in hs:
            s.smooth(l, 2)
        ok = ok and check(f"smooth L{l} post", l, V)
    check("end", 0, V)
    if not a.cmp:
        a.cmp = 1
        for l in range(1, nlev):
            check(f"end: f of L{l}", l, F)
        a.cmp = 0
    n0, n1 = hs[0].vcycle(), hs[1].vcycle()
    print("  whole-cycle norms", n0, n1, "fields equal:", bool((hs[0].get(0, V) == hs[1].get(0, V)).all()))
