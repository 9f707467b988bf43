#!/usr/bin/env python
"""Where do two sweeps of the tuned path differ from the one-thread-per-point kernels?  (debug aid)
    python tools/debug_smooth2.py [--n 511] [--opt name=value ...]"""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as entry  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, nargs="+", default=[511])
ap.add_argument("--mode", type=int, default=0)
ap.add_argument("--opt", action="append", default=[])
a = ap.parse_args()
n = a.n[0] if len(a.n) == 1 else tuple(a.n)
gs = entry.load_package()
V, F = 0, 1
res = []
for kern in (0, 1):
    s = gs.GpuSolve(gs.make_params(n, mode=a.mode, pre=2, post=2, maxiter=1, tol=1e-30))
    s.set_option("kernels", kern)
    s.set_option("graph", 0)
    for o in a.opt:
        k, v = o.split("=")
        s.set_option(k, int(v))
    shape = s.shape(0)
    rs = np.random.RandomState(7)
    v = np.zeros(shape)
    v[1:-1, 1:-1, 1:-1] = rs.uniform(-0.05, 0.05, size=tuple(d - 2 for d in shape))
    s.set(0, V, v)
    s.smooth(0, 2)
    out = s.get(0, V).copy()
    r0 = s.vcycle()      # exercises the sweep+norm variant at the end of the cycle
    r1 = s.vcycle()
    res.append((out, r0, r1, s.get(0, V).copy()))
    s.close()
for name, i in (("smooth(0,2)", 0), ("after 2 V-cycles", 3)):
    d = res[0][i] != res[1][i]
    print(name, "mismatches:", int(d.sum()), "of", d.size)
    if d.any():
        z, y, x = np.nonzero(d)
        print("  z:", np.unique(z)[:40], "\n  y:", np.unique(y)[:40], "\n  x:", np.unique(x)[:40])
        print("  max abs diff", float(np.abs(res[0][i] - res[1][i]).max()))
print("norms simple", res[0][1], res[0][2], "tuned", res[1][1], res[1][2])
