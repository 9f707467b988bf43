#!/usr/bin/env python
"""One V-cycle with plain launches (no CUDA graph) — the command profiled under ncu.
    python tools/one_cycle.py [--n 511] [--mode 0] [--pre 2] [--post 2] [--cycles 1] [--opt name=value ...]"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as entry  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=511)
ap.add_argument("--mode", type=int, default=0)
ap.add_argument("--pre", type=int, default=2)
ap.add_argument("--post", type=int, default=2)
ap.add_argument("--cycles", type=int, default=1)
ap.add_argument("--graph", type=int, default=0)
ap.add_argument("--opt", action="append", default=[])
ap.add_argument("--no-newton-step", action="store_true", help="mode 2: cycle on the zero Newton iterate (same kernels, no inner solve first)")
a = ap.parse_args()
gs = entry.load_package()
s = gs.GpuSolve(gs.make_params(a.n, mode=a.mode, pre=a.pre, post=a.post, maxiter=1, tol=1e-30))
s.set_option("graph", a.graph)
for o in a.opt:
    k, v = o.split("=")
    s.set_option(k, int(v))
if a.mode == 2 and not a.no_newton_step:
    print("newton step:", s.newton_step())
for _ in range(a.cycles):
    print("norm", s.vcycle())
print("launches", s.launches)
