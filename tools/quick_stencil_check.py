#!/usr/bin/env python
"""Few-second check of the single-sweep TMA kernels (stencil_tma): V(3,3) cycles and odd sweep counts, tuned vs
one-thread-per-point kernels, bit for bit.   python tools/quick_stencil_check.py"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as entry  # noqa: E402

gs = entry.load_package()
ok = True
for n in (63, (100, 90, 47), 127):
    out = []
    for kern in (0, 1):
        s = gs.GpuSolve(gs.make_params(n, mode=0, pre=3, post=3, maxiter=1, tol=1e-30))
        s.set_option("kernels", kern)
        s.set_option("graph", 0)
        rs = np.random.RandomState(5)
        sh = s.shape(0)
        v = np.zeros(sh)
        v[1:-1, 1:-1, 1:-1] = rs.uniform(-0.05, 0.05, size=tuple(d - 2 for d in sh))
        s.set(0, 0, v)
        s.smooth(0, 3)
        norms = [s.vcycle() for _ in range(3)]
        r = s.residual(0)
        out.append((s.get(0, 0).copy(), norms, r))
        s.close()
    same = bool(np.array_equal(out[0][0], out[1][0]))
    print(n, "fields equal", same, "norm rel", abs(out[0][2] - out[1][2]) / out[0][2])
    ok = ok and same
print("OK" if ok else "MISMATCH")
sys.exit(0 if ok else 1)
