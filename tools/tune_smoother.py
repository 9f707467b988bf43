#!/usr/bin/env python
"""Rank the smoother kernel variants of kernels_fast.cu on the GPU box: per-sweep time (CUDA events on
the library stream), GB/s at 24 B/LUP (32 in Newton mode), and a bit-for-bit check against the simple kernel.
    python tools/tune_smoother.py [--n 511] [--mode 0] [--variants 0,1,2,...] [--zchunks 32,64,128]"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as entry  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=511)
    ap.add_argument("--mode", type=int, default=0)
    ap.add_argument("--variants", default="0,1,11,12,13,2,21,22,23,24,25")
    ap.add_argument("--zchunks", default="64")
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    import numpy as np
    import torch
    gs = entry.load_package()
    p = gs.make_params(args.n, mode=args.mode, pre=2, post=2, maxiter=1, tol=1e-30)
    s = gs.GpuSolve(p)
    stream = torch.cuda.ExternalStream(s.stream)
    s.vcycle()  # realistic field values
    v0 = s.get(0, gs.V)
    (nx, ny, nz), _ = s.level_info(0)
    pts = nx * ny * nz
    bpp = 32 if args.mode == 2 else 24
    ref = None
    rows = []
    for var in [int(x) for x in args.variants.split(",")]:
        for zc in [int(x) for x in args.zchunks.split(",")]:
            if var == 0 and zc != int(args.zchunks.split(",")[0]):
                continue
            s.set_option("kernels", 0 if var == 0 else 1)
            s.set_option("smooth_variant", var)
            s.set_option("zchunk", zc)
            s.set(0, gs.V, v0)
            s.smooth(0, 2)
            out = s.get(0, gs.V)
            if ref is None:
                ref = out
            same = bool(np.array_equal(out, ref))
            s.smooth(0, 3)
            s.synchronize()
            best, tot = 1e9, 0.0
            for _ in range(3):
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record(stream)
                s.smooth(0, args.reps)
                b.record(stream)
                s.synchronize()
                ms = a.elapsed_time(b) / args.reps
                best = min(best, ms)
                tot += ms
            row = {"variant": var, "zchunk": zc, "ms_best": best, "ms_mean": tot / 3, "gbs": pts * bpp / best / 1e6,
                   "glups": pts / best / 1e6, "bitwise_equal_to_simple": same}
            rows.append(row)
            print(json.dumps(row), flush=True)
    if args.out:
        with open(args.out, "w") as f:
            json.dump({"n": args.n, "mode": args.mode, "rows": rows}, f, indent=1)


if __name__ == "__main__":
    main()
