#!/usr/bin/env python
"""Per-level / per-operator device time of the V-cycle (CUDA events on the library stream, graph off), with
the achieved GB/s against each operator's algorithmic bytes.
    python tools/profile_levels.py [--n 511] [--mode 0] [--pre 2] [--post 2] [--reps 10] [--out file.json]"""
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
    ap.add_argument("--pre", type=int, default=2)
    ap.add_argument("--post", type=int, default=2)
    ap.add_argument("--reps", type=int, default=10)
    ap.add_argument("--out", default="")
    ap.add_argument("--opt", action="append", default=[], help="name=value passed to gs_set_option")
    args = ap.parse_args()
    gs = entry.load_package()
    p = gs.make_params(args.n, mode=args.mode, pre=args.pre, post=args.post, maxiter=1, tol=1e-30)
    s = gs.GpuSolve(p)
    for o in args.opt:
        k, v = o.split("=")
        s.set_option(k, int(v))
    if args.mode == 2:
        s.newton_step()
    for _ in range(3):
        s.vcycle()
    s.set_option("timing", 1)
    s.reset_timing()
    for _ in range(args.reps):
        s.vcycle()
    sweep = 32 if args.mode == 2 else 24
    rr = 25 if args.mode == 2 else 17
    rows, total = [], 0.0
    nlev = s.num_levels
    for l in range(nlev):
        (nx, ny, nz), _ = s.level_info(l)
        pts = nx * ny * nz
        t = s.level_timing(l)
        for op, (ms, launches) in t.items():
            if launches == 0 and ms == 0:
                continue
            ms /= args.reps
            launches /= args.reps
            if op == "smooth":
                nsw = (args.pre + args.post) if l < nlev - 1 else (args.pre + args.post)
                byt = sweep * pts * nsw
            elif op == "residual":
                byt = (sweep - 8) * pts
            elif op == "restrict":
                byt = rr * pts
            elif op == "prolong":
                byt = 17 * pts
            else:
                byt = 0
            rows.append({"level": l, "n": [nx, ny, nz], "op": op, "ms": ms, "launches": launches, "alg_bytes": byt,
                         "gbs": byt / ms / 1e6 if ms > 0 else None})
            total += ms
    s.set_option("timing", 0)
    # whole cycle, graph on
    import torch
    stream = torch.cuda.ExternalStream(s.stream)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.vcycle()
    a.record(stream)
    for _ in range(args.reps):
        s.vcycle()
    b.record(stream)
    s.synchronize()
    cyc = a.elapsed_time(b) / args.reps
    print(f"n={args.n} mode={args.mode} V({args.pre},{args.post}): sum of operator times {total:.3f} ms, graph V-cycle {cyc:.3f} ms")
    for r in rows:
        g = f"{r['gbs']:8.0f} GB/s" if r["gbs"] else ""
        print(f"  L{r['level']} {str(r['n']):18s} {r['op']:9s} {r['ms']:8.4f} ms  x{r['launches']:.0f}  {g}")
    if args.out:
        with open(args.out, "w") as f:
            json.dump({"n": args.n, "mode": args.mode, "pre": args.pre, "post": args.post, "vcycle_ms_graph": cyc,
                       "sum_ops_ms": total, "rows": rows}, f, indent=1)


if __name__ == "__main__":
    main()
