#!/usr/bin/env python
"""Multi-GPU parity: the z-slab solver under torchrun must give the single-GPU fields bit for bit.
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 \
        tools/check_multigpu.py [--grid 127] [--mode 0] [--cycles 3]
Rank 0 prints one line per check and exits non-zero on a mismatch."""
import argparse
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as entry  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--grid", type=int, nargs="+", default=[127])
    ap.add_argument("--mode", type=int, default=0)
    ap.add_argument("--pre", type=int, default=2)
    ap.add_argument("--post", type=int, default=2)
    ap.add_argument("--cycles", type=int, default=3)
    args = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    gs = entry.load_package()
    box = [gs.comm_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(box, src=0)
    n = tuple(args.grid) if len(args.grid) == 3 else args.grid[0]
    p = gs.make_params(n, mode=args.mode, pre=args.pre, post=args.post, maxiter=args.cycles, tol=1e-30)
    slab = gs.GpuSolve(p, device=local, rank=rank, world=world, nccl_id=box[0])
    zb, ze, rep = gs.partition(p, world, rank, 0)
    bad = 0
    field = gs.NEWTONV if args.mode == 2 else gs.V
    norms = []
    if args.mode == 2:
        for _ in range(args.cycles):
            norms.append(slab.newton_step()[0])
    else:
        for _ in range(args.cycles):
            norms.append(slab.vcycle())
    got = slab.get(0, field)          # collective: every rank receives the whole array
    got1 = slab.get(1, field if args.mode == 2 else gs.V)
    if rank == 0:
        one = gs.GpuSolve(p, device=local)
        one.set_option("fuse_rv", 0)   # slab handles form R(v) with the separate, bit-exact restriction (gs_api.cu do_restrict_residual)
        ref = []
        if args.mode == 2:
            for _ in range(args.cycles):
                ref.append(one.newton_step()[0])
        else:
            for _ in range(args.cycles):
                ref.append(one.vcycle())
        want, want1 = one.get(0, field), one.get(1, field if args.mode == 2 else gs.V)
        same = bool(np.array_equal(got, want)) and bool(np.array_equal(got1, want1))
        nerr = max(abs(a - b) / abs(b) for a, b in zip(norms, ref))
        print(f"n={n} mode={args.mode} world={world} slabs from level {rep} replicated: fields bitwise {'OK' if same else 'MISMATCH'} "
              f"(max abs diff {float(np.max(np.abs(got - want))):.3e}); norms rel diff {nerr:.2e}: {norms[-1]!r} vs {ref[-1]!r}", flush=True)
        if not same or nerr > 1e-12:
            bad = 1
        one.close()
    slab.close()
    t = torch.tensor([bad], device="cuda")
    dist.all_reduce(t)
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(int(t.item() != 0))


if __name__ == "__main__":
    main()
