#!/usr/bin/env python
"""Under torchrun: time the ghost-plane refresh alone on every partitioned level of the bench grid (both transports).
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29521 tools/time_exchange.py [--grid 1023]"""
import argparse
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as entry  # noqa: E402
import bench  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--grid", type=int, nargs="+", default=None)
a = ap.parse_args()
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
gs = entry.load_package()
n = tuple(a.grid) if a.grid and len(a.grid) == 3 else ((a.grid[0],) * 3 if a.grid else bench.default_grid(world))
for p2p in (1, 0):
    os.environ["GPUSOLVE_P2P"] = str(p2p)
    box = [gs.comm_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(box, src=0)
    p = gs.make_params(n, mode=0, pre=2, post=2, maxiter=1, tol=1e-30)
    s = gs.GpuSolve(p, device=local, rank=rank, world=world, nccl_id=box[0])
    s.vcycle()
    rep = gs.partition(p, world, rank, 0)[2]
    for level in range(rep):
        dims = s.level_dims(level) if hasattr(s, "level_dims") else None
        for depth in (2, 1):
            ms = s.time_halo_exchange(level, depth, 50)
            t = torch.tensor([ms], device="cuda", dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            if rank == 0:
                shp = s.shape(level)
                mb = depth * shp[-1] * shp[-2] * 8 / 1e6   # logical bytes per direction and neighbour (padding not counted)
                print(f"world {world} grid {n} transport {'peer' if p2p else 'nccl'} level {level} depth {depth}: {1e3 * t.item():8.1f} us per exchange, "
                      f"{mb:7.2f} MB per direction -> {mb / t.item() / 1e3 * 1e3:7.1f} GB/s per direction", flush=True)
    s.close()
dist.barrier()
dist.destroy_process_group()
