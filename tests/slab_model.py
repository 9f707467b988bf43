"""Host-side model of the z-slab V-cycle data flow (DESIGN.md "Multi-GPU"), run by tests/test_slab_gloo.py with
one process per slab over the gloo backend.  Arithmetic is numpy fp64 in the reference's operation order
(src/cpu/CpuSolver.cpp:45-83, :141-179), so a slab result has to equal the whole-grid oracle bit for bit; what is
under test is the partition (gs_partition), the lazy two-plane ghost refresh and the norm reduction, i.e. the
protocol gs_api.cu drives with NCCL on the GPUs."""
import numpy as np

CANON = (6.0, -1.0, -1.0, -1.0, -1.0, -1.0, -1.0)


def jacobi_sweep(v, f, h, omega, zb, ze, c=CANON):
    """one damped-Jacobi sweep of the interior points of LOCAL planes [zb, ze) of v ([z,y,x], halo included);
    returns (new array, sum r^2 over the updated points)"""
    hh = h * h
    alpha = hh / c[0]
    C = v[zb:ze, 1:-1, 1:-1]
    s = 0.0 + c[0] * C
    s = s + c[1] * v[zb:ze, 1:-1, 2:]
    s = s + c[2] * v[zb:ze, 1:-1, :-2]
    s = s + c[3] * v[zb:ze, 2:, 1:-1]
    s = s + c[4] * v[zb:ze, :-2, 1:-1]
    s = s + c[5] * v[zb + 1:ze + 1, 1:-1, 1:-1]
    s = s + c[6] * v[zb - 1:ze - 1, 1:-1, 1:-1]
    r = f[zb:ze, 1:-1, 1:-1] - s / hh
    out = v.copy()
    out[zb:ze, 1:-1, 1:-1] = C + omega * (alpha * r)
    return out, float(np.sum(r * r))


class Slab:
    """global planes [z0, z0 + nzl) of a level: the owned planes [zlo, zhi) plus up to two ghost planes per side"""
    GHOST = 2

    def __init__(self, full_v, full_f, zlo, zhi):
        Z = full_v.shape[0]
        self.Z, self.zlo, self.zhi = Z, zlo, zhi
        self.z0 = max(0, zlo - self.GHOST)
        self.z1 = min(Z, zhi + self.GHOST)
        self.v = full_v[self.z0:self.z1].copy()
        self.f = full_f[self.z0:self.z1].copy()
        self.gv = self.GHOST  # ghost planes of v that are current

    def local(self, zg):
        return zg - self.z0

    def owned_interior(self):
        return self.local(max(1, self.zlo)), self.local(min(self.Z - 1, self.zhi))

    def exchange(self, dist, rank, world, depth):
        """refresh `depth` ghost planes of v on both sides from the neighbouring slabs (gs_comm_halo)"""
        import torch
        ops, bufs = [], []
        if rank > 0:
            send = torch.from_numpy(np.ascontiguousarray(self.v[self.local(self.zlo):self.local(self.zlo) + depth]))
            recv = torch.empty_like(send)
            ops += [dist.P2POp(dist.isend, send, rank - 1), dist.P2POp(dist.irecv, recv, rank - 1)]
            bufs.append(("lo", recv))
        if rank < world - 1:
            send = torch.from_numpy(np.ascontiguousarray(self.v[self.local(self.zhi) - depth:self.local(self.zhi)]))
            recv = torch.empty_like(send)
            ops += [dist.P2POp(dist.isend, send, rank + 1), dist.P2POp(dist.irecv, recv, rank + 1)]
            bufs.append(("hi", recv))
        for w in dist.batch_isend_irecv(ops):
            w.wait()
        for side, recv in bufs:
            if side == "lo":
                self.v[self.local(self.zlo) - depth:self.local(self.zlo)] = recv.numpy()
            else:
                self.v[self.local(self.zhi):self.local(self.zhi) + depth] = recv.numpy()
        self.gv = depth

    def smooth2(self, h, omega):
        """two sweeps from ghosts two planes deep without an exchange in between (kx_smooth2 on a slab): the first
        sweep also updates the inner ghost plane, redundantly with its owner"""
        assert self.gv >= 2
        ib, ie = self.owned_interior()
        lo = ib - 1 if self.zlo > 1 else ib
        hi = ie + 1 if self.zhi < self.Z - 1 else ie
        mid, _ = jacobi_sweep(self.v, self.f, h, omega, lo, hi)
        self.v, sq = jacobi_sweep(mid, self.f, h, omega, ib, ie)
        self.gv = 0
        return sq

    def smooth1(self, h, omega):
        assert self.gv >= 1
        ib, ie = self.owned_interior()
        self.v, sq = jacobi_sweep(self.v, self.f, h, omega, ib, ie)
        self.gv = 0
        return sq

    def owned(self):
        return self.v[self.local(self.zlo):self.local(self.zhi)]


def worker(rank, world, port, n, seed, out_path):
    """gloo process `rank`: slab sweeps of a seeded level-0 problem; rank 0 gathers the slabs and saves them"""
    import os
    import sys
    import torch
    import torch.distributed as dist
    here = os.path.dirname(os.path.abspath(__file__))
    for p in (os.path.dirname(here), here):
        if p not in sys.path:
            sys.path.insert(0, p)
    import __graft_entry__ as entry
    from cases import seeded_field
    gs = entry.load_package()
    gs.load()
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    try:
        # the communicator id travels like bench.py sends the NCCL id: rank 0 makes it, everybody receives it
        box = [bytes(range(128)) if rank == 0 else None]
        dist.broadcast_object_list(box, src=0)
        assert box[0] == bytes(range(128))
        p = gs.make_params(n)
        zlo, zhi, rep = gs.partition(p, world, rank, 0)
        assert rep >= 1, "level 0 must be partitioned for this test"
        shape = (n + 2, n + 2, n + 2)
        v = seeded_field(shape, seed)
        f = seeded_field(shape, seed + 1, scale=1.0)
        h = 1.0 / (n + 1)
        s = Slab(v, f, zlo, zhi)
        s.smooth2(h, 0.8)                      # ghosts current from the initial copy
        s.exchange(dist, rank, world, 2)       # lazy refresh: once, two planes deep, before the next reader
        s.smooth2(h, 0.8)
        s.exchange(dist, rank, world, 1)
        sq = s.smooth1(h, 0.8)                 # its residual is the norm of the iterate after four sweeps
        t = torch.tensor([sq], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        ms = torch.tensor([float(rank + 1)], dtype=torch.float64)   # bench.py: max over ranks of the device time
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        assert ms.item() == float(world)
        parts = [None] * world
        dist.all_gather_object(parts, (zlo, zhi, s.owned()))
        if rank == 0:
            parts.sort(key=lambda x: x[0])
            assert parts[0][0] == 0 and parts[-1][1] == n + 2
            for a, b in zip(parts, parts[1:]):
                assert a[1] == b[0], "slabs must tile the level"
            np.savez(out_path, v=np.concatenate([x[2] for x in parts], axis=0), sumsq=t.item())
    finally:
        dist.destroy_process_group()
