"""Host-side model of the z-slab V-cycle data flow (DESIGN.md "Multi-GPU"), run by tests/test_slab_gloo.py with
one process per slab over the gloo backend.  Arithmetic is numpy fp64 in the reference's operation order
(src/cpu/CpuSolver.cpp:45-83, :141-179), so a slab result has to equal the whole-grid oracle bit for bit; what is
under test is the partition (gs_partition), the lazy two-plane ghost refresh and the norm reduction, i.e. the
protocol gs_api.cu drives with NCCL on the GPUs."""
import numpy as np

CANON = (6.0, -1.0, -1.0, -1.0, -1.0, -1.0, -1.0)


def _tap_sum(v, zb, ze, c):
    C = v[zb:ze, 1:-1, 1:-1]
    s = 0.0 + c[0] * C
    s = s + c[1] * v[zb:ze, 1:-1, 2:]
    s = s + c[2] * v[zb:ze, 1:-1, :-2]
    s = s + c[3] * v[zb:ze, 2:, 1:-1]
    s = s + c[4] * v[zb:ze, :-2, 1:-1]
    s = s + c[5] * v[zb + 1:ze + 1, 1:-1, 1:-1]
    s = s + c[6] * v[zb - 1:ze - 1, 1:-1, 1:-1]
    return C, s


def _residual(v, f, h, zb, ze, c, mode, gamma, aux):
    """(centre values, r, Jacobi denominator or None) on the interior of LOCAL planes [zb, ze):
    src/cpu/CpuSolver.cpp:56-74 and :161-171 in the operation order of gs_device.cuh finish_residual"""
    hh = h * h
    C, s = _tap_sum(v, zb, ze, c)
    s = s / hh
    den = None
    if mode == 1:
        ex = np.exp(C)
        s = s + (gamma * C) * ex
        den = c[0] / hh + (gamma * (1 + C)) * ex
    elif mode == 2:
        A = aux[zb:ze, 1:-1, 1:-1]
        ex = np.exp(A)
        g1 = gamma * (1 + A)
        s = s + (g1 * C) * ex
        den = c[0] / hh + g1 * ex
    return C, f[zb:ze, 1:-1, 1:-1] - s, den


def jacobi_sweep(v, f, h, omega, zb, ze, c=CANON, mode=0, gamma=1.0, aux=None):
    """one damped-Jacobi sweep of the interior points of LOCAL planes [zb, ze) of v ([z,y,x], halo included);
    returns (new array, sum r^2 over the updated points)"""
    C, r, den = _residual(v, f, h, zb, ze, c, mode, gamma, aux)
    out = v.copy()
    if mode == 0:
        alpha = (h * h) / c[0]
        out[zb:ze, 1:-1, 1:-1] = C + omega * (alpha * r)
    else:
        out[zb:ze, 1:-1, 1:-1] = C + omega * (r / den)
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


# ---------------------------------------------------------------------------------------------------------------
# whole V-cycle on slabs (linear mode): the data flow of gs_api.cu enqueue_vcycle with world > 1 — partitioned fine
# levels, the first replicated level gathered from the slabs, everything below computed by every rank
# ---------------------------------------------------------------------------------------------------------------
def residual_planes(v, f, h, zb, ze, c=CANON, mode=0, gamma=1.0, aux=None):
    """r = f - A v on the interior points of LOCAL planes [zb, ze), zero elsewhere (src/cpu/CpuSolver.cpp:45-83)"""
    r = np.zeros_like(v)
    r[zb:ze, 1:-1, 1:-1] = _residual(v, f, h, zb, ze, c, mode, gamma, aux)[1]
    return r


def fas_rhs_planes(restv, h, zb, ze, gamma, c=CANON):
    """A(restV)/hh + gamma*restV*exp(restV) on the interior of LOCAL planes [zb, ze) (src/cpu/CpuSolver.cpp:182-209)"""
    C, s = _tap_sum(restv, zb, ze, c)
    s = s / (h * h)
    return s + (gamma * C) * np.exp(C)


def restrict_planes(fine, z0f, cb, ce, cshape):
    """full weighting onto GLOBAL coarse planes [cb, ce) from a fine array whose local plane 0 is global plane z0f
    (src/cpu/CpuSolver.cpp:211-238: running sum, x offset outermost, z offset innermost)"""
    Yc, Xc = cshape
    X = 2 * np.arange(1, Xc - 1)
    Y = 2 * np.arange(1, Yc - 1)
    Z = 2 * np.arange(cb, ce) - z0f
    w = (0.5, 1.0, 0.5)
    acc = 0.0
    for ii in (-1, 0, 1):
        for jj in (-1, 0, 1):
            for kk in (-1, 0, 1):
                fac = 0.125 * w[ii + 1] * w[jj + 1] * w[kk + 1]
                acc = acc + fac * fine[np.ix_(Z + kk, Y + jj, X + ii)]
    out = np.zeros((ce - cb, Yc, Xc))
    out[:, 1:-1, 1:-1] = acc
    return out


def _interp_axis(a, axis, n):
    """coarse index -> fine indices 1..n along `axis`: identity on even, 0.5*a + 0.5*b on odd fine indices"""
    idx = np.arange(1, n + 1)
    lo = np.take(a, idx // 2, axis=axis)
    hi = np.take(a, idx // 2 + 1, axis=axis)
    shape = [1, 1, 1]
    shape[axis] = n
    odd = (idx % 2 == 1).reshape(shape)
    return np.where(odd, 0.5 * lo + 0.5 * hi, lo)


def prolong_planes(coarse, z0c, zb, ze, nyf, nxf):
    """e on GLOBAL fine planes [zb, ze) (interior x, y) from a coarse array whose local plane 0 is global plane z0c:
    nested averages x -> y -> z (src/cpu/CpuSolver.cpp:240-290 in closed form)"""
    ax = _interp_axis(coarse, 2, nxf)
    ay = _interp_axis(ax, 1, nyf)
    zg = np.arange(zb, ze)
    lo = ay[zg // 2 - z0c]
    hi = ay[np.minimum(zg // 2 + 1 - z0c, ay.shape[0] - 1)]
    odd = (zg % 2 == 1).reshape(-1, 1, 1)
    return np.where(odd, 0.5 * lo + 0.5 * hi, lo)


class MgLevel:
    """one level on one rank: a slab with two ghost planes per side (partitioned) or the whole array (replicated)"""

    def __init__(self, n, zlo, zhi, replicated):
        self.nz, self.ny, self.nx = n
        self.Z = self.nz + 2
        self.h = 1.0 / (self.ny + 1)
        self.replicated = replicated
        self.zlo, self.zhi = (0, self.Z) if replicated else (zlo, zhi)
        self.z0 = 0 if replicated else max(0, zlo - 2)
        self.z1 = self.Z if replicated else min(self.Z, zhi + 2)
        shape = (self.z1 - self.z0, self.ny + 2, self.nx + 2)
        self.v = np.zeros(shape)
        self.f = np.zeros(shape)
        self.restV = np.zeros(shape)   # mode 1
        self.nv = np.zeros(shape)      # mode 2: the frozen Newton iterate of the level
        self.gv = 2
        self.ib = max(1, self.zlo) - self.z0          # owned interior planes, local
        self.ie = min(self.nz + 1, self.zhi) - self.z0


class SlabMultigrid:
    def __init__(self, dist, rank, world, n, levels, rep, pre=2, post=2, omega=0.8, mode=0, gamma=1.0):
        """levels: [(zlo, zhi)] per level from gs_partition; rep: first replicated level"""
        self.dist, self.rank, self.world = dist, rank, world
        self.mode, self.gamma = mode, gamma
        self.pre, self.post, self.omega = pre, post, omega
        self.lev = []
        dims = (n, n, n) if isinstance(n, int) else tuple(n)[::-1]   # (nz, ny, nx)
        for l, (zlo, zhi) in enumerate(levels):
            self.lev.append(MgLevel(dims, zlo, zhi, l >= rep))
            dims = tuple(d // 2 for d in dims)
        self.exchanges = 0

    def part(self, l):
        return self.world > 1 and not self.lev[l].replicated

    def exchange(self, l, a, depth):
        """gs_comm_halo: `depth` planes each way with the neighbours that exist"""
        import torch
        L, dist, rank = self.lev[l], self.dist, self.rank
        lo, hi = L.zlo - L.z0, L.zhi - L.z0
        assert hi - lo >= depth
        ops, bufs = [], []
        if rank + 1 < self.world and L.zhi < L.Z:
            send = torch.from_numpy(np.ascontiguousarray(a[hi - depth:hi]))
            recv = torch.empty_like(send)
            ops += [dist.P2POp(dist.isend, send, rank + 1), dist.P2POp(dist.irecv, recv, rank + 1)]
            bufs.append((slice(hi, hi + depth), recv))
        if rank > 0 and L.zlo > 0:
            send = torch.from_numpy(np.ascontiguousarray(a[lo:lo + depth]))
            recv = torch.empty_like(send)
            ops += [dist.P2POp(dist.isend, send, rank - 1), dist.P2POp(dist.irecv, recv, rank - 1)]
            bufs.append((slice(lo - depth, lo), recv))
        if ops:
            for w in dist.batch_isend_irecv(ops):
                w.wait()
        for sl, recv in bufs:
            a[sl] = recv.numpy()
        self.exchanges += 1

    def need_v_halo(self, l, depth):
        L = self.lev[l]
        if not self.part(l) or L.gv >= depth:
            return
        d = 2 if L.zhi - L.zlo >= 2 else depth
        self.exchange(l, L.v, d)
        L.gv = d

    def smooth(self, l, iters):
        L = self.lev[l]
        while iters > 0:
            self.need_v_halo(l, 2 if iters >= 2 else 1)
            if iters >= 2 and (not self.part(l) or L.gv >= 2):
                # two sweeps per pass: the first also updates the ghost plane next to the slab (if it is interior)
                lo = L.ib - 1 if (self.part(l) and L.zlo > 1) else L.ib
                hi = L.ie + 1 if (self.part(l) and L.zhi < L.Z - 1) else L.ie
                kw = dict(mode=self.mode, gamma=self.gamma, aux=L.nv)
                mid, _ = jacobi_sweep(L.v, L.f, L.h, self.omega, lo, hi, **kw)
                L.v, _ = jacobi_sweep(mid, L.f, L.h, self.omega, L.ib, L.ie, **kw)
                iters -= 2
            else:
                L.v, _ = jacobi_sweep(L.v, L.f, L.h, self.omega, L.ib, L.ie, mode=self.mode, gamma=self.gamma, aux=L.nv)
                iters -= 1
            L.gv = 0

    def coarse_range(self, l):
        """GLOBAL coarse planes of level l+1 this rank computes when restricting from level l (gs_api.cu coarse_range)"""
        F, C = self.lev[l], self.lev[l + 1]
        if self.part(l) and C.replicated:
            return max(1, (F.zlo + 1) // 2), min(C.nz + 1, (F.zhi + 1) // 2)
        return C.ib + C.z0, C.ie + C.z0

    def restrict_residual(self, l):
        F, C = self.lev[l], self.lev[l + 1]
        cb, ce = self.coarse_range(l)

        def launch(b, e):
            """fused residual+restriction onto GLOBAL coarse planes [b, e): reads fine v planes 2b-2 .. 2e"""
            rb = max(1, 2 * b - 1) - F.z0
            re = min(F.nz + 1, 2 * e) - F.z0
            r = residual_planes(F.v, F.f, F.h, rb, re, mode=self.mode, gamma=self.gamma, aux=F.nv)
            C.f[b - C.z0:e - C.z0] = restrict_planes(r, F.z0, b, e, (C.ny + 2, C.nx + 2))

        if self.part(l):
            self.need_v_halo(l, 2)   # the lowest owned coarse plane needs r one fine plane below the slab
        launch(cb, ce)
        self.finish_coarse(l, C.f)

    def finish_coarse(self, l, coarse):
        """gs_api.cu finish_coarse: gather on the first replicated level, else one ghost plane per side"""
        C = self.lev[l + 1]
        if self.part(l) and C.replicated:
            self.gather_planes(l, coarse)
        elif self.part(l + 1):
            self.exchange(l + 1, coarse, 1)

    def fas_coarse_rhs(self, l):
        """gs_api.cu do_fas_coarse_rhs (src/cpu/CpuSolver.cpp:106-112): restV = v_coarse = R v; f_coarse += A(restV)"""
        F, C = self.lev[l], self.lev[l + 1]
        self.need_v_halo(l, 1)
        cb, ce = self.coarse_range(l)
        got = restrict_planes(F.v, F.z0, cb, ce, (C.ny + 2, C.nx + 2))
        C.restV[cb - C.z0:ce - C.z0] = got
        C.v[cb - C.z0:ce - C.z0] = got
        self.finish_coarse(l, C.restV)
        self.finish_coarse(l, C.v)
        C.gv = 1
        s = fas_rhs_planes(C.restV, C.h, C.ib, C.ie, self.gamma)
        C.f[C.ib:C.ie, 1:-1, 1:-1] = C.f[C.ib:C.ie, 1:-1, 1:-1] + s
        if self.part(l + 1):
            self.exchange(l + 1, C.f, 1)   # f changed on the owned planes only

    def gather_planes(self, l, coarse):
        """gs_comm_gather_planes: every rank broadcasts the coarse planes it computed"""
        import torch
        parts = [None] * self.world
        cb, ce = self.coarse_range(l)
        self.dist.all_gather_object(parts, (cb, ce, coarse[cb:ce].copy()))
        for b, e, a in parts:
            if e > b:
                coarse[b:e] = a

    def prolong_correct(self, l):
        F, C = self.lev[l], self.lev[l + 1]
        self.need_v_halo(l + 1, 1)   # the interpolation reads one coarse plane outside the slab
        if self.mode == 1:
            C.v = C.v - C.restV      # all local planes; restV carries one current ghost plane
            C.gv = min(C.gv, 1)
        zb, ze = F.ib + F.z0, F.ie + F.z0
        e = prolong_planes(C.v, C.z0, zb, ze, F.ny, F.nx)
        F.v[F.ib:F.ie, 1:-1, 1:-1] = F.v[F.ib:F.ie, 1:-1, 1:-1] + e
        F.gv = 0

    def vcycle(self):
        """gs_api.cu enqueue_vcycle, linear mode; returns the sum of squares of the final level-0 residual"""
        import torch
        last = len(self.lev) - 1
        for i in range(last):
            self.smooth(i, self.pre)
            self.restrict_residual(i)
            C = self.lev[i + 1]
            if self.mode != 1:
                C.v[...] = 0.0
                C.gv = 2
            else:
                self.fas_coarse_rhs(i)
        self.smooth(last, self.pre + self.post)
        for i in range(last, 0, -1):
            self.prolong_correct(i - 1)
            self.smooth(i - 1, self.post)
        L = self.lev[0]
        self.need_v_halo(0, 1)
        r = residual_planes(L.v, L.f, L.h, L.ib, L.ie, mode=self.mode, gamma=self.gamma, aux=L.nv)
        t = torch.tensor([float(np.sum(r * r))], dtype=torch.float64)
        if self.part(0):
            self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM)
        return t.item()


def vcycle_worker(rank, world, port, n, seed, cycles, out_path, mode=0):
    """gloo process `rank`: `cycles` V(2,2) cycles of a seeded level-0 problem on slabs; rank 0 saves the gathered v"""
    import os
    import sys
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
        p = gs.make_params(n)
        nlev = int(np.floor(np.log(n) / np.log(2.0))) + 1
        levels, rep = [], 0
        for l in range(nlev):
            zlo, zhi, rep = gs.partition(p, world, rank, l)
            levels.append((zlo, zhi))
        mg = SlabMultigrid(dist, rank, world, n, levels, rep, mode=mode)
        shape = (n + 2, n + 2, n + 2)
        v = seeded_field(shape, seed)
        f = seeded_field(shape, seed + 1, scale=1.0)
        L = mg.lev[0]
        L.v[...] = v[L.z0:L.z1]
        L.f[...] = f[L.z0:L.z1]
        if mode == 2:   # a frozen Newton iterate on every level, ghost planes included (gs_set_field)
            for l, M in enumerate(mg.lev):
                M.nv[...] = seeded_field((M.Z, M.ny + 2, M.nx + 2), seed + 10 + l)[M.z0:M.z1]
        sums = [mg.vcycle() for _ in range(cycles)]
        parts = [None] * world
        dist.all_gather_object(parts, (L.zlo, L.zhi, L.v[L.zlo - L.z0:L.zhi - L.z0]))
        if rank == 0:
            parts.sort(key=lambda x: x[0])
            np.savez(out_path, v=np.concatenate([x[2] for x in parts], axis=0), sums=np.array(sums), rep=rep,
                     exchanges=mg.exchanges)
    finally:
        dist.destroy_process_group()
