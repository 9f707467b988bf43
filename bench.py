#!/usr/bin/env python
"""bench.py — GpuSolve V-cycle hot path on B200: V-cycles/s at 513^3 fp64 (config n=511, linear
V(2,2), BASELINE.json configs[1]) with the fine-level smoother against the HBM roofline.

  python bench.py --gpus N --steps K --warmup W            the CUDA backend (C ABI, lib/libgpusolve_cuda.so)
  python bench.py --impl reference ...                      the reference's own CPU path on the host cores

A step is one V(2,2) cycle of the hot path (including the per-cycle residual norm the reference's
solve loop reads back).  One JSON line on stdout (rank 0).

N = 1: the 513^3 grid.  N > 1 (one process per GPU under torchrun): WEAK scaling by default — the same lattice
points per GPU (511x511x1023, 511x1023x1023, 1023^3 at N = 2, 4, 8), split into z-slabs; `value` is then the whole
job's throughput in 513^3-EQUIVALENT V-cycles/s (cycles/s on the grid x its lattice points / those of 513^3) and the
metric string says so.  `--grid 1023` gives the strong-scaling run of BASELINE.json configs[4]; every N > 1 line also
carries a `strong_1023` sub-record (1023^3 on N GPUs against 1023^3 on one GPU of the same box) and a `slab_parity`
record (slab run == one-GPU run bit for bit, checked before anything is timed).

What the fractions mean (VERDICT r01 "weak" 3): `roofline.frac` is a BANDWIDTH fraction — the compulsory bytes of one
pass of the dominant kernel (read v, read f, write v'' once: 24 B per point, 32 in mode 2) / its time / the measured
copy peak — and is <= ~1.  That a pass applies two Jacobi sweeps shows up in `glups` and `lup_equivalent_gbs`
(24 B per lattice UPDATE / time), not in `frac`.  `vcycle.frac_of_peak` is per GPU and counts the passes launched.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry  # noqa: E402


def interior_points(n, levels=None):
    """interior points per level of the reference hierarchy (src/cpu/CpuGridData.cpp:19-29)"""
    import math
    if isinstance(n, int):
        n = (n, n, n)
    nlev = int(math.floor(math.log(min(n)) / math.log(2.0))) + 1
    out, d = [], list(n)
    for _ in range(nlev):
        out.append(d[0] * d[1] * d[2])
        d = [x // 2 for x in d]
    return out


def vcycle_algorithmic_bytes(n, pre, post, mode=0):
    """B_V = sum_{l<L-1} P_l [24(v1+v2) + 17 + 17] + P_{L-1} 24 (v1+v2) + 24 P_0 (mode 2: 32/25).
    One pass per operator the REFERENCE applies — the figure the survey's bound is quoted on."""
    P = interior_points(n)
    sweep = 32 if mode == 2 else 24
    rr = 25 if mode == 2 else 17
    b = sum(p * (sweep * (pre + post) + rr + 17) for p in P[:-1]) + P[-1] * sweep * (pre + post) + sweep * P[0]
    return b


def pass_level_bytes(n, mode, level_launches):
    """Compulsory traffic of the passes actually LAUNCHED in one cycle: every smoother / residual-norm launch reads v
    and f (and the Newton coefficient) and writes v once (24 / 32 B per point however many sweeps it applies), the
    fused residual+restriction 17 / 25 B per fine point, prolongation+correction 17.  level_launches[l] = {op:
    launches per cycle}."""
    P = interior_points(n)
    sweep = 32 if mode == 2 else 24
    rr = 25 if mode == 2 else 17
    b = 0.0
    for l, ops in enumerate(level_launches):
        b += P[l] * sweep * (ops.get("smooth", 0) + ops.get("residual", 0))
        b += P[l] * rr * min(ops.get("restrict", 0), 1)
        b += P[l] * 17 * min(ops.get("prolong", 0), 1)
    return b


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return json.load(f).get("hbm_gbs", 6650.0), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md: 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons while the timed region runs"""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.rows = []
        self.proc = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(gpu_index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), line.strip()))

    def stop(self, t0, t1):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], None, set()
        for ts, line in self.rows:
            parts = [p.strip() for p in line.split(",")]
            if len(parts) < 8:
                continue
            if t0 - 0.05 <= ts <= t1 + 0.15:
                try:
                    sm.append(float(parts[1]))
                    mx = float(parts[2])
                except ValueError:
                    continue
                for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), parts[4:8]):
                    if val.lower().startswith("active"):
                        reasons.add(name)
        if not sm:  # region shorter than the sampling period: take every sample seen
            for ts, line in self.rows:
                parts = [p.strip() for p in line.split(",")]
                try:
                    sm.append(float(parts[1]))
                    mx = float(parts[2])
                except (ValueError, IndexError):
                    pass
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons), "samples": len(sm)}


# ---------------------------------------------------------------------------------------------------------------------
# the reference's CPU path on the host cores (oracle/_ref when built, else the C port)
# ---------------------------------------------------------------------------------------------------------------------
def host_ram_gb():
    try:
        with open("/proc/meminfo") as f:
            for line in f:
                if line.startswith("MemAvailable:"):
                    return int(line.split()[1]) / 1e6
    except OSError:
        pass
    return 0.0


def cpu_grid_ram_gb(n):
    """six (n+2)^3 arrays per level (src/cpu/CpuGridData.cpp:32-39) plus newtonF and slack"""
    return 8.0 * (n[0] + 2) * (n[1] + 2) * (n[2] + 2) * 6 * 1.15 / 1e9


def cpu_sample_grid(n, steps_total, cores, budget_s, per_cycle_511_8cores=3.6):
    """the largest grid, obtained by halving the target like the hierarchy does, whose CPU run fits the time budget
    (one 511^3 V(2,2) cycle of the LTO-built reference is ~3.6 s on 8 cores)"""
    n = tuple(n)
    est = per_cycle_511_8cores * (n[0] * n[1] * n[2]) / 511.0 ** 3 * 8.0 / max(cores, 1) * steps_total
    while min(n) > 63 and est > budget_s:
        n = tuple((d + 1) // 2 - 1 for d in n)
        est /= 8.0
    return n


def cpu_oracle():
    orc = entry.load_oracle()
    kind = "reference" if orc.available("reference") else "port"
    if kind == "port" and not orc.available("port"):
        orc.build(ref=False)
    return orc, kind


def run_cpu(n, pre, post, steps, warm, budget_s, mode=0):
    """Times the reference's CPU V-cycle on all host threads.  Never shrinks the grid silently: if `steps + warm` cycles
    of grid n do not fit the budget it first runs FEWER cycles of the stated grid (at least one warm-up and one timed),
    and only if even that does not fit (or the grid does not fit host memory) it runs a smaller grid and says so
    (`extrapolated`).  Returns a dict."""
    orc, kind = cpu_oracle()
    cores = os.cpu_count()
    n = tuple(n)
    # calibrate the cost per lattice point on a grid that takes a fraction of a second
    cal = cpu_sample_grid(n, 2, cores, 1.5)
    o = orc.Oracle(orc.make_params(cal, mode=mode, pre=pre, post=post, maxiter=1, tol=1e-30), kind)
    o.set_threads(cores)
    o.vcycle()
    t0 = time.time()
    o.vcycle()
    per_pt = (time.time() - t0) / sum(interior_points(cal))
    o.close()
    est = per_pt * sum(interior_points(n))            # seconds per cycle on the stated grid
    sn, run_steps, run_warm, extrapolated = n, steps, warm, False
    fits_ram = cpu_grid_ram_gb(n) < 0.8 * host_ram_gb()
    if not fits_ram or est * 2 > budget_s:
        extrapolated = True
        while min(sn) > 63 and (cpu_grid_ram_gb(sn) >= 0.8 * host_ram_gb() or per_pt * sum(interior_points(sn)) * 2 > budget_s):
            sn = tuple((d + 1) // 2 - 1 for d in sn)
        est = per_pt * sum(interior_points(sn))
    total = max(2, int(budget_s / max(est, 1e-9)))
    if run_steps + run_warm > total:
        run_warm = max(1, min(run_warm, total // 4)) if warm > 0 else (1 if total > 1 else 0)
        run_steps = max(1, min(steps, total - run_warm))
    o = orc.Oracle(orc.make_params(sn, mode=mode, pre=pre, post=post, maxiter=1, tol=1e-30), kind)
    o.set_threads(cores)
    for _ in range(run_warm):
        o.vcycle()
    t0 = time.time()
    for _ in range(run_steps):
        o.vcycle()
    dt = time.time() - t0
    o.close()
    scale = sum(interior_points(sn)) / sum(interior_points(n))
    gname = "x".join(map(str, sn))
    desc = (f"{run_steps} timed V({pre},{post}) cycle(s) after {run_warm} warm-up of GpuSolve-cpu ({kind}, -O3 -flto, OpenMP) at n={gname} "
            f"on {cores} threads: {dt / run_steps:.3f} s per cycle")
    if extrapolated:
        desc += (f"; EXTRAPOLATED to n={'x'.join(map(str, n))} by lattice points (x{scale:.5f}): the stated grid needs "
                 f"{cpu_grid_ram_gb(n):.0f} GB / ~{per_pt * sum(interior_points(n)):.0f} s per cycle on this host")
    return {"cycles_per_s": run_steps / dt * scale, "kind": kind, "cores": cores, "sample": desc, "extrapolated": extrapolated,
            "grid_run": list(sn), "steps_run": run_steps, "warmup_run": run_warm, "s_per_cycle_run": dt / run_steps}


def run_reference_arm(args, emit):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    n, pre, post = resolve_grid(args), args.pre, args.post
    r = run_cpu(n, pre, post, args.steps, args.warmup, budget_s=150.0)
    value = r["cycles_per_s"] * work_ratio(n)
    line = {
        "impl": "reference", "metric": metric_name(args, n), "value": value, "unit": "V-cycles/s",
        "n_gpus": args.gpus, "steps": r["steps_run"], "warmup": r["warmup_run"], "steps_requested": args.steps, "warmup_requested": args.warmup,
        # wall time per step of what actually ran (so steps x ms_per_step is the real timed region)
        "ms_per_step": 1e3 * r["s_per_cycle_run"], "extrapolated": r["extrapolated"], "grid_run": r["grid_run"],
        "higher_is_better": True, "scaling": "weak" if args.grid is None else "strong", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic (analytic RHS)", "config": workload_config(args),
        "cpu_baseline": {"value": value, "unit": "V-cycles/s", "cores": r["cores"], "kind": r["kind"], "sample": r["sample"]},
        "e2e": {"value": value, "unit": "V-cycles/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)
    return 0


# ---------------------------------------------------------------------------------------------------------------------
# workload
# ---------------------------------------------------------------------------------------------------------------------
REF_GRID = (511, 511, 511)   # BASELINE.json configs[1]: "513^3" = 511 interior points per axis


def default_grid(gpus):
    """N = 1: the 513^3 grid the metric is quoted on.  N > 1: the same lattice points PER GPU (weak scaling): the grid
    doubles along z, then y, then x, which at N = 8 is the 1025^3 grid of BASELINE.json configs[4]."""
    g = [511, 511, 511]
    axis, k = 2, gpus
    while k > 1:
        g[axis] = 2 * g[axis] + 1
        axis = (axis - 1) % 3
        k //= 2
    return tuple(g)


def resolve_grid(args):
    if args.grid is None:
        return default_grid(args.gpus)
    if len(args.grid) == 1:
        return (args.grid[0],) * 3
    if len(args.grid) != 3:
        raise SystemExit("--grid takes one or three integers")
    return tuple(args.grid)


def work_ratio(n):
    """lattice points of all levels relative to the 513^3 reference grid: 1 for the headline grid"""
    return sum(interior_points(n)) / sum(interior_points(REF_GRID))


def metric_name(args, n):
    base = f"V-cycles/s at 513^3 fp64, linear V({args.pre},{args.post})"
    if tuple(n) == REF_GRID:
        return base
    arr = "x".join(str(d + 2) for d in n)
    kind = "weak scaling" if args.grid is None else "grid given by --grid"
    return (f"513^3-equivalent V-cycles/s fp64, linear V({args.pre},{args.post}) ({kind}: {arr} arrays on {args.gpus} GPU(s), "
            f"cycles/s x {work_ratio(n):.3f})")


def workload_config(args):
    n = resolve_grid(args)
    arr = "x".join(str(d + 2) for d in n)
    cfg = {"workload": f"linear Poisson 7-point, n={'x'.join(map(str, n))} interior ({arr} arrays), V({args.pre},{args.post}) cycles, "
                       f"fp64, omega 0.8",
           "grid": list(n), "mode": 0, "pre": args.pre, "post": args.post,
           "l2": "inputs larger than L2 (each level-0 array is %.2f GB)" % ((n[0] + 2) * (n[1] + 2) * (n[2] + 2) * 8 / 1e9),
           "partition": "single GPU" if args.gpus == 1 else f"z-slabs over {args.gpus} GPUs, ghost planes over NVLink, coarse levels replicated"}
    if n != REF_GRID:
        cfg["value_normalisation"] = ("value = V-cycles/s on this grid x %.4f (its lattice points over all levels / those of the 513^3 "
                                      "grid): whole-job throughput in 513^3-equivalent V-cycles/s" % work_ratio(n))
    return cfg


class Job:
    """torch.distributed plumbing of one bench process"""

    def __init__(self):
        import torch
        self.torch = torch
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        torch.cuda.set_device(self.local)
        self.dist = None
        if self.world > 1:
            import torch.distributed as dist
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local))
            self.dist = dist
        self.gs = entry.load_package()
        self.gs.load()  # raises if the CUDA library is missing: no fallback

    def nccl_id(self):
        if self.dist is None:
            return None
        box = [self.gs.comm_unique_id() if self.rank == 0 else None]
        self.dist.broadcast_object_list(box, src=0)
        return box[0]

    def solver(self, n, mode=0, pre=2, post=2, slab=True):
        p = self.gs.make_params(n, mode=mode, pre=pre, post=post, maxiter=1, tol=1e-30)
        if slab and self.world > 1:
            return self.gs.GpuSolve(p, device=self.local, rank=self.rank, world=self.world, nccl_id=self.nccl_id())
        return self.gs.GpuSolve(p, device=self.local)

    def barrier(self, *solvers):
        if self.dist is not None:
            self.dist.barrier()
        self.torch.cuda.synchronize()
        for s in solvers:
            s.synchronize()

    def max_over_ranks(self, x):
        if self.dist is None:
            return x
        t = self.torch.tensor([x], device="cuda", dtype=self.torch.float64)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def time_cycles(self, solver, steps, warm, step=None):
        """CUDA events on the library's stream around `steps` calls, barrier + synchronize on both sides, max over ranks;
        returns (ms per step, last return value, kernels launched)"""
        torch = self.torch
        step = step or solver.vcycle
        stream = torch.cuda.ExternalStream(solver.stream, device=torch.device("cuda", self.local))
        for _ in range(warm):
            step()
        self.barrier(solver)
        l0 = solver.launches
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record(stream)
        out = None
        for _ in range(steps):
            out = step()
        ev1.record(stream)
        self.barrier(solver)
        return self.max_over_ranks(ev0.elapsed_time(ev1)) / steps, out, solver.launches - l0


def per_level_profile(solver, reps):
    """per-operator CUDA events (graph off): ({op: {ms_per_cycle, launches_per_cycle}}, [per level {op: launches}],
    [per level {op: ms}], sweeps per cycle on level 0)"""
    solver.set_option("timing", 1)
    solver.reset_timing()
    for _ in range(reps):
        solver.vcycle()
    allt = solver.timing()
    nlev = solver.num_levels
    ll, lm = [], []
    for l in range(nlev):
        t = solver.level_timing(l)
        ll.append({k: v[1] / reps for k, v in t.items() if v[1]})
        lm.append({k: v[0] / reps for k, v in t.items() if v[1]})
    sweeps0 = solver.level_sweeps(0) / reps
    solver.set_option("timing", 0)
    per_op = {k: {"ms_per_cycle": v[0] / reps, "launches_per_cycle": v[1] / reps} for k, v in allt.items()}
    return per_op, ll, lm, sweeps0


def slab_parity(job, n=255, cycles=2):
    """Before anything is timed: the slab solver must give the ONE-GPU fields bit for bit (Jacobi is order independent
    and every rank runs the same kernels on its planes) and the norms to 1e-12 (only the summation tree differs) — all
    three modes, with GPUSOLVE_SLAB_MIN_POINTS=0 so that several levels are partitioned and (world >= 3) there are ranks
    with a neighbour on both sides.  Raises SystemExit on a mismatch."""
    import numpy as np
    gs = job.gs
    old = os.environ.get("GPUSOLVE_SLAB_MIN_POINTS")
    os.environ["GPUSOLVE_SLAB_MIN_POINTS"] = "0"
    rec = {"grid": n, "cycles": cycles, "world": job.world, "modes": [], "bitwise": True, "max_norm_rel_diff": 0.0, "partitioned_levels": None}
    try:
        for mode in (0, 1, 2):
            pre = 3 if mode == 2 else 2
            slab = job.solver(n, mode=mode, pre=pre, post=pre)
            p = gs.make_params(n, mode=mode, pre=pre, post=pre, maxiter=1, tol=1e-30)
            rec["partitioned_levels"] = gs.partition(p, job.world, job.rank, 0)[2]
            field = gs.NEWTONV if mode == 2 else gs.V
            step = (lambda s: s.newton_step()[0]) if mode == 2 else (lambda s: s.vcycle())
            norms = [step(slab) for _ in range(cycles)]
            got = [slab.get(l, field if (mode == 2 and l == 0) else gs.V) for l in (0, 1)]   # collective
            ok, nerr = True, 0.0
            if job.rank == 0:
                one = job.solver(n, mode=mode, pre=pre, post=pre, slab=False)
                # FAS: a one-GPU handle folds R(v) into the fused residual+restriction pass (equal to the separate restriction
                # within rounding, not bit for bit); slab handles never do, so the bitwise comparison is made with it off
                one.set_option("fuse_rv", 0)
                ref = [step(one) for _ in range(cycles)]
                want = [one.get(l, field if (mode == 2 and l == 0) else gs.V) for l in (0, 1)]
                ok = all(bool(np.array_equal(a, b)) for a, b in zip(got, want))
                nerr = max(abs(a - b) / abs(b) for a, b in zip(norms, ref))
                one.close()
            slab.close()
            bad = job.max_over_ranks(0.0 if (ok and nerr <= 1e-12) else 1.0)
            rec["modes"].append(mode)
            rec["bitwise"] = rec["bitwise"] and ok
            rec["max_norm_rel_diff"] = max(rec["max_norm_rel_diff"], nerr)
            if bad:
                if job.rank == 0:
                    sys.stderr.write(f"slab parity FAILED: mode {mode} world {job.world} n {n}: fields bitwise {ok}, norm rel diff {nerr:.3e}\n")
                raise SystemExit(3)
    finally:
        if old is None:
            os.environ.pop("GPUSOLVE_SLAB_MIN_POINTS", None)
        else:
            os.environ["GPUSOLVE_SLAB_MIN_POINTS"] = old
    return rec


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="cuda", choices=["cuda", "reference"])
    ap.add_argument("--grid", type=int, nargs="+", default=None,
                    help="interior points per axis, one or three integers (511 -> 513^3 arrays; 1023 -> 1025^3); default: 511^3 per GPU")
    ap.add_argument("--pre", type=int, default=2)
    ap.add_argument("--post", type=int, default=2)
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the modes / small / strong-scaling sub-records")
    ap.add_argument("--no-parity", action="store_true", help="skip the slab-vs-one-GPU check at N > 1")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "cuda" else args.warmup

    # libraries (NCCL's version banner, torchrun) print to stdout: keep fd 1 for the one JSON line
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)

    def emit(line):
        sys.stdout.flush()
        os.write(real_stdout, (json.dumps(line) + "\n").encode())

    if args.impl == "reference":
        return run_reference_arm(args, emit)

    import numpy as np
    job = Job()
    gs, rank, world = job.gs, job.rank, job.world
    if args.gpus != world and world > 1:
        args.gpus = world
    peak, peak_src = measured_peaks()

    parity = None
    if world > 1 and not args.no_parity:
        parity = slab_parity(job)

    n, pre, post = resolve_grid(args), args.pre, args.post
    wr = work_ratio(n)
    solver = job.solver(n, 0, pre, post)

    # ---- whole-job throughput: inputs resident in HBM ----
    for _ in range(args.warmup):
        solver.vcycle()
    job.barrier(solver)
    sampler = ClockSampler(job.local) if rank == 0 else None
    t_wall0 = time.time()
    ms_per_step, norm, launches = job.time_cycles(solver, args.steps, 0)
    t_wall1 = time.time()
    clocks = sampler.stop(t_wall0, t_wall1) if sampler else None
    value = 1e3 / ms_per_step * wr

    # ---- dominant kernel (fine-level two-sweep smoother pass) inside the V-cycle, per-operator CUDA events ----
    reps = max(3, min(args.steps, 10))
    per_op, ll, lm, sweeps0 = per_level_profile(solver, reps)
    P = interior_points(n)
    sm_n = ll[0].get("smooth", 0.0)
    launch_ms = lm[0].get("smooth", 0.0) / max(sm_n, 1e-9)
    per_launch = sweeps0 / max(sm_n, 1e-9)
    # the post-smoothing pass may carry the prolongation + correction (fuse_pc): then it also reads the coarse iterate
    # (1 B per fine point) and there is no separate prolongation launch on level 0
    fused_pc = per_launch > 1.5 and ll[0].get("prolong", 0.0) == 0.0 and len(P) > 1
    pass_bytes = (25.0 if fused_pc else 24.0) * P[0] / world   # read v, read f, write v'' once per launch (this rank's slab)
    achieved = pass_bytes / (launch_ms * 1e-3) / 1e9
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    if os.path.exists(tpath) and n == REF_GRID:
        with open(tpath) as f:
            traffic = json.load(f).get(("smooth2pc_n%d" if fused_pc else "smooth2_n%d" if per_launch > 1.5 else "smooth_n%d") % n[0], None)
    kname = ("smooth2u_tma<PC>: prolongation + correction + two damped-Jacobi sweeps in one pass" if fused_pc else
             "smooth2u_tma: two damped-Jacobi sweeps per pass (temporal blocking, z-march unrolled over the ring period)" if per_launch > 1.5
             else "stencil_tma: one damped-Jacobi sweep per pass")
    # the other level-0 two-sweep pass of a cycle: the one that also reduces the residual norm (GS_OP_RESIDUAL)
    nr_n = ll[0].get("residual", 0.0)
    norm_ms = lm[0].get("residual", 0.0) / max(nr_n, 1e-9)
    norm_pass = {"kernel": "smooth2u_tma<NORM>: the cycle's residual norm + the next cycle's two pre-smoothing sweeps", "ms_per_launch": norm_ms,
                 "achieved": 24.0 * P[0] / world / max(norm_ms * 1e-3, 1e-12) / 1e9,
                 "frac": 24.0 * P[0] / world / max(norm_ms * 1e-3, 1e-12) / 1e9 / peak} if nr_n else None
    # the third level-0 pass of a cycle: residual + restriction (17 B per fine point: v and f read, 1/8 of a point written)
    rr_n = ll[0].get("restrict", 0.0)
    rr_ms = lm[0].get("restrict", 0.0)   # all restriction-side launches of level 0 (one in mode 0)
    rr_pass = {"kernel": "resid_restrict_tma: residual + full-weighting restriction, r never stored", "ms_per_cycle": rr_ms,
               "launches_per_cycle": rr_n, "achieved": 17.0 * P[0] / world / max(rr_ms * 1e-3, 1e-12) / 1e9,
               "frac": 17.0 * P[0] / world / max(rr_ms * 1e-3, 1e-12) / 1e9 / peak} if rr_n and len(P) > 1 else None
    roofline = {"bound": "hbm", "kernel": "fine-level smoother (level 0, inside the V-cycle) - " + kname, "achieved": achieved, "peak": peak,
                "unit": "GB/s", "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": pass_bytes, "ms_per_launch": launch_ms, "sweeps_per_launch": per_launch,
                "glups": sweeps0 * P[0] / max(lm[0].get("smooth", 0.0) * 1e-3, 1e-12) / 1e9,
                "lup_equivalent_gbs": 24.0 * P[0] / world * per_launch / (launch_ms * 1e-3) / 1e9,
                "dram_gbs": (traffic / (launch_ms * 1e-3) / 1e9) if traffic else None,
                "dram_frac": (traffic / (launch_ms * 1e-3) / 1e9 / peak) if traffic else None,
                "norm_pass": norm_pass, "resid_restrict_pass": rr_pass,
                "share_of_cycle": launch_ms * sm_n / max(sum(sum(x.values()) for x in lm), 1e-12),
                "note": "achieved = compulsory bytes of ONE PASS (v and f read, v'' written once: 24 B per point of this GPU's slab, 25 when "
                        "the pass also carries the prolongation) / average launch time, CUDA events inside the V-cycle; frac is a bandwidth "
                        "fraction.  The pass applies sweeps_per_launch Jacobi sweeps: glups and lup_equivalent_gbs (24 B per lattice "
                        "UPDATE, SURVEY 8d) carry that"}
    if fused_pc and world == 1:
        # what the fused pass replaces, measured in the same run: prolongation + correction and the two post-sweeps as two passes
        try:
            solver.set_option("fuse_pc", 0)
            ums, _, _ = job.time_cycles(solver, max(3, min(args.steps, 10)), 3)
            _, ull, ulm, _ = per_level_profile(solver, 3)
            solver.set_option("fuse_pc", 1)
            u_s = ulm[0].get("smooth", 0.0) / max(ull[0].get("smooth", 1.0), 1e-9)
            u_p = ulm[0].get("prolong", 0.0)
            roofline["replaces"] = {
                "passes": "prolong_correct_zmarch (17 B per point) + smooth2u_tma (24 B per point), option fuse_pc = 0",
                "prolong_ms": u_p, "smooth2_ms": u_s, "smooth2_frac": 24.0 * P[0] / max(u_s * 1e-3, 1e-12) / 1e9 / peak,
                "prolong_frac": 17.0 * P[0] / max(u_p * 1e-3, 1e-12) / 1e9 / peak,
                "vcycle_ms_unfused": ums, "vcycle_ms_fused": ms_per_step,
                "note": "the fused pass moves 25 B per point where the two it replaces move 41; it is latency-bound (coarse-value loads, "
                        "one step less TMA lead), so its own bandwidth fraction is lower than either of theirs while the cycle is faster"}
        except Exception as ex:  # never cost the line
            roofline["replaces"] = {"error": repr(ex)}
            try:
                solver.set_option("fuse_pc", 1)
            except Exception:
                pass
    vbytes = vcycle_algorithmic_bytes(n, pre, post)
    launched_bytes = pass_level_bytes(n, 0, ll)

    # ---- end to end through the C ABI with HOST buffers: H2D of the step's right-hand side from pinned
    #      memory (gs_set_field), the V-cycle, D2H of the step's residual norm ----
    e2e = None
    if not args.no_e2e:
        torch = job.torch
        shape = solver.shape(0)
        f_host = torch.empty(shape, dtype=torch.float64).pin_memory()
        f_np = f_host.numpy()
        f_np[...] = solver.get(0, gs.F)
        k = max(2, min(args.steps, 5))
        for _ in range(2):
            solver.set(0, gs.F, f_np)
            solver.vcycle()
        job.barrier(solver)
        t0 = time.perf_counter()
        for _ in range(k):
            solver.set(0, gs.F, f_np)      # host -> device, inside the timed region
            norm = solver.vcycle()         # device -> host: the cycle's residual norm
        job.barrier(solver)
        dt = job.max_over_ranks(time.perf_counter() - t0)
        nbytes = int(np.prod(shape)) * 8
        e2e = {"value": k / dt * wr, "unit": "V-cycles/s", "h2d_bytes_per_step": nbytes, "d2h_bytes_per_step": 8,
               "steps": k, "ms_per_step": 1e3 * dt / k, "handles": 1,
               "note": "gs_set_field(F) from pinned host memory + gs_vcycle + norm read-back per step; the result read back is the "
                       "8-byte residual norm (the reference's solve contract), never the field"}
        if world == 1:
            # Two independent problems in flight (two handles, one host thread each): every step still pays its own
            # 1.08 GB H2D, V-cycle and norm D2H, but one handle's cycle runs under the other's copy, so the step rate is
            # the PCIe link's, not link + cycle.  This is the headline e2e; the one-handle figure stays beside it.
            second = job.solver(n, 0, pre, post)
            g_host = torch.empty(shape, dtype=torch.float64).pin_memory()
            g_np = g_host.numpy()
            g_np[...] = f_np
            pair = [(solver, f_np), (second, g_np)]
            for s_, h_ in pair:
                for _ in range(2):
                    s_.set(0, gs.F, h_)
                    s_.vcycle()
            job.barrier(solver, second)
            k2 = max(3, min(args.steps, 6))
            norms = [None, None]

            def drive(i):
                s_, h_ = pair[i]
                for _ in range(k2):
                    s_.set(0, gs.F, h_)
                    norms[i] = s_.vcycle()

            th = [threading.Thread(target=drive, args=(i,)) for i in range(2)]
            t0 = time.perf_counter()
            for t in th:
                t.start()
            for t in th:
                t.join()
            job.barrier(solver, second)
            dt2 = time.perf_counter() - t0
            second.close()
            del g_host
            ok2 = all(x is not None and x == x and x >= 0.0 for x in norms)   # both handles returned a residual norm
            if ok2 and 2 * k2 / dt2 > e2e["value"]:
                e2e = {"value": 2 * k2 / dt2 * wr, "unit": "V-cycles/s", "h2d_bytes_per_step": nbytes, "d2h_bytes_per_step": 8,
                       "steps": 2 * k2, "ms_per_step": 1e3 * dt2 / (2 * k2), "handles": 2,
                       "one_handle": {"value": e2e["value"], "ms_per_step": e2e["ms_per_step"]},
                       "h2d_gbs": nbytes * 2 * k2 / dt2 / 1e9,
                       "note": "two handles driven by two host threads (independent right-hand sides): per step gs_set_field(F) from "
                               "pinned host memory (1.08 GB H2D) + gs_vcycle + the 8-byte norm D2H; one handle's cycle runs under the "
                               "other's copy, so the rate is bound by the PCIe link (h2d_gbs).  one_handle = the same loop on one handle"}
        del f_host
    solver.close()

    extras = {}
    if not args.no_extras:
        # ---- strong scaling on 1023^3 (BASELINE.json configs[4]) ----
        big = (1023, 1023, 1023)
        t1 = None
        if rank == 0:
            one = job.solver(big, 0, 2, 2, slab=False)
            for _ in range(3):
                one.vcycle()
            one.synchronize()
            torch = job.torch
            st = torch.cuda.ExternalStream(one.stream, device=torch.device("cuda", job.local))
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(st)
            for _ in range(10):
                one.vcycle()
            e1.record(st)
            one.synchronize()
            t1 = e0.elapsed_time(e1) / 10
            one.close()
        if world == 1:
            extras["strong_ref_1023"] = {"ms_per_cycle": t1, "grid": list(big), "cycles": 10, "note": "1023^3 (1025^3 arrays) on this one GPU"}
        else:
            t1 = job.max_over_ranks(t1 if t1 is not None else 0.0)
            s = job.solver(big, 0, 2, 2)
            tn, _, _ = job.time_cycles(s, 10, 3)
            s.close()
            extras["strong_1023"] = {"ms_per_cycle": tn, "ms_per_cycle_1gpu": t1, "efficiency": t1 / (world * tn), "speedup": t1 / tn,
                                     "grid": list(big), "cycles": 10,
                                     "note": "same 1023^3 grid on N GPUs vs on GPU 0 alone in this run (CUDA events, max over ranks)"}
        # ---- BASELINE.json configs[2], [3] at 513^3 and the shipped config at 129^3, one GPU ----
        if world == 1:
          try:
              extras["modes"] = {}
              for mode, (mpre, mpost) in ((1, (2, 2)), (2, (3, 3))):
                  s = job.solver(REF_GRID, mode, mpre, mpost)
                  if mode == 2:
                      s.newton_step()      # a realistic Newton iterate on every level
                  ms, _, ml = job.time_cycles(s, 10, 3)
                  pop, mll, mlm, msw = per_level_profile(s, 5)
                  s.close()
                  sweep = 32 if mode == 2 else 24
                  spn = mll[0].get("smooth", 0.0)
                  sp_ms = mlm[0].get("smooth", 0.0) / max(spn, 1e-9)
                  lb = pass_level_bytes(REF_GRID, mode, mll)
                  extras["modes"][str(mode)] = {
                      "config": f"mode {mode} ({'FAS, gamma=1' if mode == 1 else 'Newton-linearised, precomputed coefficient'}), 513^3, V({mpre},{mpost})",
                      "ms_per_cycle": ms, "vcycles_per_s": 1e3 / ms, "launches_per_cycle": ml / 10,
                      "dominant_kernel": "smooth2u_tma<%s> level 0" % ("KM_NONLINEAR" if mode == 1 else "KM_NEWTON_COEF"),
                      "pass_ms": sp_ms, "pass_bytes": sweep * P[0], "pass_frac_of_hbm_peak": sweep * P[0] / (sp_ms * 1e-3) / 1e9 / peak,
                      "sweeps_per_pass": msw / max(spn, 1e-9), "glups": msw * P[0] / max(mlm[0].get("smooth", 1e-9) * 1e-3, 1e-12) / 1e9,
                      "vcycle_pass_level_bytes": lb, "vcycle_frac_of_peak": lb / (ms * 1e-3) / 1e9 / peak, "per_operator": pop,
                      "binds": ("instruction issue, not memory: exp + divide per point per sweep (gs_exp / gs_div: a degree-13 polynomial and "
                                "a Newton-refined reciprocal without the library's special-case scaffolding): 65 % of the fp64 pipe, 64 % of the "
                                "issue slots, DRAM 30 %, dependency waits the largest stall (ncu: profiles/r02aa_ncu_modes_summary.txt, "
                                "smooth2u_tma<1,...>; with the library calls 55 % / 68 % / 27 %, profiles/r02a_ncu_modes_summary.txt)" if mode == 1 else
                                "latency and the divide (no pipe saturated): DRAM 49 %, issue 62 %, fp64 41 % for the two-sweep pass, 32 B per "
                                "point with the precomputed coefficient and one divide per point per sweep; the single-sweep pass of V(3,3) "
                                "reaches DRAM 79 % (ncu: profiles/r02aa_ncu_modes_summary.txt, smooth2u_tma<3,...> / stencil_tma<3,...>)")}
              # the shipped examples/data-2nd_order.conf (BASELINE configs[0]): Newton, 127^3, V(3,3) — ms per Newton iteration
              p = gs.make_params(127, mode=2, pre=3, post=3, maxiter=10, tol=1e-5)
              s = gs.GpuSolve(p, device=job.local)
              s.newton_step()
              s.synchronize()
              t0 = time.perf_counter()
              inner = 0
              for _ in range(3):
                  inner += s.newton_step()[1]
              s.synchronize()
              gpu_ms = (time.perf_counter() - t0) / 3 * 1e3
              s.close()
              small = {"n": 127, "mode": 2, "pre": 3, "post": 3, "ms_per_newton_iter": gpu_ms, "inner_cycles_per_iter": inner / 3,
                       "note": "examples/data-2nd_order.conf as shipped; wall clock over 3 Newton iterations after one warm-up (includes the host's "
                               "norm read-backs)"}
              if not args.no_cpu:
                  orc, kind = cpu_oracle()
                  o = orc.Oracle(orc.make_params(127, mode=2, pre=3, post=3, maxiter=10, tol=1e-5), kind)
                  o.set_threads(os.cpu_count())
                  o.newton_begin()
                  import numpy as _np

                  def cpu_step():
                      o.newton_compF()
                      o.set(0, orc.V, _np.zeros(o.shape(0)))
                      c = o.newton_find_error()
                      o.newton_compF()
                      return c
                  cpu_step()
                  t0 = time.perf_counter()
                  cin = sum(cpu_step() for _ in range(3))
                  small["cpu_ms_per_newton_iter"] = (time.perf_counter() - t0) / 3 * 1e3
                  small["cpu_inner_cycles_per_iter"] = cin / 3
                  small["cpu_kind"], small["cpu_cores"] = kind, os.cpu_count()
                  o.close()
              extras["small"] = small
          except Exception as ex:  # a sub-record must never cost the line
            extras["extras_error"] = repr(ex)

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        r = run_cpu(n, pre, post, 3, 1, budget_s=30.0)
        cpu = {"value": r["cycles_per_s"] * wr, "unit": "V-cycles/s", "cores": r["cores"], "kind": r["kind"], "sample": r["sample"],
               "extrapolated": r["extrapolated"]}

    if rank == 0:
        line = {
            "metric": metric_name(args, n),
            "value": value, "unit": "V-cycles/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak" if args.grid is None else "strong",
            "vs_baseline": None, "dtype": "f64",
            "data": "synthetic (analytic RHS, src/cpu/CpuGridData.cpp:47-58)", "config": workload_config(args),
            "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks,
            "vcycle": {"algorithmic_bytes": vbytes, "pass_level_bytes": launched_bytes,
                       "achieved_gbs_per_gpu": launched_bytes / world / (ms_per_step * 1e-3) / 1e9,
                       "frac_of_peak": launched_bytes / world / (ms_per_step * 1e-3) / 1e9 / peak,
                       "survey_bound_gbs_per_gpu": vbytes / world / (ms_per_step * 1e-3) / 1e9,
                       "per_operator": per_op, "per_level_ms": lm,
                       "glups_smoother_equivalent": (pre + post) * sum(P) / (ms_per_step * 1e-3) / 1e9,
                       "note": "pass_level_bytes = compulsory traffic of the passes launched per cycle (a two-sweep pass counts once); "
                               "frac_of_peak = that / world / time / peak, i.e. per GPU; algorithmic_bytes is SURVEY 8d's B_V (one pass per "
                               "reference operator)"},
            "final_residual_norm": norm, "vcycles_per_s_on_this_grid": 1e3 / ms_per_step, "work_ratio_vs_513": wr,
        }
        if parity is not None:
            line["slab_parity"] = parity
        line.update(extras)
        emit(line)
    if job.dist is not None:
        job.dist.barrier()
        job.dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
