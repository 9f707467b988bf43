#!/usr/bin/env python
"""bench.py — GpuSolve V-cycle hot path on B200: V-cycles/s at 513^3 fp64 (config n=511, linear
V(2,2), BASELINE.json configs[1]) with the fine-level smoother against the HBM roofline.

  python bench.py --gpus N --steps K --warmup W            the CUDA backend (C ABI, lib/libgpusolve_cuda.so)
  python bench.py --impl reference ...                      the reference's own CPU path on the host cores

A step is one V(2,2) cycle of the hot path (including the per-cycle residual norm the reference's
solve loop reads back).  One JSON line on stdout (rank 0).  N > 1: one process per GPU under
torchrun, the grid is split into z-slabs (strong scaling of the same grid), NCCL halo exchange.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
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
    """B_V = sum_{l<L-1} P_l [24(v1+v2) + 17 + 17] + P_{L-1} 24 (v1+v2) + 24 P_0 (mode 2: 32/25)"""
    P = interior_points(n)
    sweep = 32 if mode == 2 else 24
    rr = 25 if mode == 2 else 17
    b = sum(p * (sweep * (pre + post) + rr + 17) for p in P[:-1]) + P[-1] * sweep * (pre + post) + sweep * P[0]
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


def cpu_sample_grid(n, steps_total, cores, budget_s):
    """the largest grid, obtained by halving the target like the hierarchy does, whose CPU run fits the budget
    (one 511^3 V(2,2) cycle is ~9 s on 8 cores)"""
    n = tuple(n)
    est = 9.0 * (n[0] * n[1] * n[2]) / 511.0 ** 3 * 8.0 / max(cores, 1) * steps_total
    while min(n) > 63 and est > budget_s:
        n = tuple((d + 1) // 2 - 1 for d in n)
        est /= 8.0
    return n


def run_cpu(n, pre, post, steps, warm, budget_s):
    """the reference's CPU V-cycle (oracle/_ref when built, else the C port) on all host threads; returns
    (V-cycles/s on grid n, kind, cores, sample description)"""
    orc = entry.load_oracle()
    kind = "reference" if orc.available("reference") else "port"
    if kind == "port" and not orc.available("port"):
        orc.build(ref=False)
    cores = os.cpu_count()
    sn = cpu_sample_grid(n, steps + warm, cores, budget_s)
    o = orc.Oracle(orc.make_params(sn, mode=0, pre=pre, post=post, maxiter=1, tol=1e-30), kind)
    o.set_threads(cores)
    for _ in range(warm):
        o.vcycle()
    t0 = time.time()
    for _ in range(steps):
        o.vcycle()
    dt = time.time() - t0
    o.close()
    scale = sum(interior_points(sn)) / sum(interior_points(tuple(n)))
    desc = (f"{steps} V({pre},{post}) cycle(s) of GpuSolve-cpu ({kind}) at n={'x'.join(map(str, sn))} on {cores} OpenMP threads: "
            f"{dt / steps:.2f} s per cycle" + ("" if tuple(sn) == tuple(n) else f"; throughput scaled by lattice points x{scale:.5f} "
                                               f"to n={'x'.join(map(str, n))}"))
    return steps / dt * scale, kind, cores, desc


def run_reference_arm(args, emit):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    n, pre, post = resolve_grid(args), args.pre, args.post
    steps, warm = args.steps, args.warmup
    v, kind, cores, sample = run_cpu(n, pre, post, steps, warm, budget_s=150.0)
    value = v * work_ratio(n)
    line = {
        "impl": "reference", "metric": "V-cycles/s at 513^3 fp64, linear V(2,2)", "value": value, "unit": "V-cycles/s",
        "n_gpus": args.gpus, "steps": steps, "warmup": warm, "ms_per_step": 1e3 / v,
        "higher_is_better": True, "scaling": "weak" if args.grid is None else "strong", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic (analytic RHS)", "config": workload_config(args),
        "cpu_baseline": {"value": value, "unit": "V-cycles/s", "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": "V-cycles/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)
    return 0


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


def workload_config(args):
    n = resolve_grid(args)
    arr = "x".join(str(d + 2) for d in n)
    cfg = {"workload": f"linear Poisson 7-point, n={'x'.join(map(str, n))} interior ({arr} arrays), V({args.pre},{args.post}) cycles, "
                       f"fp64, omega 0.8",
           "grid": list(n), "mode": 0, "pre": args.pre, "post": args.post,
           "l2": "inputs larger than L2 (each level-0 array is %.2f GB)" % ((n[0] + 2) * (n[1] + 2) * (n[2] + 2) * 8 / 1e9),
           "partition": "single GPU" if args.gpus == 1 else f"z-slabs over {args.gpus} GPUs, NCCL halo exchange, coarse levels replicated"}
    if n != REF_GRID:
        cfg["value_normalisation"] = ("value = V-cycles/s on this grid x %.4f (its lattice points over all levels / those of the 513^3 "
                                      "grid): whole-job throughput in 513^3-equivalent V-cycles/s" % work_ratio(n))
    return cfg


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
    import torch
    gs = entry.load_package()
    gs.load()  # raises if the CUDA library is missing: no fallback

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.gpus != world and world > 1:
        args.gpus = world
    dist = None
    torch.cuda.set_device(local_rank)
    nccl_id = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        box = [gs.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(box, src=0)
        nccl_id = box[0]

    n, pre, post = resolve_grid(args), args.pre, args.post
    wr = work_ratio(n)
    p = gs.make_params(n, mode=0, pre=pre, post=post, maxiter=1, tol=1e-30)
    solver = gs.GpuSolve(p, device=local_rank, rank=rank, world=world, nccl_id=nccl_id)
    stream = torch.cuda.ExternalStream(solver.stream, device=torch.device("cuda", local_rank))

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()
        solver.synchronize()

    # ---- whole-job throughput: inputs resident in HBM ----
    for _ in range(args.warmup):
        solver.vcycle()
    barrier()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    l0 = solver.launches
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t_wall0 = time.time()
    ev0.record(stream)
    norm = 0.0
    for _ in range(args.steps):
        norm = solver.vcycle()
    ev1.record(stream)
    barrier()
    t_wall1 = time.time()
    ms = ev0.elapsed_time(ev1)
    launches = solver.launches - l0
    clocks = sampler.stop(t_wall0, t_wall1) if sampler else None
    if dist is not None:
        t = torch.tensor([ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    ms_per_step = ms / args.steps
    value = 1e3 / ms_per_step * wr

    # ---- dominant kernel (fine-level smoother sweep) inside the V-cycle, per-operator CUDA events ----
    solver.set_option("timing", 1)
    solver.reset_timing()
    reps = max(3, min(args.steps, 10))
    for _ in range(reps):
        solver.vcycle()
    lt = solver.level_timing(0)
    allt = solver.timing()
    solver.set_option("timing", 0)
    sm_ms, sm_n = lt["smooth"]
    sweeps = solver.level_sweeps(0)          # Jacobi sweeps those launches applied (2 per launch of smooth2_tma)
    P = interior_points(n)
    per_launch = sweeps / max(sm_n, 1)
    launch_ms = sm_ms / max(sm_n, 1)
    launch_bytes = 24.0 * P[0] / world * per_launch      # 24 B per lattice update
    peak, peak_src = measured_peaks()
    achieved = launch_bytes / (launch_ms * 1e-3) / 1e9
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    if os.path.exists(tpath) and n == REF_GRID:
        with open(tpath) as f:
            traffic = json.load(f).get(("smooth2_n%d" if per_launch > 1.5 else "smooth_n%d") % n[0], None)
    kname = ("smooth2u_tma: two damped-Jacobi sweeps per pass (temporal blocking, z-march unrolled over the ring period)" if per_launch > 1.5
             else "stencil_tma: one damped-Jacobi sweep per pass")
    roofline = {"bound": "hbm", "kernel": "fine-level smoother (level 0, inside the V-cycle) - " + kname, "achieved": achieved, "peak": peak,
                "unit": "GB/s", "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": launch_bytes, "ms_per_launch": launch_ms, "sweeps_per_launch": per_launch,
                "glups": sweeps * P[0] / (sm_ms * 1e-3) / 1e9,
                "dram_gbs": (traffic / (launch_ms * 1e-3) / 1e9) if traffic else None,
                "note": "achieved = algorithmic bytes (24 B per lattice update x updates per launch) / launch time; a temporally blocked "
                        "launch does two updates per point for one pass over v and f, so achieved may exceed the copy peak while the DRAM "
                        "traffic (ncu, `traffic`) stays at ~25 B per point"}
    vbytes = vcycle_algorithmic_bytes(n, pre, post)
    per_op = {k: {"ms_per_cycle": v[0] / reps, "launches_per_cycle": v[1] / reps} for k, v in allt.items()}

    # ---- end to end through the C ABI with HOST buffers: H2D of the step's right-hand side from pinned
    #      memory (gs_set_field), the V-cycle, D2H of the step's residual norm ----
    e2e = None
    if not args.no_e2e:
        shape = solver.shape(0)
        f_host = torch.empty(shape, dtype=torch.float64).pin_memory()
        f_np = f_host.numpy()
        f_np[...] = solver.get(0, gs.F)
        k = max(2, min(args.steps, 5))
        for _ in range(2):
            solver.set(0, gs.F, f_np)
            solver.vcycle()
        barrier()
        t0 = time.perf_counter()
        for _ in range(k):
            solver.set(0, gs.F, f_np)      # host -> device, inside the timed region
            norm = solver.vcycle()         # device -> host: the cycle's residual norm
        barrier()
        dt = time.perf_counter() - t0
        if dist is not None:
            t = torch.tensor([dt], device="cuda", dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dt = float(t.item())
        e2e = {"value": k / dt * wr, "unit": "V-cycles/s", "h2d_bytes_per_step": int(np.prod(shape)) * 8 // 1, "d2h_bytes_per_step": 8,
               "steps": k, "note": "gs_set_field(F) from pinned host memory + gs_vcycle + norm read-back per step"}
        del f_host

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        v, kind, cores, sample = run_cpu(n, pre, post, 1, 0, budget_s=30.0)
        cpu = {"value": v * wr, "unit": "V-cycles/s", "cores": cores, "kind": kind, "sample": sample}

    if rank == 0:
        line = {
            "metric": "V-cycles/s at 513^3 fp64, linear V(2,2)" if (pre, post) == (2, 2) else f"V-cycles/s at 513^3 fp64, linear V({pre},{post})",
            "value": value, "unit": "V-cycles/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak" if args.grid is None else "strong",
            "vs_baseline": None, "dtype": "f64",
            "data": "synthetic (analytic RHS, src/cpu/CpuGridData.cpp:47-58)", "config": workload_config(args),
            "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks,
            "vcycle": {"algorithmic_bytes": vbytes, "achieved_gbs": vbytes / (ms_per_step * 1e-3) / 1e9,
                       "frac_of_peak": vbytes / (ms_per_step * 1e-3) / 1e9 / peak, "per_operator": per_op,
                       "glups_smoother_equivalent": (pre + post) * sum(P) / (ms_per_step * 1e-3) / 1e9},
            "final_residual_norm": norm, "vcycles_per_s_on_this_grid": 1e3 / ms_per_step, "work_ratio_vs_513": wr,
        }
        emit(line)
    solver.close()
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
