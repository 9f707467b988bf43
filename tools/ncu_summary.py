#!/usr/bin/env python
"""Summarise an .ncu-rep or its `--page raw --csv` export (read here, no GPU): one line per launch with the roofline-relevant counters.
    python tools/ncu_summary.py gpurun_out/prof.ncu-rep [min_us]"""
import csv
import io
import re
import subprocess
import sys

rep = sys.argv[1]
min_us = float(sys.argv[2]) if len(sys.argv) > 2 else 0.0
out = open(rep).read() if rep.endswith(".csv") else subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
h, units = rows[0], rows[1]
cols = {
    "time_us": "gpu__time_duration.sum",
    "dram_rd": "dram__bytes_read.sum", "dram_wr": "dram__bytes_write.sum",
    "dram%": "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "l2hit%": "lts__t_sector_hit_rate.pct",
    "l1%": "l1tex__throughput.avg.pct_of_peak_sustained_active",
    "sm%": "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "warps%": "sm__warps_active.avg.pct_of_peak_sustained_active",
    "regs": "launch__registers_per_thread", "grid": "launch__grid_size", "block": "launch__block_size",
    "smem_dyn": "launch__shared_mem_per_block_dynamic",
    "fp64%": "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
    "lsu_mem%": "l1tex__data_pipe_lsu_wavefronts_mem_shared.avg.pct_of_peak_sustained_elapsed",
    "issue%": "sm__inst_issued.avg.pct_of_peak_sustained_active",
}
idx = {k: h.index(v) for k, v in cols.items() if v in h}
ki = h.index("Kernel Name")


def num(r, k):
    if k not in idx:
        return None
    try:
        v = float(r[idx[k]].replace(",", ""))
    except ValueError:
        return None
    u = units[idx[k]]
    scale = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1, "ms": 1e3, "us": 1, "ns": 1e-3, "s": 1e6}
    if k in ("dram_rd", "dram_wr"):
        return v * scale.get(u, 1)
    if k == "time_us":
        return v * scale.get(u, 1)
    return v


for r in rows[2:]:
    t = num(r, "time_us")
    if t is None or t < min_us:
        continue
    name = re.sub(r"\(.*", "", r[ki])
    name = re.sub(r".*::", "", name)[:46]
    rd, wr = num(r, "dram_rd") or 0, num(r, "dram_wr") or 0
    parts = [f"{name:46s} {t:9.1f}us grid={int(num(r,'grid') or 0):6d}x{int(num(r,'block') or 0):4d} regs={int(num(r,'regs') or 0):3d}",
             f"dram rd {rd/1e6:8.1f} wr {wr/1e6:8.1f} MB ({(rd+wr)/t/1e3:6.0f} GB/s)"]
    for k in ("dram%", "l2hit%", "l1%", "lsu_mem%", "sm%", "fp64%", "issue%", "warps%"):
        v = num(r, k)
        if v is not None:
            parts.append(f"{k}={v:5.1f}")
    print("  ".join(parts))
