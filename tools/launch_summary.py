#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: launches, total time and share per kernel.
    python tools/launch_summary.py gpurun_out/launches.csv"""
import csv
import re
import sys
from collections import defaultdict

rows = [r for r in csv.reader(open(sys.argv[1], errors="replace")) if len(r) > 10 and r[0].isdigit()]
tot = defaultdict(lambda: [0, 0.0])
for r in rows:
    name = re.sub(r"\(.*", "", r[4])
    name = re.sub(r".*::", "", name)
    t = float(r[-1].replace(",", ""))
    unit = r[-2]
    t *= {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(unit, 1e-3)
    tot[name][0] += 1
    tot[name][1] += t
total = sum(v[1] for v in tot.values())
for name, (n, t) in sorted(tot.items(), key=lambda kv: -kv[1][1]):
    print(f"{name:48s} launches={n:4d} total_us={t:10.1f} share={t / total:.3f}")
