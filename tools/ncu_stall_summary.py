#!/usr/bin/env python
"""Warp-stall breakdown per launch from a `--page raw --csv` export of an ncu --set full capture (read here, no GPU):
average stalled warps per issue-active cycle by reason.   python tools/ncu_stall_summary.py raw.csv"""
import csv
import re
import sys

rows = list(csv.reader(open(sys.argv[1])))
h = rows[0]
ki = h.index("Kernel Name")
cols = [(i, re.sub(r".*issue_stalled_(.*)_per_issue_active.ratio", r"\1", n)) for i, n in enumerate(h)
        if n.startswith("smsp__average_warps_issue_stalled_") and n.endswith("_per_issue_active.ratio") and "not_issued" not in n]
for r in rows[2:]:
    vals = []
    for i, n in cols:
        try:
            vals.append((float(r[i]), n))
        except ValueError:
            pass
    tot = sum(v for v, _ in vals) or 1.0
    name = re.sub(r"\(.*", "", r[ki])
    name = re.sub(r".*::", "", name)[:44]
    print(f"{name:44s} " + "  ".join(f"{n} {100 * v / tot:.0f}%" for v, n in sorted(vals, reverse=True)[:6]))
