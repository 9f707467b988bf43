#!/usr/bin/env python
"""Where the warps of one kernel in an .ncu-rep stall (read here, no GPU): stall-reason totals and the SASS lines
holding most samples.   python tools/ncu_stalls.py rep.ncu-rep <kernel regex> [launch index] [min percent]"""
import csv
import io
import subprocess
import sys

rep, rx = sys.argv[1], sys.argv[2]
skip = sys.argv[3] if len(sys.argv) > 3 else "0"
minpct = float(sys.argv[4]) if len(sys.argv) > 4 else 1.2
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + rx, "--launch-skip", skip,
                      "--launch-count", "1"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
print(rows[0][1][:100])
h = rows[1]
ia, isamp = h.index("Source"), h.index("# Samples")
st = [i for i, c in enumerate(h) if c.startswith("stall_") and "Not Issued" not in c]
body = [r for r in rows[2:] if len(r) > isamp and r[isamp].isdigit()]
tot = sum(int(r[isamp]) for r in body)
agg = {}
for r in body:
    for j in st:
        if r[j].isdigit():
            agg[h[j][6:]] = agg.get(h[j][6:], 0) + int(r[j])
print("samples", tot, " ".join(f"{k} {100 * v / tot:.1f}%" for k, v in sorted(agg.items(), key=lambda t: -t[1])[:9]))
for i, r in enumerate(body):
    n = int(r[isamp])
    if n > tot * minpct / 100:
        ss = sorted([(int(r[j]), h[j][6:]) for j in st if r[j].isdigit() and int(r[j]) > 0], reverse=True)[:2]
        print(f"{i:5d} {100 * n / tot:5.1f}% {r[ia].strip()[:60]:60s} {ss}")
