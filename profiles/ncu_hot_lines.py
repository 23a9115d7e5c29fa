#!/usr/bin/env python
"""Per-CUDA-line instruction and stall-sample shares of one kernel:
   ncu -i rep --page source --csv --print-source cuda,sass --kernel-name regex:<k> > f.csv; python ncu_hot_lines.py f.csv [top]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Line No")
hdr = rows[hi]
i_ex, i_s = hdr.index("Instructions Executed"), hdr.index("Warp Stall Sampling (All Samples)")
data = []
for r in rows[hi + 1:]:
    if r and r[0].isdigit() and len(r) > i_ex and r[i_ex].isdigit():
        data.append((int(r[i_ex]), int(r[i_s]) if r[i_s].isdigit() else 0, int(r[0]), r[1].strip()))
tot, ts = sum(d[0] for d in data) or 1, sum(d[1] for d in data) or 1
print(rows[1][1] if len(rows) > 1 else "", "warp-instr", tot, "samples", ts)
for e, s, ln, src in sorted(data, reverse=True)[: int(sys.argv[2]) if len(sys.argv) > 2 else 30]:
    print(f"{100 * e / tot:5.1f}% exec {100 * s / ts:5.1f}% samp  L{ln:<5d} {src[:105]}")
