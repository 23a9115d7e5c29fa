#!/usr/bin/env python
"""Hottest SASS lines of one kernel from `ncu -i rep --page source --csv ...` output (stdin or file)."""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]
i_src, i_s, i_ex = hdr.index("Source"), hdr.index("Warp Stall Sampling (All Samples)"), hdr.index("Instructions Executed")
data = [(int(r[i_s]), int(r[i_ex]), k, r[i_src].strip()) for k, r in enumerate(rows[2:]) if len(r) > i_ex and r[i_s].isdigit()]
tot = sum(d[0] for d in data) or 1
totex = sum(d[1] for d in data) or 1
print(rows[0][1], "samples", tot, "warp-instr", totex)
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
for s, e, k, src in sorted(data, reverse=True)[:top]:
    print(f"{100 * s / tot:5.1f}% samp  exec {e:9d}  #{k:4d} {src[:100]}")
print("--- most executed")
for s, e, k, src in sorted(data, key=lambda d: -d[1])[:10]:
    print(f"{100 * e / totex:5.1f}% exec  samp {100 * s / tot:5.1f}%  #{k:4d} {src[:100]}")
