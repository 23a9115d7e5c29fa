#!/usr/bin/env python
"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel name."""
import collections
import csv
import sys

lines = [l for l in open(sys.argv[1]) if not l.startswith("==")]
rows = list(csv.DictReader(lines))
f = {"ns": 1e-6, "us": 1e-3, "ms": 1.0}
agg = collections.OrderedDict()
for x in rows:
    try:
        v = float(x["Metric Value"].replace(",", "")) * f.get(x["Metric Unit"], 1.0)
    except (ValueError, KeyError):
        continue
    a = agg.setdefault(x["Kernel Name"][:72], [0, 0.0])
    a[0] += 1
    a[1] += v
tot = sum(a[1] for a in agg.values())
for n, a in sorted(agg.items(), key=lambda kv: -kv[1][1])[: int(sys.argv[2]) if len(sys.argv) > 2 else 40]:
    print(f"{a[1]:9.3f} ms {a[0]:4d}x {a[1] / a[0] * 1e3:9.1f} us/launch {100 * a[1] / tot:5.1f}%  {n}")
print(f"total {tot:.3f} ms over {len(rows)} launches")
