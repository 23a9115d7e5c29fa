#!/usr/bin/env python
"""Per-kernel counts of the SASS mnemonics that prove the Blackwell-native paths (B200_PROFILING.md):
UTC*MMA (tcgen05.mma), LDTM / STTM (tcgen05.ld / st), UTCBAR (tcgen05.commit), UBLKCP (cp.async.bulk: the TMA engine's
bulk mode), UTMALDG / UTMASTG (tensor TMA), SYNCS (mbarrier), HMMA (legacy mma.sync -- must be 0).
    python profiles/sass_mnemonics.py explainn_b200/lib/libexplainn_b200.so > profiles/r02_sass_mnemonics.txt"""
import collections
import re
import subprocess
import sys

out = subprocess.run(["cuobjdump", "-sass", sys.argv[1]], capture_output=True, text=True).stdout
pat = re.compile(r"\b(UTC[A-Z]*MMA|LDTM|STTM|UTCBAR|UBLKCP|UTMALDG|UTMASTG|UTCCP|HMMA|SYNCS)\b")
cur, counts = None, collections.OrderedDict()
for line in out.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().split("(")[0]
        counts[cur] = collections.Counter()
        continue
    if cur:
        for mm in pat.findall(line):
            counts[cur][mm] += 1
keys = ["UTCHMMA", "UTCQMMA", "LDTM", "STTM", "UTCBAR", "UBLKCP", "UTMALDG", "UTMASTG", "SYNCS", "HMMA"]
tot = collections.Counter()
print(f"{'kernel':58s} " + " ".join(f"{k:>8s}" for k in keys))
for name, c in counts.items():
    if sum(c.values()) == 0:
        continue
    tot.update(c)
    print(f"{name[:58]:58s} " + " ".join(f"{c.get(k, 0):8d}" for k in keys))
print(f"{'TOTAL':58s} " + " ".join(f"{tot.get(k, 0):8d}" for k in keys))
