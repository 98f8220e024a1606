#!/usr/bin/env python
"""Summarise an ncu launch list (--metrics gpu__time_duration.sum --csv): per kernel count, total
time and share.  python tools/launch_summary.py launches.csv [> summary.txt]"""
import csv, re, sys
from collections import defaultdict

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 14 and r[12] == "gpu__time_duration.sum"]
agg = defaultdict(lambda: [0, 0.0])
for r in rows:
    name = r[4]
    name = re.sub(r"^void ", "", name)
    name = re.sub(r"\(.*$", "", name)
    if name.startswith("at::") or name.startswith("cub::") or "at::native" in name:
        name = re.sub(r"<.*$", "", name)
    agg[name][0] += 1
    agg[name][1] += float(r[14]) * 1e-6
tot = sum(v[1] for v in agg.values())
ours = sum(v[1] for k, v in agg.items() if k.startswith("pcc::"))
print(f"launches {len(rows)}  total {tot:.2f} ms  pcc:: kernels {ours:.2f} ms ({100 * ours / tot:.1f} %)")
print(f"{'kernel':90s} {'n':>6s} {'ms':>10s} {'share':>7s}")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{k[:90]:90s} {v[0]:6d} {v[1]:10.3f} {100 * v[1] / tot:6.2f}%")
