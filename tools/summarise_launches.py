#!/usr/bin/env python3
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel:
launches, total / mean device time and share.  usage: summarise_launches.py in.csv > out.md"""
import csv
import re
import sys
from collections import defaultdict

rows = []
with open(sys.argv[1], newline="") as f:
    lines = [l for l in f if l.startswith('"')]
for r in csv.DictReader(lines):
    if r.get("Metric Name") == "gpu__time_duration.sum":
        rows.append((r["Kernel Name"], float(r["Metric Value"].replace(",", "")), r["Metric Unit"]))
agg = defaultdict(lambda: [0, 0.0])
for name, v, unit in rows:
    ns = v * {"ns": 1.0, "us": 1e3, "ms": 1e6, "s": 1e9}.get(unit, 1.0)
    short = re.sub(r"\(.*", "", name)
    short = re.sub(r"^void ", "", short)
    agg[short][0] += 1
    agg[short][1] += ns
tot = sum(v[1] for v in agg.values())
print("| kernel | launches | total ms | mean us | share |")
print("|---|---:|---:|---:|---:|")
for k, (n, ns) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print("| `%s` | %d | %.3f | %.1f | %.1f%% |" % (k[:110], n, ns / 1e6, ns / n / 1e3, 100 * ns / tot))
print("\ntotal: %d launches, %.1f ms of device time (cold-cache, serialised under ncu: compare shares)" % (len(rows), tot / 1e6))
