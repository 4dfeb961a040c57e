#!/usr/bin/env python
"""Per-kernel totals of an `ncu --metrics gpu__time_duration.sum --csv` launch list.
usage: launch_summary.py launches.csv ["header line"]"""
import collections
import csv
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 14 and r[0].isdigit()]
tot = collections.OrderedDict()
for r in rows:
    name, ns = r[4], float(r[14].replace(",", ""))
    n, t = tot.get(name, (0, 0.0))
    tot[name] = (n + 1, t + ns)
allns = sum(t for _, t in tot.values())
if len(sys.argv) > 2:
    print(sys.argv[2])
print(f"{'kernel':92s} {'launches':>8s} {'ms':>10s} {'share':>7s}")
for name, (n, t) in sorted(tot.items(), key=lambda kv: -kv[1][1]):
    print(f"{name[:90]:92s} {n:8d} {t / 1e6:10.3f} {100 * t / allns:6.1f}%")
