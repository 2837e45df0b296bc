#!/usr/bin/env python3
"""Per-kernel share of the step from an `ncu --metrics gpu__time_duration.sum --csv` launch list. usage: ncu_launch_shares.py launches.csv"""
import csv
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
h = rows[0]
ki, vi, ui = h.index("Kernel Name"), h.index("Metric Value"), h.index("Metric Unit")
tot = {}
for r in rows[1:]:
    if r[h.index("Metric Name")] != "gpu__time_duration.sum":
        continue
    us = float(r[vi].replace(",", "")) / (1e3 if r[ui] in ("ns", "nsecond") else 1.0)
    name = r[ki].split("(")[0].replace("void ", "")[:130]
    t = tot.setdefault(name, [0, 0.0])
    t[0] += 1
    t[1] += us
allus = sum(v[1] for v in tot.values())
print("# ncu --metrics gpu__time_duration.sum --clock-control none, `python bench.py --steps 3 --warmup 3` (cold-cache, serialised launches: shares, not absolutes)")
for name, (n, us) in sorted(tot.items(), key=lambda kv: -kv[1][1]):
    print("%6.2f%% %6d launches %11.1f us total %9.2f us avg  %s" % (100 * us / allus, n, us, us / n, name))
