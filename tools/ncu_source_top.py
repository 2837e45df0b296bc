#!/usr/bin/env python3
"""Top stall locations of `ncu -i X.ncu-rep --page source --csv`: usage ncu_source_top.py source.csv [N]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
N = int(sys.argv[2]) if len(sys.argv) > 2 else 40
h = rows[1]
si, ni, ei = h.index("Source"), h.index("# Samples"), h.index("Instructions Executed")
stall = [i for i, n in enumerate(h) if n.startswith("stall_") and "Not Issued" not in n]
data = []
for r in rows[2:]:
    try:
        s = int(r[ni])
    except Exception:
        continue
    top = sorted(((int(r[i] or 0), h[i][6:]) for i in stall), reverse=True)[:2]
    data.append((s, r[ei], r[si], top))
tot = sum(d[0] for d in data)
texe = sum(int(d[1] or 0) for d in data)
print("total samples", tot, "warp instructions", texe)
agg = {}
for r in rows[2:]:
    for i in stall:
        try:
            agg[h[i]] = agg.get(h[i], 0) + int(r[i] or 0)
        except Exception:
            pass
print({k[6:]: "%.1f%%" % (100.0 * v / max(tot, 1)) for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:9]})
for d in sorted(data, key=lambda x: -x[0])[:N]:
    print("%5.2f%% %10s  %-70s %s" % (100 * d[0] / tot, d[1], d[2][:70], d[3]))
