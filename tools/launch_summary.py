#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: launches, total, average and share per kernel."""
import csv
import sys
from collections import OrderedDict

path, title = sys.argv[1], (sys.argv[2] if len(sys.argv) > 2 else "")
rows = [r for r in csv.reader(open(path, errors="replace")) if len(r) > 5]
hdr = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
cols = rows[hdr]
kn, mv, mu = cols.index("Kernel Name"), cols.index("Metric Value"), cols.index("Metric Unit")
agg = OrderedDict()
for r in rows[hdr + 1:]:
    if len(r) <= mv:
        continue
    v = float(r[mv].replace(",", ""))
    v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(r[mu], 1e-3)
    n, t = agg.get(r[kn], (0, 0.0))
    agg[r[kn]] = (n + 1, t + v)
total = sum(t for _, t in agg.values())
print(f"# ncu launch list summary{': ' + title if title else ''}")
print("# ncu --metrics gpu__time_duration.sum --clock-control none; cold-cache serialised times: compare SHARES")
print(f"{'kernel':64s} {'n':>5s} {'total_us':>10s} {'avg_us':>9s}  share")
for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{k[:64]:64s} {n:5d} {t:10.1f} {t / n:9.1f} {100 * t / total:5.1f}%")
