#!/usr/bin/env python3
"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel: count, total, share, mean."""
import collections
import csv
import sys

rows = list(csv.reader(l for l in open(sys.argv[1]) if l.startswith('"')))
h = rows[0]
ik, iv = h.index("Kernel Name"), h.index("Metric Value")
only = sys.argv[2] if len(sys.argv) > 2 else ""
agg = collections.defaultdict(lambda: [0, 0.0])
for r in rows[1:]:
    try:
        v = float(r[iv].replace(",", ""))
    except ValueError:
        continue
    if only and only not in r[ik]:
        continue
    name = r[ik].split("(")[0][-70:]
    agg[name][0] += 1
    agg[name][1] += v
tot = sum(v[1] for v in agg.values())
print(f"{'n':>5s} {'total_us':>11s} {'share':>6s} {'mean_us':>9s}  kernel")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{v[0]:5d} {v[1] / 1e3:11.1f} {100 * v[1] / tot:5.1f}% {v[1] / v[0] / 1e3:9.1f}  {k}")
