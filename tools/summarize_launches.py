#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel:
count, total/mean device time and share.  usage: summarize_launches.py launches.csv"""
import collections
import csv
import sys

lines = [ln for ln in open(sys.argv[1]) if ln.startswith('"')]
rows = list(csv.DictReader(lines))
agg = collections.OrderedDict()
for r in rows:
    name = r["Kernel Name"].split("(")[0]
    grid, block = r["Grid Size"], r["Block Size"]
    # launches of the full-size batches keep their own line; the many small grids of the
    # configs[4] replay (one per batch) are folded into one line per kernel
    g0 = int(grid.strip("()").split(",")[0])
    key = (name, grid if g0 >= 1000 or g0 in (128, 148) else "(small batches)", block)
    agg.setdefault(key, []).append(float(r["Metric Value"]) / 1e6)
tot = sum(sum(v) for v in agg.values())
print("| kernel | grid | block | launches | total ms | mean ms | share |")
print("|---|---|---|---|---|---|---|")
for (name, grid, block), v in agg.items():
    print(f"| {name} | {grid} | {block} | {len(v)} | {sum(v):.3f} | {sum(v)/len(v):.3f} | {100*sum(v)/tot:.1f}% |")
