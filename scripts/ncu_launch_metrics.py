#!/usr/bin/env python
"""Print per-launch metrics of an `ncu --csv --metrics ...` log, one line per launch.
usage: ncu_launch_metrics.py log.csv [min_grid_size]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
min_grid = int(sys.argv[2]) if len(sys.argv) > 2 else 0
start = next(i for i, r in enumerate(rows) if r and r[0] == "ID")
h = rows[start]
ii, ki, mi, vi = h.index("ID"), h.index("Kernel Name"), h.index("Metric Name"), h.index("Metric Value")
cur, names = {}, {}
for r in rows[start + 1:]:
    if len(r) > vi:
        key = r[mi].split(".")[0].replace("smsp__", "").replace("sm__", "").replace("launch__", "")
        cur.setdefault(r[ii], {})[key] = float(r[vi].replace(",", ""))
        names[r[ii]] = r[ki][:28]
for k, m in cur.items():
    if m.get("grid_size", min_grid) >= min_grid:
        print(k, names[k], {a: round(v, 1) for a, v in m.items()})
