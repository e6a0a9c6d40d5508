#!/usr/bin/env python
"""profiles/int_count.json <- one ncu metrics CSV (ncu --csv --metrics ... of scripts/profile_targets.py or bench.py).
usage: python scripts/update_int_count.py <csv> --key k_step_8 --units 1048576 [--kernel 'k_step<8'] [--which last]
Stores, for `key`: integer thread-instructions per unit (smsp__sass_thread_inst_executed_op_integer_pred_on.sum /
units), all thread-instructions per unit, ALU-pipe warp instructions and DRAM bytes per launch, the duration under
ncu, and the digest of the sources of the CURRENT tree (run it right after the profiled build came back)."""
import argparse
import csv
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from gymothelloenv_b200 import build  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("csv")
ap.add_argument("--key", required=True)
ap.add_argument("--units", type=float, required=True)
ap.add_argument("--kernel", default="")
ap.add_argument("--which", default="last", choices=["first", "last", "mean"])
ap.add_argument("--nodes-per-ply", type=float, default=None)
a = ap.parse_args()
rows = [r for r in csv.reader(open(a.csv)) if len(r) > 10]
hdr = rows[0]
ki, mi, vi, ii = hdr.index("Kernel Name"), hdr.index("Metric Name"), hdr.index("Metric Value"), hdr.index("ID")
launches = {}
for r in rows[1:]:
    if a.kernel and a.kernel not in r[ki]:
        continue
    d = launches.setdefault(int(r[ii]), {"kernel": r[ki]})
    try:
        d[r[mi]] = float(r[vi].replace(",", ""))
    except ValueError:
        pass
ids = sorted(launches)
if not ids:
    raise SystemExit("no launch of %r in %s" % (a.kernel, a.csv))
pick = [ids[0]] if a.which == "first" else ([ids[-1]] if a.which == "last" else ids)


def avg(name):
    v = [launches[i][name] for i in pick if name in launches[i]]
    return sum(v) / len(v) if v else None


ints = avg("smsp__sass_thread_inst_executed_op_integer_pred_on.sum")
alls = avg("smsp__thread_inst_executed.sum")
entry = {
    "kernel": launches[pick[0]]["kernel"], "units_per_launch": a.units,
    "int_thread_inst_per_unit": ints / a.units, "thread_inst_per_unit": alls / a.units if alls else None,
    "warp_inst_per_launch": avg("smsp__inst_executed.sum"), "alu_pipe_warp_inst_per_launch": avg("sm__inst_executed_pipe_alu.sum"),
    "fma_pipe_warp_inst_per_launch": avg("sm__inst_executed_pipe_fma.sum"),
    "alu_pipe_active_pct": avg("sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active"),
    "issue_active_pct": avg("smsp__issue_active.avg.pct_of_peak_sustained_active"),
    "dram_bytes_per_launch": (avg("dram__bytes_read.sum") or 0) + (avg("dram__bytes_write.sum") or 0),
    "registers": avg("launch__registers_per_thread"), "ncu_duration_ns": avg("gpu__time_duration.sum"),
    "source": os.path.relpath(os.path.abspath(a.csv), ROOT), "source_digest": build.source_digest(),
}
if a.nodes_per_ply:
    entry["nodes_per_ply"] = a.nodes_per_ply
path = os.path.join(ROOT, "profiles", "int_count.json")
data = json.load(open(path)) if os.path.isfile(path) else {}
data[a.key] = entry
json.dump(data, open(path, "w"), indent=1, sort_keys=True)
print(a.key, json.dumps(entry))
