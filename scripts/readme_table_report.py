#!/usr/bin/env python
"""README.md:44-50 table on the GPU next to the C oracle and the README's own cells -> JSON lines + a markdown table.
usage (GPU box): python scripts/readme_table_report.py > gpurun_out/<tag>/readme_table.jsonl"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tests import readme_table as rt  # noqa: E402

rows = []
cpu = rt.all_cpu_tallies()
for (b, w) in sorted(rt.README):
    r = rt.cell(b, w, cpu=cpu[(b, w)])
    rows.append(r)
    print(json.dumps(r), flush=True)
print("| black \\\\ white | GPU W/D/L % (N) | C oracle W/D/L (N) | chi2 (df 2) | README | z of README cell |", file=sys.stderr)
print("|---|---|---|---|---|---|", file=sys.stderr)
for r in rows:
    print("| %s vs %s | %s (%d) | %s (%d%s) | %.2f | %s | %s |" % (
        r["black"], r["white"], " / ".join("%.1f" % x for x in r["gpu_wdl_pct"]), r["gpu_games"],
        " / ".join(str(x) for x in r["cpu_oracle_wdl"]), r["cpu_oracle_games"], ", sampled" if r["cpu_sampled"] else "",
        r["chi2_df2"], " / ".join(str(x) for x in r["readme_wdl"]), " / ".join("%+.1f" % z for z in r["readme_z"])),
        file=sys.stderr)
