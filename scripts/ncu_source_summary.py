#!/usr/bin/env python
"""Summarise an ncu report's source page: instructions executed per CUDA source line
(top N) and per SASS opcode, for the first profiled kernel matching --kernel.
usage: ncu_source_summary.py report.ncu-rep [--kernel k_step] [--top 40]"""
import argparse, collections, csv, io, re, subprocess, sys

ap = argparse.ArgumentParser()
ap.add_argument("report"); ap.add_argument("--kernel", default=""); ap.add_argument("--top", type=int, default=40)
a = ap.parse_args()
txt = subprocess.run(["ncu", "-i", a.report, "--page", "source", "--csv", "--print-source", "sass,cuda"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
per_line = collections.Counter(); per_op = collections.Counter(); line_src = {}
cur_file = None; hdr = None; total = 0; seen_kernels = 0; active = False
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        cur_file = r[1].split("/")[-1]; continue
    if r[0] == "Function Name":
        continue
    if r[0] == "Line No":
        hdr = r; continue
    if r[0] == "Kernel Name":
        seen_kernels += 1; active = (a.kernel in r[1]) and seen_kernels == 1; continue
    if hdr is None or len(r) < 8:
        continue
    if r[0] != "":      # a CUDA source line
        cur_line = (cur_file, r[0]); line_src[cur_line] = r[1].strip()[:110]; continue
    try:
        n = int(r[7])
    except ValueError:
        continue
    if seen_kernels > 1:
        continue
    per_line[cur_line] += n; total += n
    m = re.match(r"\s*(@!?U?P\d+\s+)?([A-Z0-9_.]+)", r[3])
    if m:
        per_op[m.group(2).split(".")[0]] += n
print("total warp-instructions executed: %d" % total)
print("\n== per opcode")
for op, n in per_op.most_common(25):
    print("%-10s %10d %5.1f%%" % (op, n, 100.0 * n / total))
print("\n== per source line (top %d)" % a.top)
for k, n in per_line.most_common(a.top):
    print("%6.2f%% %9d  %s:%s  %s" % (100.0 * n / total, n, k[0], k[1], line_src.get(k, "")))
