#!/usr/bin/env python
"""Bucket an ncu source page of k_step by code region and pipe.
usage: ncu_bucket_summary.py report.ncu-rep [--kernel k_step]"""
import argparse, collections, csv, io, re, subprocess

ap = argparse.ArgumentParser()
ap.add_argument("report"); ap.add_argument("--kernel", default="k_step")
a = ap.parse_args()
ALU = {"LOP3", "SHF", "ISETP", "SEL", "IADD3", "PRMT", "LEA", "VIADD", "PLOP3", "IABS", "IMNMX", "VIMNMX", "FLO", "ICMP", "SGXT", "BMSK", "MOV", "UISETP"}
XU = {"BREV", "POPC", "MUFU"}
BUCKETS = [  # (file, first line, last line, name)  -- line ranges of the CURRENT sources
    ("oth_core.cuh", 183, 237, "legal moves"), ("oth_core.cuh", 239, 314, "flips"),
    ("oth_core.cuh", 316, 358, "nth_bit / bit utils"), ("oth_core.cuh", 360, 415, "philox + bounded"),
    ("oth_core.cuh", 84, 124, "BB ops (select/popc/test)"), ("oth_core.cuh", 417, 515, "env_step / reset / random_pick"),
    ("oth_kernels.cuh", 61, 74, "spread4 / lds_u8"), ("oth_kernels.cuh", 141, 213, "nibble_bytes + fast8_full"),
    ("oth_kernels.cuh", 347, 440, "stage + emit_dense"), ("oth_kernels.cuh", 441, 470, "load/store env"),
]
txt = subprocess.run(["ncu", "-i", a.report, "--page", "source", "--csv", "--print-source", "sass,cuda"],
                     capture_output=True, text=True).stdout
tot = collections.Counter(); alu = collections.Counter(); fma = collections.Counter(); xu = collections.Counter(); lsu = collections.Counter()
cur_file = None; hdr = None; seen = 0; cur = None
for r in csv.reader(io.StringIO(txt)):
    if not r: continue
    if r[0] == "File Path": cur_file = r[1].split("/")[-1]; continue
    if r[0] == "Line No": hdr = r; continue
    if r[0] == "Kernel Name": seen += 1; continue
    if r[0] == "Function Name" or hdr is None or len(r) < 8: continue
    if r[0] != "": cur = (cur_file, int(r[0]) if r[0].isdigit() else -1); continue
    if seen > 1: continue
    try: n = int(r[7])
    except ValueError: continue
    m = re.match(r"\s*(@!?U?P\d+\s+)?([A-Z0-9_.]+)", r[3])
    op = m.group(2).split(".")[0] if m else "?"
    name = "other (%s)" % cur[0]
    for f, lo, hi, nm in BUCKETS:
        if cur[0] == f and lo <= cur[1] <= hi: name = nm; break
    tot[name] += n
    if op in ALU: alu[name] += n
    elif op in ("IMAD", "FFMA", "FMUL", "FADD", "IMUL"): fma[name] += n
    elif op in XU: xu[name] += n
    elif op in ("LDS", "STS", "LDG", "STG", "LDC", "LDCU", "ATOMG"): lsu[name] += n
T = sum(tot.values())
print("| region | all instr | ALU pipe | FMA pipe | XU | LSU/const |\n|---|---|---|---|---|---|")
for k, n in tot.most_common():
    print("| %s | %.1f%% | %.1f%% | %.1f%% | %.1f%% | %.1f%% |" % (k, 100 * n / T, 100 * alu[k] / T, 100 * fma[k] / T, 100 * xu[k] / T, 100 * lsu[k] / T))
print("| TOTAL | 100%% | %.1f%% | %.1f%% | %.1f%% | %.1f%% |" % tuple(100 * sum(c.values()) / T for c in (alu, fma, xu, lsu)))
