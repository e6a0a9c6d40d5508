#!/usr/bin/env python
"""Turn one gpurun_out/<tag>/ session into committed evidence under profiles/<tag>_*.
usage: python scripts/profile_summary.py <tag>"""
import collections, csv, io, json, os, subprocess, sys

tag = sys.argv[1]
src = os.path.join("gpurun_out", tag)
os.makedirs("profiles", exist_ok=True)
out = []

def launches(fname="launches.csv", title="## ncu launch list (gpu__time_duration.sum, ns; cold-cache, serialised: compare shares)\n",
             cmd="command: python bench.py --steps 20 --warmup 3 --no-cpu-baseline --e2e-steps 3 (first 400 launches)\n"):
    p = os.path.join(src, fname)
    if not os.path.isfile(p):
        return
    rows = list(csv.reader(open(p)))
    for i, r in enumerate(rows):
        if r and r[0] == "ID":
            hdr, start = r, i + 1
            break
    ki, vi = hdr.index("Kernel Name"), hdr.index("Metric Value")
    agg = collections.defaultdict(list)
    for r in rows[start:]:
        if len(r) > vi:
            agg[r[ki][:70]].append(float(r[vi].replace(",", "")))
    tot = sum(sum(v) for v in agg.values())
    out.append(title)
    out.append(cmd)
    out.append("| kernel | launches | total ns | avg ns | share |\n|---|---|---|---|---|")
    for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
        out.append("| `%s` | %d | %.0f | %.0f | %.1f%% |" % (k, len(v), sum(v), sum(v) / len(v), 100 * sum(v) / tot))
    out.append("")

WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "launch__registers_per_thread", "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__cycles_elapsed.avg",
        "l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum", "l1tex__t_requests_pipe_lsu_mem_global_op_st.sum",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smsp__sass_thread_inst_executed_op_integer_pred_on.sum",
        "launch__occupancy_limit_registers", "launch__grid_size", "launch__block_size"]

def full(name):
    rep = os.path.join(src, name + ".ncu-rep")
    if not os.path.isfile(rep):
        return
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    hdr, units = rows[0], rows[1]
    out.append("## ncu --set full: %s (one row per captured launch)\n" % name)
    traffic = None
    for r in rows[2:]:
        out.append("kernel `%s`\n" % r[hdr.index("Kernel Name")])
        out.append("| metric | value | unit |\n|---|---|---|")
        vals = {}
        for w in WANT:
            if w in hdr:
                i = hdr.index(w)
                vals[w] = (r[i], units[i])
                out.append("| %s | %s | %s |" % (w, r[i], units[i]))
        out.append("")
        try:
            def b(k):
                v, u = vals[k]
                return float(v.replace(",", "")) * {"Mbyte": 1e6, "Gbyte": 1e9, "Kbyte": 1e3, "byte": 1}[u]
            traffic = b("dram__bytes_read.sum") + b("dram__bytes_write.sum")
        except Exception:
            pass
    s = subprocess.run([sys.executable, "scripts/ncu_source_summary.py", rep, "--top", "30"], capture_output=True, text=True).stdout
    out.append("### source-page summary (instructions executed per opcode / CUDA line)\n\n```\n" + s + "```\n")
    return traffic

out.append("# profile summary %s\n" % tag)
for f in ("gpu.txt",):
    p = os.path.join(src, f)
    if os.path.isfile(p):
        out.append("```\n" + open(p).read() + "```\n")
for f in ("bench.json", "bench_ref.json"):
    p = os.path.join(src, f)
    if os.path.isfile(p) and os.path.getsize(p):
        out.append("## %s\n\n```json\n%s```\n" % (f, open(p).read()))
launches()
launches("launches_timed_region.csv",
         "## ncu launch list of the TIMED REGION only (cudaProfilerStart/Stop around the K steps; ns per launch)\n",
         "command: ncu --profile-from-start off ... python bench.py --steps 20 --warmup 3 --no-cpu-baseline --e2e-steps 3 "
         "--profile-timed-region (every launch between the two CUDA events that bracket the K timed steps)\n")
traffic = full("prof_step")
open(os.path.join("profiles", tag + "_summary.md"), "w").write("\n".join(out))
if traffic:
    json.dump({"k_step_8_bytes_per_launch": traffic, "source": "profiles/%s_summary.md (ncu --set full, dram__bytes_read+write)" % tag,
               "envs_per_launch": 1 << 20}, open("profiles/traffic.json", "w"))
for f in ("launches.csv", "launches_timed_region.csv"):
    p = os.path.join(src, f)
    if os.path.isfile(p):
        open(os.path.join("profiles", tag + "_" + f), "w").write(open(p).read())
print("wrote profiles/%s_summary.md" % tag, "traffic", traffic)
