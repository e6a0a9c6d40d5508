#!/bin/bash
# One GPU-box session for kernel A/B work: GPU parity tests on the main build, then bench.py on the main build and on
# every variant library (scripts/build_variant.py), then ncu of the main build's step kernel.
# Usage (under gpurun, repo root):  bash scripts/gpu_ab.sh <tag> [variant ...]
set -u
TAG=$1; shift
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=name,driver_version,clocks.max.sm,clocks.max.mem,power.limit --format=csv > $OUT/gpu.txt 2>&1
echo "== pytest -m gpu"; timeout 1200 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -8 $OUT/pytest_gpu.log
BENCH="python bench.py --steps 2000 --warmup 100 --no-cpu-baseline --e2e-steps 3"
echo "== bench main"; timeout 300 $BENCH > $OUT/bench_main.json 2> $OUT/bench_main.err; echo "rc=$?"
python - $OUT/bench_main.json main <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print(sys.argv[2], "ms_per_step %.5f value %.4g frac %.4f" % (d["ms_per_step"], d["value"], d["roofline"]["frac"]))
except Exception as e: print(sys.argv[2], "FAILED", e)
PY
for v in "$@"; do
  echo "== bench $v"; OTH_LIB_DIR=_lib_$v timeout 300 $BENCH > $OUT/bench_$v.json 2> $OUT/bench_$v.err; echo "rc=$?"
  python - $OUT/bench_$v.json $v <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print(sys.argv[2], "ms_per_step %.5f value %.4g frac %.4f" % (d["ms_per_step"], d["value"], d["roofline"]["frac"]))
except Exception as e: print(sys.argv[2], "FAILED", e)
PY
done
SHORT="python bench.py --steps 20 --warmup 3 --no-cpu-baseline --e2e-steps 3"
echo "== ncu counters of the step kernel (main build)"
ncu --clock-control none -k regex:k_step -s 30 -c 3 --csv --log-file $OUT/ncu_step_metrics.csv --metrics gpu__time_duration.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum,smsp__sass_thread_inst_executed_op_integer_pred_on.sum,sm__inst_executed_pipe_alu.sum,sm__inst_executed_pipe_fma.sum,sm__inst_executed_pipe_xu.sum,sm__inst_executed_pipe_lsu.sum,sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active,sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active,dram__bytes_read.sum,dram__bytes_write.sum,launch__registers_per_thread,l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum $SHORT > $OUT/ncu_metrics.log 2>&1
echo "ncu metrics rc=$?"
echo "== ncu full (main build)"
ncu --set full --clock-control none --import-source on -k regex:k_step -s 40 -c 1 -o $OUT/prof_step $SHORT > $OUT/ncu_full.log 2>&1
echo "ncu full rc=$?"
ls -la $OUT
