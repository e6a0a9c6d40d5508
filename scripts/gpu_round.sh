#!/bin/bash
# One GPU-box session: smoke, GPU parity tests, bench, ncu launch list + full capture of k_step.
# Usage (from the repo root, under gpurun):  bash scripts/gpu_round.sh [tag]
set -u
TAG=${1:-r01}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=name,driver_version,clocks.max.sm,clocks.max.mem,power.limit --format=csv > $OUT/gpu.txt 2>&1
echo "== smoke" ; python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke.log 2>&1 ; echo "smoke rc=$?" ; tail -3 $OUT/smoke.log
echo "== pytest -m gpu" ; timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1 ; echo "pytest rc=$?" ; tail -15 $OUT/pytest_gpu.log
echo "== bench" ; timeout 600 python bench.py > $OUT/bench.json 2> $OUT/bench.err ; echo "bench rc=$?" ; cat $OUT/bench.json ; tail -5 $OUT/bench.err
echo "== bench reference arm" ; timeout 300 python bench.py --impl reference --steps 5 --warmup 1 > $OUT/bench_ref.json 2> $OUT/bench_ref.err ; echo "rc=$?" ; cat $OUT/bench_ref.json
SHORT="python bench.py --steps 20 --warmup 3 --no-cpu-baseline --e2e-steps 3"
echo "== ncu launch list"
$SHORT > $OUT/short_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches.csv $SHORT > $OUT/ncu_launches.log 2>&1
echo "ncu launches rc=$?"
echo "== ncu launch list of the TIMED REGION only (cudaProfilerStart/Stop around the K steps)"
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file $OUT/launches_timed_region.csv $SHORT --profile-timed-region > $OUT/ncu_launches_timed.log 2>&1
echo "ncu timed-region launches rc=$?"
echo "== ncu full k_step"
$SHORT > $OUT/short_plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_step -s 60 -c 3 -o $OUT/prof_step $SHORT > $OUT/ncu_full.log 2>&1
echo "ncu full rc=$?"
ls -la $OUT
