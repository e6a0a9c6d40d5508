#!/bin/bash
# Round-2 GPU session: parity suite, the driver's own bench commands, every --config once, ncu counters.
set -u
TAG=$1
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=name,driver_version,clocks.max.sm,clocks.max.mem,power.limit --format=csv > $OUT/gpu.txt 2>&1
echo "== pytest -m gpu"; timeout 1500 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -6 $OUT/pytest_gpu.log
echo "== bench (driver command)"; timeout 600 python bench.py --gpus 1 --steps 20 --warmup 5 > $OUT/bench_driver.json 2> $OUT/bench_driver.err; echo "rc=$?"; tail -3 $OUT/bench_driver.err
echo "== bench (2000 steps)"; timeout 600 python bench.py --steps 2000 --warmup 100 --no-cpu-baseline > $OUT/bench_2000.json 2> $OUT/bench_2000.err; echo "rc=$?"
echo "== bench reference arm (driver command)"; timeout 600 python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > $OUT/bench_ref.json 2> $OUT/bench_ref.err; echo "rc=$?"; tail -3 $OUT/bench_ref.err
for C in sizes ppo_feed tournament scalar; do
  echo "== bench --config $C"; timeout 900 python bench.py --config $C --steps 20 --warmup 5 --cpu-seconds 6 > $OUT/bench_$C.json 2> $OUT/bench_$C.err; echo "rc=$?"; tail -3 $OUT/bench_$C.err
done
python - $OUT <<'PY'
import json,sys,glob,os
for f in sorted(glob.glob(sys.argv[1]+"/bench_*.json")):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        r=d.get("roofline") or {}
        print(os.path.basename(f), "value %.4g %s ms/step %s frac %s int_frac %s e2e %.4g cpu %s" % (d["value"], d["unit"], d.get("ms_per_step"), r.get("frac"), r.get("int_frac"), (d.get("e2e") or {}).get("value", float("nan")), (d.get("cpu_baseline") or {}).get("value")))
    except Exception as e:
        print(os.path.basename(f), "FAILED", e)
PY
echo "== ncu counters"; bash scripts/gpu_counts.sh $TAG
ls $OUT
