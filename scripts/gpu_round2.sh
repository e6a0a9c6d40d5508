#!/bin/bash
# Round-2 GPU session: parity suite, the driver's own bench commands, every --config once, ncu counters.
set -u
TAG=$1
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=name,driver_version,clocks.max.sm,clocks.max.mem,power.limit --format=csv > $OUT/gpu.txt 2>&1
echo "== smoke"; python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 $OUT/smoke.log
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
echo "== ncu launch list of the timed region (headline)"
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file $OUT/launches_timed_region.csv python bench.py --steps 20 --warmup 3 --no-cpu-baseline --e2e-steps 3 --profile-timed-region > $OUT/ncu_launches_timed.log 2>&1; echo "rc=$?"
echo "== ncu launch list (whole short run)"
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $OUT/launches.csv python bench.py --steps 20 --warmup 3 --no-cpu-baseline --e2e-steps 3 > $OUT/ncu_launches.log 2>&1; echo "rc=$?"
echo "== ncu full of the headline kernel"
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:k_step -s 19 -c 1 -o $OUT/prof_step_8 python scripts/profile_targets.py step --size 8 > $OUT/ncu_full.log 2>&1; echo "rc=$?"
echo "== ncu full of the maximin-2 tournament kernel"
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:k_rollout -s 1 -c 1 -o $OUT/prof_rollout_gm2 python scripts/profile_targets.py rollout --pairing greedy_vs_maximin2 --games 262144 > $OUT/ncu_full_rollout.log 2>&1; echo "rc=$?"
echo "== search bench"; python scripts/bench_search.py > $OUT/search.jsonl 2> $OUT/search.err; echo "rc=$?"
echo "== ncu counters"; bash scripts/gpu_counts.sh $TAG
ls $OUT
