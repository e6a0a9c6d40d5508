#!/bin/bash
# A/B of step-kernel schedules and variants.  usage (under gpurun): bash scripts/gpu_ab2.sh <tag>
set -u
TAG=$1; OUT=gpurun_out/$TAG; mkdir -p $OUT
echo "== pytest -m gpu"; timeout 1500 python -m pytest ${PYTEST_ARGS:-tests} -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 $OUT/pytest_gpu.log
H="python bench.py --steps 2000 --warmup 100 --no-cpu-baseline --e2e-steps 3"
Z="python bench.py --config sizes --steps 300 --warmup 20 --no-cpu-baseline --e2e-steps 3"
run() {  # name, env assignments..., -- command
  local name=$1; shift
  env "$@" > $OUT/$name.json 2> $OUT/$name.err
  python - $OUT/$name.json $name <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    if "sizes" in d: print(sys.argv[2], {k:(round(v["ms_per_step"]*1e3,2), round(v["hbm_frac"],3)) for k,v in d["sizes"].items()})
    else: print(sys.argv[2], "us/step %.2f frac %.4f" % (d["ms_per_step"]*1e3, d["roofline"]["frac"]))
except Exception as e: print(sys.argv[2], "FAILED", e)
PY
}
run head_main X=1 $H
run head_driver X=1 python bench.py --gpus 1 --steps 20 --warmup 5 --no-cpu-baseline --e2e-steps 3
run sizes_main X=1 $Z
