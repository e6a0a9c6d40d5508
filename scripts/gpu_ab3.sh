#!/bin/bash
# A/B of library variants (gymothelloenv_b200/_lib_<name>) on headline, sizes and tournaments.
# usage (under gpurun): bash scripts/gpu_ab3.sh <tag> <variant> [<variant> ...]     ("main" = _lib)
set -u
TAG=$1; shift; OUT=gpurun_out/$TAG; mkdir -p $OUT
for V in "$@"; do
  D=_lib_$V; [ $V = main ] && D=_lib
  export OTH_LIB_DIR=$D
  python bench.py --steps 2000 --warmup 100 --no-cpu-baseline --e2e-steps 3 > $OUT/head_$V.json 2> $OUT/head_$V.err
  python bench.py --config sizes --steps 300 --warmup 20 --no-cpu-baseline --e2e-steps 3 > $OUT/sizes_$V.json 2> $OUT/sizes_$V.err
  python bench.py --config tournament --steps 5 --warmup 2 --no-cpu-baseline > $OUT/tour_$V.json 2> $OUT/tour_$V.err
  python - $OUT $V <<'PY'
import json,sys
out,v=sys.argv[1],sys.argv[2]
def last(f):
    return json.loads(open(f).read().strip().splitlines()[-1])
try:
    d=last("%s/head_%s.json"%(out,v)); print(v, "headline us/step %.2f frac %.4f clocks %s" % (d["ms_per_step"]*1e3, d["roofline"]["frac"], d["clocks"]))
except Exception as e: print(v, "headline FAILED", e)
try:
    d=last("%s/sizes_%s.json"%(out,v)); print(v, "sizes", {k:(round(x["ms_per_step"]*1e3,2), round(x["hbm_frac"],3)) for k,x in d["sizes"].items()})
except Exception as e: print(v, "sizes FAILED", e)
try:
    d=last("%s/tour_%s.json"%(out,v)); print(v, "tournaments", {k:"%.4g"%x["games_per_s"] for k,x in d["pairings"].items()})
except Exception as e: print(v, "tournament FAILED", e)
PY
done
