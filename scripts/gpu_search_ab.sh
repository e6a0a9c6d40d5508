#!/bin/bash
# Search / tournament validation + timing after a change of the policies.  usage (under gpurun): bash scripts/gpu_search_ab.sh <tag>
set -u
TAG=$1; OUT=gpurun_out/$TAG; mkdir -p $OUT
nvidia-smi --query-gpu=name,driver_version,clocks.max.sm --format=csv,noheader > $OUT/gpu.txt
echo "== pytest (search, tournament, self-play, teacher/student, readme table, scalar api, no-alloc)"
timeout 1500 python -m pytest tests/test_gpu_search.py tests/test_gpu_tournament.py tests/test_gpu_selfplay.py \
    tests/test_gpu_teacher_student.py tests/test_gpu_readme_table.py tests/test_gpu_scalar_api.py tests/test_gpu_dqn_feed.py \
    tests/test_zz_no_alloc.py -m gpu -x -q > $OUT/pytest_gpu.log 2>&1
echo "pytest rc=$?"; tail -5 $OUT/pytest_gpu.log
for p in greedy_vs_maximin2 greedy_vs_maximin3; do
  python bench.py --config tournament --pairing $p --steps 5 --warmup 2 --no-cpu-baseline > $OUT/bench_tournament_$p.json 2> $OUT/bench_tournament_$p.err
  python - $OUT/bench_tournament_$p.json <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    for k,v in d["pairings"].items(): print(k, "games/s %.4g" % v["games_per_s"], "plies/s %.4g" % v["outer_plies_per_s"], "WDL", v["black_wdl_pct"])
except Exception as e: print("FAILED", e)
PY
done
python scripts/bench_search.py > $OUT/bench_search.json 2> $OUT/bench_search.err; tail -3 $OUT/bench_search.json | cut -c1-600
