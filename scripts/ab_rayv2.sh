set -u
OUT=gpurun_out/r01w; mkdir -p $OUT
SHORT="python bench.py --steps 1500 --warmup 100 --no-cpu-baseline --e2e-steps 3"
echo "== A (HEAD) pytest"; timeout 1200 python -m pytest tests -m gpu -x -q > $OUT/pytest_A.log 2>&1; echo rc=$?; tail -3 $OUT/pytest_A.log
echo "== A bench"; $SHORT > $OUT/bench_A.json 2>$OUT/bench_A.err; python - <<'PY'
import json; d=json.loads(open('gpurun_out/r01w/bench_A.json').read().strip().splitlines()[-1]); print('A', d['ms_per_step'], d['roofline']['frac'])
PY
python scripts/bench_configs.py > $OUT/configs_A.jsonl 2>&1; tail -12 $OUT/configs_A.jsonl | cut -c1-300
cp gymothelloenv_b200/_lib_rayv2/libothello_b200.so gymothelloenv_b200/_lib/libothello_b200.so
echo "== B (RAY_V2) pytest"; timeout 1200 python -m pytest tests -m gpu -x -q > $OUT/pytest_B.log 2>&1; echo rc=$?; tail -3 $OUT/pytest_B.log
echo "== B bench"; $SHORT > $OUT/bench_B.json 2>$OUT/bench_B.err; python - <<'PY'
import json; d=json.loads(open('gpurun_out/r01w/bench_B.json').read().strip().splitlines()[-1]); print('B', d['ms_per_step'], d['roofline']['frac'])
PY
python scripts/bench_configs.py > $OUT/configs_B.jsonl 2>&1; tail -12 $OUT/configs_B.jsonl | cut -c1-300
