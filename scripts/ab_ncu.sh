# usage: bash scripts/ab_ncu.sh <tag> <variant> [<variant> ...]   compares _lib (A) with _lib_<variant> on k_step<8>:
# bench.py (CUDA events, 1500 back-to-back steps) and an ncu metric list of three launches each.
set -u
OUT=gpurun_out/$1; mkdir -p $OUT; shift
M=gpu__time_duration.sum,smsp__inst_executed.sum,sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active,sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active,launch__registers_per_thread,sm__warps_active.avg.pct_of_peak_sustained_active,smsp__issue_active.avg.pct_of_peak_sustained_active,dram__bytes_write.sum,dram__bytes_read.sum,sm__inst_executed_pipe_lsu.sum,sm__inst_executed_pipe_alu.sum,sm__inst_executed_pipe_fma.sum,smsp__warp_issue_stalled_math_pipe_throttle_per_warp_active.pct,smsp__warp_issue_stalled_long_scoreboard_per_warp_active.pct,smsp__warp_issue_stalled_not_selected_per_warp_active.pct,smsp__warp_issue_stalled_wait_per_warp_active.pct
SHORT="python bench.py --steps 40 --warmup 5 --no-cpu-baseline --e2e-steps 3"
cp gymothelloenv_b200/_lib/libothello_b200.so /tmp/lib_A.so
for V in A "$@"; do
  if [ $V = A ]; then cp /tmp/lib_A.so gymothelloenv_b200/_lib/libothello_b200.so; else cp gymothelloenv_b200/_lib_$V/libothello_b200.so gymothelloenv_b200/_lib/libothello_b200.so; fi
  python bench.py --steps 1500 --warmup 100 --no-cpu-baseline --e2e-steps 3 > $OUT/bench_$V.json 2> $OUT/bench_$V.err
  python -c "import json;d=json.loads(open('$OUT/bench_$V.json').read().strip().splitlines()[-1]);print('$V bench', d['ms_per_step'], d['roofline']['frac'])"
  ncu --metrics $M --clock-control none -k regex:k_step -s 20 -c 3 --csv --log-file $OUT/ncu_$V.csv $SHORT > $OUT/ncu_$V.log 2>&1
  python - <<PY
import csv
rows=[r for r in csv.reader(open('$OUT/ncu_$V.csv')) if len(r)>10]
h=rows[0]; mi=h.index('Metric Name'); vi=h.index('Metric Value'); ii=h.index('ID')
first=rows[1][ii]
print('$V', '; '.join('%s=%s' % (r[mi].split('.')[0].replace('smsp__warp_issue_stalled_','stall_').replace('_per_warp_active',''), r[vi]) for r in rows[1:] if r[ii]==first))
PY
done
cp /tmp/lib_A.so gymothelloenv_b200/_lib/libothello_b200.so
