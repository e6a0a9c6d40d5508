#!/bin/bash
# ncu counters (one steady launch each) of every kernel bench.py reports a roofline for.  Under gpurun:
#   bash scripts/gpu_counts.sh <tag>          -> gpurun_out/<tag>/counts_*.csv + counts_*.json
set -u
TAG=$1
OUT=gpurun_out/$TAG
mkdir -p $OUT
M=gpu__time_duration.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum,smsp__sass_thread_inst_executed_op_integer_pred_on.sum,sm__inst_executed_pipe_alu.sum,sm__inst_executed_pipe_fma.sum,sm__inst_executed_pipe_xu.sum,sm__inst_executed_pipe_lsu.sum,sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active,sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active,dram__bytes_read.sum,dram__bytes_write.sum,launch__registers_per_thread,l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum,smsp__thread_inst_executed_per_inst_executed.ratio
for S in 8 6 10 12; do
  python scripts/profile_targets.py step --size $S > $OUT/counts_step_$S.json 2>$OUT/counts_step_$S.err || { echo "plain run S=$S failed"; continue; }
  SKIP=$((2 * S + 3))
  ncu --clock-control none --kernel-name-base demangled -k regex:k_step -s $SKIP -c 2 --csv --log-file $OUT/counts_step_$S.csv --metrics $M python scripts/profile_targets.py step --size $S > /dev/null 2>$OUT/counts_step_${S}_ncu.err
  echo "step S=$S rc=$?"
done
for P in rand_vs_rand rand_vs_greedy greedy_vs_maximin2 greedy_vs_maximin3; do
  G=262144; [ $P = greedy_vs_maximin3 ] && G=65536
  python scripts/profile_targets.py rollout --pairing $P --games $G > $OUT/counts_rollout_$P.json 2>$OUT/counts_rollout_$P.err || { echo "plain rollout $P failed"; continue; }
  ncu --clock-control none --kernel-name-base demangled -k regex:k_rollout -s 1 -c 1 --csv --log-file $OUT/counts_rollout_$P.csv --metrics $M python scripts/profile_targets.py rollout --pairing $P --games $G > /dev/null 2>$OUT/counts_rollout_${P}_ncu.err
  echo "rollout $P rc=$?"; cat $OUT/counts_rollout_$P.json
done
