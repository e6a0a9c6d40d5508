#!/bin/bash
# Multi-GPU check (default 2 GPUs; second argument = N) of the driver's launch form (gpurun --gpus 2): headline, tournament (NCCL all-reduce of the tally),
# and the reference arm under torchrun (rank 0 alone works).
set -u
TAG=$1; NG=${2:-2}; OUT=gpurun_out/$TAG; mkdir -p $OUT
run() {
  local N=$1 name=$2; shift 2
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) bench.py --gpus $N "$@" > $OUT/$name.json 2> $OUT/$name.err
  echo "$name rc=$?"; grep '^{' $OUT/$name.json | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('  n_gpus', d.get('n_gpus'), 'value %.4g' % d['value'], d['unit'], 'e2e %.4g' % (d.get('e2e') or {}).get('value', 0), d.get('clocks'))
for k,v in (d.get('pairings') or {}).items(): print('   ', k, '%.4g games/s' % v['games_per_s'], v['black_wdl_pct'])
"
}
run $NG n${NG}_headline --steps 20 --warmup 5
run $NG n${NG}_tournament --config tournament --steps 3 --warmup 1

