#!/bin/bash
# Multi-GPU session (gpurun --gpus 8): the driver's launch form for N = 8 / 4 / 2.
set -u
TAG=$1; OUT=gpurun_out/$TAG; mkdir -p $OUT
run() {  # N name args...
  local N=$1 name=$2; shift 2
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) bench.py --gpus $N "$@" > $OUT/$name.json 2> $OUT/$name.err
  echo "$name rc=$?"
  python - $OUT/$name.json <<'PY'
import json,sys
try:
    d=json.loads([l for l in open(sys.argv[1]).read().splitlines() if l.startswith("{")][-1])
    r=d.get("roofline") or {}
    extra = {k:(round(v["ms_per_step"]*1e3,2), "%.3g" % v["env_steps_per_s"]) for k,v in d["sizes"].items()} if "sizes" in d else ""
    e=d.get("e2e") or {}
    print("  n_gpus %d value %.4g %s ms/step %.5f frac %s | e2e %.4g (%.1f GB/s/GPU, %s of ceiling) %s" % (d["n_gpus"], d["value"], d["unit"], d["ms_per_step"], r.get("frac"), e.get("value", 0), e.get("host_link_gbs_per_gpu", 0), e.get("frac_of_host_copy_ceiling"), extra))
    if d.get("e2e_packed"): p=d["e2e_packed"]; print("  e2e_packed %.4g (%.1f GB/s/GPU, %s of ceiling)" % (p["value"], p["host_link_gbs_per_gpu"], p.get("frac_of_host_copy_ceiling")))
    if "pairings" in d:
        for k,v in d["pairings"].items(): print("  ", k, "%.4g games/s" % v["games_per_s"], v["black_wdl_pct"])
except Exception as ex: print("  FAILED", ex)
PY
}
run 8 n8_headline --steps 200 --warmup 20
run 8 n8_sizes --config sizes --steps 100 --warmup 10
run 8 n8_tournament --config tournament --steps 3 --warmup 1
run 4 n4_sizes --config sizes --steps 100 --warmup 10
run 2 n2_sizes --config sizes --steps 100 --warmup 10
run 8 n8_ppo_feed --config ppo_feed --steps 50 --warmup 10
ls $OUT
