#!/usr/bin/env python
"""BASELINE configs[4]: board sizes 6 / 10 / 12 (and 8), 2^20 random-policy envs PER GPU, lockstep step + int8 obs +
legal mask with auto-reset, on every rank of one box (run under torchrun; weak scaling, no data-path collective).
One JSON line per size on rank 0: aggregate env steps/s over the max-over-ranks device time.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29551 \\
        scripts/bench_sizes_sharded.py
"""
import json, os, sys
import torch
import torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gymothelloenv_b200 import VecOthelloEnv

B, K = 1 << 20, 300
rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
dev = torch.device("cuda", local)
torch.cuda.set_device(dev)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
_pk = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
PEAK = float(json.load(open(_pk))["hbm_gbs"]) if os.path.isfile(_pk) else 6650.0
for S in (6, 8, 10, 12):
    env = VecOthelloEnv(B, board_size=S, device=dev, seed=0, auto_reset=True, obs_layout="board", obs_dtype=torch.int8,
                        env_id_base=rank * B)
    env.reset()
    for _ in range(2 * S):
        env.step(policy="random")
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(dev)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(K):
        env.step(policy="random")
    b.record()
    torch.cuda.synchronize(dev)
    t = torch.tensor([a.elapsed_time(b) * 1e-3], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    sec = float(t.item())
    W = (S * S + 63) // 64
    bytes_per = 2 * (24 * W + 8) + 2 * S * S + 15
    if rank == 0:
        print(json.dumps({"config": "configs[4]", "board_size": S, "n_gpus": world, "envs_per_gpu": B, "steps": K,
                          "us_per_step_max_over_ranks": round(sec / K * 1e6, 2),
                          "env_steps_per_s": world * B * K / sec, "bytes_per_env_step": bytes_per,
                          "hbm_frac_per_gpu": round(bytes_per * B * K / sec / 1e9 / PEAK, 3)}), flush=True)
    del env
if world > 1:
    dist.destroy_process_group()
