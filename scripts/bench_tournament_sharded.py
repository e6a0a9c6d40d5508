#!/usr/bin/env python
"""BASELINE configs[2]: greedy vs maximin-2 / maximin-3 tournaments of 16,777,216 games sharded over the ranks of one
box (run under torchrun; one process per GPU, contiguous block of games per rank, Philox keyed by the global game id,
ONE all-reduce of the 4-word tally over NCCL at the end).  Prints one JSON line per pairing on rank 0: games/s and
outer plies/s over the max-over-ranks device time, and black's W/D/L.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29541 \\
        scripts/bench_tournament_sharded.py [games]
"""
import json, os, sys
import torch
import torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gymothelloenv_b200.tournament import play_tournament

games = int(sys.argv[1]) if len(sys.argv) > 1 else 16777216
rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
dev = torch.device("cuda", local)
torch.cuda.set_device(dev)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
for black, white in (("greedy", "maximin2"), ("maximin2", "greedy"), ("greedy", "maximin3"), ("maximin3", "greedy"),
                     ("rand", "greedy")):
    play_tournament(black, white, num_games=world * 4096, seed=1, device=dev)          # warm-up (+ NCCL init)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(dev)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    res = play_tournament(black, white, num_games=games, initial_rand_steps=10, seed=0, device=dev)
    b.record()
    torch.cuda.synchronize(dev)
    t = torch.tensor([a.elapsed_time(b) * 1e-3], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    sec = float(t.item())
    if rank == 0:
        g = res["games"]
        print(json.dumps({"config": "configs[2] sharded tournament", "black": black, "white": white, "games": g,
                          "n_gpus": world, "seconds_max_over_ranks": sec, "games_per_s": g / sec,
                          "outer_plies_per_s": res["plies"] / sec,
                          "black_wdl_pct": [round(100.0 * res[k] / g, 3) for k in ("wins", "draws", "losses")],
                          "includes": "rollout kernels + the NCCL all-reduce of the 4-word tally + its host read"}),
              flush=True)
if world > 1:
    dist.destroy_process_group()
