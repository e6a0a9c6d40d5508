#!/usr/bin/env python
"""bench.py -- the hot path of gymothelloenv_b200 on the configs BASELINE.json names.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--config NAME]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

--config headline (default)  configs[1]: 8x8 lockstep batch of 2^20 random-policy envs per GPU.  One "step" = ONE
      lockstep ply of the whole batch = ONE launch of k_step<8>: in-kernel random policy (Philox) + rules step +
      auto-reset + the declared outputs (state, int8 obs (B,8,8), dense legal mask (B,64), reward, done,
      player_turn, winner, action).  Envs are staggered over game phases.  `value` = env steps/s with everything
      resident in HBM, all ranks (weak scaling, no data-path collective); `e2e` = the same step through HOST buffers.
--config sizes        configs[4]: board 6 / 10 / 12 (multi-word bitboards), 2^20 envs per GPU
--config ppo_feed     configs[3]: 65,536 envs, planes4 float32 obs + mask (the PPO rollout feed)
--config tournament   configs[2] (+ configs[0] protocol): fused greedy vs maximin-2/3 tournaments, 2^21 games per GPU
--config scalar       configs[0] through the scalar Gym drop-in API, next to the reference's wall time

Timing: W warm-up steps, >= 0.3 s of untimed steps, then 5 runs of EXACTLY K steps each between barrier +
synchronize, CUDA events on the launch stream, max over ranks; ms_per_step = median run / K (all runs are printed).
`--impl reference` = the reference's own CPU code (oracle/_ref, see oracle/make_ref.py) on all host cores, for the
same config.  Prints ONE JSON line.
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

CONFIGS = ("headline", "sizes", "ppo_feed", "tournament", "scalar")


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--warmup", type=int, default=100)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="headline", choices=CONFIGS)
    ap.add_argument("--num-envs", type=int, default=1 << 20, help="envs (games for --config tournament) per GPU")
    ap.add_argument("--board-size", type=int, default=10, help="--config sizes: which size `value` reports")
    ap.add_argument("--pairing", default="greedy_vs_maximin2", help="--config tournament: which pairing `value` reports")
    ap.add_argument("--repeats", type=int, default=5, help="timed runs of K steps; the median is reported")
    ap.add_argument("--e2e-steps", type=int, default=30)
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="bounded CPU-baseline sample per worker")
    ap.add_argument("--cpu-maximin3", action="store_true", help="--config tournament: also sample maximin-3 on the CPU")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--graph", action="store_true", help="headline: replay a one-step CUDA graph per timed step")
    ap.add_argument("--profile-timed-region", action="store_true",
                    help="cudaProfilerStart/Stop around the first timed run of K steps (ncu --profile-from-start off)")
    return ap.parse_args()


def main():
    a = parse()
    if a.impl == "reference":
        from benchlib import reference_arm
        return reference_arm.run(a)
    if a.config == "headline":
        from benchlib import headline
        return headline.run(a)
    if a.config == "sizes":
        from benchlib import configs
        return configs.run_sizes(a)
    if a.config == "ppo_feed":
        from benchlib import configs
        return configs.run_ppo_feed(a)
    if a.config == "tournament":
        from benchlib import configs2
        return configs2.run_tournament(a)
    from benchlib import configs2
    return configs2.run_scalar(a)


if __name__ == "__main__":
    sys.exit(main())
