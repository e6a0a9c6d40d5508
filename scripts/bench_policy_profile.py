#!/usr/bin/env python
"""One launch of the batched policy kernel per (kind, depth) on mid-game positions, for ncu.
usage: ncu ... python scripts/bench_policy_profile.py [num_envs] [plies]"""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gymothelloenv_b200 import VecOthelloEnv
B = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 16
plies = int(sys.argv[2]) if len(sys.argv) > 2 else 24
env = VecOthelloEnv(B, board_size=8, seed=0, auto_reset=True)
env.reset()
for _ in range(plies):
    env.step(policy="random")
print("mean mobility", float(env.legal_mask.sum(1).float().mean()))
for d in (2, 3, 4):
    env.policy_actions("maximin", d)
torch.cuda.synchronize()
