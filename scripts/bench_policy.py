#!/usr/bin/env python
"""Times the batched policy kernel (VecOthelloEnv.policy_actions) on mid-game 8x8 positions.
usage: python scripts/bench_policy.py [num_envs]"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gymothelloenv_b200 import VecOthelloEnv  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 18
env = VecOthelloEnv(B, board_size=8, seed=0, auto_reset=True)
env.reset()
for _ in range(24):                       # mid-game: ~24 random plies in
    env.step(policy="random")
for kind, depth, reps in (("greedy", 1, 20), ("maximin", 2, 10), ("maximin", 3, 3), ("maximin", 4, 1)):
    env.policy_actions(kind, depth)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        acts = env.policy_actions(kind, depth)
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / reps
    print(json.dumps({"policy": "%s-%d" % (kind, depth), "envs": B, "ms": ms, "decisions_per_s": B / (ms * 1e-3),
                      "checksum": int(acts.sum().item())}))
