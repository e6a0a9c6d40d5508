#!/usr/bin/env python
"""us per step of the 8x8 step kernel at small batch sizes (int8 board + mask = the fast kernel; planes4 float32 = the
PPO feed through the general kernel), in-kernel random policy, CUDA-graph replay of 20 steps so that launch latency
does not hide the kernel.  usage (GPU box): [OTH_LIB_DIR=_lib_<variant>] python scripts/bench_small_batches.py"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from gymothelloenv_b200 import VecOthelloEnv  # noqa: E402

for layout, dtype in (("board", torch.int8), ("planes4", torch.float32)):
    for B in (4096, 16384, 32768, 65536, 98304, 131072, 262144):
        env = VecOthelloEnv(B, board_size=8, seed=0, auto_reset=True, obs_layout=layout, obs_dtype=dtype)
        env.reset()
        for _ in range(20):
            env.step(policy="random")
        g = env.capture_steps("random", 20)
        for _ in range(5):
            g.replay()
        torch.cuda.synchronize()
        best = 1e9
        for _ in range(5):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(10):
                g.replay()
            b.record()
            torch.cuda.synchronize()
            best = min(best, a.elapsed_time(b) / 200.0)
        print(json.dumps({"layout": layout, "envs": B, "us_per_step": best * 1e3, "env_steps_per_s": B / (best * 1e-3)}), flush=True)
