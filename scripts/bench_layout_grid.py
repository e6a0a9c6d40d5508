#!/usr/bin/env python
"""Fraction of measured HBM bandwidth for every (board size, obs layout, dtype): finds slow writer paths.
usage: python scripts/bench_layout_grid.py [num_envs]"""
import json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gymothelloenv_b200 import VecOthelloEnv
B = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 18
PLANES = {"board": 1, "board+mask": 2, "planes4": 4, "planes3": 3}
for S in range(4, 17):
    row = {"S": S}
    for layout, P in PLANES.items():
        for dname, dt, es in (("i8", torch.int8, 1), ("f32", torch.float32, 4)):
            env = VecOthelloEnv(B, board_size=S, seed=0, auto_reset=True, obs_layout=layout, obs_dtype=dt)
            env.reset()
            for _ in range(S * 2):
                env.step(policy="random")
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(50):
                env.step(policy="random")
            b.record()
            torch.cuda.synchronize()
            us = a.elapsed_time(b) * 20.0
            W = (S * S + 63) // 64
            bytes_per = 2 * (24 * W + 8) + P * S * S * es + S * S + 15
            row["%s/%s" % (layout, dname)] = [round(us, 1), round(bytes_per * B / (us * 1e-6) / 1e9 / 6454.6, 2)]
            del env
    print(json.dumps(row))
