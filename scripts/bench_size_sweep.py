#!/usr/bin/env python
"""us per lockstep step for every board size and two output layouts (random in-kernel policy, auto-reset).
usage: python scripts/bench_size_sweep.py [num_envs]"""
import json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gymothelloenv_b200 import VecOthelloEnv
_pk = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
PEAK = float(json.load(open(_pk))["hbm_gbs"]) if os.path.isfile(_pk) else 6650.0
B = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 18
for S in range(4, 17):
    row = {"S": S, "envs": B}
    for name, layout, dt in (("board_i8", "board", torch.int8), ("planes4_f32", "planes4", torch.float32)):
        env = VecOthelloEnv(B, board_size=S, seed=0, auto_reset=True, obs_layout=layout, obs_dtype=dt)
        env.reset()
        for _ in range(S * 2):
            env.step(policy="random")
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(100):
            env.step(policy="random")
        b.record()
        torch.cuda.synchronize()
        us = a.elapsed_time(b) * 10.0
        planes, es = (1, 1) if layout == "board" else (4, 4)
        W = (S * S + 63) // 64
        bytes_per = 2 * (24 * W + 8) + planes * S * S * es + S * S + 15
        row[name + "_us"] = round(us, 2)
        row[name + "_hbm_frac"] = round(bytes_per * B / (us * 1e-6) / 1e9 / PEAK, 3)
        del env
    print(json.dumps(row))
