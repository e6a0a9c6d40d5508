#!/usr/bin/env python
"""How do the two phases of k_step<8> add up?  Times, on 2^20 8x8 envs (random in-kernel policy, auto-reset):
  full      step + int8 obs + legal mask (the headline launch)
  rules     the same step with NO dense outputs (state + scalars only)
  writer    oth_observe alone (int8 obs + legal mask from the current state)
If full ~= rules + writer the ALU-heavy rules phase and the FMA/LSU-heavy writer phase do not overlap inside the
SM (each warp runs them back to back and the resident warps do not mix phases well enough); if full ~= max(...) they do.
usage: python scripts/bench_phase_split.py [num_envs]"""
import json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gymothelloenv_b200 import VecOthelloEnv

B = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20


def timed(fn, reps=300, warm=20):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) * 1e3 / reps


def mk(layout, mask):
    env = VecOthelloEnv(B, board_size=8, seed=0, auto_reset=True, obs_layout=layout, obs_dtype=torch.int8,
                        with_legal_mask=mask)
    env.reset()
    for _ in range(40):
        env.step(policy="random")
    return env


full = mk("board", True)
rules = mk("none", False)
row = {"envs": B,
       "full_us": round(timed(lambda: full.step(policy="random")), 2),
       "rules_only_us": round(timed(lambda: rules.step(policy="random")), 2),
       "writer_only_us": round(timed(lambda: full.observe()), 2)}
row["sum_us"] = round(row["rules_only_us"] + row["writer_only_us"], 2)
print(json.dumps(row))
