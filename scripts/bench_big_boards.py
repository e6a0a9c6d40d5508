#!/usr/bin/env python
"""Greedy / maximin decisions on multi-word boards (10x10, 12x12, 16x16): fused tournaments and the batched policy call.
usage (GPU box): [OTH_LIB_DIR=_lib_<variant>] python scripts/bench_big_boards.py > gpurun_out/<tag>/big_<variant>.jsonl"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from gymothelloenv_b200 import VecOthelloEnv  # noqa: E402
from gymothelloenv_b200.tournament import rollout  # noqa: E402


def timed(fn):
    fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    out = fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b), out


for S, games in ((10, 1 << 18), (12, 1 << 17), (16, 1 << 15)):
    for black, white, g in (("rand", "greedy", games), ("greedy", "maximin2", games // 8)):
        ms, (tally, _) = timed(lambda: rollout(g, black, white, board_size=S, initial_rand_steps=10, seed=3))
        t = [int(x) for x in tally.tolist()]
        print(json.dumps({"S": S, "pairing": "%s_vs_%s" % (black, white), "games": g, "ms": ms,
                          "games_per_s": g / (ms * 1e-3), "plies_per_s": t[3] / (ms * 1e-3), "tally": t}), flush=True)
    env = VecOthelloEnv(1 << 16, board_size=S, seed=1, auto_reset=True)
    env.reset()
    for _ in range(S * S // 3):
        env.step(policy="random")
    for name, depth in (("greedy", 1), ("maximin", 2)):
        ms, acts = timed(lambda: env.policy_actions(name, depth))
        print(json.dumps({"S": S, "call": "policy_actions(%s, %d)" % (name, depth), "positions": 1 << 16, "ms": ms,
                          "decisions_per_s": (1 << 16) / (ms * 1e-3), "checksum": int(acts.sum())}), flush=True)
