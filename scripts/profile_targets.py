#!/usr/bin/env python
"""A few launches of ONE kernel for ncu, plus a JSON line with the units the launch processed.
usage: python scripts/profile_targets.py step --size 10 [--envs 1048576] [--layout board --dtype int8]
       python scripts/profile_targets.py rollout --pairing greedy_vs_maximin2 [--games 262144]
ncu: --kernel-name-base demangled -k regex:'k_step<10' -s <skip> -c 1 ...   (the LAST launches are the steady ones)"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from gymothelloenv_b200 import VecOthelloEnv  # noqa: E402
from gymothelloenv_b200.tournament import rollout  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("target", choices=["step", "rollout"])
ap.add_argument("--size", type=int, default=8)
ap.add_argument("--envs", type=int, default=1 << 20)
ap.add_argument("--layout", default="board")
ap.add_argument("--dtype", default="int8")
ap.add_argument("--pairing", default="greedy_vs_maximin2")
ap.add_argument("--games", type=int, default=1 << 18)
ap.add_argument("--launches", type=int, default=6)
a = ap.parse_args()
if a.target == "step":
    env = VecOthelloEnv(a.envs, board_size=a.size, seed=0, auto_reset=True, obs_layout=a.layout,
                        obs_dtype=torch.int8 if a.dtype == "int8" else torch.float32)
    env.reset()
    # mid-game mix WITHOUT thousands of launches under ncu: stagger by running 2*S plies with auto-reset
    for _ in range(2 * a.size + a.launches):
        env.step(policy="random")
    torch.cuda.synchronize()
    print(json.dumps({"target": "step", "size": a.size, "units_per_launch": a.envs, "launches": 2 * a.size + a.launches}))
else:
    black, white = a.pairing.split("_vs_")
    for _ in range(2):
        tally, _ = rollout(a.games, black, white, board_size=a.size, initial_rand_steps=10, seed=0)
    torch.cuda.synchronize()
    t = [int(x) for x in tally.tolist()]            # every rollout() call returns a fresh tally: ONE launch's plies
    print(json.dumps({"target": "rollout", "pairing": a.pairing, "games": a.games, "units_per_launch": t[3],
                      "plies_per_launch": t[3], "launches": 2}))
