#!/usr/bin/env python
"""A/B of the two maximin searches inside oth_rollout (lane-cooperative vs per-thread DFS), 8x8.
usage: python scripts/bench_rollout_ab.py"""
import json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gymothelloenv_b200.tournament import rollout
for black, white, games in (("greedy", "maximin2", 1 << 20), ("greedy", "maximin3", 1 << 18), ("greedy", "maximin4", 1 << 14),
                            ("maximin4", "rand", 1 << 14), ("maximin3", "maximin3", 1 << 17)):
    res = {}
    for name, kw in (("coop", {}), ("per_thread", {"search_per_thread": True})):
        rollout(2048, black, white, seed=1, **kw)
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        tally, _ = rollout(games, black, white, board_size=8, initial_rand_steps=10, seed=0, **kw)
        b.record()
        torch.cuda.synchronize()
        res[name] = (a.elapsed_time(b) * 1e-3, tally.tolist())
    assert res["coop"][1] == res["per_thread"][1], (black, white, res)
    print(json.dumps({"black": black, "white": white, "games": games, "coop_s": res["coop"][0], "per_thread_s": res["per_thread"][0],
                      "coop_speedup": res["per_thread"][0] / res["coop"][0], "tally_identical": True, "tally": res["coop"][1]}))
