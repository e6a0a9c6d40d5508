#!/usr/bin/env python
"""Positions scored per outer ply of the tournament pairings (README protocol: first <= 10 plies random), measured by
playing the games in lockstep on VecOthelloEnv and asking oth_search (full width, the same moves as maximin) for its
node count at every maximin decision.  Writes `nodes_per_ply` into profiles/int_count.json for the rollout kernels.
usage (GPU box): python scripts/count_search_nodes.py [games]   (prints JSON lines; run scripts/apply_search_nodes.py here)"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from gymothelloenv_b200 import VecOthelloEnv  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
for name, black, white in (("greedy_vs_maximin2", ("greedy", 1), ("maximin", 2)), ("greedy_vs_maximin3", ("greedy", 1), ("maximin", 3)),
                           ("maximin2_vs_greedy", ("maximin", 2), ("greedy", 1)), ("maximin3_vs_greedy", ("maximin", 3), ("greedy", 1))):
    n = B if "3" not in name else B // 4
    env = VecOthelloEnv(n, board_size=8, seed=0, initial_rand_steps=10)
    env.reset()
    nodes = torch.zeros((), dtype=torch.int64, device=env.device)
    decisions = torch.zeros((), dtype=torch.int64, device=env.device)
    plies = torch.zeros((), dtype=torch.int64, device=env.device)
    for t in range(70):
        live = ~env.terminated()
        if not bool(live.any()):
            break
        turn = env.player_turn.clone()
        cnt, mx = env.rand_steps()
        opening = cnt < mx                                        # the env overrides these plies with random moves
        acts = torch.zeros((n,), dtype=torch.int64, device=env.device)
        for colour, (kind, depth) in ((-1, black), (1, white)):
            sel = live & (turn == colour)
            if kind == "greedy":
                a = env.policy_actions("greedy")
            else:
                a, nd = env.search_actions(depth, prune=False, cooperative=False, with_nodes=True)
                use = sel & ~opening
                nodes += nd[use].sum()
                decisions += use.sum()
            acts = torch.where(sel, a, acts)
        plies += live.sum()
        env.step(acts)
    print(json.dumps({"pairing": name, "games": n, "plies": int(plies), "maximin_decisions": int(decisions),
                      "positions_scored": int(nodes), "nodes_per_ply": int(nodes) / int(plies),
                      "nodes_per_maximin_decision": int(nodes) / max(1, int(decisions))}), flush=True)
