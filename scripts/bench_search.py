#!/usr/bin/env python
"""Decisions/s and positions scored per decision of oth_search on mid-game 8x8 positions: full-width maximin vs
alpha-beta, per thread vs lane-cooperative, next to the round-1 maximin kernels (oth_policy).
usage (GPU box): python scripts/bench_search.py [num_positions] > gpurun_out/<tag>/search.jsonl"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from gymothelloenv_b200 import VecOthelloEnv  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 16
env = VecOthelloEnv(B, board_size=8, seed=0, auto_reset=True)
env.reset()
for t in range(24):                                      # mid-game: 24 random plies from the start
    env.step(policy="random")


def timed(fn, reps):
    fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        out = fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps, out


ref = {}
for depth in (2, 3, 4, 5, 6):
    n = B if depth <= 4 else (B // 8 if depth == 5 else B // 64)
    sub = VecOthelloEnv(n, board_size=8, seed=0)
    st = env.get_state()
    sub.set_state({k: (v[:, :n] if v.dim() == 2 else v[:n]) if isinstance(v, torch.Tensor) else v for k, v in st.items()})
    reps = 3 if depth <= 4 else 1
    rows = []
    if depth <= 4:
        ms, a = timed(lambda: sub.policy_actions("maximin", depth), reps)
        ref[depth] = a.clone()
        rows.append(("round-1 maximin kernel (oth_policy, lane-cooperative, full width)", ms, None, a))
    for name, prune, coop in (("full width, per thread", False, False), ("alpha-beta, per thread", True, False),
                              ("alpha-beta, lane-cooperative", True, True)):
        if not prune and depth >= 6:
            continue
        ms, (a, nodes) = timed(lambda: sub.search_actions(depth, prune=prune, cooperative=coop, with_nodes=True), reps)
        rows.append((name, ms, float(nodes.sum().item()) / n, a))
    base = rows[0][3]
    for name, ms, npd, a in rows:
        print(json.dumps({"depth": depth, "positions": n, "variant": name, "ms": ms, "decisions_per_s": n / (ms * 1e-3),
                          "positions_scored_per_decision": npd, "same_moves_as_first_row": bool(torch.equal(a, base))}), flush=True)
