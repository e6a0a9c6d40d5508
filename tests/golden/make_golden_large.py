#!/usr/bin/env python
"""Golden vectors for the LARGE boards (13x13 .. 16x16: 3-4 words per colour) from the LIVE reference.

    python tests/golden/make_golden_large.py      # writes trajectories_large.npz, policies_large.npz

Same recorders as make_golden.py (imported from it), kept in separate files so the 4..12 fixtures stay
byte-identical.  tests/golden_io.py merges both sets, so every test that replays the golden episodes or the
policy choices covers these sizes too.  The reference accepts any board_size (othello.py:222-254); 16 is the
largest this build instantiates (256 cells = 4 x 64-bit words).
"""
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import make_golden as mg  # noqa: E402

SIZES = [13, 14, 15, 16]


def gen_trajectories():
    eps = []
    for S in SIZES:
        for seed, disk in ((0, 0), (1, 1)):
            eps.append(mg.play_episode(S, 1, disk, seed, 0, 0))
        eps.append(mg.play_episode(S, 0, 1, 10, 15, 0))      # invalid => forfeited turn
        eps.append(mg.play_episode(S, 1, 0, 11, 2, 0))       # invalid => sudden death
        eps.append(mg.play_episode(S, 1, 0, 20, 0, 1))       # possible_actions_in_obs
    flat = {"num_episodes": np.array(len(eps))}
    for i, e in enumerate(eps):
        for k, v in e.items():
            flat["e%d_%s" % (i, k)] = v
    np.savez_compressed(os.path.join(HERE, "trajectories_large.npz"), **flat)
    print("trajectories_large: %d episodes, %d plies" % (len(eps), sum(len(e["act"]) for e in eps)))


def gen_policies():
    flat = {}
    for S in SIZES:
        t0 = time.time()
        pos = mg.sample_positions(S, 2, 16, 2000 + S)
        rows = []
        for i, (board, turn) in enumerate(pos):
            r = mg.policy_choices(S, board, turn, {1, 2} if i < 6 else {1})
            rows.append((r["greedy"], r["mm1"], r["mm2"], r["mm3"], r["mm4"]))
        flat["s%d_board" % S] = np.stack([p[0] for p in pos])
        flat["s%d_turn" % S] = np.array([p[1] for p in pos], dtype=np.int8)
        flat["s%d_choice" % S] = np.array(rows, dtype=np.int16)
        print("policies S=%d: %d positions (%.1fs)" % (S, len(pos), time.time() - t0))
    flat["sizes"] = np.array(SIZES, dtype=np.int32)
    np.savez_compressed(os.path.join(HERE, "policies_large.npz"), **flat)


if __name__ == "__main__":
    which = sys.argv[1:] or ["traj", "policies"]
    if "traj" in which:
        gen_trajectories()
    if "policies" in which:
        gen_policies()
