"""Loaders for the committed golden vectors (tests/golden, produced by make_golden.py
from the live reference)."""
import json
import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

_EP_KEYS = ("cfg", "act", "rew", "done", "board", "turn", "legal", "obs", "winner", "cnt", "state4")


def _load_episodes(fname):
    z = np.load(os.path.join(GOLDEN, fname))
    eps = []
    for i in range(int(z["num_episodes"])):
        e = {}
        for k in _EP_KEYS:
            name = "e%d_%s" % (i, k)
            if name in z.files:
                e[k] = z[name]
        S, sudden, disk, seed, p_inv, mask_in_obs = [int(x) for x in e["cfg"]]
        e.update(S=S, sudden=sudden, disk=disk, seed=seed, p_invalid=p_inv, mask_in_obs=mask_in_obs)
        eps.append(e)
    return eps


def load_trajectories():
    """Board sizes 4..12 (make_golden.py) followed by 13..16 (make_golden_large.py)."""
    return _load_episodes("trajectories.npz") + _load_episodes("trajectories_large.npz")


def load_policies():
    out = {}
    for fname in ("policies.npz", "policies_large.npz"):
        z = np.load(os.path.join(GOLDEN, fname))
        for S in z["sizes"]:
            S = int(S)
            out[S] = dict(board=z["s%d_board" % S], turn=z["s%d_turn" % S], choice=z["s%d_choice" % S])
    return out


def load_wrappers():
    with open(os.path.join(GOLDEN, "wrappers.json")) as f:
        return json.load(f)


def load_perft():
    with open(os.path.join(GOLDEN, "perft.json")) as f:
        return {int(k): v for k, v in json.load(f).items()}


def load_selfplay():
    """tests/golden/selfplay.npz (make_golden_selfplay.py): list of cases, each a dict with cfg fields,
    `protagonist` (colour per game, shared by all workers), `test` = (games, wins) and per-call arrays."""
    z = np.load(os.path.join(GOLDEN, "selfplay.npz"))
    cases = []
    for c in range(int(z["num_cases"])):
        pre = "c%d_" % c
        S, B, G, disk, seed, tg = [int(x) for x in z[pre + "cfg"]]
        d = dict(S=S, B=B, games=G, disk=bool(disk), seed=seed, test_games=tg,
                 protagonist=z[pre + "protagonist"], test=tuple(int(x) for x in z[pre + "test"]))
        for k in ("game", "obs", "actions", "rewards", "dones", "types", "choices", "outputs", "masks", "bad_masks"):
            d[k] = z[pre + k]
        cases.append(d)
    return cases
