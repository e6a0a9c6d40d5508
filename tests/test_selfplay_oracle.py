"""CPU: the self-play protocol restatement (oracle/selfplay_oracle.py) against the golden
vectors recorded from the live reference's PPOEnvs + forked subproc_worker processes."""
import numpy as np
import pytest

from oracle.selfplay_oracle import SelfPlayOracle, rule_base_games
from tests import golden_io
from tests.selfplay_fake import FakePolicy

CASES = golden_io.load_selfplay()
KEYS = ("obs", "actions", "rewards", "dones", "types", "choices", "outputs")


@pytest.mark.parametrize("c", range(len(CASES)))
def test_selfplay_oracle_reproduces_reference_ppoenvs(c):
    case = CASES[c]
    S, B = case["S"], case["B"]
    orc = SelfPlayOracle(B, S, disk_reward=case["disk"])
    pol = FakePolicy(S)
    k = 0
    n_pass_emits = 0
    for g in range(case["games"]):
        orc.reset(np.full(B, case["protagonist"][g]))
        while True:
            out = orc.step(pol)
            assert int(case["game"][k]) == g
            for key in KEYS:
                assert np.array_equal(out[key], case[key][k]), (c, g, k, key)
            # masks / bad_masks (envs.py:198-199)
            assert np.array_equal(case["masks"][k], 1.0 - out["dones"].astype(np.float32))
            assert (case["bad_masks"][k] == 1.0).all()
            n_pass_emits += int(((out["types"] == 0) & (out["dones"] == 0) & (out["rewards"] == 0)).sum())
            k += 1
            if out["dones"].all():
                break
    assert k == len(case["game"])


def test_golden_covers_both_colours_and_pass_runs():
    colours = set()
    for case in CASES:
        colours.update(int(x) for x in case["protagonist"][:case["games"]])
    assert colours == {-1, 1}


@pytest.mark.parametrize("c", [i for i, cs in enumerate(CASES) if cs["test_games"]])
def test_rule_base_game_wins_match_reference(c):
    case = CASES[c]
    S, B = case["S"], case["B"]
    per = case["test_games"] // B
    prot = np.repeat(case["protagonist"][case["games"]:case["games"] + per, None], B, axis=1)
    wins = rule_base_games(FakePolicy(S), "greedy", per, S, prot, disk_reward=case["disk"])
    assert case["test"] == (per * B, int(wins.sum()))
