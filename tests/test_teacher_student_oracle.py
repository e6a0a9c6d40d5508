"""The teacher/student protocol oracle against the live reference's recorded tensors (CPU)."""
import os

import numpy as np
import pytest

from oracle.teacher_student_oracle import TeacherStudentOracle
from tests.selfplay_fake import FakePolicy

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = np.load(os.path.join(HERE, "golden", "teacher_student.npz"))
STUDENT_SALT = 4321
WIN_PAIRS = [({"rand": 0.625, "greedy": 0.25}, {"rand": 0.5, "greedy": 0.125}),
             ({"rand": 0.5, "greedy": 0.5}, {"rand": 0.75, "greedy": 0.125}),
             ({"rand": 0.0, "greedy": 0.0}, {"rand": 0.0, "greedy": 0.0})]
KEYS = ("who", "obs", "actions", "rewards", "dones", "types", "choices", "outputs")


def policies_for(S, teacher):
    t, s = FakePolicy(S, 0), FakePolicy(S, STUDENT_SALT)
    return {teacher: t, -teacher: s}


def golden_case(c):
    pre = "c%d_" % c
    S, B, E, disk = (int(x) for x in GOLD[pre + "cfg"])
    return S, B, [int(t) for t in GOLD[pre + "teachers"]], bool(disk), {k: GOLD[pre + k] for k in KEYS + ("episode",)}


def test_live_reference_never_blocked():
    for c in range(int(GOLD["num_cases"])):
        assert tuple(GOLD["c%d_blocked" % c]) == (-1, -1, 0)


@pytest.mark.parametrize("c", range(int(GOLD["num_cases"])))
def test_oracle_reproduces_live_reference(c):
    S, B, teachers, disk, g = golden_case(c)
    orc = TeacherStudentOracle(B, S, disk_reward=disk)
    k = 0
    for ep, teacher in enumerate(teachers):
        orc.reset(teacher, *WIN_PAIRS[ep])
        pol = policies_for(S, teacher)
        n_calls = int((g["episode"] == ep).sum())
        for _ in range(n_calls):
            out = orc.step(pol)
            for key in KEYS:
                assert np.array_equal(out[key], g[key][k]), (c, ep, k, key)
            k += 1
    assert k == len(g["episode"])


def test_golden_covers_the_hard_cases():
    """forced-pass runs (same colour twice, stale saved state), both end-of-game orders, 'over' rows,
    teacher on both colours, a non-zero teacher reward"""
    passes = overs = stale = 0
    ends = set()
    for c in range(int(GOLD["num_cases"])):
        S, B, teachers, disk, g = golden_case(c)
        ty, dn, ob = g["types"], g["dones"], g["obs"]
        for i in range(B):
            for k in range(1, len(ty)):
                if g["episode"][k] != g["episode"][k - 1]:
                    continue
                if ty[k][i] == ty[k - 1][i] and ty[k][i] < 2 and not dn[k - 1][i]:
                    passes += 1
                    stale += int(np.array_equal(ob[k][i], ob[k - 1][i]))
                if dn[k][i] and not dn[k - 1][i] and ty[k][i] < 2:
                    ends.add(int(ty[k][i]))
            overs += int((ty[:, i] == 2).sum())
        assert set(np.unique(g["who"])) == {0, 1}
    assert passes >= 10 and stale >= 5 and overs >= 5 and ends == {0, 1}
    r = GOLD["c0_rewards"]
    assert (np.abs(r - 0.25) < 1e-7).any() and (np.abs(r - 0.125) < 1e-7).any()
