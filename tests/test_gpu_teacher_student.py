"""GPU parity of the teacher/student vec-env against the live reference's recorded
tensors (tests/golden/teacher_student.npz) and against the pinned CPU restatement on larger batches."""
import numpy as np
import pytest
import torch

from oracle.teacher_student_oracle import TeacherStudentOracle
from tests.selfplay_fake import FakePolicy
from tests.test_teacher_student_oracle import GOLD, KEYS, STUDENT_SALT, WIN_PAIRS, golden_case

pytestmark = pytest.mark.gpu


def _mk(B, S, disk, legacy=True, **kw):
    from gymothelloenv_b200 import PPOTeacherStudentEnvs
    return PPOTeacherStudentEnvs(B, None, FakePolicy(S, 0), FakePolicy(S, STUDENT_SALT), "cuda:0", board_size=S,
                                 num_disk_as_reward=disk, legacy_stale_state=legacy, **kw)


def _as_numpy(ret):
    who, obs, actions, rewards, dones, infos, outputs, masks, bad = ret
    return dict(who=who.who.cpu().numpy(), obs=obs.cpu().numpy().astype(np.int8),
                actions=actions.cpu().numpy().reshape(-1).astype(np.int64),
                rewards=rewards.cpu().numpy().reshape(-1).astype(np.float32),
                dones=dones.cpu().numpy().astype(np.uint8), types=infos.type.cpu().numpy(),
                choices=infos.choices_mask.cpu().numpy().astype(np.uint8),
                outputs=outputs.cpu().numpy().astype(np.float32),
                masks=masks.cpu().numpy().reshape(-1), bad_masks=bad.cpu().numpy().reshape(-1))


@pytest.mark.parametrize("c", range(int(GOLD["num_cases"])))
def test_device_reproduces_live_reference(c):
    S, B, teachers, disk, g = golden_case(c)
    envs = _mk(B, S, disk)
    k = 0
    for ep, teacher in enumerate(teachers):
        envs.reset(teacher, *WIN_PAIRS[ep])
        for _ in range(int((g["episode"] == ep).sum())):
            out = _as_numpy(envs.step(None, None))
            for key in KEYS:
                assert np.array_equal(out[key], g[key][k]), (c, ep, k, key)
            assert np.array_equal(out["masks"], GOLD["c%d_masks" % c][k])
            assert np.array_equal(out["bad_masks"], GOLD["c%d_bad_masks" % c][k])
            k += 1
    assert k == len(g["episode"])


@pytest.mark.parametrize("S,disk,legacy", [(8, True, True), (8, False, False), (6, True, True), (10, False, True),
                                           (5, True, False), (12, True, True), (15, True, True)])
def test_device_vs_oracle_large_batch(S, disk, legacy):
    B = 320
    envs = _mk(B, S, disk, legacy)
    orc = TeacherStudentOracle(B, S, disk_reward=disk, legacy_stale_state=legacy)
    n_rows = {0: 0, 1: 0, 2: 0}
    for ep, teacher in enumerate((1, -1)):
        wp = WIN_PAIRS[ep]
        envs.reset(teacher, *wp)
        orc.reset(teacher, *wp)
        pol = {teacher: FakePolicy(S, 0), -teacher: FakePolicy(S, STUDENT_SALT)}
        for call in range(3 * S * S):
            want = orc.step(pol)
            got = _as_numpy(envs.step(None, None))
            for key in KEYS:
                assert np.array_equal(got[key], want[key]), (S, ep, call, key)
            assert envs.policy_calls == orc.policy_calls            # same colours served, in the same order
            for t in (0, 1, 2):
                n_rows[t] += int((want["types"] == t).sum())
            if (want["types"] == 2).all():
                break
        assert (want["types"] == 2).all()
    assert min(n_rows.values()) > 0


def test_infos_and_who_iterate_like_the_reference():
    S, B = 6, 16
    envs = _mk(B, S, True)
    envs.reset(-1, *WIN_PAIRS[0])
    who, obs, actions, rewards, dones, infos, outputs, masks, bad = envs.step(None, None)
    assert list(who) == ["teacher"] * B                             # black tuple first, teacher is black
    assert [d["type"] for d in infos] == ["black_transition"] * B
    assert infos[0]["choices"] == torch.nonzero(infos.choices_mask[0]).flatten().tolist()
    who2 = envs.step(None, None)[0]
    assert list(who2) == ["student"] * B and who2[3] == "student" and bool((~who2.is_teacher).all())
    with pytest.raises(ValueError):
        envs.reset(0, *WIN_PAIRS[0])


def test_rule_base_test_uses_the_selected_network():
    """envs.py:142-184: teacher=True plays the teacher's network, else the student's (same rule_base_game)."""
    from oracle.selfplay_oracle import rule_base_games
    S, B, per = 6, 48, 2
    envs = _mk(B, S, True, test_rand_steps=0)
    prot = np.where(np.arange(per * B).reshape(per, B) % 3 == 0, -1, 1)
    for teacher_flag, salt in ((True, 0), (False, STUDENT_SALT)):
        games, wins = envs.test("greedy", per * B, None, teacher=teacher_flag, protagonists=torch.from_numpy(prot))
        want = rule_base_games(FakePolicy(S, salt), "greedy", per, S, prot, disk_reward=True)
        assert (games, wins) == (per * B, int(want.sum()))
