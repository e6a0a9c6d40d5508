"""GPU tests for the DQN feed: the epsilon-greedy action choice over the legal
moves (reference dqn.py:264-286,478-488) and the 3-plane state (run_2agent.py:29-46).

The reference draws from Python's `random`; the device uses Philox, so the explore/exploit STREAM cannot
match -- what is checked is every deterministic property (eps = 0: first argmax over the ascending legal
list incl. ties; eps = 1: an entry of the legal list), the choice an independent numpy Philox predicts, and
the distribution."""
import numpy as np
import pytest
import torch

from tests import philox_ref
from tests.test_gpu_selfplay import _random_case

pytestmark = pytest.mark.gpu
EPS_STREAM = 4


def _legal_lists(legal):
    return [np.nonzero(row)[0] for row in legal]


def _reference_exploit(q, legal):
    """dqn.py:271-273: possible_moves[argmax(q[0][possible_moves])] -- first occurrence on ties"""
    out = np.zeros(len(q), dtype=np.int64)
    for i, pm in enumerate(_legal_lists(legal)):
        if len(pm):
            out[i] = pm[int(torch.argmax(torch.from_numpy(q[i][pm])))]
    return out


@pytest.mark.parametrize("S", [8, 6, 10, 12, 5, 13, 16])
def test_epsilon_greedy_exploit_explore_and_philox_prediction(S):
    from gymothelloenv_b200 import epsilon_greedy
    B, seed, base, counter = 6000, 77, 1 << 33, 5
    q, legal, words = _random_case(B, S, 100 + S)
    q = np.round(q)                                    # many exact ties: first occurrence must win
    qd, wd = torch.from_numpy(q).cuda(), torch.from_numpy(words).cuda()
    has = legal.any(1)
    n = legal.sum(1)
    gid = base + np.arange(B, dtype=np.int64)
    r0 = philox_ref.draw(seed, gid, EPS_STREAM, np.full(B, 2 * counter))
    r1 = philox_ref.draw(seed, gid, EPS_STREAM, np.full(B, 2 * counter + 1))
    u = (r0.astype(np.int64) >> 8).astype(np.float32) * np.float32(1.0 / 16777216.0)
    want_exploit = _reference_exploit(q, legal)
    lists = _legal_lists(legal)
    want_explore = np.array([lists[i][philox_ref.bounded(r1[i], n[i])] if n[i] else 0 for i in range(B)], dtype=np.int64)
    for eps in (0.0, 1.0, 0.3):
        a, qc, ex = epsilon_greedy(qd, wd, S, eps, seed=seed, counter=counter, env_id_base=base)
        a, qc, ex = a.cpu().numpy(), qc.cpu().numpy(), ex.cpu().numpy()
        explore = (np.float32(eps) >= u) & has         # dqn.py:266
        assert np.array_equal(ex, explore)
        want = np.where(explore, want_explore, want_exploit)
        assert np.array_equal(a, want)
        assert (a[~has] == 0).all() and (qc[~has] == 0).all()
        assert legal[np.arange(B)[has], a[has]].all()
        assert np.array_equal(qc[has], q[np.arange(B)[has], a[has]])
    # eps = 0 explores only when u == 0 exactly (the reference's `epsilon >= random.random()` has the same edge)
    assert (u > 0).all()


def test_epsilon_greedy_distribution_and_counters():
    from gymothelloenv_b200 import epsilon_greedy
    S, B, eps = 8, 400000, 0.25
    rng = np.random.RandomState(1)
    row = rng.randn(64).astype(np.float32)
    leg = np.zeros(64, dtype=bool)
    leg[[2, 9, 33, 34, 60, 63]] = True
    best = np.nonzero(leg)[0][np.argmax(row[leg])]
    word = np.uint64(sum(1 << int(i) for i in np.nonzero(leg)[0]))
    qd = torch.from_numpy(np.tile(row, (B, 1))).cuda()
    wd = torch.from_numpy(np.full((1, B), word, dtype=np.uint64).view(np.int64)).cuda()
    a0, _, e0 = epsilon_greedy(qd, wd, S, eps, seed=9, counter=0)
    a1, _, _ = epsilon_greedy(qd, wd, S, eps, seed=9, counter=1)
    a0b, _, _ = epsilon_greedy(qd, wd, S, eps, seed=9, counter=0)
    assert torch.equal(a0, a0b) and not torch.equal(a0, a1)
    a0, e0 = a0.cpu().numpy(), e0.cpu().numpy()
    tol = 5 * np.sqrt(0.25 / B) + 1e-3
    assert abs(e0.mean() - eps) < tol
    counts = np.bincount(a0, minlength=64)
    assert counts[~leg].sum() == 0
    want = np.full(int(leg.sum()), eps / leg.sum())
    want[list(np.nonzero(leg)[0]).index(best)] += 1 - eps
    assert np.abs(counts[leg] / B - want).max() < tol
    assert (a0[~e0] == best).all()


def test_epsilon_schedule_follows_the_reference_agent():
    """dqn.py:196-201,266,283-284: all-random until initial_replay_size, then linear annealing to final_epsilon."""
    from gymothelloenv_b200 import VecOthelloEnv, EpsilonGreedy
    env = VecOthelloEnv(2048, board_size=8, device="cuda:0", seed=3)
    env.reset()
    agent = EpsilonGreedy(env, anealing_steps=10, initial_epsilon=1.0, final_epsilon=0.1, initial_replay_size=3)
    q = torch.randn(2048, 64, device="cuda")
    eps_ref, t_ref, trace = 1.0, 0, []
    for step in range(20):
        a, qc, ex = agent.get_action(q)
        if t_ref < 3:
            assert bool(ex.all())                      # self.t < initial_replay_size: always random
        if eps_ref > 0.1 and t_ref >= 3:
            eps_ref -= (1.0 - 0.1) / 10
        trace.append(agent.epsilon)
        assert abs(agent.epsilon - eps_ref) < 1e-12
        mask = env.legal_mask
        assert bool(mask[torch.arange(2048, device="cuda"), a].all())
        agent.tick()
        t_ref += 1
    # Ground truth = the reference's own update rule in floating point.  Its `epsilon > final_epsilon` guard
    # sees 0.10000000000000026 after ten decrements and subtracts once more, so with these parameters the
    # reference settles at 0.01, BELOW final_epsilon; the mirror must do the same, not clamp.
    assert trace[0] == 1.0 and trace[-1] == eps_ref
    assert abs(eps_ref - 0.01) < 1e-9 and eps_ref < agent.final_epsilon
    frac = float(ex.float().mean())                    # the last call ran with that settled epsilon
    assert abs(frac - eps_ref) < 5 * np.sqrt(eps_ref * (1 - eps_ref) / 2048) + 1e-3
    a, _, ex = agent.get_action_at_test(q)
    assert 0.02 < float(ex.float().mean()) < 0.09      # dqn.py:480: 0.05


@pytest.mark.parametrize("S", [8, 6, 10])
def test_planes3_state_is_run_2agent_make_state(S):
    """run_2agent.py:29-46 on the mover-perspective obs: [black in {0,1}, white in {0,1}, turn (0 black, 1 white)]."""
    from gymothelloenv_b200 import VecOthelloEnv
    B = 512
    env3 = VecOthelloEnv(B, board_size=S, device="cuda:0", seed=5, obs_layout="planes3", obs_dtype=torch.float32)
    envb = VecOthelloEnv(B, board_size=S, device="cuda:0", seed=5, obs_layout="board", obs_dtype=torch.int8)
    o3, ob = env3.reset(), envb.reset()
    was_terminated = np.zeros(B, dtype=bool)           # envs finished BEFORE the last step report zeros
    compared_terminal = 0
    for _ in range(S * S):
        obs = ob.cpu().numpy().astype(np.int64)
        turn = envb.player_turn.cpu().numpy()
        want = np.zeros((B, 3, S, S), dtype=np.float32)
        blk = np.where(turn[:, None, None] == -1, obs == 1, obs == -1)
        wht = np.where(turn[:, None, None] == -1, obs == -1, obs == 1)
        want[:, 0], want[:, 1] = blk, wht
        want[:, 2] = (turn == 1)[:, None, None]
        live = ~was_terminated
        got = o3.cpu().numpy()
        assert np.array_equal(got[live], want[live])
        assert (got[~live] == 0).all()
        now_terminated = envb.terminated().cpu().numpy().astype(bool)
        compared_terminal += int((now_terminated & live).sum())
        was_terminated = now_terminated
        o3 = env3.step(policy="random")[0]
        ob = envb.step(policy="random")[0]
    assert compared_terminal == B                      # every game's terminal board was compared once
