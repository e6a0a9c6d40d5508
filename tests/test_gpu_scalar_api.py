"""GPU parity of the drop-in scalar API (gymothelloenv_b200.othello / simple_policies /
util) against games recorded from the live reference: same constructor arguments, same
seeds => same numpy RandomState streams => identical action streams, rewards, boards."""
import numpy as np
import pytest

from tests import golden_io

pytestmark = pytest.mark.gpu

WRAP = golden_io.load_wrappers()
EPISODES = golden_io.load_trajectories()


def _make_policy(kind, seed):
    from gymothelloenv_b200 import simple_policies
    if kind == "rand":
        return simple_policies.RandomPolicy(seed=seed)
    if kind == "greedy":
        return simple_policies.GreedyPolicy()
    return simple_policies.MaxiMinPolicy(int(kind[7:]))


@pytest.mark.parametrize("idx", range(len(WRAP["two_agent"])))
def test_two_agent_games_reproduce_reference(idx):
    """run_2agent.py:92-203 protocol (fixed colours) through SimpleOthelloEnv + policies."""
    from gymothelloenv_b200 import othello, util
    cfg = WRAP["two_agent"][idx]
    S = cfg["S"]
    env = othello.SimpleOthelloEnv(board_size=S, seed=cfg["seed"], initial_rand_steps=cfg["init_rand"],
                                   num_disk_as_reward=bool(cfg["disk_reward"]))
    pol = {-1: _make_policy(cfg["black"], cfg["seed"]), 1: _make_policy(cfg["white"], cfg["seed"])}
    for g, want in enumerate(cfg["games"]):
        obs = env.reset()
        for p in pol.values():
            p.reset(env)
        assert env.max_rand_steps == want["max_rand_steps"]
        done, t = False, 0
        while not done:
            turn = env.player_turn
            state = util.make_state(obs, env)
            a = int(pol[turn].get_action(state))
            obs, reward, done, info = env.step(a)
            ctx = "cfg %d game %d ply %d" % (idx, g, t)
            assert info is None
            assert a == want["requested"][t], ctx
            assert turn == want["turn"][t], ctx
            assert reward == want["reward"][t] and done == bool(want["done"][t]), ctx
            t += 1
        assert t == len(want["requested"])
        assert env.env.board_state.ravel().tolist() == want["final_board"]
        w, b = env.env.count_disks()
        assert (int(w), int(b)) == (want["white"], want["black"]) and env.env.winner == want["winner"]


@pytest.mark.parametrize("idx", range(len(WRAP["othello_env"])))
def test_othello_env_with_embedded_opponent(idx):
    """OthelloEnv (othello.py:96-214): opponent auto-plays inside reset/step, reward sign flip."""
    from gymothelloenv_b200 import othello, simple_policies
    cfg = WRAP["othello_env"][idx]
    opp = simple_policies.RandomPolicy(seed=cfg["seed"] + 100)
    kw = {"white_policy": opp} if cfg["protagonist"] == -1 else {"black_policy": opp}
    env = othello.OthelloEnv(protagonist=cfg["protagonist"], board_size=cfg["S"],
                             initial_rand_steps=cfg["init_rand"], seed=cfg["seed"], **kw)
    prot_rnd = np.random.RandomState(cfg["seed"] + 7)
    for want in cfg["games"]:
        obs = env.reset()
        assert env.max_rand_steps == want["max_rand_steps"]
        assert np.asarray(obs).ravel().tolist() == want["reset_obs"]
        done, t = False, 0
        while not done:
            a = int(env.possible_moves[prot_rnd.randint(0, len(env.possible_moves))])
            obs, reward, done, _ = env.step(a)
            assert a == want["requested"][t]
            assert reward == want["reward"][t] and done == bool(want["done"][t])
            assert int(np.sum(obs)) == want["obs_sum"][t]
            assert list(env.possible_moves) == want["moves_after"][t]
            t += 1
        assert np.asarray(obs).ravel().tolist() == want["final_obs"]
    with pytest.raises(AssertionError):
        env.env.set_player_turn(-env.protagonist)
        env.env.terminated = False
        env.step(0)                                   # othello.py:177: not the protagonist's turn


@pytest.mark.parametrize("idx", [i for i, e in enumerate(EPISODES) if e["S"] in (5, 8, 10) and e["seed"] in (0, 10, 20, 21)])
def test_base_env_replays_golden(idx):
    from gymothelloenv_b200 import othello, util
    e = EPISODES[idx]
    env = othello.OthelloBaseEnv(board_size=e["S"], sudden_death_on_invalid_move=bool(e["sudden"]),
                                 num_disk_as_reward=bool(e["disk"]),
                                 possible_actions_in_obs=bool(e["mask_in_obs"]), mute=True)
    assert env.possible_moves == []                   # before the first reset (othello.py:242)
    obs = env.reset()
    assert obs.dtype == np.int64 and np.array_equal(obs.reshape(e["obs"][0].shape), e["obs"][0])
    assert env.action_space.n == e["S"] ** 2 and env.observation_space.shape == obs.shape
    for t, a in enumerate(e["act"]):
        obs, reward, done, info = env.step(int(a))
        assert info is None and isinstance(reward, int) and isinstance(done, bool)
        assert reward == e["rew"][t] and done == bool(e["done"][t])
        assert np.array_equal(obs.reshape(e["obs"][t + 1].shape), e["obs"][t + 1])
        assert np.array_equal(env.board_state, e["board"][t + 1]) and env.player_turn == e["turn"][t + 1]
        assert env.possible_moves == np.nonzero(e["legal"][t + 1])[0].tolist()
        assert env.winner == e["winner"][t + 1] and env.terminated == bool(e["done"][t])
        assert tuple(int(x) for x in env.count_disks()) == tuple(e["cnt"][t + 1])
        if "state4" in e and t % 3 == 0:
            st = util.make_state(obs, env)
            assert st.dtype == np.float64 and np.array_equal(st, e["state4"][t + 1])
            if not done:
                assert np.array_equal(util.undo_state(st, env.player_turn), obs)
    with pytest.raises(ValueError, match="Game has terminated!"):
        env.step(0)


def test_state_injection_and_helpers():
    """set_board_state / set_player_turn / get_possible_actions / update_board / copy_env,
    the calls the reference's planners make (simple_policies.py:12-18,78-84,135-138)."""
    from gymothelloenv_b200 import othello, simple_policies
    e = next(x for x in EPISODES if x["S"] == 8 and x["seed"] == 1 and not x["mask_in_obs"] and x["p_invalid"] == 0)
    env = othello.OthelloBaseEnv(board_size=8, mute=True)
    env.reset()
    clone = simple_policies.copy_env(env)
    for t in (5, 20, 41):
        board, turn = e["board"][t], int(e["turn"][t])
        legal = np.nonzero(e["legal"][t])[0].tolist()
        clone.reset()
        before = list(clone.possible_moves)
        clone.set_board_state(board * turn, perspective=turn)      # as simple_policies.py:79-80
        assert np.array_equal(clone.board_state, board)
        assert clone.possible_moves == before                       # untouched until set_player_turn
        clone.set_player_turn(turn)
        assert clone.possible_moves == legal and clone.player_turn == turn
        assert clone.get_possible_actions() == legal
        assert clone.get_possible_actions(board * turn) == legal    # explicit mover-perspective board
        a = int(e["act"][t])
        clone.update_board(a)                                       # flips only, turn unchanged
        assert np.array_equal(clone.board_state, e["board"][t + 1]) and clone.player_turn == turn
        # a 3-d input takes plane 0 (othello.py:382-383)
        clone.set_board_state(np.stack([board, np.zeros_like(board)]), perspective=1)
        assert np.array_equal(clone.board_state, board)
    assert clone.determine_winner(sudden_death=True) == -clone.player_turn and clone.terminated
    with pytest.raises(ValueError):
        clone.step(0)


def test_board_size_is_clamped_and_bounds_checked():
    from gymothelloenv_b200 import othello, OthelloB200Error
    env = othello.OthelloBaseEnv(board_size=2, mute=True)
    assert env.board_size == 4 and env.reset().shape == (4, 4)       # othello.py:230
    with pytest.raises((ValueError, OthelloB200Error)):
        othello.OthelloBaseEnv(board_size=40)


def test_greedy_equals_maximin1_on_scalar_api():
    """README.md:54: Greedy and Maximin-1 are the same policy."""
    from gymothelloenv_b200 import othello, simple_policies, util
    env = othello.SimpleOthelloEnv(board_size=8, seed=3, initial_rand_steps=0)
    g, m = simple_policies.GreedyPolicy(), simple_policies.MaxiMinPolicy(1)
    rnd = np.random.RandomState(0)
    obs = env.reset()
    g.reset(env); m.reset(env)
    done = False
    while not done:
        assert g.get_action(util.make_state(obs, env)) == m.get_action(None)
        obs, _, done, _ = env.step(env.possible_moves[rnd.randint(0, len(env.possible_moves))])
