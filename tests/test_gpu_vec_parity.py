"""GPU parity: the CUDA engine (through the C ABI, via VecOthelloEnv) against
 (a) the golden vectors recorded from the live reference, and
 (b) the pinned C oracle on large seeded random action streams,
bit-exact on board, turn, cached legal set, obs, reward, done, winner, make_state planes.
"""
import numpy as np
import pytest
import torch

from oracle import oracle
from tests import golden_io, philox_ref

pytestmark = pytest.mark.gpu


def _env(**kw):
    from gymothelloenv_b200 import VecOthelloEnv
    return VecOthelloEnv(device="cuda:0", **kw)


def _dev_snapshot(env):
    torch.cuda.synchronize()
    return dict(
        board=env.board_state().cpu().numpy(),
        turn=env.player_turn.cpu().numpy().astype(np.int32),
        legal=env.legal_mask.cpu().numpy().astype(np.uint8),
        obs=env.obs.cpu().numpy() if env.obs is not None else None,
        reward=env.reward.cpu().numpy(),
        done=env.done.cpu().numpy().astype(np.uint8),
        winner=env.winner.cpu().numpy().astype(np.int32),
        term=env.terminated().cpu().numpy().astype(np.uint8),
    )


# --------------------------------------------------------------------------- golden replay
EPISODES = golden_io.load_trajectories()
GROUPS = {}
for _e in EPISODES:
    GROUPS.setdefault((_e["S"], _e["sudden"], _e["disk"], _e["mask_in_obs"]), []).append(_e)


@pytest.mark.parametrize("key", sorted(GROUPS))
def test_golden_replay_lockstep(key):
    """All golden episodes of one config stepped in lockstep in ONE VecOthelloEnv;
    finished episodes keep receiving (ignored) actions -- the no-op path."""
    S, sudden, disk, mask_in_obs = key
    eps = GROUPS[key]
    B = len(eps)
    T = max(len(e["act"]) for e in eps)
    env = _env(num_envs=B, board_size=S, sudden_death_on_invalid_move=bool(sudden), num_disk_as_reward=bool(disk),
               obs_layout="board+mask" if mask_in_obs else "board")
    env.reset()
    snap = _dev_snapshot(env)
    for i, e in enumerate(eps):
        assert np.array_equal(snap["board"][i], e["board"][0])
        assert np.array_equal(snap["legal"][i], e["legal"][0])
        assert np.array_equal(snap["obs"][i].reshape(e["obs"][0].shape), e["obs"][0])
        assert snap["turn"][i] == -1 and snap["done"][i] == 0
    if not mask_in_obs:
        st4 = env.observe("planes4", torch.int8).cpu().numpy()
        for i, e in enumerate(eps):
            assert np.array_equal(st4[i], e["state4"][0])
    for t in range(T):
        acts = np.array([int(e["act"][t]) if t < len(e["act"]) else 0 for e in eps], dtype=np.int64)
        env.step(torch.from_numpy(acts).cuda())
        snap = _dev_snapshot(env)
        st4 = None if mask_in_obs else env.observe("planes4", torch.int8).cpu().numpy()
        st4f = None if mask_in_obs else env.observe("planes4", torch.float32).cpu().numpy()
        for i, e in enumerate(eps):
            n = len(e["act"])
            if t < n:
                ctx = "S=%d ep=%d t=%d" % (S, i, t)
                assert np.array_equal(snap["board"][i], e["board"][t + 1]), ctx
                assert snap["turn"][i] == e["turn"][t + 1], ctx
                assert np.array_equal(snap["legal"][i], e["legal"][t + 1]), ctx
                assert np.array_equal(snap["obs"][i].reshape(e["obs"][t + 1].shape), e["obs"][t + 1]), ctx
                assert snap["reward"][i] == e["rew"][t], ctx
                assert snap["done"][i] == e["done"][t], ctx
                if e["done"][t]:
                    assert snap["winner"][i] == e["winner"][t + 1], ctx
                if st4 is not None:
                    assert np.array_equal(st4[i], e["state4"][t + 1]), ctx
                    assert np.array_equal(st4f[i], e["state4"][t + 1].astype(np.float32)), ctx
            else:   # finished: no-op, done stays 1, reward 0, zero obs, board unchanged
                assert snap["done"][i] == 1 and snap["reward"][i] == 0
                assert not snap["obs"][i].any()
                assert np.array_equal(snap["board"][i], e["board"][n])


# --------------------------------------------------------------------------- oracle fuzz
def _sample_actions(rng, legal, p_invalid, N):
    """Uniform legal action per env from a dense mask; with prob p_invalid any integer."""
    B = legal.shape[0]
    score = rng.random(legal.shape) * legal
    a = score.argmax(1).astype(np.int64)
    if p_invalid > 0:
        bad = rng.random(B) < p_invalid
        junk = rng.integers(-3, N + 3, size=B)
        a = np.where(bad, junk, a)
    return a


def _state_legal(env):
    """The CACHED possible_moves set held in the state (not the per-step mask output)."""
    words = env._legal.cpu().numpy().astype(np.uint64)          # [W, B]
    bits = ((words[:, :, None] >> np.arange(64, dtype=np.uint64)[None, None, :]) & np.uint64(1)).astype(np.uint8)
    return np.transpose(bits, (1, 0, 2)).reshape(words.shape[1], -1)[:, :env.num_cells]


def _compare(env, orc, live_before, ctx):
    """live_before: envs that were NOT finished before the call (None = all).  For a
    finished env the step is a no-op whose dense outputs (obs, mask) are zero, while the
    cached list in the state stays what the reference's possible_moves attribute holds."""
    snap = _dev_snapshot(env)
    board, legal, turn, term, winner, n_moves = orc.export()
    assert np.array_equal(snap["board"], board), ctx
    assert np.array_equal(snap["term"], term.astype(np.uint8)), ctx
    assert np.array_equal(snap["turn"], turn), ctx
    assert np.array_equal(_state_legal(env), legal), ctx
    want_mask = legal.copy()
    if live_before is not None:
        want_mask[~live_before] = 0
    assert np.array_equal(snap["legal"], want_mask), ctx
    return snap


FUZZ = [(8, 1, 0, 0.0), (8, 1, 1, 0.02), (8, 0, 0, 0.2), (4, 1, 0, 0.02), (5, 0, 1, 0.1), (6, 1, 1, 0.02),
        (7, 1, 0, 0.02), (9, 0, 0, 0.1), (10, 1, 1, 0.01), (11, 1, 0, 0.02), (12, 1, 1, 0.01), (12, 0, 0, 0.1),
        (13, 1, 1, 0.01), (14, 0, 0, 0.1), (15, 1, 0, 0.02), (16, 1, 1, 0.01), (16, 0, 0, 0.1)]


@pytest.mark.parametrize("S,sudden,disk,p_invalid", FUZZ)
def test_random_streams_vs_oracle(S, sudden, disk, p_invalid):
    B = 2048 + 77   # ragged last warp / block
    N = S * S
    rng = np.random.default_rng(1234 + S)
    env = _env(num_envs=B, board_size=S, sudden_death_on_invalid_move=bool(sudden), num_disk_as_reward=bool(disk),
               obs_layout="board+mask")
    orc = oracle.OracleVec(B, S, sudden_death=sudden, disk_reward=disk)
    env.reset()
    _compare(env, orc, None, "reset")
    steps = 0
    for t in range(N + 30):
        _, legal, _, term, _, _ = orc.export()
        acts = _sample_actions(rng, legal, p_invalid, N)
        env.step(torch.from_numpy(acts).cuda())
        live = term == 0
        reward, done, _ = orc.step(acts)
        snap = _compare(env, orc, live, "S=%d t=%d" % (S, t))
        assert np.array_equal(snap["reward"].astype(np.int32), reward)
        assert np.array_equal(snap["done"], done)
        want_obs = orc.observe(with_mask=True)
        want_obs[~live] = 0            # finished before this call: zero obs (no-op contract)
        assert np.array_equal(snap["obs"], want_obs)
        steps += int(live.sum())
        if not live.any():
            break
    assert steps > B * (N - 10) * (0.5 if p_invalid else 0.9) or sudden


@pytest.mark.parametrize("S", [6, 8, 10, 5, 7, 9, 11, 12, 4, 13, 14, 15, 16])
def test_auto_reset_and_all_layouts(S):
    """auto_reset: the terminal transition reports reward/done/winner, the state restarts
    in the same call; every obs layout x dtype equals the oracle's observation."""
    B, N = 1024 + 5, S * S
    rng = np.random.default_rng(7 + S)
    env = _env(num_envs=B, board_size=S, auto_reset=True, obs_layout="planes4", obs_dtype=torch.float32)
    orc = oracle.OracleVec(B, S)
    env.reset()
    finished = 0
    for t in range(2 * N):
        _, legal, _, _, _, _ = orc.export()
        acts = _sample_actions(rng, legal, 0.0, N)
        env.step(torch.from_numpy(acts).cuda())
        reward, done, _ = orc.step(acts)
        _, _, _, _, owinner, _ = orc.export()
        snap = _dev_snapshot(env)
        assert np.array_equal(snap["reward"].astype(np.int32), reward)
        assert np.array_equal(snap["done"], done)
        assert np.array_equal(snap["winner"][done == 1], owinner[done == 1])
        finished += int(done.sum())
        orc.reset_done()
        board, olegal, turn, term, _, _ = orc.export()
        assert not snap["term"].any()
        assert np.array_equal(snap["board"], board) and np.array_equal(snap["turn"], turn)
        assert np.array_equal(snap["legal"], olegal)
        assert np.array_equal(snap["obs"], orc.make_state().astype(np.float32))
        if t % 7 == 0:
            for layout, want in (("board", orc.observe(False)[:, 0]), ("board+mask", orc.observe(True)),
                                 ("planes4", orc.make_state()), ("planes3", orc.make_state()[:, :3]),
                                 ("absolute", board)):
                for dt in (torch.int8, torch.float32):
                    got = env.observe(layout, dt).cpu().numpy()
                    assert got.dtype == (np.int8 if dt == torch.int8 else np.float32)
                    assert np.array_equal(got, want.astype(got.dtype)), (layout, dt, t)
    assert finished > B


# --------------------------------------------------------------------------- policies
@pytest.mark.parametrize("S,depths", [(8, (1, 2, 3)), (6, (1, 2, 3, 4)), (4, (1, 2, 3, 4)), (5, (2, 3)), (10, (1, 2)),
                                      (12, (1, 2)), (9, (2,)), (7, (3,)), (11, (2,)), (13, (1, 2)), (14, (2,)), (15, (1,)),
                                      (16, (1,))])
def test_policy_choices_vs_oracle(S, depths):
    """greedy / maximin-k pick the same action index as the oracle on positions reached by
    random play (incl. forced passes and near-terminal boards)."""
    B, N = (384 if S <= 12 else 96), S * S     # the oracle's cell-walk maximin is the slow side on large boards
    rng = np.random.default_rng(99 + S)
    env = _env(num_envs=B, board_size=S)
    orc = oracle.OracleVec(B, S)
    env.reset()
    checked = 0
    for t in range(N):
        _, legal, _, term, _, _ = orc.export()
        if t % 5 == 2 or t >= N - 12:
            got_g = env.policy_actions("greedy").cpu().numpy()
            assert np.array_equal(got_g, orc.policy("greedy")), "greedy S=%d t=%d" % (S, t)
            for d in depths:
                if d >= 3 and S >= 8 and t % 10 != 2:
                    continue
                sub = slice(0, 96 if d >= 3 else B)
                got = env.policy_actions("maximin", d).cpu().numpy()
                want = orc.policy("maximin", d) if d < 3 else None
                if want is None:
                    o2 = oracle.OracleVec(96, S)
                    board, _, turn, oterm, _, _ = orc.export()
                    live = oterm[sub] == 0
                    o2.inject(board[sub], np.where(turn[sub] == 0, 1, turn[sub]))
                    want = o2.policy("maximin", d)
                    assert np.array_equal(got[sub][live], want[live]), "maximin-%d S=%d t=%d" % (d, S, t)
                else:
                    assert np.array_equal(got, want), "maximin-%d S=%d t=%d" % (d, S, t)
                checked += 1
        acts = _sample_actions(rng, legal, 0.0, N)
        env.step(torch.from_numpy(acts).cuda())
        orc.step(acts)
        if (term == 1).all():
            break
    assert checked > 0


def test_golden_policy_choices():
    """Choices recorded from the reference's GreedyPolicy / MaxiMinPolicy(1..4)."""
    for S, g in golden_io.load_policies().items():
        P = len(g["turn"])
        env = _env(num_envs=P, board_size=S)
        env.reset()
        env.set_board_state(torch.from_numpy(g["board"]), perspective=1, turn=torch.from_numpy(g["turn"]))
        assert np.array_equal(env.policy_actions("greedy").cpu().numpy(), g["choice"][:, 0])
        for d in (1, 2, 3, 4):
            want = g["choice"][:, d]
            sel = want >= 0
            if sel.any():
                got = env.policy_actions("maximin", d).cpu().numpy()
                assert np.array_equal(got[sel], want[sel]), "S=%d maximin-%d" % (S, d)


# --------------------------------------------------------------------------- in-kernel random policy
@pytest.mark.parametrize("S", [8, 6, 10])
def test_fused_random_policy_matches_philox_and_oracle(S):
    """step(policy='random'): every applied action is (a) the one an independent numpy
    Philox4x32-10 predicts from (seed, global env id, draw counter) and the oracle's legal
    list, and (b) legal; the resulting states match the oracle."""
    B, N, seed, base = 1500, S * S, 0xABCDEF0123, 1 << 33
    env = _env(num_envs=B, board_size=S, seed=seed, env_id_base=base, auto_reset=True)
    orc = oracle.OracleVec(B, S)
    env.reset()
    ctr = np.zeros(B, dtype=np.int64)
    gid = base + np.arange(B, dtype=np.int64)
    for t in range(N + 20):
        _, legal, _, _, _, n_moves = orc.export()
        r = philox_ref.draw(seed, gid, 0, ctr)
        idx = philox_ref.bounded(r, n_moves)
        order = np.cumsum(legal, axis=1) - 1              # rank of each legal cell
        want = ((order == idx[:, None]) & (legal == 1)).argmax(1)
        env.step(policy="random")
        got = env.actions_taken.cpu().numpy()
        assert np.array_equal(got, want), "t=%d" % t
        ctr += 1
        reward, done, valid = orc.step(got)
        assert valid.all()
        orc.reset_done()
        board, olegal, turn, _, _, _ = orc.export()
        assert np.array_equal(env.board_state().cpu().numpy(), board)
        assert np.array_equal(env.legal_mask.cpu().numpy().astype(np.uint8), olegal)
        assert np.array_equal(env.done.cpu().numpy().astype(np.uint8), done)


def test_random_opening_budget_and_override():
    """initial_rand_steps: max_rand_steps = 2*U{0..init/2} per reset (othello.py:62-64);
    while rand_step_cnt < max the submitted action is replaced by a uniform legal move."""
    B, S, seed = 4096, 8, 42
    env = _env(num_envs=B, board_size=S, seed=seed, initial_rand_steps=10)
    env.reset()
    cnt, mx = [x.cpu().numpy() for x in env.rand_steps()]
    assert (cnt == 0).all() and set(np.unique(mx)) == {0, 2, 4, 6, 8, 10}
    want_mx = 2 * philox_ref.bounded(philox_ref.draw(seed, np.arange(B), 0, np.zeros(B, dtype=np.int64)), 6)
    assert np.array_equal(mx, want_mx)
    freq = np.bincount(mx // 2, minlength=6) / B
    assert np.abs(freq - 1 / 6).max() < 0.03
    orc = oracle.OracleVec(B, S)
    for t in range(12):
        alive = ~env.terminated().cpu().numpy()
        bogus = torch.full((B,), -1, dtype=torch.int64, device="cuda")  # never a legal action
        env.step(bogus)
        got = env.actions_taken.cpu().numpy()
        overridden = alive & (t < mx)
        spent = alive & ~(t < mx)
        assert (got[spent] == -1).all() and (got[~alive] == -1).all() and (got[overridden] >= 0).all()
        _, _, valid = orc.step(got)
        assert valid[overridden].all()
        done = env.done.cpu().numpy()
        assert done[spent].all()                # bogus move => sudden death once the budget is spent
        assert np.array_equal(env.board_state().cpu().numpy(), orc.export()[0])
    cnt2, _ = [x.cpu().numpy() for x in env.rand_steps()]
    assert np.array_equal(cnt2, np.minimum(mx, 12))


def test_get_set_state_roundtrip():
    env = _env(num_envs=300, board_size=8, seed=3)
    env.reset()
    for _ in range(20):
        env.step(policy="random")
    saved = env.get_state()
    a = [env.step(policy="random")[0].clone() for _ in range(10)]
    env.set_state(saved)
    b = [env.step(policy="random")[0].clone() for _ in range(10)]
    assert all(torch.equal(x, y) for x, y in zip(a, b))


def test_host_buffer_step_matches_device_step():
    B = 1000
    e1 = _env(num_envs=B, board_size=8)
    e2 = _env(num_envs=B, board_size=8)
    e1.reset(); e2.reset()
    acts = e1.policy_actions("greedy")
    e1.step(acts)
    obs_h, rew_h, done_h, mask_h = e2.step_host(acts.cpu().pin_memory())
    assert torch.equal(obs_h, e1.obs.cpu()) and torch.equal(mask_h, e1._mask_u8.cpu())
    assert torch.equal(rew_h, e1.reward.cpu()) and torch.equal(done_h, e1._done_u8.cpu())
    h2d, d2h = e2.host_bytes_per_step()
    assert h2d == 8 * B and d2h == B * (64 + 64 + 4 + 1)
    # narrow host actions (int16 / int8 / int32) cross PCIe at their own width and play the same ply
    for dt, width in ((torch.int16, 2), (torch.int8, 1), (torch.int32, 4)):
        acts = e1.policy_actions("greedy")
        acts[::9] = -1                                         # an invalid action survives the narrowing
        e1.step(acts)
        obs_h, rew_h, done_h, mask_h = e2.step_host(acts.to(dt).cpu().pin_memory())
        assert torch.equal(obs_h, e1.obs.cpu()) and torch.equal(mask_h, e1._mask_u8.cpu())
        assert torch.equal(rew_h, e1.reward.cpu()) and torch.equal(done_h, e1._done_u8.cpu())
        assert e2.host_bytes_per_step()[0] == width * B
    with pytest.raises(TypeError):
        e2.step_host(torch.zeros(B, dtype=torch.uint8).pin_memory())


def test_perft_through_the_engine():
    """perft via set_board_state + step on a growing batch."""
    want = golden_io.load_perft()
    for S, depth in ((8, 5), (6, 5), (4, 6)):
        env = _env(num_envs=1, board_size=S)
        env.reset()
        boards = env.board_state().cpu().numpy()
        turns = env.player_turn.cpu().numpy()
        for d in range(1, depth + 1):
            e = _env(num_envs=len(boards), board_size=S)
            e.reset()
            e.set_board_state(torch.from_numpy(boards), perspective=1, turn=torch.from_numpy(turns))
            legal = e.legal_mask.cpu().numpy()
            term = e.terminated().cpu().numpy()
            # terminal positions reached earlier count as leaves at every deeper level
            src, act = np.nonzero(legal & ~term[:, None])
            n_leaf_terminal = int(((legal.sum(1) == 0) | term).sum())
            e2 = _env(num_envs=len(src), board_size=S)
            e2.reset()
            e2.set_board_state(torch.from_numpy(boards[src]), perspective=1, turn=torch.from_numpy(turns[src]))
            e2.step(torch.from_numpy(act.astype(np.int64)).cuda())
            nb = e2.board_state().cpu().numpy()
            nt = e2.player_turn.cpu().numpy()
            assert len(src) + n_leaf_terminal == want[S][d - 1], (S, d)
            if n_leaf_terminal:
                keep_b = boards[(legal.sum(1) == 0) | term]
                keep_t = turns[(legal.sum(1) == 0) | term]
                nb = np.concatenate([nb, keep_b]); nt = np.concatenate([nt, keep_t])
            boards, turns = nb, nt


def test_captured_step_graph_replays_the_same_plies():
    """VecOthelloEnv.capture_steps: replaying a CUDA graph of k fused policy+step launches plays exactly the plies
    k eager calls play (the Philox position lives in the env state, not in launch arguments)."""
    kw = dict(num_envs=3000, board_size=8, seed=9, auto_reset=True, initial_rand_steps=4)
    eager, graphed = _env(**kw), _env(**kw)
    eager.reset(); graphed.reset()
    g = graphed.capture_steps("random", steps=4)
    for _ in range(20):
        eager.step(policy="random")
    for _ in range(5):
        g.replay()
    torch.cuda.synchronize()
    for name in ("_black", "_white", "_legal", "_meta", "_ctr"):
        assert torch.equal(getattr(eager, name), getattr(graphed, name)), name
    assert torch.equal(eager.obs, graphed.obs) and torch.equal(eager._mask_u8, graphed._mask_u8)
    with pytest.raises(ValueError):
        graphed.capture_steps("maximin")
