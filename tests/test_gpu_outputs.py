"""Output plumbing of oth_step / oth_observe: the packed 16-byte scalar record against the separate arrays, the
compile-time fast kernel against the general one, finished envs, seed(), and the packed host path."""
import ctypes as C

import numpy as np
import pytest
import torch

from gymothelloenv_b200 import VecOthelloEnv, _cabi

pytestmark = pytest.mark.gpu


def _raw_step(env, out, policy_src=_cabi.ACT_RANDOM, actions=None):
    lib = _cabi.load()
    _cabi.check(lib.oth_step(env._state(), env.num_envs, env.board_size, env._flags(), policy_src, _cabi.ptr(actions),
                             env.seed_value, env.env_id_base, env.initial_rand_steps, out,
                             C.c_void_p(torch.cuda.current_stream(env.device).cuda_stream)), "oth_step")


@pytest.mark.parametrize("S,B", [(8, 1024), (8, 1000), (6, 512), (10, 384), (7, 256), (12, 256), (4, 128), (14, 128), (16, 128)])
def test_packed_record_equals_separate_arrays_and_fast_equals_general(S, B):
    """Two envs on the same seed: A steps through VecOthelloEnv (packed record only -> the FAST kernel when B is a
    multiple of 128 and S is even), B through the raw C ABI with EVERY separate array plus the packed record (the
    general kernel).  States, dense outputs and all scalars must agree every ply, and the record's fields must equal
    the separate arrays."""
    dev = torch.device("cuda", 0)
    a = VecOthelloEnv(B, board_size=S, device=dev, seed=11, initial_rand_steps=4, auto_reset=True)
    b = VecOthelloEnv(B, board_size=S, device=dev, seed=11, initial_rand_steps=4, auto_reset=True)
    a.reset()
    b.reset()
    N = S * S
    obs = torch.zeros((B, S, S), dtype=torch.int8, device=dev)
    mask = torch.zeros((B, N), dtype=torch.uint8, device=dev)
    reward = torch.zeros(B, dtype=torch.float32, device=dev)
    done = torch.zeros(B, dtype=torch.uint8, device=dev)
    turn = torch.zeros(B, dtype=torch.int8, device=dev)
    winner = torch.zeros(B, dtype=torch.int8, device=dev)
    act = torch.zeros(B, dtype=torch.int64, device=dev)
    packed = torch.zeros((B, 16), dtype=torch.uint8, device=dev)
    out = _cabi.OthOutputs(_cabi.ptr(obs), _cabi.ptr(mask), _cabi.ptr(reward), _cabi.ptr(done), _cabi.ptr(turn),
                           _cabi.ptr(winner), _cabi.ptr(act), _cabi.OBS_BOARD, _cabi.DTYPE_I8, _cabi.ptr(packed))
    finished = 0
    for t in range(N + 10):
        a.step(policy="random")
        _raw_step(b, out)
        torch.cuda.synchronize()
        for name in ("_black", "_white", "_legal", "_meta", "_ctr"):
            assert torch.equal(getattr(a, name), getattr(b, name)), (name, t)
        assert torch.equal(a.obs, obs) and torch.equal(a._mask_u8, mask), t
        assert torch.equal(a.reward, reward) and torch.equal(a._done_u8, done), t
        assert torch.equal(a.player_turn, turn) and torch.equal(a.winner, winner) and torch.equal(a.actions_taken, act), t
        assert torch.equal(packed.view(torch.float32)[:, 0], reward)
        assert torch.equal(packed[:, 4], done) and torch.equal(packed.view(torch.int8)[:, 5], turn)
        assert torch.equal(packed.view(torch.int8)[:, 6], winner) and torch.equal(packed.view(torch.int64)[:, 1], act)
        assert int(packed[:, 7].max()) == 0
        finished += int(done.sum())
    assert finished >= B                      # every env ended at least one game (auto-reset keeps them going)


def test_winner_stays_readable_on_finished_envs():
    """without auto_reset, stepping a finished env is a no-op that must keep reporting its winner."""
    env = VecOthelloEnv(512, board_size=6, seed=3)
    env.reset()
    first = {}
    for t in range(60):
        env.step(policy="random")
        d, w = env.done.cpu().numpy(), env.winner.cpu().numpy().copy()
        for i in np.flatnonzero(d):
            if i in first:
                assert w[i] == first[i], (t, i)
            else:
                first[i] = w[i]
        if d.all():
            break
    assert env.done.all() and len(first) == 512
    meta_w = ((env._meta >> _cabi.META_WINNER_SHIFT) & 3).cpu().numpy()
    expect = np.where(meta_w == 1, 1, np.where(meta_w == 2, -1, 0))
    for _ in range(3):
        env.step(policy="random")
        assert np.array_equal(env.winner.cpu().numpy(), expect)
        assert float(env.reward.abs().sum()) == 0.0 and bool(env.done.all())
        assert int(env.obs.abs().sum()) == 0
    env.refresh_outputs()                     # oth_observe writes the same winner through the same record
    assert np.array_equal(env.winner.cpu().numpy(), expect)
    assert set(np.unique(expect)) <= {-1, 0, 1} and (expect != 0).any()


def test_seed_then_reset_reproduces_a_fresh_env():
    def play(env, n=40):
        out = []
        for _ in range(n):
            env.step(policy="random")
            out.append(env.actions_taken.clone())
        return torch.stack(out)
    fresh = VecOthelloEnv(256, board_size=8, seed=77, initial_rand_steps=6, auto_reset=True)
    fresh.reset()
    ref = play(fresh)
    env = VecOthelloEnv(256, board_size=8, seed=5, initial_rand_steps=6, auto_reset=True)
    env.reset()
    play(env, 17)                             # mid-game, counters advanced
    env.seed(77)
    env.reset()
    assert torch.equal(play(env), ref)


@pytest.mark.parametrize("S", [8, 10])
def test_step_host_packed_expands_to_the_dense_outputs(S):
    B = 640
    dense = VecOthelloEnv(B, board_size=S, seed=2, auto_reset=True)
    pk = VecOthelloEnv(B, board_size=S, seed=2, auto_reset=True)
    dense.reset()
    pk.reset()
    acts = torch.empty((B,), dtype=torch.int16, pin_memory=True)
    for t in range(S * S + 5):
        dense.step(policy="random")                       # defines the action stream
        acts.copy_(dense.actions_taken.to(torch.int16))
        torch.cuda.synchronize()
        got = pk.step_host(acts, packed=True)
        obs, mask = pk.expand_packed_host(got)
        assert np.array_equal(obs, dense.obs.cpu().numpy()), t
        assert np.array_equal(mask, dense.legal_mask.cpu().numpy()), t
        assert np.array_equal(got["reward"].numpy(), dense.reward.cpu().numpy())
        assert np.array_equal(got["done"].numpy(), dense._done_u8.cpu().numpy())
        sub = [3, 77, 639]
        o2, m2 = pk.expand_packed_host(got, sub)
        assert np.array_equal(o2, obs[sub]) and np.array_equal(m2, mask[sub])
    h2d, d2h = pk.host_bytes_per_step()
    assert h2d == 2 * B and d2h == B * (24 * pk.words + 6)


@pytest.mark.parametrize("S,B", [(8, 256), (8, 200), (6, 128), (10, 128), (7, 96)])
def test_held_envs_sit_a_ply_out(S, B):
    """actions[i] == OTH_ACTION_HOLD: env i keeps its state and re-emits its outputs (reward 0, done 0) while the
    others move.  Per-env replay: env i of the held run after its k-th real move equals env i of a run in which every
    env moves on every ply, after k plies (same start, same deterministic action rule)."""
    free = VecOthelloEnv(B, board_size=S, seed=4, initial_rand_steps=0)
    held = VecOthelloEnv(B, board_size=S, seed=4, initial_rand_steps=0)
    free.reset()
    held.reset()
    keys = ("black", "white", "legal", "meta")

    def snap(env):                                            # everything a step produces, (B, ...) per field
        st = env.get_state()
        d = {k: (st[k].t() if st[k].dim() == 2 else st[k]).cpu().numpy().copy() for k in keys}
        d.update(obs=env.obs.cpu().numpy().copy(), mask=env.legal_mask.cpu().numpy().copy(),
                 done=env.done.cpu().numpy().copy(), reward=env.reward.cpu().numpy().copy(),
                 turn=env.player_turn.cpu().numpy().copy())
        return d
    hist_a, hist = [], []
    for t in range(S * S):
        a = free.policy_actions("greedy")                     # a deterministic function of the position
        free.step(a)
        hist_a.append(a.cpu().numpy().copy())
        hist.append(snap(free))
    hist_a = np.stack(hist_a)
    hist = {k: np.stack([h[k] for h in hist]) for k in hist[0]}      # (T, B, ...)
    rng = np.random.RandomState(S)
    hold = torch.full((B,), _cabi.ACTION_HOLD, dtype=torch.int64, device="cuda")
    moves = np.zeros(B, dtype=np.int64)                       # real moves each env of the held run has made
    prev = snap(held)
    for t in range(4 * S * S):
        sit = rng.rand(B) < 0.5
        term0 = held.terminated().cpu().numpy().astype(bool)
        a = held.policy_actions("greedy")
        held.step(torch.where(torch.from_numpy(sit).cuda(), hold, a))
        cur = snap(held)
        s_idx, m_idx = np.flatnonzero(sit & ~term0), np.flatnonzero(~sit & ~term0)
        for k in keys + ("obs", "mask", "turn"):              # held: nothing moved, outputs re-emitted
            assert np.array_equal(cur[k][s_idx], prev[k][s_idx]), (k, t)
        assert not cur["done"][s_idx].any() and not cur["reward"][s_idx].any()
        assert np.array_equal(a.cpu().numpy()[m_idx], hist_a[moves[m_idx], m_idx])
        for k in hist:                                        # moved: exactly the free run's next ply of that env
            assert np.array_equal(cur[k][m_idx], hist[k][moves[m_idx], m_idx]), (k, t)
        moves[m_idx] += 1
        prev = cur
        if cur["done"].all() or held.terminated().all():
            break
    assert bool(held.terminated().all()) and moves.min() > 4


def test_the_stub_in_integration_md_runs_as_written():
    """INTEGRATION.md section B is the binding a maintainer of the reference would copy: execute it verbatim."""
    import os
    import re
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    text = open(os.path.join(root, "INTEGRATION.md")).read()
    section = text[text.index("## B. The ctypes stub"):text.index("## C. Batched use")]
    code = re.search(r"```python\n(.*?)```", section, flags=re.S).group(1)
    code = code.replace('"gymothelloenv_b200/_lib/libothello_b200.so"', repr(_cabi.LIB_PATH))
    scope = {}
    exec(compile(code, "INTEGRATION.md:B", "exec"), scope)
    torch.cuda.synchronize()
    assert scope["mask"][0].nonzero().flatten().tolist() == [18, 20, 34]              # white's replies to 19 (oracle)
    assert float(scope["reward"][0]) == 0.0 and not bool(scope["done"][0]) and int(scope["turn"][0]) == 1
    assert int(scope["obs"][0].sum()) == -3                                          # mover white: 1 own, 4 opponent
