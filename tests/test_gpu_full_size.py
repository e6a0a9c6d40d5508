"""GPU: the BASELINE workloads at their FULL sizes -- configs[1] (2^20 8x8 envs in one launch, in-kernel random
policy, auto-reset, int8 obs + legal mask), configs[4] (6x6 / 10x10 / 12x12 at 2^20 envs) and configs[3] (65,536 envs,
planes4 float32 obs) -- checked through size-independent properties and a sampled oracle replay.

The oracle cannot replay 2^20 x 70 plies in test time, so at this size the test checks
  (1) invariants of EVERY env after EVERY step, computed on the device with torch from the raw state: colours
      disjoint, cached legal set on empty cells only, obs == mover-perspective expansion of the bitboards
      (othello.py:363-369), legal mask == expansion of the cached list, exactly one disc more per ply
      (othello.py:391-410 places one disc; flips only recolour) or the 4-disc start position after an auto-reset,
      non-zero reward only on terminal transitions;
  (2) sharding invariance: two 2^19-env halves with env_id_base 0 / 2^19 reach bit-identical states (the Philox
      streams are keyed by the global env id);
  (3) bit-exact parity against the pinned C oracle on a random sample of 4096 of the 2^20 envs, replaying the
      actions the kernel drew: board, turn, cached legal set, reward, done after every ply.
"""
import numpy as np
import pytest
import torch

from oracle import oracle

pytestmark = pytest.mark.gpu

SAMPLE = 4096
CASES = [  # (board size, envs, obs layout, obs dtype)
    (8, 1 << 20, "board", torch.int8),          # configs[1]
    (6, 1 << 20, "board", torch.int8),          # configs[4]
    (10, 1 << 20, "board", torch.int8),
    (12, 1 << 20, "board", torch.int8),
    (8, 65536, "planes4", torch.float32),       # configs[3]
]


def _bits(words, n):
    """int64 [W, B] -> int8 [B, n], bit a of the W-word board = cell a"""
    sh = torch.arange(64, device=words.device, dtype=torch.int64)
    b = ((words.unsqueeze(2) >> sh) & 1).to(torch.int8)            # [W, B, 64]
    return b.permute(1, 0, 2).reshape(words.shape[1], -1)[:, :n]


@pytest.mark.parametrize("S,B,layout,dtype", CASES)
def test_full_size_batch_invariants_sharding_and_sampled_oracle(S, B, layout, dtype):
    from gymothelloenv_b200 import VecOthelloEnv
    dev, N, steps = "cuda:0", S * S, S * S + 6
    kw = dict(seed=3, auto_reset=True, obs_layout=layout, obs_dtype=dtype)
    env = VecOthelloEnv(B, S, dev, **kw)
    halves = [VecOthelloEnv(B // 2, S, dev, env_id_base=k * (B // 2), **kw) for k in (0, 1)]
    env.reset()
    for h in halves:
        h.reset()
    rng = np.random.default_rng(5 + S)
    idx_np = np.sort(rng.choice(B, size=SAMPLE, replace=False))
    idx = torch.from_numpy(idx_np).to(dev)
    orc = oracle.OracleVec(SAMPLE, S)
    prev_cnt = torch.full((B,), 4, dtype=torch.int64, device=dev)
    finished = 0
    for t in range(steps):
        env.step(policy="random")
        for h in halves:
            h.step(policy="random")
        black, white, legal = env._black, env._white, env._legal
        # (1) invariants of every env
        assert not bool((black & white).any()), t
        assert not bool((legal & (black | white)).any()), t
        b8, w8, l8 = _bits(black, N), _bits(white, N), _bits(legal, N)
        mover_white = (env.player_turn == 1).unsqueeze(1)
        if layout == "board":                                           # othello.py:363-369
            want_obs = torch.where(mover_white, w8 - b8, b8 - w8)
            assert torch.equal(env.obs.reshape(B, N), want_obs), t
        else:                                                           # util.py:48-74 (legal plane needs >= 2 moves)
            lp = l8 * (l8.sum(1, keepdim=True) >= 2).to(torch.int8)
            want_obs = torch.stack([b8, w8, mover_white.to(torch.int8).expand(B, N), lp], dim=1).to(dtype)
            assert torch.equal(env.obs.reshape(B, 4, N), want_obs), t
        assert torch.equal(env._mask_u8, l8.to(torch.uint8)), t
        assert bool(((env.player_turn == 1) | (env.player_turn == -1)).all()), t
        cnt = (b8.sum(1) + w8.sum(1)).to(torch.int64)
        done = env.done
        assert bool(torch.where(done, cnt == 4, cnt == prev_cnt + 1).all()), t
        assert not bool((env.reward != 0)[~done].any()), t
        prev_cnt = cnt
        finished += int(done.sum().item())
        # (3) the sampled envs against the oracle, on the actions the kernel drew
        acts = env.actions_taken[idx].cpu().numpy()
        reward, odone, valid = orc.step(acts)
        assert valid.all(), t
        assert np.array_equal(reward.astype(np.float32), env.reward[idx].cpu().numpy()), t
        assert np.array_equal(odone, env._done_u8[idx].cpu().numpy()), t
        orc.reset_done()
        oboard, olegal, oturn, oterm, _, _ = orc.export()
        dboard = (w8[idx] - b8[idx]).cpu().numpy().reshape(SAMPLE, S, S)      # absolute colours, WHITE = +1
        assert np.array_equal(dboard, oboard), t
        assert np.array_equal(l8[idx].cpu().numpy().astype(np.uint8), olegal), t
        assert np.array_equal(env.player_turn[idx].cpu().numpy().astype(np.int32), oturn), t
        assert not oterm.any()
    assert finished > B // 2                    # most envs completed a game and restarted inside the window
    # (2) sharding invariance at full size
    for name in ("_black", "_white", "_legal", "_meta", "_ctr"):
        whole = getattr(env, name)
        parts = torch.cat([getattr(h, name) for h in halves], dim=-1)
        assert torch.equal(whole, parts), name


@pytest.mark.parametrize("black,white,games", [("greedy", "maximin2", 1 << 21), ("maximin3", "greedy", 1 << 19),
                                               ("rand", "greedy", 1 << 22)])
def test_full_size_tournament_shards_add_up(black, white, games):
    """configs[2] / configs[0] at (per-GPU) full size: the tally of one 2^21-game launch equals the sum of the tallies of
    its two halves launched with their global env-id offsets -- what makes the 8-GPU all-reduce exact --
    and every game is accounted for (W + D + L = games, plies within a game's bounds)."""
    from gymothelloenv_b200.tournament import rollout
    whole, _ = rollout(games, black, white, board_size=8, initial_rand_steps=10, seed=11, device="cuda:0")
    parts = [rollout(games // 2, black, white, board_size=8, initial_rand_steps=10, seed=11, device="cuda:0",
                     env_id_base=k * (games // 2))[0] for k in (0, 1)]
    whole = whole.cpu().numpy()
    assert np.array_equal(whole, parts[0].cpu().numpy() + parts[1].cpu().numpy())
    assert int(whole[:3].sum()) == games
    assert 50 * games < int(whole[3]) <= 60 * games + games        # <= 60 placements + forced passes are not plies
