"""GPU parity of the fused tournament kernel (oth_rollout): every recorded game is
re-derived on the host from (seed, global env id) with an independent numpy Philox and
the C oracle's rules + policies -- action by action -- and W/D/L rates are compared with
the oracle's CPU tournament and the README table (README.md:44-50)."""
import math

import numpy as np
import pytest
import torch

from oracle import oracle
from tests import philox_ref

pytestmark = pytest.mark.gpu


def _rollout(*a, **kw):
    from gymothelloenv_b200.tournament import rollout
    return rollout(*a, device="cuda:0", **kw)


def _replay_games(S, black, white, init_rand, seed, base, rec, games):
    """Host re-derivation of `games` single-game envs; returns nothing, asserts equality."""
    from gymothelloenv_b200.tournament import parse_policy
    spec = {-1: parse_policy(black), 1: parse_policy(white)}
    trace = rec["trace"].cpu().numpy()
    result = rec["result"].cpu().numpy()
    plies = rec["plies"].cpu().numpy()
    counts = rec["counts"].cpu().numpy()
    orc = oracle.OracleVec(games, S)
    gid = base + np.arange(games, dtype=np.int64)
    ctr = np.zeros(games, dtype=np.int64)
    budget = np.zeros(games, dtype=np.int64)
    if init_rand // 2 > 0:
        budget = 2 * philox_ref.bounded(philox_ref.draw(seed, gid, 0, ctr), init_rand // 2 + 1)
        ctr += 1
    ply = np.zeros(games, dtype=np.int64)
    for t in range(trace.shape[1]):
        board, legal, turn, term, winner, n_moves = orc.export()
        live = term == 0
        if not live.any():
            break
        want = np.full(games, -1, dtype=np.int64)
        # random plies: opening budget or a random policy for the side to move
        kinds = np.array([spec[int(x)][0] for x in turn])
        is_rand = live & ((ply < budget) | np.isin(kinds, ("rand", "random")))
        r = philox_ref.draw(seed, gid, 0, ctr)
        idx = philox_ref.bounded(r, np.maximum(n_moves, 1))
        order = np.cumsum(legal, axis=1) - 1
        pick = ((order == idx[:, None]) & (legal == 1)).argmax(1)
        want[is_rand] = pick[is_rand]
        ctr += is_rand
        for colour in (-1, 1):
            kind, depth = spec[colour]
            sel = live & ~is_rand & (turn == colour)
            if sel.any():
                acts = orc.policy("greedy" if kind == "greedy" else "maximin", depth)
                want[sel] = acts[sel]
        got = trace[:, t].astype(np.int64)
        assert np.array_equal(got[live], want[live]), "ply %d" % t
        assert (got[~live] == -1).all()
        orc.step(np.where(live, want, 0))
        ply += live
    board, _, _, term, winner, _ = orc.export()
    assert term.all()
    assert np.array_equal(result, winner) and np.array_equal(plies, ply)
    assert np.array_equal(counts[:, 0], (board == -1).reshape(games, -1).sum(1))
    assert np.array_equal(counts[:, 1], (board == 1).reshape(games, -1).sum(1))


ROLLOUT_CASES = [
    (8, "rand", "greedy", 10, 512), (8, "greedy", "rand", 10, 512), (8, "greedy", "maximin2", 10, 192),
    (8, "maximin3", "greedy", 6, 40), (8, "rand", "rand", 0, 700), (6, "maximin3", "maximin2", 4, 128),
    (6, "maximin4", "rand", 0, 64), (10, "greedy", "rand", 10, 160), (12, "rand", "maximin2", 8, 48),
    (5, "maximin2", "greedy", 4, 200), (4, "maximin4", "maximin3", 2, 100), (9, "greedy", "greedy", 10, 64),
    (8, "maximin2", "maximin3", 10, 96), (8, "rand", "maximin4", 10, 33), (7, "maximin3", "maximin3", 2, 130),
    (13, "greedy", "rand", 10, 64), (14, "rand", "greedy", 6, 64), (16, "greedy", "maximin2", 10, 8), (15, "rand", "rand", 0, 96),
]


def _check_rollout(S, black, white, init_rand, games, **kw):
    seed, base = 12345 + S, 7_000_000_000
    tally, rec = _rollout(games, black, white, board_size=S, initial_rand_steps=init_rand, seed=seed,
                          env_id_base=base, record=True, trace_len=S * S + 8, **kw)
    torch.cuda.synchronize()
    _replay_games(S, black, white, init_rand, seed, base, rec, games)
    res = rec["result"].cpu().numpy()
    w, d, l, plies = tally.tolist()
    assert (w, d, l) == (int((res == -1).sum()), int((res == 0).sum()), int((res == 1).sum()))
    assert plies == int(rec["plies"].sum().item())


@pytest.mark.parametrize("S,black,white,init_rand,games", ROLLOUT_CASES)
def test_rollout_games_rederived_on_host(S, black, white, init_rand, games):
    """Every action of every game against the oracle's policies (ragged last warps, mixed depths in one warp,
    lane-cooperative maximin for S <= 8, per-thread DFS for S > 8)."""
    _check_rollout(S, black, white, init_rand, games)


@pytest.mark.parametrize("S,black,white,init_rand,games",
                         [c for c in ROLLOUT_CASES if "maximin" in c[1] + c[2] and c[0] <= 8])
def test_cooperative_search_fallback_rederived_on_host(S, black, white, init_rand, games):
    """OTH_F_SEARCH_SMALL_LISTS: the cooperative search may keep only 3 root moves / 12 pairs, so most decisions
    overflow and take the per-thread DFS fallback (depth 4: also mid-way through the root's moves)."""
    _check_rollout(S, black, white, init_rand, games, search_small_lists=True)


@pytest.mark.parametrize("black,white,games", [("greedy", "maximin2", 4099), ("maximin3", "rand", 1500),
                                               ("maximin4", "maximin2", 200), ("maximin3", "maximin3", 777)])
def test_cooperative_and_per_thread_search_play_identical_games(black, white, games):
    """OTH_F_SEARCH_PER_THREAD: the per-thread DFS and the lane-cooperative search are two implementations of
    MaxiMinPolicy.search; every recorded game (result, plies, final counts, action trace) must coincide."""
    kw = dict(board_size=8, initial_rand_steps=10, seed=99, env_id_base=123456789, record=True, trace_len=72)
    ta, ra = _rollout(games, black, white, **kw)
    tb, rb = _rollout(games, black, white, search_per_thread=True, **kw)
    assert ta.tolist() == tb.tolist()
    for k in ("result", "plies", "counts", "trace"):
        assert torch.equal(ra[k], rb[k]), k


def test_sharding_invariance_and_multi_game_envs():
    """Global env ids key the RNG: any split of the id range plays the same games."""
    S, seed = 8, 99
    _, whole = _rollout(1000, "greedy", "rand", board_size=S, seed=seed, env_id_base=0, record=True)
    _, a = _rollout(400, "greedy", "rand", board_size=S, seed=seed, env_id_base=0, record=True)
    _, b = _rollout(600, "greedy", "rand", board_size=S, seed=seed, env_id_base=400, record=True)
    for k in ("result", "plies", "counts"):
        assert torch.equal(whole[k], torch.cat([a[k], b[k]]))
    tally, rec = _rollout(300, "rand", "rand", board_size=6, seed=5, games_per_env=7, record=True)
    w, d, l, plies = tally.tolist()
    assert w + d + l == 2100 and plies == int(rec["plies"].sum().item())
    assert len(set(rec["plies"].view(300, 7)[0].tolist())) >= 1
    # consecutive games of one env differ (the draw counter runs on across games)
    counts = rec["counts"].view(300, 7, 2).cpu().numpy()
    assert (counts[:, 0] != counts[:, 1]).any()


def _rates(t):
    n = float(sum(t[:3]))
    return [x / n for x in t[:3]], n


@pytest.mark.parametrize("black,white,readme", [("rand", "greedy", (38, 1, 61)), ("greedy", "rand", (61, 5, 34)),
                                                ("rand", "rand", (48, 6, 46)), ("greedy", "maximin2", (27, 3, 70))])
def test_tallies_match_oracle_tournament_and_readme(black, white, readme):
    """Large-N GPU W/D/L rate vs (a) the C oracle playing the same protocol on the CPU and
    (b) the README's 100-game cell, each inside its own binomial interval (4 sigma)."""
    from gymothelloenv_b200.tournament import parse_policy, play_tournament
    n_gpu = 200_000
    res = play_tournament(black, white, num_games=n_gpu, board_size=8, initial_rand_steps=10, seed=2024, device="cuda:0")
    assert res["games"] == n_gpu and 55 < res["plies"] / n_gpu <= 60.5
    p_gpu = [res["wins"] / n_gpu, res["draws"] / n_gpu, res["losses"] / n_gpu]
    n_cpu = 600 if "maximin" in (black + white) else 4000
    _, t = oracle.tournament(8, 7, n_cpu, black=parse_policy(black), white=parse_policy(white), init_rand=10)
    p_cpu, _ = _rates(t)
    for k in range(3):
        sigma = math.sqrt(max(p_gpu[k] * (1 - p_gpu[k]), 1e-4) * (1.0 / n_cpu + 1.0 / n_gpu))
        assert abs(p_gpu[k] - p_cpu[k]) < 4 * sigma + 1e-3, (k, p_gpu, p_cpu)
    for k in range(3):
        sigma = math.sqrt(max(p_gpu[k] * (1 - p_gpu[k]), 1e-4) / 100.0)
        assert abs(p_gpu[k] - readme[k] / 100.0) < 4 * sigma + 0.01, (k, p_gpu, readme)


def test_greedy_and_maximin1_tournaments_are_identical():
    """README.md:54: the Greedy and Maximin-1 rows/columns are the same numbers."""
    a, _ = _rollout(5000, "greedy", "rand", seed=1)
    b, _ = _rollout(5000, "maximin1", "rand", seed=1)
    assert a.tolist() == b.tolist()
