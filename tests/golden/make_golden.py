#!/usr/bin/env python
"""Generate the committed golden vectors from the LIVE reference (build container only).

    python tests/golden/make_golden.py          # rewrites tests/golden/*.npz / *.json

The reference (omurammm/GymOthelloEnv @ /root/reference) is imported unmodified through
``oracle/ref_shim.py``.  Nothing here is imported by the product.  The reference ships
no tests or fixtures of its own, so these files -- outputs of the
reference itself on seeded inputs -- are what pins the oracle and the CUDA path.

Files written
  trajectories.npz  replayed action streams through OthelloBaseEnv (othello.py:217-501)
                    for board sizes 4..12: board, turn, cached legal list, obs, reward,
                    done, winner, disc counts and util.make_state planes after every ply,
                    including invalid actions (both sudden-death modes), passes,
                    double-pass and full-board terminals, both reward modes.
  policies.npz      GreedyPolicy / MaxiMinPolicy(1..3) choices (simple_policies.py:57-163)
                    on positions sampled from random games.
  wrappers.json     SimpleOthelloEnv / OthelloEnv seeded games with RandomPolicy /
                    GreedyPolicy (othello.py:21-214, simple_policies.py:21-44), i.e. the
                    host-side numpy RandomState streams, plus a run_2agent-style
                    tournament tally (run_2agent.py:92-203, fixed colours).
  perft.json        perft counts through set_player_turn + step.
"""
import json
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle import ref_shim  # noqa: E402

othello, simple_policies, util = ref_shim.load_reference()

SIZES = [4, 5, 6, 7, 8, 9, 10, 11, 12]


def dense_legal(env):
    n = env.board_size ** 2
    m = np.zeros(n, dtype=np.uint8)
    if len(env.possible_moves):
        m[np.asarray(env.possible_moves, dtype=np.int64)] = 1
    return m


def snapshot(env, obs, rec):
    rec["board"].append(np.array(env.board_state, dtype=np.int8, copy=True))
    rec["turn"].append(int(env.player_turn))
    rec["legal"].append(dense_legal(env))
    o = np.array(obs, dtype=np.int8, copy=True)
    if o.ndim == 2:
        o = o[None]
    rec["obs"].append(o)
    rec["winner"].append(int(env.winner))
    w, b = env.count_disks()
    rec["cnt"].append((int(w), int(b)))
    if not env.possible_actions_in_obs:
        rec["state4"].append(np.array(util.make_state(np.array(obs, copy=True), env), dtype=np.int8))


def play_episode(S, sudden, disk_reward, seed, p_invalid_pct, mask_in_obs):
    env = othello.OthelloBaseEnv(board_size=S, sudden_death_on_invalid_move=bool(sudden),
                                 num_disk_as_reward=bool(disk_reward),
                                 possible_actions_in_obs=bool(mask_in_obs), mute=True)
    rnd = np.random.RandomState(seed)
    rec = {k: [] for k in ("board", "turn", "legal", "obs", "winner", "cnt", "state4", "act", "rew", "done")}
    obs = env.reset()
    snapshot(env, obs, rec)
    n = env.board_size ** 2
    guard = 0
    while not env.terminated and guard < 4 * n:
        guard += 1
        # Appendix B recipe: one randint per ply picks among the cached legal list.
        a = env.possible_moves[rnd.randint(0, len(env.possible_moves))]
        if p_invalid_pct and rnd.randint(0, 100) < p_invalid_pct:
            k = rnd.randint(0, 10)
            a = [-1, n, n + 7, 1000][k] if k < 4 else int(rnd.randint(0, n))
        obs, reward, done, info = env.step(a)
        assert info is None
        rec["act"].append(int(a))
        rec["rew"].append(int(reward))
        rec["done"].append(int(bool(done)))
        snapshot(env, obs, rec)
    out = {
        "cfg": np.array([S, sudden, disk_reward, seed, p_invalid_pct, mask_in_obs], dtype=np.int32),
        "act": np.array(rec["act"], dtype=np.int16),
        "rew": np.array(rec["rew"], dtype=np.int16),
        "done": np.array(rec["done"], dtype=np.uint8),
        "board": np.stack(rec["board"]),
        "turn": np.array(rec["turn"], dtype=np.int8),
        "legal": np.stack(rec["legal"]),
        "obs": np.stack(rec["obs"]),
        "winner": np.array(rec["winner"], dtype=np.int8),
        "cnt": np.array(rec["cnt"], dtype=np.int16),
    }
    if rec["state4"]:
        out["state4"] = np.stack(rec["state4"])
    return out


def gen_trajectories():
    eps = []
    for S in SIZES:
        for seed in (0, 1, 2):
            for disk in (0, 1):
                eps.append(play_episode(S, 1, disk, seed, 0, 0))
        for seed in (10, 11):
            for disk in (0, 1):
                eps.append(play_episode(S, 0, disk, seed, 15, 0))   # invalid => forfeited turn
                eps.append(play_episode(S, 1, disk, seed, 4, 0))    # invalid => sudden death
        eps.append(play_episode(S, 1, 0, 20, 0, 1))                 # possible_actions_in_obs
        eps.append(play_episode(S, 0, 1, 21, 10, 1))
    flat = {"num_episodes": np.array(len(eps))}
    for i, e in enumerate(eps):
        for k, v in e.items():
            flat["e%d_%s" % (i, k)] = v
    np.savez_compressed(os.path.join(HERE, "trajectories.npz"), **flat)
    plies = sum(len(e["act"]) for e in eps)
    print("trajectories: %d episodes, %d plies" % (len(eps), plies))


def sample_positions(S, games, every, seed0):
    """Positions (env objects are rebuilt later) from uniformly random games."""
    out = []
    for g in range(games):
        env = othello.OthelloBaseEnv(board_size=S, mute=True)
        rnd = np.random.RandomState(seed0 + g)
        env.reset()
        ply = 0
        while not env.terminated:
            if ply % every == (g % every):
                out.append((np.array(env.board_state, dtype=np.int8, copy=True), int(env.player_turn)))
            env.step(env.possible_moves[rnd.randint(0, len(env.possible_moves))])
            ply += 1
    return out


def policy_choices(S, board, turn, depths):
    env = othello.OthelloBaseEnv(board_size=S, mute=True)
    env.reset()
    env.set_board_state(board, perspective=othello.WHITE_DISK)   # absolute board
    env.set_player_turn(turn)
    assert len(env.possible_moves) > 0
    obs = np.array(env.get_observation(), copy=True)
    res = {}
    g = simple_policies.GreedyPolicy()
    g.reset(env)
    res["greedy"] = int(g.get_action(util.make_state(obs, env)))
    for d in (1, 2, 3, 4):
        if d in depths:
            m = simple_policies.MaxiMinPolicy(d)
            m.reset(env)
            res["mm%d" % d] = int(m.get_action(None))
        else:
            res["mm%d" % d] = -1
    return res


def gen_policies():
    flat = {}
    plan = {
        # S: (games, every, n_depth3, n_depth4)
        4: (6, 1, 10 ** 9, 10 ** 9),
        5: (6, 2, 10 ** 9, 40),
        6: (8, 3, 60, 6),
        8: (12, 4, 40, 2),
        9: (3, 8, 6, 0),
        10: (4, 8, 8, 0),
        12: (3, 12, 4, 0),
    }
    for S, (games, every, n3, n4) in plan.items():
        t0 = time.time()
        pos = sample_positions(S, games, every, 1000 + S)
        rows = []
        for i, (board, turn) in enumerate(pos):
            depths = {1, 2}
            if i < n3:
                depths.add(3)
            if i < n4:
                depths.add(4)
            r = policy_choices(S, board, turn, depths)
            rows.append((r["greedy"], r["mm1"], r["mm2"], r["mm3"], r["mm4"]))
        flat["s%d_board" % S] = np.stack([p[0] for p in pos])
        flat["s%d_turn" % S] = np.array([p[1] for p in pos], dtype=np.int8)
        flat["s%d_choice" % S] = np.array(rows, dtype=np.int16)   # greedy, mm1, mm2, mm3, mm4
        print("policies S=%d: %d positions (%.1fs)" % (S, len(pos), time.time() - t0))
    flat["sizes"] = np.array(sorted(plan), dtype=np.int32)
    np.savez_compressed(os.path.join(HERE, "policies.npz"), **flat)


def make_policy(kind, seed):
    if kind == "rand":
        return simple_policies.RandomPolicy(seed=seed)
    if kind == "greedy":
        return simple_policies.GreedyPolicy()
    if kind.startswith("maximin"):
        return simple_policies.MaxiMinPolicy(int(kind[7:]))
    raise ValueError(kind)


def two_agent_games(S, black_kind, white_kind, games, init_rand, seed, disk_reward):
    """run_2agent.py:92-203 with fixed colours (black = 'row' policy of the README table)
    and the 4-plane util.make_state fed to the policies."""
    with ref_shim.muted():
        env = othello.SimpleOthelloEnv(board_size=S, seed=seed, initial_rand_steps=init_rand,
                                       num_disk_as_reward=bool(disk_reward))
        pol = {-1: make_policy(black_kind, seed), 1: make_policy(white_kind, seed)}
        out = []
        for g in range(games):
            obs = env.reset()
            for p in pol.values():
                p.reset(env)
            rec = {"max_rand_steps": int(env.max_rand_steps), "requested": [], "turn": [], "reward": [],
                   "done": []}
            done = False
            while not done:
                turn = int(env.player_turn)
                state = util.make_state(np.array(obs, copy=True), env)
                a = int(pol[turn].get_action(state))
                obs, reward, done, _ = env.step(a)
                rec["requested"].append(a)
                rec["turn"].append(turn)
                rec["reward"].append(int(reward))
                rec["done"].append(int(bool(done)))
            w, b = env.env.count_disks()
            rec["final_board"] = np.array(env.env.board_state, dtype=np.int8).ravel().tolist()
            rec["white"], rec["black"] = int(w), int(b)
            rec["winner"] = int(env.env.winner)
            # reward of the last ply is from the mover's perspective; for black:
            rec["black_reward"] = int(reward) * (1 if turn == -1 else -1)
            out.append(rec)
    return out


def othello_env_games(S, protagonist, opp_kind, games, init_rand, seed):
    """OthelloEnv (othello.py:96-214) with an embedded RandomPolicy opponent (the only
    scripted opponent that works inside OthelloEnv in this fork); the
    protagonist draws from its own RandomState over env.possible_moves."""
    with ref_shim.muted():
        opp = make_policy(opp_kind, seed + 100)
        kw = {"white_policy": opp} if protagonist == -1 else {"black_policy": opp}
        env = othello.OthelloEnv(protagonist=protagonist, board_size=S, initial_rand_steps=init_rand,
                                 seed=seed, **kw)
        prot_rnd = np.random.RandomState(seed + 7)
        out = []
        for g in range(games):
            obs = env.reset()
            rec = {"max_rand_steps": int(env.max_rand_steps),
                   "reset_obs": np.array(obs, dtype=np.int8).ravel().tolist(),
                   "requested": [], "reward": [], "done": [], "obs_sum": [], "moves_after": []}
            done = False
            while not done:
                a = int(env.possible_moves[prot_rnd.randint(0, len(env.possible_moves))])
                obs, reward, done, _ = env.step(a)
                rec["requested"].append(a)
                rec["reward"].append(int(reward))
                rec["done"].append(int(bool(done)))
                rec["obs_sum"].append(int(np.sum(obs)))
                rec["moves_after"].append([int(x) for x in env.possible_moves])
            rec["final_obs"] = np.array(obs, dtype=np.int8).ravel().tolist()
            out.append(rec)
    return out


def gen_wrappers():
    doc = {"two_agent": [], "othello_env": []}
    for (S, bk, wk, games, init, seed, disk) in [
        (8, "rand", "rand", 4, 10, 0, 0),
        (8, "rand", "greedy", 10, 10, 0, 0),      # BASELINE config 1 protocol, first 10 games
        (8, "greedy", "rand", 4, 10, 3, 1),
        (8, "greedy", "maximin2", 2, 10, 0, 0),
        (6, "maximin2", "rand", 3, 6, 5, 1),
        (6, "maximin3", "greedy", 2, 4, 1, 0),
        (10, "rand", "greedy", 2, 10, 2, 0),
        (5, "rand", "maximin1", 3, 4, 9, 1),
    ]:
        t0 = time.time()
        games_out = two_agent_games(S, bk, wk, games, init, seed, disk)
        doc["two_agent"].append({"S": S, "black": bk, "white": wk, "init_rand": init, "seed": seed,
                                 "disk_reward": disk, "games": games_out})
        print("two_agent S=%d %s vs %s: %d games (%.1fs)" % (S, bk, wk, games, time.time() - t0))
    for (S, prot, games, init, seed) in [(8, -1, 3, 10, 0), (8, 1, 3, 10, 1), (6, 1, 3, 4, 2), (10, -1, 2, 0, 3)]:
        doc["othello_env"].append({"S": S, "protagonist": prot, "opponent": "rand", "init_rand": init,
                                   "seed": seed, "games": othello_env_games(S, prot, "rand", games, init, seed)})
    with open(os.path.join(HERE, "wrappers.json"), "w") as f:
        json.dump(doc, f, separators=(",", ":"))


def perft(env, depth):
    """Leaves reached by legal moves; a terminal position is a leaf."""
    if depth == 0 or env.terminated:
        return 1
    board = np.array(env.board_state, copy=True)
    turn = env.player_turn
    total = 0
    for mv in list(env.possible_moves):
        child = othello.OthelloBaseEnv(board_size=env.board_size, mute=True)
        child.reset()
        child.set_board_state(board, perspective=othello.WHITE_DISK)
        child.set_player_turn(turn)
        child.step(mv)
        total += perft(child, depth - 1)
    return total


def gen_perft():
    doc = {}
    for S, maxd in ((4, 8), (6, 6), (8, 6), (10, 4)):
        env = othello.OthelloBaseEnv(board_size=S, mute=True)
        vals = []
        for d in range(1, maxd + 1):
            env.reset()
            vals.append(perft(env, d))
        doc[str(S)] = vals
        print("perft S=%d: %s" % (S, vals))
    with open(os.path.join(HERE, "perft.json"), "w") as f:
        json.dump(doc, f)


if __name__ == "__main__":
    which = sys.argv[1:] or ["traj", "policies", "wrappers", "perft"]
    if "traj" in which:
        gen_trajectories()
    if "policies" in which:
        gen_policies()
    if "wrappers" in which:
        gen_wrappers()
    if "perft" in which:
        gen_perft()
