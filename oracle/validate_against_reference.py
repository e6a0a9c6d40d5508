"""TEST INFRASTRUCTURE -- live differential fuzz of the oracles against the imported reference.

    python oracle/validate_against_reference.py [--games 6] [--sizes 4,5,...,16] [--prebuilt]

For every board size: random games (with a share of invalid actions, both sudden-death modes, both reward modes)
are stepped through the reference's OthelloBaseEnv, the C oracle (oracle/othello_oracle.c) and the Python port
(oracle/rules_py.py) side by side; after every ply board, side to move, cached possible_moves, reward and done must
be identical.  On sampled positions the reference's GreedyPolicy / MaxiMinPolicy(1..3) choices must equal the C
oracle's.  tests/test_oracle_vs_live_reference.py runs a short version wherever the reference is importable
(the live tree in the build container, or the oracle/_ref build).  Never imported by the product package."""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import oracle, ref_shim, rules_py  # noqa: E402


def fuzz_rules(S, games, seed, prebuilt=False):
    """-> number of plies compared"""
    othello, _, _ = ref_shim.load_reference(prefer_prebuilt=prebuilt)
    rng = np.random.RandomState(seed)
    plies = 0
    for g in range(games):
        sudden, disk = bool(g & 1), bool(g & 2)
        p_bad = 0.0 if g % 3 == 0 else 0.08
        ref = othello.OthelloBaseEnv(S, sudden_death_on_invalid_move=sudden, num_disk_as_reward=disk, mute=True)
        ref.reset()
        orc = oracle.OracleVec(1, S, sudden_death=sudden, disk_reward=disk)
        port = rules_py.PyOthello(S, sudden_death_on_invalid_move=sudden, num_disk_as_reward=disk)
        n = ref.board_size
        done = False
        while not done:
            board, legal, turn, term, winner, _ = orc.export()
            assert np.array_equal(board[0], np.asarray(ref.board_state)), (S, g, plies)
            assert np.array_equal(port.grid, np.asarray(ref.board_state))
            assert turn[0] == ref.player_turn == port.turn
            assert np.flatnonzero(legal[0]).tolist() == list(ref.possible_moves) == list(port.moves), (S, g, plies)
            moves = list(ref.possible_moves)
            if moves and rng.rand() >= p_bad:
                a = moves[rng.randint(len(moves))]
            else:
                a = int(rng.randint(0, n * n))
            _, r_ref, done, _ = ref.step(a)
            r_orc, d_orc, _ = orc.step(np.array([a]))
            _, r_port, d_port, _ = port.step(a)
            assert int(r_ref) == int(r_orc[0]) == int(r_port), (S, g, plies, a)
            assert bool(done) == bool(d_orc[0]) == bool(d_port)
            plies += 1
        board, legal, turn, term, winner, _ = orc.export()
        assert np.array_equal(board[0], np.asarray(ref.board_state)) and winner[0] == ref.winner == port.winner
    return plies


def fuzz_policies(S, positions, depths, seed, prebuilt=False):
    """-> number of (position, policy) choices compared"""
    othello, sp, util = ref_shim.load_reference(prefer_prebuilt=prebuilt)
    rng = np.random.RandomState(seed)
    checked = 0
    with ref_shim.muted():
        env = othello.SimpleOthelloEnv(board_size=S, seed=seed)
        pols = [("greedy", 1, sp.GreedyPolicy())] + [("maximin", d, sp.MaxiMinPolicy(d)) for d in depths]
        while checked < positions * len(pols):
            obs = env.reset()
            for _, _, p in pols:
                p.reset(env)
            done, t = False, 0
            while not done:
                if t % 4 == 1:
                    orc = oracle.OracleVec(1, S)
                    orc.inject(np.asarray(env.env.board_state, dtype=np.int8)[None], np.array([env.player_turn]))
                    state = util.make_state(np.array(obs, copy=True), env)
                    for kind, d, p in pols:
                        want = int(p.get_action(state))
                        got = int(orc.policy(kind, d)[0])
                        assert want == got, (S, kind, d, want, got)
                        checked += 1
                moves = env.possible_moves
                obs, _, done, _ = env.step(moves[rng.randint(len(moves))])
                t += 1
    return checked


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--games", type=int, default=6)
    ap.add_argument("--sizes", default=",".join(str(s) for s in range(4, 17)))
    ap.add_argument("--prebuilt", action="store_true", help="use the oracle/_ref build instead of the live tree")
    a = ap.parse_args()
    total = 0
    for S in [int(x) for x in a.sizes.split(",")]:
        n = fuzz_rules(S, a.games, 1000 + S, a.prebuilt)
        depths = (1, 2, 3) if S <= 6 else ((1, 2) if S <= 8 else (1,))
        c = fuzz_policies(S, 6 if S <= 8 else 2, depths, 2000 + S, a.prebuilt)
        print("S=%2d  %5d plies identical (reference / C oracle / port), %3d policy choices identical" % (S, n, c))
        total += n
    print("OK: %d plies" % total)


if __name__ == "__main__":
    main()
