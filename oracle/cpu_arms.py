"""TEST / BASELINE INFRASTRUCTURE -- the CPU arms bench.py times next to the GPU numbers.

Each arm is the reference's own code (oracle/_ref, built by oracle/make_ref.py; `kind: "reference"`) driven by the
harness SURVEY 8(c) prescribes, on all host cores: ONE fork pool for the whole arm, numpy and the reference imported
in the parent, every worker plays for `seconds` of its own wall clock and the rate is sum(units) / max(worker wall).
Without oracle/_ref the playout arms fall back to the port oracle/rules_py.py (`kind: "port"`), and say so.

  playout(S)            OthelloBaseEnv(S, mute=True) + uniform random legal moves     configs[1], configs[4]
  playout_make_state    the same + util.make_state(obs, env) every ply                 configs[3]
  two_agent(b, w)       SimpleOthelloEnv + policies, the run_2agent.py:119-167 loop    configs[0], configs[2]
Never imported by the product package.
"""
import multiprocessing as mp
import os
import time

import numpy as np

_REF = None            # (othello, simple_policies, util) loaded once in the parent, inherited by the forked workers


def reference_modules():
    global _REF
    if _REF is None:
        from oracle import ref_shim
        try:
            _REF = ref_shim.load_reference(prefer_prebuilt=True)
        except Exception:
            _REF = False
    return _REF or None


def kind():
    return "reference" if reference_modules() else "port"


# ------------------------------------------------------------------------------- workers
def _playout_worker(args):
    S, seed, seconds, with_state = args
    ref = reference_modules()
    rng = np.random.RandomState(seed)
    t0 = time.perf_counter()
    steps = games = 0
    if ref:
        othello, _, util = ref
        env = othello.OthelloBaseEnv(S, mute=True)
        while time.perf_counter() - t0 < seconds:
            obs = env.reset()
            done = False
            while not done:
                if with_state:
                    util.make_state(np.array(obs, copy=True), env)
                moves = env.possible_moves
                obs, _, done, _ = env.step(moves[rng.randint(len(moves))])
                steps += 1
            games += 1
    else:
        from oracle import rules_py
        env = rules_py.PyOthello(S)
        while time.perf_counter() - t0 < seconds:
            env.reset()
            done = False
            while not done:
                if with_state:
                    rules_py.make_state(env)
                moves = env.moves
                _, _, done = env.step(moves[rng.randint(len(moves))])[:3]
                steps += 1
            games += 1
    return steps, games, time.perf_counter() - t0


def _make_policy(sp, spec, seed):
    name, depth = spec
    if name in ("rand", "random"):
        return sp.RandomPolicy(seed=seed)
    if name == "greedy":
        return sp.GreedyPolicy()
    return sp.MaxiMinPolicy(depth)


def _two_agent_worker(args):
    """run_2agent.py:119-167 with fixed colours: black policy vs white policy on SimpleOthelloEnv, util.make_state
    fed to the policies, policy.reset(env) every game.  Plays until `seconds` are over but finishes the game in
    progress (a maximin-3 game takes seconds), min_games at least."""
    S, black, white, init_rand, seed, seconds, min_games = args
    import contextlib
    import io
    othello, sp, util = reference_modules()
    tally = [0, 0, 0]
    plies = games = 0
    t0 = time.perf_counter()
    with contextlib.redirect_stdout(io.StringIO()):
        env = othello.SimpleOthelloEnv(board_size=S, seed=seed, initial_rand_steps=init_rand)
        pol = {-1: _make_policy(sp, black, seed), 1: _make_policy(sp, white, seed + 1)}
        while games < min_games or time.perf_counter() - t0 < seconds:
            obs = env.reset()
            for p in pol.values():
                p.reset(env)
            done = False
            while not done:
                turn = int(env.player_turn)
                state = util.make_state(np.array(obs, copy=True), env)
                obs, reward, done, _ = env.step(int(pol[turn].get_action(state)))
                plies += 1
            w = int(env.env.winner)
            tally[0 if w < 0 else (1 if w == 0 else 2)] += 1
            games += 1
    return plies, games, time.perf_counter() - t0, tally


# ------------------------------------------------------------------------------- arms
class CpuArms(object):
    """One pool for every arm of a bench run."""

    def __init__(self, procs=None):
        reference_modules()                                  # import before forking
        self.procs = procs or os.cpu_count() or 1
        self.pool = mp.get_context("fork").Pool(self.procs)
        self.pool.map(_noop, range(self.procs))              # workers are up before anything is timed

    def close(self):
        self.pool.close()
        self.pool.join()

    def playout(self, S, seconds, with_state=False, seed=1000):
        res = self.pool.map(_playout_worker, [(S, seed + i, seconds, with_state) for i in range(self.procs)])
        wall = max(r[2] for r in res)
        steps, games = sum(r[0] for r in res), sum(r[1] for r in res)
        return {"value": steps / wall, "unit": "env_steps/s", "cores": self.procs, "kind": kind(),
                "sample": "%.0f s of uniformly random %dx%d games%s on %d processes (%s), %d games, %d steps"
                          % (seconds, S, S, " + util.make_state per ply" if with_state else "", self.procs,
                             _what(), games, steps)}

    def two_agent(self, S, black, white, init_rand, seconds, min_games=1, seed=0):
        if not reference_modules():
            return None                                      # the policies exist in the reference only
        res = self.pool.map(_two_agent_worker, [(S, black, white, init_rand, seed + 17 * i, seconds, min_games)
                                                for i in range(self.procs)])
        wall = max(r[2] for r in res)
        plies, games = sum(r[0] for r in res), sum(r[1] for r in res)
        tally = [sum(r[3][k] for r in res) for k in range(3)]
        return {"value": games / wall, "unit": "games/s", "plies_per_s": plies / wall, "cores": self.procs,
                "kind": "reference", "games": games, "black_wdl": tally,
                "sample": "SAMPLED: %d games of %s vs %s (%dx%d, %d random opening plies) in %.1f s on %d processes, "
                          "reference SimpleOthelloEnv + policies, run_2agent.py:119-167 loop"
                          % (games, _name(black), _name(white), S, S, init_rand, wall, self.procs)}


def _noop(i):
    return i


def _what():
    return "reference OthelloBaseEnv, oracle/_ref" if reference_modules() else "port oracle/rules_py.py"


def _name(spec):
    return spec[0] if spec[0] != "maximin" else "maximin-%d" % spec[1]
