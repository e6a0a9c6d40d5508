"""Scripted policies with the reference's API (``simple_policies.py``): ``reset(env)``,
``get_action(obs)``, ``get_test_action(obs)``, ``seed(s)``.

RandomPolicy keeps the host ``numpy.random.RandomState`` stream of the reference
(``simple_policies.py:21-44``) so seeded games are reproducible against it.  Greedy and
MaxiMin do NOT clone-and-step Python envs (``simple_policies.py:69-92,111-155``): they
call the batched policy kernel (``oth_policy``) on the env's device state -- flip
popcounts / depth-k minimax in registers -- and return the same action index
(first-occurrence tie-breaks included).
"""
import numpy as np

from . import _cabi
from .othello import OthelloBaseEnv

WHITE_DISK = 1
BLACK_DISK = -1
PROTAGONIST_TURN = 1
OPPONENT_TURN = -1


def copy_env(env, mute_env=True):
    """simple_policies.py:12-18: a fresh env with the same rules configuration."""
    new_env = env.__class__(board_size=env.board_size,
                            sudden_death_on_invalid_move=env.sudden_death_on_invalid_move, mute=mute_env)
    new_env.reset()
    return new_env


def undo_state(state, player_turn):
    """simple_policies.py:46-54 / util.py:77-85: plane state -> mover-perspective board."""
    assert int((player_turn + 1) / 2) == int(state[2][0][0])
    if player_turn == -1:
        return state[0] - state[1]
    return state[1] - state[0]


class _PolicyBase(object):
    def __init__(self):
        self.env = None

    def reset(self, env):
        self.env = env.env if hasattr(env, 'env') else env      # simple_policies.py:27-31

    def get_test_action(self, obs):
        return self.get_action(obs)

    def _base_env(self):
        if not isinstance(self.env, OthelloBaseEnv):
            raise TypeError("policy.reset(env) must be called with a gymothelloenv_b200 env first")
        return self.env


class RandomPolicy(_PolicyBase):
    """Uniform over env.possible_moves with its own RandomState (simple_policies.py:21-44)."""

    def __init__(self, seed=0):
        super(RandomPolicy, self).__init__()
        self.rnd = np.random.RandomState(seed=seed)

    def seed(self, seed):
        self.rnd = np.random.RandomState(seed=seed)

    def get_action(self, obs):
        possible_moves = self.env.possible_moves
        ix = self.rnd.randint(0, len(possible_moves))
        return possible_moves[ix]


def _device_pick(base_env, kind, depth, obs_board=None):
    """One oth_policy launch on the B=1 device state of `base_env` -> python int."""
    vec = base_env._vec
    if obs_board is not None:
        # GreedyPolicy scores moves on the board decoded from `obs`, not on the env's own
        # board (simple_policies.py:70,79-80).  They coincide in every reference driver; if a
        # caller passes a different position, evaluate that one in a scratch env that keeps the
        # env's cached move list (the list the reference iterates over, :75).
        cur = base_env.board_state * base_env.player_turn
        if not np.array_equal(np.asarray(obs_board, dtype=np.int64), cur):
            scratch = base_env._scratch()
            scratch.set_board_state(np.asarray(obs_board, dtype=np.int8).reshape(1, -1),
                                    perspective=base_env.player_turn, turn=base_env.player_turn,
                                    recompute_legal=True)
            legal_now = set(scratch.possible_moves(0))
            assert all(m in legal_now for m in base_env.possible_moves)      # simple_policies.py:82
            scratch._legal.copy_(vec._legal)
            vec = scratch
    if vec is base_env._vec:
        return base_env.device_pick(kind, depth)             # one launch, the action lands in pinned memory
    return int(vec.policy_actions(kind, depth)[0].item())


class GreedyPolicy(_PolicyBase):
    """Greed is good (simple_policies.py:57-95): the move that leaves the mover with the
    most discs, lowest index on ties."""

    def get_action(self, obs):
        env = self._base_env()
        board = undo_state(obs, env.player_turn)
        return _device_pick(env, "greedy", 1, obs_board=board)


class MaxiMinPolicy(_PolicyBase):
    """Depth-k maximin on own disc count (simple_policies.py:98-163).  Depths 1..4 search every node as the reference
    does; depths 5..8 (the reference accepts any depth) run the alpha-beta search, which returns the same move.
    The `obs` argument is ignored, as in the reference (:157-163)."""

    def __init__(self, max_search_depth=1):
        super(MaxiMinPolicy, self).__init__()
        self.max_search_depth = max_search_depth
        if not (1 <= int(max_search_depth) <= _cabi.MAX_SEARCH_DEPTH):
            raise ValueError("max_search_depth must be in 1..%d in this build" % _cabi.MAX_SEARCH_DEPTH)

    def get_action(self, obs):
        env = self._base_env()
        if env.terminated or not env.possible_moves:
            return None                                           # search() returns (count, None) at a leaf
        return _device_pick(env, "maximin", int(self.max_search_depth))


class AlphaBetaPolicy(MaxiMinPolicy):
    """MaxiMinPolicy with alpha-beta pruning (the reference's README.md:55-56: "can be improved by pruning (not
    included)"): the SAME move as MaxiMinPolicy(depth) -- full window at the root, first-occurrence ties -- for a
    fraction of the positions searched."""

    def get_action(self, obs):
        env = self._base_env()
        if env.terminated or not env.possible_moves:
            return None
        return _device_pick(env, "alphabeta", int(self.max_search_depth))


class HumanPolicy(object):
    """simple_policies.py:166-176"""

    def __init__(self, board_size):
        self.board_size = board_size

    def reset(self, env):
        pass

    def get_action(self, obs):
        return int(input('Enter action index:'))
