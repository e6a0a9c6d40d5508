"""Helpers with the reference's ``util.py`` surface: ``make_state``, ``undo_state``,
``create_policy``.

``make_state`` (``util.py:48-74``) is the 4-plane encoding [black, white, turn, legal]
the PPO/DQN drivers feed their networks; here the planes are written by the engine
(obs layout ``planes4``), including the reference's quirk that the legal plane stays
empty when exactly one move exists (``util.py:55``).
"""
import numpy as np
import torch

from .simple_policies import AlphaBetaPolicy, GreedyPolicy, HumanPolicy, MaxiMinPolicy, RandomPolicy, undo_state  # noqa: F401


def create_policy(policy_type='rand', board_size=8, seed=0, search_depth=1, agent_name='', device=None):
    """util.py:28-45 (learner-backed kinds 'ppo' / 'rainbow' / 'dqn' are out of scope)."""
    if policy_type == 'rand':
        return RandomPolicy(seed=seed)
    if policy_type == 'greedy':
        return GreedyPolicy()
    if policy_type == 'alphabeta':
        return AlphaBetaPolicy(search_depth)
    if policy_type == 'maximin':
        return MaxiMinPolicy(search_depth)
    if policy_type == 'human':
        return HumanPolicy(board_size)
    raise ValueError("policy_type %r is not provided by gymothelloenv_b200 "
                     "(rand | greedy | maximin | alphabeta | human)" % (policy_type,))


def make_state(obs, env):
    """util.py:48-74 -> float64 (4,S,S) = [black, white, turn plane, legal plane].

    `env` is an OthelloBaseEnv / SimpleOthelloEnv / OthelloEnv of this package.  `obs`
    must be the env's current observation (that is how every reference driver calls it);
    a different board is encoded through a scratch env with the env's turn and cached
    move list, exactly what the reference's arithmetic would produce."""
    base = env.env if hasattr(env, 'env') else env
    board = np.asarray(obs)
    if board.ndim > 2:
        raise ValueError("make_state expects the (S,S) observation (possible_actions_in_obs=False)")
    vec = base._vec
    cur = base.board_state * base.player_turn
    if np.array_equal(board.astype(np.int64), cur):
        return base.current_planes().astype(np.float64)      # written by the device with the env's last call
    else:
        scratch = base._scratch()
        scratch.set_board_state(board.astype(np.int8).reshape(1, -1), perspective=base.player_turn,
                                turn=base.player_turn, recompute_legal=False)
        scratch._legal.copy_(vec._legal)
        vec = scratch
    planes = vec.observe("planes4", torch.int8)
    return planes[0].cpu().numpy().astype(np.float64)
