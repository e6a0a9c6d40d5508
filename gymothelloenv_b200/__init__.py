"""gymothelloenv_b200 -- B200-native batched Othello environment.

Drop-in for the env hot path of omurammm/GymOthelloEnv: the reference's class API
(``othello.py``, ``simple_policies.py``, ``util.py``) is mirrored by the same-named
modules in this package; all rules/policy compute runs in hand-written sm_100a CUDA
kernels behind a C ABI (``include/othello_b200.h``).  No CPU fallback.
"""
from . import _cabi
from ._cabi import OthelloB200Error
from .vec_env import VecOthelloEnv, shard_range
from .envs import Envs, PPOEnvs, PPOTeacherStudentEnvs
from .sampling import masked_categorical, LegalMoveSampler, epsilon_greedy, EpsilonGreedy, evaluate_actions, ChoiceStore

BLACK_DISK = -1   # othello.py:10
NO_DISK = 0       # othello.py:11
WHITE_DISK = 1    # othello.py:12

__all__ = ["VecOthelloEnv", "Envs", "PPOEnvs", "PPOTeacherStudentEnvs", "masked_categorical", "LegalMoveSampler", "evaluate_actions", "ChoiceStore", "epsilon_greedy", "EpsilonGreedy", "OthelloB200Error", "shard_range", "BLACK_DISK", "NO_DISK", "WHITE_DISK"]
