"""TEST INFRASTRUCTURE -- ctypes binding of oracle/othello_oracle.c (the CPU oracle).

Importers allowed: tests/, __graft_entry__.smoke(), bench.py (cpu_baseline and
--impl reference legs).  The product package ``gymothelloenv_b200`` must never import
this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SRC = os.path.join(_HERE, "othello_oracle.c")
_LIB = os.path.join(_HERE, "_build", "libothello_oracle.so")

_lib = None


def build(force=False):
    """gcc -O2 -shared; rebuilt when the source is newer than the library."""
    if (not force and os.path.isfile(_LIB) and os.path.getmtime(_LIB) >= os.path.getmtime(_SRC)):
        return _LIB
    os.makedirs(os.path.dirname(_LIB), exist_ok=True)
    tmp = _LIB + ".%d.tmp" % os.getpid()
    subprocess.check_call(["gcc", "-O2", "-std=c11", "-fPIC", "-shared", "-Wall", "-o", tmp, _SRC])
    os.replace(tmp, _LIB)
    return _LIB


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(build())
        L = _lib
        L.orc_sizeof_env.restype = C.c_int
        L.orc_max_board_size.restype = C.c_int
        L.orc_perft.restype = C.c_int64
        L.orc_perft.argtypes = [C.c_void_p, C.c_int]
        L.orc_random_playouts.restype = C.c_int64
        L.orc_random_playouts.argtypes = [C.c_int, C.c_uint64, C.c_int64, C.c_int, C.c_void_p]
        L.orc_tournament.restype = C.c_int64
        L.orc_tournament.argtypes = [C.c_int, C.c_uint64, C.c_int64, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                     C.c_void_p]
        L.orc_vec_init.argtypes = [C.c_void_p, C.c_int64, C.c_int, C.c_int, C.c_int]
        L.orc_vec_step.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_vec_reset_done.argtypes = [C.c_void_p, C.c_int64]
        L.orc_vec_export.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_vec_observe.argtypes = [C.c_void_p, C.c_int64, C.c_int, C.c_void_p, C.c_void_p]
        L.orc_vec_policy.argtypes = [C.c_void_p, C.c_int64, C.c_int, C.c_int, C.c_void_p]
        L.orc_vec_inject.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


class OracleVec(object):
    """B independent oracle envs stepped in lockstep (all state lives in C structs)."""

    def __init__(self, num_envs, board_size=8, sudden_death=True, disk_reward=False):
        L = lib()
        self.B = int(num_envs)
        self.S = min(max(4, int(board_size)), L.orc_max_board_size())
        self.N = self.S * self.S
        self._buf = np.zeros(self.B * L.orc_sizeof_env(), dtype=np.uint8)
        L.orc_vec_init(_p(self._buf), self.B, self.S, int(bool(sudden_death)), int(bool(disk_reward)))

    def step(self, actions):
        """-> (reward int32[B], done uint8[B], was_valid uint8[B]); terminated envs are no-ops."""
        a = np.ascontiguousarray(actions, dtype=np.int64)
        assert a.shape == (self.B,)
        reward = np.zeros(self.B, dtype=np.int32)
        done = np.zeros(self.B, dtype=np.uint8)
        valid = np.zeros(self.B, dtype=np.uint8)
        lib().orc_vec_step(_p(self._buf), self.B, _p(a), _p(reward), _p(done), _p(valid))
        return reward, done, valid

    def reset_done(self):
        lib().orc_vec_reset_done(_p(self._buf), self.B)

    def export(self):
        """-> board int8[B,S,S] (white=+1), legal uint8[B,N] (cached list), turn, terminated, winner, n_moves"""
        board = np.zeros((self.B, self.S, self.S), dtype=np.int8)
        legal = np.zeros((self.B, self.N), dtype=np.uint8)
        meta = np.zeros((self.B, 4), dtype=np.int32)
        lib().orc_vec_export(_p(self._buf), self.B, _p(board), _p(legal), _p(meta))
        return board, legal, meta[:, 0].copy(), meta[:, 1].copy(), meta[:, 2].copy(), meta[:, 3].copy()

    def observe(self, with_mask=False):
        obs = np.zeros((self.B, 2 if with_mask else 1, self.S, self.S), dtype=np.int8)
        lib().orc_vec_observe(_p(self._buf), self.B, int(with_mask), _p(obs), None)
        return obs

    def make_state(self):
        st = np.zeros((self.B, 4, self.S, self.S), dtype=np.int8)
        lib().orc_vec_observe(_p(self._buf), self.B, 0, None, _p(st))
        return st

    def policy(self, kind, depth=1):
        """kind: 'greedy' | 'maximin' -> int64[B] (-1 for terminated envs)"""
        out = np.zeros(self.B, dtype=np.int64)
        lib().orc_vec_policy(_p(self._buf), self.B, 1 if kind == "greedy" else 2, int(depth), _p(out))
        return out

    def inject(self, board, turn):
        b = np.ascontiguousarray(board, dtype=np.int8).reshape(self.B, self.N)
        t = np.ascontiguousarray(turn, dtype=np.int8).reshape(self.B)
        lib().orc_vec_inject(_p(self._buf), self.B, _p(b), _p(t))

    def perft(self, index, depth):
        sz = lib().orc_sizeof_env()
        return int(lib().orc_perft(C.c_void_p(self._buf.ctypes.data + index * sz), int(depth)))


def random_playouts(board_size, seed, games, with_obs=False):
    """Single-thread uniformly-random playouts; -> (env_steps, [black W, D, L])"""
    tally = np.zeros(3, dtype=np.int64)
    steps = lib().orc_random_playouts(int(board_size), int(seed), int(games), int(with_obs), _p(tally))
    return int(steps), tally.tolist()


_KIND = {"rand": 0, "random": 0, "greedy": 1, "maximin": 2}


def tournament(board_size, seed, games, black=("greedy", 1), white=("rand", 1), init_rand=10):
    tally = np.zeros(3, dtype=np.int64)
    steps = lib().orc_tournament(int(board_size), int(seed), int(games), _KIND[black[0]], int(black[1]),
                                 _KIND[white[0]], int(white[1]), int(init_rand), _p(tally))
    return int(steps), tally.tolist()
