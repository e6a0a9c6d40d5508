"""ctypes binding of libothello_b200.so (include/othello_b200.h).

There is NO fallback: if the CUDA library is missing or a call fails, this raises.
PyTorch is used only to own device memory and streams; the library itself has no
torch dependency (plain pointers and sizes cross the boundary).
"""
import ctypes as C
import os

_PKG = os.path.dirname(os.path.abspath(__file__))
# OTH_LIB_DIR (development only): directory name of an A/B variant build made by scripts/build_variant.py
_VARIANT = os.environ.get("OTH_LIB_DIR")
LIB_PATH = os.path.join(_PKG, _VARIANT or "_lib", "libothello_b200.so")

ABI_VERSION = 2
MIN_BOARD_SIZE = 4
MAX_BOARD_SIZE = 16

# flags (OTH_F_*)
F_SUDDEN_DEATH = 0x01
F_DISK_REWARD = 0x02
F_AUTO_RESET = 0x04
F_LEGACY_SINGLE_MOVE_PLANE = 0x08
F_ZERO_IF_TERMINATED = 0x10
F_SP_FRESH_STATE = 0x20
F_SEARCH_SMALL_LISTS = 0x40
F_SEARCH_PER_THREAD = 0x80

# self-play phase bits (OTH_SP_*) and categorical modes (OTH_CAT_*)
SP_PENDING, SP_EMITTED, SP_FINISHED, SP_PREV_P = 0x01, 0x02, 0x04, 0x08
# teacher/student phase bits and tuple types (OTH_TS_*)
TS_EMITTED, TS_FINISHED, TS_PEND_BLACK, TS_PEND_WHITE = 0x01, 0x02, 0x04, 0x08
TS_PREV_BLACK, TS_PREV_WHITE, TS_QUEUED, TS_QUEUED_WHITE = 0x10, 0x20, 0x40, 0x80
TS_TYPE_BLACK, TS_TYPE_WHITE, TS_TYPE_OVER = 0, 1, 2
CAT_SAMPLE, CAT_MODE, CAT_EVALUATE = 0, 1, 2

OBS_NONE, OBS_BOARD, OBS_BOARD_MASK, OBS_PLANES4, OBS_PLANES3, OBS_ABSOLUTE = 0, 1, 2, 3, 4, 5
OBS_PLANES = {OBS_NONE: 0, OBS_BOARD: 1, OBS_BOARD_MASK: 2, OBS_PLANES4: 4, OBS_PLANES3: 3, OBS_ABSOLUTE: 1}
DTYPE_I8, DTYPE_F32 = 0, 1
ACT_EXTERNAL, ACT_RANDOM, ACT_GREEDY = 0, 1, 2
ACTION_HOLD = -(1 << 63)          # OTH_ACTION_HOLD: the env sits this ply out
POLICY_RANDOM, POLICY_GREEDY, POLICY_MAXIMIN, POLICY_ALPHABETA = 0, 1, 2, 3
MAX_SEARCH_DEPTH = 8
SEARCH_PRUNE, SEARCH_COOPERATIVE = 1, 2

META_TURN_WHITE = 0x1
META_TERMINATED = 0x2
META_WINNER_SHIFT = 2
META_RANDCNT_SHIFT = 8
META_RANDMAX_SHIFT = 16

EXPORTED_SYMBOLS = (
    "oth_abi_version", "oth_last_error", "oth_words_per_board", "oth_reset", "oth_observe", "oth_step",
    "oth_policy", "oth_rollout", "oth_set_board", "oth_update_board", "oth_int_alu_probe",
    "oth_selfplay_new_game", "oth_selfplay_begin", "oth_selfplay_advance", "oth_masked_categorical",
    "oth_epsilon_greedy", "oth_ts_new_game", "oth_ts_begin", "oth_ts_advance", "oth_rollout_noalloc",
    "oth_masked_logprob_forward", "oth_masked_logprob_backward", "oth_search",
)
NOALLOC_MAX_GAMES = 1 << 21


class OthState(C.Structure):
    _fields_ = [("black", C.c_void_p), ("white", C.c_void_p), ("legal", C.c_void_p), ("meta", C.c_void_p),
                ("ctr", C.c_void_p)]


class OthOutputs(C.Structure):
    _fields_ = [("obs", C.c_void_p), ("legal_mask", C.c_void_p), ("reward", C.c_void_p), ("done", C.c_void_p),
                ("player_turn", C.c_void_p), ("winner", C.c_void_p), ("actions_out", C.c_void_p),
                ("obs_layout", C.c_int32), ("obs_dtype", C.c_int32), ("packed", C.c_void_p)]


class OthSelfPlay(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in (
        "protagonist", "phase", "sav_black", "sav_white", "sav_legal", "sav_meta", "sav_choices", "tr_action",
        "tr_reward", "tr_done", "tr_type", "tr_outputs", "pend_outputs", "pend_action")]


class OthTeacherStudent(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in (
        "phase", "sav_black", "sav_white", "sav_legal", "sav_meta", "sav_choices", "sav_action", "sav_outputs",
        "queued_reward", "tr_black", "tr_white", "tr_legal", "tr_choices", "tr_meta", "tr_action", "tr_reward",
        "tr_done", "tr_type", "tr_who", "tr_outputs")]


class OthPolicySpec(C.Structure):
    _fields_ = [("kind", C.c_int32), ("depth", C.c_int32)]


class OthelloB200Error(RuntimeError):
    pass


_lib = None


def load():
    """Loads the shared library (once). Raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.isfile(LIB_PATH):
        raise OthelloB200Error(
            "libothello_b200.so is missing (%s). Build it with `python -m gymothelloenv_b200.build` "
            "(needs nvcc; targets sm_100a). There is no CPU fallback." % LIB_PATH)
    if not _VARIANT:
        # a library built from other sources than the ones on disk has produced wrong conclusions before: refuse it
        from . import build as _build
        stamp = LIB_PATH + ".digest"
        have = open(stamp).read().strip() if os.path.isfile(stamp) else None
        want = _build.source_digest()
        if have != want:
            raise OthelloB200Error(
                "libothello_b200.so is stale: built from sources with digest %s, the tree has %s. Rebuild with "
                "`python -m gymothelloenv_b200.build`." % (have and have[:12], want[:12]))
    L = C.CDLL(LIB_PATH)
    L.oth_abi_version.restype = C.c_int
    L.oth_last_error.restype = C.c_char_p
    L.oth_words_per_board.restype = C.c_int
    L.oth_words_per_board.argtypes = [C.c_int]
    L.oth_reset.restype = C.c_int
    L.oth_reset.argtypes = [OthState, C.c_int64, C.c_int, C.c_uint64, C.c_int64, C.c_int, C.c_int, C.c_void_p,
                            C.c_void_p]
    L.oth_observe.restype = C.c_int
    L.oth_observe.argtypes = [OthState, C.c_int64, C.c_int, C.c_int, OthOutputs, C.c_void_p]
    L.oth_step.restype = C.c_int
    L.oth_step.argtypes = [OthState, C.c_int64, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_uint64, C.c_int64,
                           C.c_int, OthOutputs, C.c_void_p]
    L.oth_policy.restype = C.c_int
    L.oth_policy.argtypes = [OthState, C.c_int64, C.c_int, OthPolicySpec, C.c_uint64, C.c_uint64, C.c_int64,
                             C.c_void_p, C.c_void_p]
    L.oth_search.restype = C.c_int
    L.oth_search.argtypes = [OthState, C.c_int64, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
    L.oth_rollout.restype = C.c_int
    L.oth_rollout.argtypes = [C.c_int64, C.c_int, C.c_int, OthPolicySpec, OthPolicySpec, C.c_int, C.c_int,
                              C.c_uint64, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                              C.c_int, C.c_void_p]
    L.oth_rollout_noalloc.restype = C.c_int
    L.oth_rollout_noalloc.argtypes = [C.c_int64, C.c_int, C.c_int, OthPolicySpec, OthPolicySpec, C.c_int, C.c_int,
                                      C.c_uint64, C.c_int64, C.c_void_p, C.c_void_p]
    L.oth_set_board.restype = C.c_int
    L.oth_set_board.argtypes = [OthState, C.c_int64, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int,
                                C.c_void_p]
    L.oth_update_board.restype = C.c_int
    L.oth_update_board.argtypes = [OthState, C.c_int64, C.c_int, C.c_void_p, C.c_void_p]
    L.oth_int_alu_probe.restype = C.c_int
    L.oth_int_alu_probe.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_uint64), C.c_void_p]
    L.oth_selfplay_new_game.restype = C.c_int
    L.oth_selfplay_new_game.argtypes = [OthState, OthSelfPlay, C.c_int64, C.c_int, C.c_uint64, C.c_int64, C.c_int,
                                        C.c_uint64, C.c_void_p, C.c_void_p]
    L.oth_selfplay_begin.restype = C.c_int
    L.oth_selfplay_begin.argtypes = [OthState, OthSelfPlay, C.c_int64, C.c_int, C.c_int, OthOutputs, C.c_void_p,
                                     C.c_void_p]
    L.oth_selfplay_advance.restype = C.c_int
    L.oth_selfplay_advance.argtypes = [OthState, OthSelfPlay, C.c_int64, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                       C.c_void_p, C.c_void_p, C.c_uint64, C.c_int64, C.c_int, OthOutputs,
                                       C.c_void_p, C.c_void_p]
    L.oth_masked_categorical.restype = C.c_int
    L.oth_masked_categorical.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_int, C.c_void_p, C.c_uint64,
                                         C.c_uint64, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]
    L.oth_masked_logprob_forward.restype = C.c_int
    L.oth_masked_logprob_forward.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_void_p, C.c_void_p,
                                             C.c_void_p, C.c_void_p]
    L.oth_masked_logprob_backward.restype = C.c_int
    L.oth_masked_logprob_backward.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_void_p, C.c_void_p,
                                              C.c_void_p, C.c_void_p, C.c_void_p]
    L.oth_epsilon_greedy.restype = C.c_int
    L.oth_epsilon_greedy.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_float, C.c_uint64, C.c_uint64,
                                     C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    L.oth_ts_new_game.restype = C.c_int
    L.oth_ts_new_game.argtypes = [OthState, OthTeacherStudent, C.c_int64, C.c_int, C.c_uint64, C.c_int64, C.c_int,
                                  C.c_void_p]
    L.oth_ts_begin.restype = C.c_int
    L.oth_ts_begin.argtypes = [OthState, OthTeacherStudent, C.c_int64, C.c_int, C.c_int, C.c_int, C.c_float, OthOutputs,
                               C.c_void_p, C.c_void_p]
    L.oth_ts_advance.restype = C.c_int
    L.oth_ts_advance.argtypes = [OthState, OthTeacherStudent, C.c_int64, C.c_int, C.c_int, C.c_int, C.c_float, C.c_int,
                                 C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_int64, C.c_int,
                                 OthOutputs, C.c_void_p, C.c_void_p]
    if L.oth_abi_version() != ABI_VERSION:
        raise OthelloB200Error("libothello_b200.so ABI version mismatch")
    _lib = L
    return L


def check(rc, what):
    if rc != 0:
        msg = load().oth_last_error()
        raise OthelloB200Error("%s failed (code %d): %s" % (what, rc, msg.decode() if msg else "?"))


def words_per_board(board_size):
    n = board_size * board_size
    return (n + 63) // 64


def ptr(t):
    """device pointer of a torch tensor (or None)"""
    return None if t is None else C.c_void_p(t.data_ptr())
