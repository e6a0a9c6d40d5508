"""CPU: the error behaviour of the C ABI (include/othello_b200.h, SURVEY 8(b) "Errors" row).

Every entry point checks its arguments BEFORE any CUDA call, returns an OTH_ERR_* code, never throws across the
boundary and leaves a message for oth_last_error() (thread-local).  These calls return from the argument checks, so
they need no GPU: the fake pointers (aligned non-zero integers) are never dereferenced.  Zero-sized batches are a
successful no-op.  The Python host turns a non-zero code into OthelloB200Error (gymothelloenv_b200/_cabi.check).
"""
import ctypes as C
import threading

import pytest

from gymothelloenv_b200 import _cabi

ERR_BAD_ARG, ERR_BAD_BOARD_SIZE, ERR_NULL_POINTER = 1001, 1002, 1003
FAKE = 0x10000          # aligned, non-null, never dereferenced: every call below fails (or no-ops) before a launch


@pytest.fixture(scope="module")
def lib():
    return _cabi.load()


def _state(null=None):
    f = {n: FAKE for n in ("black", "white", "legal", "meta", "ctr")}
    if null:
        f[null] = None
    return _cabi.OthState(**f)


def _outputs(**kw):
    f = dict(obs=FAKE, legal_mask=FAKE, reward=FAKE, done=FAKE, player_turn=FAKE, winner=FAKE, actions_out=FAKE,
             obs_layout=_cabi.OBS_BOARD, obs_dtype=_cabi.DTYPE_I8)
    f.update(kw)
    return _cabi.OthOutputs(**f)


def _spec(kind=_cabi.POLICY_GREEDY, depth=1):
    return _cabi.OthPolicySpec(kind, depth)


def _selfplay(null=None):
    f = {n: FAKE for n, _ in _cabi.OthSelfPlay._fields_}
    if null:
        f[null] = None
    return _cabi.OthSelfPlay(**f)


def _ts(null=None):
    f = {n: FAKE for n, _ in _cabi.OthTeacherStudent._fields_}
    if null:
        f[null] = None
    return _cabi.OthTeacherStudent(**f)


def _err(lib):
    return lib.oth_last_error().decode()


def test_header_error_codes_match():
    import os
    import re
    hdr = open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "include", "othello_b200.h")).read()
    codes = dict(re.findall(r"#define\s+(OTH_ERR_\w+)\s+(\d+)", hdr))
    assert int(codes["OTH_ERR_BAD_ARG"]) == ERR_BAD_ARG
    assert int(codes["OTH_ERR_BAD_BOARD_SIZE"]) == ERR_BAD_BOARD_SIZE
    assert int(codes["OTH_ERR_NULL_POINTER"]) == ERR_NULL_POINTER
    assert int(codes["OTH_ERR_CUDA"]) == 2000


@pytest.mark.parametrize("S", [-1, 0, 3, 17, 64])
def test_unsupported_board_size_everywhere(lib, S):
    """The C ABI does not clamp (the Python classes do, othello.py:230): sizes outside the build are an error."""
    assert lib.oth_words_per_board(S) == 0
    assert lib.oth_reset(_state(), 8, S, 0, 0, 0, 1, None, None) == ERR_BAD_BOARD_SIZE
    assert "board_size %d unsupported" % S in _err(lib) and "4..16" in _err(lib)
    assert lib.oth_observe(_state(), 8, S, 0, _outputs(), None) == ERR_BAD_BOARD_SIZE
    assert lib.oth_step(_state(), 8, S, 0, _cabi.ACT_RANDOM, None, 0, 0, 0, _outputs(), None) == ERR_BAD_BOARD_SIZE
    assert lib.oth_policy(_state(), 8, S, _spec(), 0, 0, 0, FAKE, None) == ERR_BAD_BOARD_SIZE
    assert lib.oth_rollout(8, S, 0, _spec(), _spec(), 0, 1, 0, 0, FAKE, None, None, None, None, 0, None) == ERR_BAD_BOARD_SIZE
    assert lib.oth_set_board(_state(), 8, S, FAKE, FAKE, FAKE, 1, 0, None) == ERR_BAD_BOARD_SIZE
    assert lib.oth_update_board(_state(), 8, S, FAKE, None) == ERR_BAD_BOARD_SIZE
    assert lib.oth_masked_categorical(FAKE, FAKE, 8, S, _cabi.CAT_SAMPLE, None, 0, 0, 0, FAKE, FAKE, None) == ERR_BAD_BOARD_SIZE
    assert lib.oth_epsilon_greedy(FAKE, FAKE, 8, S, 0.5, 0, 0, 0, FAKE, FAKE, FAKE, None) == ERR_BAD_BOARD_SIZE


def test_words_per_board_for_every_built_size(lib):
    for S in range(_cabi.MIN_BOARD_SIZE, _cabi.MAX_BOARD_SIZE + 1):
        assert lib.oth_words_per_board(S) == _cabi.words_per_board(S) == (S * S + 63) // 64


@pytest.mark.parametrize("field", ["black", "white", "legal", "meta", "ctr"])
def test_null_state_pointer(lib, field):
    assert lib.oth_reset(_state(field), 8, 8, 0, 0, 0, 1, None, None) == ERR_NULL_POINTER
    assert _err(lib) == "null state pointer"
    assert lib.oth_step(_state(field), 8, 8, 0, _cabi.ACT_RANDOM, None, 0, 0, 0, _outputs(), None) == ERR_NULL_POINTER
    assert lib.oth_policy(_state(field), 8, 8, _spec(), 0, 0, 0, FAKE, None) == ERR_NULL_POINTER


def test_num_envs_range(lib):
    assert lib.oth_reset(_state(), -1, 8, 0, 0, 0, 1, None, None) == ERR_BAD_ARG
    assert _err(lib) == "num_envs out of range"
    assert lib.oth_step(_state(), (1 << 40) + 1, 8, 0, _cabi.ACT_RANDOM, None, 0, 0, 0, _outputs(), None) == ERR_BAD_ARG
    assert lib.oth_rollout(-5, 8, 0, _spec(), _spec(), 0, 1, 0, 0, FAKE, None, None, None, None, 0, None) == ERR_BAD_ARG
    assert lib.oth_rollout(5, 8, 0, _spec(), _spec(), 0, -1, 0, 0, FAKE, None, None, None, None, 0, None) == ERR_BAD_ARG
    assert lib.oth_masked_categorical(FAKE, FAKE, -1, 8, _cabi.CAT_SAMPLE, None, 0, 0, 0, FAKE, FAKE, None) == ERR_BAD_ARG
    assert lib.oth_epsilon_greedy(FAKE, FAKE, -1, 8, 0.5, 0, 0, 0, FAKE, FAKE, FAKE, None) == ERR_BAD_ARG


def test_zero_sized_batches_are_a_successful_no_op(lib):
    """Empty input (the reference's Envs with zero workers would simply loop over nothing, envs.py:190)."""
    assert lib.oth_reset(_state(), 0, 8, 0, 0, 0, 1, None, None) == 0
    assert lib.oth_observe(_state(), 0, 8, 0, _outputs(), None) == 0
    assert lib.oth_step(_state(), 0, 8, 0, _cabi.ACT_RANDOM, None, 0, 0, 0, _outputs(), None) == 0
    assert lib.oth_policy(_state(), 0, 8, _spec(), 0, 0, 0, FAKE, None) == 0
    assert lib.oth_rollout(0, 8, 0, _spec(), _spec(), 0, 1, 0, 0, FAKE, None, None, None, None, 0, None) == 0
    assert lib.oth_rollout(64, 8, 0, _spec(), _spec(), 0, 0, 0, 0, FAKE, None, None, None, None, 0, None) == 0
    assert lib.oth_set_board(_state(), 0, 8, FAKE, FAKE, FAKE, 1, 0, None) == 0
    assert lib.oth_update_board(_state(), 0, 8, FAKE, None) == 0


def test_step_argument_checks(lib):
    st = _state()
    assert lib.oth_step(st, 8, 8, 0, 7, None, 0, 0, 0, _outputs(), None) == ERR_BAD_ARG
    assert _err(lib) == "bad action_source"
    assert lib.oth_step(st, 8, 8, 0, -1, None, 0, 0, 0, _outputs(), None) == ERR_BAD_ARG
    assert lib.oth_step(st, 8, 8, 0, _cabi.ACT_EXTERNAL, None, 0, 0, 0, _outputs(), None) == ERR_NULL_POINTER
    assert _err(lib) == "actions is null"
    for bad in (-1, 255, 1000):                        # the opening budget lives in 8 bits of meta
        assert lib.oth_step(st, 8, 8, 0, _cabi.ACT_RANDOM, None, 0, 0, bad, _outputs(), None) == ERR_BAD_ARG
        assert _err(lib) == "init_rand_steps must be in 0..254"
        assert lib.oth_reset(st, 8, 8, 0, 0, bad, 1, None, None) == ERR_BAD_ARG


def test_output_descriptor_checks(lib):
    st = _state()
    for layout in (-1, 6, 99):
        assert lib.oth_observe(st, 8, 8, 0, _outputs(obs_layout=layout), None) == ERR_BAD_ARG
        assert _err(lib) == "bad obs_layout"
    assert lib.oth_observe(st, 8, 8, 0, _outputs(obs_dtype=2), None) == ERR_BAD_ARG
    assert _err(lib) == "bad obs_dtype"
    # the dense writers store 16-byte vectors: a misaligned base is refused, not silently mis-stored
    assert lib.oth_observe(st, 8, 8, 0, _outputs(obs=FAKE + 4), None) == ERR_BAD_ARG
    assert _err(lib) == "obs must be 16-byte aligned"
    assert lib.oth_step(st, 8, 8, 0, _cabi.ACT_RANDOM, None, 0, 0, 0, _outputs(legal_mask=FAKE + 8), None) == ERR_BAD_ARG
    assert _err(lib) == "legal_mask must be 16-byte aligned"
    # layout / dtype are ignored when there is no obs buffer
    assert lib.oth_observe(st, 0, 8, 0, _outputs(obs=None, obs_layout=99, obs_dtype=99), None) == 0


def test_policy_spec_checks(lib):
    st = _state()
    assert lib.oth_policy(st, 8, 8, _spec(kind=4), 0, 0, 0, FAKE, None) == ERR_BAD_ARG
    assert _err(lib) == "bad policy kind"
    for depth in (0, -2, _cabi.MAX_SEARCH_DEPTH + 1):
        assert lib.oth_policy(st, 8, 8, _spec(_cabi.POLICY_MAXIMIN, depth), 0, 0, 0, FAKE, None) == ERR_BAD_ARG
        assert _err(lib) == "search depth must be in 1..8"
        assert lib.oth_rollout(8, 8, 0, _spec(), _spec(_cabi.POLICY_MAXIMIN, depth), 0, 1, 0, 0, FAKE, None, None, None,
                               None, 0, None) == ERR_BAD_ARG
    assert lib.oth_policy(st, 8, 8, _spec(), 0, 0, 0, None, None) == ERR_NULL_POINTER
    assert _err(lib) == "actions_out is null"
    assert lib.oth_rollout(8, 8, 0, _spec(), _spec(), 0, 1, 0, 0, FAKE, None, None, None, FAKE, 0, None) == ERR_BAD_ARG
    assert _err(lib) == "trace_stride must be positive"
    assert lib.oth_update_board(st, 8, 8, None, None) == ERR_NULL_POINTER
    # oth_search: depth range, flags, null output
    for depth in (0, _cabi.MAX_SEARCH_DEPTH + 1):
        assert lib.oth_search(st, 8, 8, depth, 0, FAKE, None, None) == ERR_BAD_ARG
        assert _err(lib) == "search depth must be in 1..8"
    assert lib.oth_search(st, 8, 8, 3, 4, FAKE, None, None) == ERR_BAD_ARG
    assert _err(lib) == "bad search flags"
    assert lib.oth_search(st, 8, 8, 3, _cabi.SEARCH_PRUNE, None, None, None) == ERR_NULL_POINTER
    assert lib.oth_search(st, 0, 8, 3, _cabi.SEARCH_PRUNE, FAKE, None, None) == 0
    # the differentiable evaluate_actions pair
    assert lib.oth_masked_logprob_forward(None, FAKE, 8, 8, FAKE, FAKE, FAKE, None) == ERR_NULL_POINTER
    assert lib.oth_masked_logprob_forward(FAKE, FAKE, 8, 8, FAKE, None, None, None) == ERR_NULL_POINTER
    assert _err(lib) == "no output requested"
    assert lib.oth_masked_logprob_backward(FAKE, FAKE, 8, 8, FAKE, FAKE, FAKE, None, None) == ERR_NULL_POINTER
    assert _err(lib) == "grad_logits is null"
    assert lib.oth_masked_logprob_backward(FAKE, FAKE, 0, 8, FAKE, FAKE, FAKE, FAKE, None) == 0


def test_sampler_argument_checks(lib):
    assert lib.oth_masked_categorical(None, FAKE, 8, 8, _cabi.CAT_SAMPLE, None, 0, 0, 0, FAKE, FAKE, None) == ERR_NULL_POINTER
    assert lib.oth_masked_categorical(FAKE, None, 8, 8, _cabi.CAT_SAMPLE, None, 0, 0, 0, FAKE, FAKE, None) == ERR_NULL_POINTER
    assert lib.oth_masked_categorical(FAKE, FAKE, 8, 8, 3, None, 0, 0, 0, FAKE, FAKE, None) == ERR_BAD_ARG
    assert _err(lib) == "bad mode"
    assert lib.oth_masked_categorical(FAKE, FAKE, 8, 8, _cabi.CAT_EVALUATE, None, 0, 0, 0, FAKE, FAKE, None) == ERR_NULL_POINTER
    assert _err(lib) == "actions_in is null"
    assert lib.oth_epsilon_greedy(None, FAKE, 8, 8, 0.5, 0, 0, 0, FAKE, FAKE, FAKE, None) == ERR_NULL_POINTER
    assert lib.oth_epsilon_greedy(FAKE, FAKE, 8, 8, 0.5, 0, 0, 0, None, FAKE, FAKE, None) == ERR_NULL_POINTER
    for eps in (-0.01, 1.5, float("nan")):
        assert lib.oth_epsilon_greedy(FAKE, FAKE, 8, 8, eps, 0, 0, 0, FAKE, FAKE, FAKE, None) == ERR_BAD_ARG
        assert _err(lib) == "epsilon must be in [0, 1]"


def test_selfplay_and_teacher_student_argument_checks(lib):
    st, out = _state(), _outputs(obs_layout=_cabi.OBS_PLANES4, obs_dtype=_cabi.DTYPE_F32)
    assert lib.oth_selfplay_new_game(st, _selfplay("phase"), 8, 8, 0, 0, 0, 0, None, None) == ERR_NULL_POINTER
    assert _err(lib) == "null self-play pointer"
    assert lib.oth_selfplay_begin(st, _selfplay(), 8, 8, 0, out, None, None) == ERR_NULL_POINTER
    assert _err(lib) == "need_count is null"
    assert lib.oth_selfplay_advance(st, _selfplay(), 8, 8, 0, None, None, None, None, 0, 0, 0, out, FAKE, None) == ERR_NULL_POINTER
    assert _err(lib) == "actions is null"
    assert lib.oth_selfplay_advance(st, _selfplay(), 8, 8, 0, FAKE, None, None, None, 0, 0, 300, out, FAKE, None) == ERR_BAD_ARG
    assert lib.oth_ts_new_game(st, _ts("tr_who"), 8, 8, 0, 0, 0, None) == ERR_NULL_POINTER
    assert _err(lib) == "null teacher/student pointer"
    assert lib.oth_ts_begin(st, _ts(), 8, 8, 0, 0, 0.0, out, FAKE, None) == ERR_BAD_ARG
    assert _err(lib) == "teacher must be -1 or +1"
    assert lib.oth_ts_begin(st, _ts(), 8, 8, 0, 1, 0.0, out, None, None) == ERR_NULL_POINTER
    assert lib.oth_ts_advance(st, _ts(), 8, 8, 0, 1, 0.0, 0, FAKE, None, None, None, 0, 0, 0, out, FAKE, None) == ERR_BAD_ARG
    assert _err(lib) == "colour must be -1 or +1"
    assert lib.oth_ts_advance(st, _ts(), 8, 8, 0, 1, 0.0, -1, None, None, None, None, 0, 0, 0, out, FAKE, None) == ERR_NULL_POINTER


def test_last_error_is_thread_local(lib):
    """SURVEY 8(b) Threading: no global mutable state -- a failure on one host thread does not overwrite the message
    another thread reads."""
    assert lib.oth_policy(_state(), 8, 8, _spec(kind=4), 0, 0, 0, FAKE, None) == ERR_BAD_ARG
    seen = {}

    def other():
        seen["before"] = _err(lib)
        lib.oth_reset(_state(), -1, 8, 0, 0, 0, 1, None, None)
        seen["after"] = _err(lib)
    t = threading.Thread(target=other)
    t.start()
    t.join()
    assert seen["before"] == "" and seen["after"] == "num_envs out of range"
    assert _err(lib) == "bad policy kind"


def test_python_host_raises_with_the_library_message(lib):
    rc = lib.oth_reset(_state(), 8, 17, 0, 0, 0, 1, None, None)
    with pytest.raises(_cabi.OthelloB200Error, match=r"oth_reset failed \(code 1002\): board_size 17 unsupported"):
        _cabi.check(rc, "oth_reset")
    _cabi.check(0, "oth_reset")
