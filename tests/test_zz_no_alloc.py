"""oth_rollout_noalloc (the allocation-free tournament entry point, static device scratch) against oth_rollout.

Collected LAST on purpose.  The entry point has no counterpart in the
reference; its contract is "the same tally as oth_rollout", so the existing entry point is the checker here
(integer counters: bit-exact).  Everything goes through the C ABI with ctypes."""
import ctypes as C

import pytest
import torch

from gymothelloenv_b200 import _cabi

RAND = (_cabi.POLICY_RANDOM, 1)
GREEDY = (_cabi.POLICY_GREEDY, 1)
MAXIMIN2 = (_cabi.POLICY_MAXIMIN, 2)


def _spec(p):
    return _cabi.OthPolicySpec(p[0], p[1])


def test_noalloc_rejects_more_games_than_its_scratch_holds():
    """Argument checks come before any CUDA call, so this runs without a GPU."""
    lib = _cabi.load()
    too_many = _cabi.NOALLOC_MAX_GAMES + 1
    rc = lib.oth_rollout_noalloc(too_many, 8, 0, _spec(RAND), _spec(RAND), 0, 1, 0, 0, C.c_void_p(16), None)
    assert rc == 1001                                            # OTH_ERR_BAD_ARG
    msg = lib.oth_last_error().decode()
    assert "exceed" in msg and str(_cabi.NOALLOC_MAX_GAMES) in msg
    rc = lib.oth_rollout_noalloc(1 << 12, 8, 0, _spec(RAND), _spec(RAND), 0, 1 << 12, 0, 0, C.c_void_p(16), None)
    assert rc == 1001                                            # B * games_per_env counts, not B alone
    assert lib.oth_rollout_noalloc(64, 8, 0, _spec(RAND), _spec(RAND), 0, 1, 0, 0, None, None) == 1003   # null tally
    assert lib.oth_rollout_noalloc(64, 3, 0, _spec(RAND), _spec(RAND), 0, 1, 0, 0, C.c_void_p(16), None) == 1002


@pytest.mark.gpu
def test_noalloc_tally_equals_oth_rollout_on_two_calls_in_a_row():
    assert torch.cuda.is_available(), "GPU test on a box without CUDA"
    lib = _cabi.load()
    dev = torch.device("cuda", 0)
    with torch.cuda.device(dev):
        stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        ref = torch.zeros((4,), dtype=torch.int64, device=dev)
        got = torch.zeros((4,), dtype=torch.int64, device=dev)
        # two calls in a row share the static scratch; the second is SMALLER and uses another board size, so any
        # record left over from the first call would be counted if the reduction read past its own range
        calls = [
            dict(B=40000 + 17, S=8, black=RAND, white=GREEDY, init=10, gpe=2, seed=5, base=0),
            dict(B=3000 + 5, S=6, black=GREEDY, white=RAND, init=4, gpe=1, seed=9, base=1 << 33),
            dict(B=2048, S=8, black=GREEDY, white=MAXIMIN2, init=10, gpe=1, seed=1, base=77),      # cooperative kernel
            dict(B=512, S=10, black=RAND, white=RAND, init=0, gpe=3, seed=2, base=0),               # multi-word boards
        ]
        for c in calls:
            ref.zero_()
            got.zero_()
            _cabi.check(lib.oth_rollout(c["B"], c["S"], 0, _spec(c["black"]), _spec(c["white"]), c["init"], c["gpe"],
                                        c["seed"], c["base"], _cabi.ptr(ref), None, None, None, None, 0, stream),
                        "oth_rollout")
            _cabi.check(lib.oth_rollout_noalloc(c["B"], c["S"], 0, _spec(c["black"]), _spec(c["white"]), c["init"],
                                                c["gpe"], c["seed"], c["base"], _cabi.ptr(got), stream),
                        "oth_rollout_noalloc")
            torch.cuda.synchronize(dev)
            r, g = ref.tolist(), got.tolist()
            assert sum(r[:3]) == c["B"] * c["gpe"] and r[3] > 0
            assert g == r, (c, r, g)
        # like oth_rollout the tally ACCUMULATES: the same call twice doubles it
        c = calls[1]
        for _ in range(2):
            _cabi.check(lib.oth_rollout_noalloc(c["B"], c["S"], 0, _spec(c["black"]), _spec(c["white"]), c["init"],
                                                c["gpe"], c["seed"], c["base"], _cabi.ptr(got), stream),
                        "oth_rollout_noalloc")
        _cabi.check(lib.oth_rollout(c["B"], c["S"], 0, _spec(c["black"]), _spec(c["white"]), c["init"], c["gpe"],
                                    c["seed"], c["base"], _cabi.ptr(ref), None, None, None, None, 0, stream),
                    "oth_rollout")
        torch.cuda.synchronize(dev)
        # got = last loop value (calls[3]) + 2 x calls[1]; ref = calls[3] + 1 x calls[1]  ->  compare the increments
        base3 = torch.zeros((4,), dtype=torch.int64, device=dev)
        c3 = calls[3]
        _cabi.check(lib.oth_rollout(c3["B"], c3["S"], 0, _spec(c3["black"]), _spec(c3["white"]), c3["init"], c3["gpe"],
                                    c3["seed"], c3["base"], _cabi.ptr(base3), None, None, None, None, 0, stream),
                    "oth_rollout")
        torch.cuda.synchronize(dev)
        inc_ref = (ref - base3).tolist()
        inc_got = (got - base3).tolist()
        assert inc_got == [2 * x for x in inc_ref], (inc_ref, inc_got)
