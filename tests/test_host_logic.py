"""CPU-only checks: the C ABI library loads and exports every declared symbol, host-side
helpers (sharding, policy parsing, argument validation), the numpy Philox used by the GPU
tests, and the multi-rank tally reduction over gloo (world_size 2)."""
import ctypes
import os
import re
import socket

import numpy as np
import pytest
import torch

from tests import philox_ref

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _lib_path():
    from gymothelloenv_b200 import _cabi, build
    if not os.path.isfile(_cabi.LIB_PATH):
        build.build()
    return _cabi.LIB_PATH


def test_library_exports_every_declared_symbol():
    from gymothelloenv_b200 import _cabi
    lib = ctypes.CDLL(_lib_path())
    header = open(os.path.join(ROOT, "include", "othello_b200.h")).read()
    declared = set(re.findall(r"^\s*(?:int|const char \*)\s*\*?\s*(oth_[a-z_0-9]+)\s*\(", header, flags=re.M))
    assert declared == set(_cabi.EXPORTED_SYMBOLS), declared ^ set(_cabi.EXPORTED_SYMBOLS)
    for name in declared:
        assert getattr(lib, name) is not None
    lib.oth_abi_version.restype = ctypes.c_int
    lib.oth_words_per_board.restype = ctypes.c_int
    assert lib.oth_abi_version() == 2                       # no GPU needed for these two
    assert [lib.oth_words_per_board(s) for s in (3, 4, 8, 9, 11, 12, 13, 14, 16, 17)] == [0, 1, 1, 2, 2, 3, 3, 4, 4, 0]


def test_header_constants_match_python_binding():
    from gymothelloenv_b200 import _cabi
    header = open(os.path.join(ROOT, "include", "othello_b200.h")).read()
    defs = {k: int(v.rstrip("u"), 0) for k, v in re.findall(r"#define\s+(OTH_[A-Z0-9_]+)\s+(0x[0-9a-fA-F]+u?|\d+)\b", header)}
    pairs = {"OTH_F_SUDDEN_DEATH": _cabi.F_SUDDEN_DEATH, "OTH_F_DISK_REWARD": _cabi.F_DISK_REWARD,
             "OTH_F_AUTO_RESET": _cabi.F_AUTO_RESET, "OTH_F_LEGACY_SINGLE_MOVE_PLANE": _cabi.F_LEGACY_SINGLE_MOVE_PLANE,
             "OTH_F_SEARCH_SMALL_LISTS": _cabi.F_SEARCH_SMALL_LISTS, "OTH_F_SEARCH_PER_THREAD": _cabi.F_SEARCH_PER_THREAD, "OTH_F_SP_FRESH_STATE": _cabi.F_SP_FRESH_STATE,
             "OTH_OBS_BOARD": _cabi.OBS_BOARD, "OTH_OBS_BOARD_MASK": _cabi.OBS_BOARD_MASK,
             "OTH_OBS_PLANES4": _cabi.OBS_PLANES4, "OTH_OBS_PLANES3": _cabi.OBS_PLANES3,
             "OTH_OBS_ABSOLUTE": _cabi.OBS_ABSOLUTE, "OTH_DTYPE_F32": _cabi.DTYPE_F32,
             "OTH_ACT_RANDOM": _cabi.ACT_RANDOM, "OTH_ACT_GREEDY": _cabi.ACT_GREEDY,
             "OTH_POLICY_MAXIMIN": _cabi.POLICY_MAXIMIN, "OTH_MAX_SEARCH_DEPTH": _cabi.MAX_SEARCH_DEPTH,
             "OTH_MAX_BOARD_SIZE": _cabi.MAX_BOARD_SIZE, "OTH_MIN_BOARD_SIZE": _cabi.MIN_BOARD_SIZE,
             "OTH_META_RANDCNT_SHIFT": _cabi.META_RANDCNT_SHIFT, "OTH_META_RANDMAX_SHIFT": _cabi.META_RANDMAX_SHIFT,
             "OTH_META_WINNER_SHIFT": _cabi.META_WINNER_SHIFT}
    for k, v in pairs.items():
        assert defs[k] == v, k
    # struct layouts: 5 pointers / 7 pointers + 2 int32 + the packed-record pointer
    assert ctypes.sizeof(_cabi.OthState) == 40 and ctypes.sizeof(_cabi.OthOutputs) == 72
    assert ctypes.sizeof(_cabi.OthPolicySpec) == 8


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_product_fails_loudly_without_cuda():
    """No CPU fallback: constructing any env without a CUDA device raises."""
    from gymothelloenv_b200 import OthelloB200Error, VecOthelloEnv, othello
    from gymothelloenv_b200.tournament import rollout
    with pytest.raises(OthelloB200Error):
        VecOthelloEnv(4)
    with pytest.raises(OthelloB200Error):
        othello.SimpleOthelloEnv()
    with pytest.raises(OthelloB200Error):
        rollout(4, "rand", "greedy")


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "gymothelloenv_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", text, flags=re.M), f
                assert "othello_oracle" not in text and "rules_py" not in text, f


def test_shard_range_partitions_exactly():
    from gymothelloenv_b200 import shard_range
    for total in (0, 1, 7, 100, 1 << 20, 16777216 + 3):
        for world in (1, 2, 3, 4, 8):
            blocks = [shard_range(total, r, world) for r in range(world)]
            assert blocks[0][0] == 0 and sum(c for _, c in blocks) == total
            for (f0, c0), (f1, _) in zip(blocks, blocks[1:]):
                assert f0 + c0 == f1
            assert max(c for _, c in blocks) - min(c for _, c in blocks) <= 1
    with pytest.raises(ValueError):
        shard_range(10, 2, 2)


def test_policy_parsing_and_validation():
    from gymothelloenv_b200.tournament import parse_policy
    from gymothelloenv_b200.vec_env import policy_spec
    assert parse_policy("rand") == ("rand", 1) and parse_policy("maximin3") == ("maximin", 3)
    assert parse_policy(("maximin", 2)) == ("maximin", 2) and parse_policy("maximin-2") == ("maximin", 2)
    assert parse_policy("greedy") == ("greedy", 1)
    assert parse_policy("alphabeta5") == ("alphabeta", 5) and parse_policy("alphabeta") == ("alphabeta", 1)
    with pytest.raises(ValueError):
        parse_policy("mcts4")
    with pytest.raises(ValueError):
        policy_spec("maximin", 9)
    with pytest.raises(ValueError):
        policy_spec("alphabeta", 0)
    assert (policy_spec("alphabeta", 6).kind, policy_spec("maximin", 8).depth) == (3, 8)
    with pytest.raises(ValueError):
        policy_spec("mcts")
    s = policy_spec("maximin", 3)
    assert (s.kind, s.depth) == (2, 3)


def test_numpy_philox_known_answers():
    """Random123 kat_vectors for philox4x32-10."""
    kat = [((0, 0, 0, 0), (0, 0), (0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8)),
           ((0xffffffff,) * 4, (0xffffffff,) * 2, (0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd)),
           ((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0),
            (0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1))]
    for ctr, key, want in kat:
        got = philox_ref.philox4x32_10(*ctr, *key)
        assert tuple(int(x) for x in got) == want
    r = philox_ref.draw(7, np.arange(1000), 0, np.zeros(1000, dtype=np.int64))
    idx = philox_ref.bounded(r, 6)
    assert idx.min() == 0 and idx.max() == 5


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _gloo_worker(rank, world, port, total, out):
    import torch.distributed as dist
    from gymothelloenv_b200.tournament import sharded_tally
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)

    def local(first, count, tally):              # stands in for the rollout kernel: a pure
        ids = np.arange(first, first + count)     # function of the GLOBAL game ids
        res = (ids * 2654435761 >> 7) % 3
        tally += torch.tensor([(res == 0).sum(), (res == 1).sum(), (res == 2).sum(), (ids % 61 + 40).sum()])

    tally, w = sharded_tally(total, local, torch.zeros(4, dtype=torch.int64))
    assert w == world
    if rank == 0:
        torch.save(tally, out)
    dist.barrier()
    dist.destroy_process_group()


def test_sharded_tally_over_gloo_world_size_2(tmp_path):
    """Same global id space => the reduced tally is independent of the world size."""
    import torch.multiprocessing as mp
    total = 10007
    out = str(tmp_path / "tally.pt")
    mp.spawn(_gloo_worker, args=(2, _free_port(), total, out), nprocs=2, join=True)
    got = torch.load(out)
    ids = np.arange(total)
    res = (ids * 2654435761 >> 7) % 3
    want = [(res == 0).sum(), (res == 1).sum(), (res == 2).sum(), (ids % 61 + 40).sum()]
    assert got.tolist() == [int(x) for x in want]


def test_build_digest_covers_every_native_source():
    """A source file outside the digest makes `build()` keep a stale library after an edit."""
    from gymothelloenv_b200 import build
    hashed = {os.path.abspath(p) for p in build._sources()}
    csrc = os.path.join(ROOT, "gymothelloenv_b200", "csrc")
    for n in os.listdir(csrc):
        assert os.path.abspath(os.path.join(csrc, n)) in hashed, n
    assert os.path.abspath(os.path.join(ROOT, "include", "othello_b200.h")) in hashed
