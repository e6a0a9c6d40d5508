"""W/D/L tournaments between scripted policies, fused on the GPU.

Replaces the reference's 100-game tally loops (``run.py:74-133``, ``run_2agent.py:92-203``)
and the evaluation games of the PPO workers (``rule_base_game``,
``ppo_run_self_play.py:371-441``): every game is played start to finish inside ONE kernel
(``oth_rollout``), boards in registers, nothing but four counters leaves the SMs.

Multi-GPU: games are split in contiguous blocks over the ranks of the
default ``torch.distributed`` group; the Philox stream of a game is keyed by its GLOBAL
env id so the set of games played is the same for any world size; the only collective
is one all-reduce (sum) of the 4-word tally -- the buffer the kernel accumulates into.
"""
import ctypes as C

import torch

from . import _cabi
from .vec_env import policy_spec, require_cuda, shard_range


def parse_policy(spec):
    """'rand' | 'greedy' | 'maximin2' | ('maximin', 3) -> (kind, depth)  (util.py:28-45 create_policy)"""
    if isinstance(spec, (tuple, list)):
        return str(spec[0]), int(spec[1]) if len(spec) > 1 else 1
    s = str(spec)
    for name in ("maximin", "alphabeta", "rand", "random", "greedy"):
        if s.startswith(name):
            tail = s[len(name):].lstrip("-_")
            return name, int(tail) if tail else 1
    raise ValueError("unknown policy %r" % (spec,))


def rollout(num_envs, black, white, board_size=8, initial_rand_steps=10, games_per_env=1, seed=0, env_id_base=0,
            device=None, num_disk_as_reward=False, record=False, trace_len=0, tally=None, search_small_lists=False,
            search_per_thread=False):
    """Launches oth_rollout for `num_envs` threads x `games_per_env` games on this device.

    Returns (tally int64[4] CUDA tensor = black wins, draws, white wins, plies, records)
    where records is None or a dict of per-game CUDA tensors (result, plies, counts, trace)."""
    dev = require_cuda(device)
    lib = _cabi.load()
    bk, bd = parse_policy(black)
    wk, wd = parse_policy(white)
    if tally is None:
        tally = torch.zeros((4,), dtype=torch.int64, device=dev)
    games = int(num_envs) * int(games_per_env)
    rec = None
    ptrs = [None, None, None, None]
    stride = 0
    if record:
        rec = {"result": torch.zeros((games,), dtype=torch.int8, device=dev),
               "plies": torch.zeros((games,), dtype=torch.int16, device=dev),
               "counts": torch.zeros((games, 2), dtype=torch.int16, device=dev)}
        ptrs[:3] = [_cabi.ptr(rec["result"]), _cabi.ptr(rec["plies"]), _cabi.ptr(rec["counts"])]
        if trace_len:
            stride = int(trace_len)
            rec["trace"] = torch.full((games, stride), -1, dtype=torch.int16, device=dev)
            ptrs[3] = _cabi.ptr(rec["trace"])
    flags = _cabi.F_SUDDEN_DEATH | (_cabi.F_DISK_REWARD if num_disk_as_reward else 0)
    if search_small_lists:                        # tests: force the cooperative search onto its per-thread fallback
        flags |= _cabi.F_SEARCH_SMALL_LISTS
    if search_per_thread:                         # A/B: per-thread DFS instead of the lane-cooperative search
        flags |= _cabi.F_SEARCH_PER_THREAD
    with torch.cuda.device(dev):
        _cabi.check(lib.oth_rollout(int(num_envs), int(board_size), flags, policy_spec(bk, bd), policy_spec(wk, wd),
                                    int(initial_rand_steps), int(games_per_env), int(seed) & 0xFFFFFFFFFFFFFFFF,
                                    int(env_id_base), _cabi.ptr(tally), ptrs[0], ptrs[1], ptrs[2], ptrs[3], stride,
                                    C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)), "oth_rollout")
    return tally, rec


def sharded_tally(num_units, local_fn, tally, group=None):
    """Splits `num_units` (envs) in contiguous blocks over the ranks of `group`, calls
    local_fn(first_global_id, count, tally) on this rank's block (it accumulates into the
    4-word `tally` tensor) and sums the tallies with ONE all-reduce.  Host-side plumbing
    only; also exercised on CPU/gloo by tests/test_host_logic.py."""
    import torch.distributed as dist
    rank, world = 0, 1
    if dist.is_available() and dist.is_initialized():
        rank, world = dist.get_rank(group), dist.get_world_size(group)
    first, count = shard_range(num_units, rank, world)
    if count > 0:
        local_fn(first, count, tally)
    if world > 1:
        dist.all_reduce(tally, op=dist.ReduceOp.SUM, group=group)
    return tally, world


def play_tournament(black, white, num_games=100, board_size=8, initial_rand_steps=10, seed=0, device=None,
                    games_per_env=1, group=None):
    """README protocol (README.md:36-50): `black` (row) vs `white` (column), fixed colours,
    first <= initial_rand_steps plies random.  Sharded over torch.distributed ranks when a
    process group is initialised.  Returns a dict with black's wins/draws/losses, plies."""
    num_envs_total = (int(num_games) + games_per_env - 1) // games_per_env
    dev = require_cuda(device)
    tally = torch.zeros((4,), dtype=torch.int64, device=dev)

    def local(first, count, t):
        rollout(count, black, white, board_size=board_size, initial_rand_steps=initial_rand_steps,
                games_per_env=games_per_env, seed=seed, env_id_base=first, device=dev, tally=t)

    tally, world = sharded_tally(num_envs_total, local, tally, group)
    w, d, l, plies = [int(x) for x in tally.tolist()]
    return {"wins": w, "draws": d, "losses": l, "games": w + d + l, "plies": plies,
            "black": str(black), "white": str(white), "world_size": world}
