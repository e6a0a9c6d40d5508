"""bench.py --config tournament | scalar (BASELINE.json configs[2] / configs[0])."""
import time

from . import common
from .configs import _cpu

PAIRINGS = {            # --pairing name -> (black, white) as (kind, depth)
    "greedy_vs_maximin2": (("greedy", 1), ("maximin", 2)),
    "maximin2_vs_greedy": (("maximin", 2), ("greedy", 1)),
    "greedy_vs_maximin3": (("greedy", 1), ("maximin", 3)),
    "maximin3_vs_greedy": (("maximin", 3), ("greedy", 1)),
    "rand_vs_greedy": (("rand", 1), ("greedy", 1)),
}


def _label(p):
    return p[0] if p[0] != "maximin" else "maximin%d" % p[1]


# ------------------------------------------------------------------------------------ configs[2]: sharded tournament
def run_tournament(args):
    """greedy vs maximin-2 / maximin-3 (both colour assignments) + rand vs greedy, README protocol (<= 10 random
    opening plies, fixed colours), 2^21 games per GPU in ONE fused launch per pairing, W/D/L all-reduced (NCCL sum of
    4 words).  A "step" = one whole tournament launch of the --pairing (default greedy_vs_maximin2); the other
    pairings are measured once each and reported under `pairings`.  value = games/s over all ranks."""
    import torch
    from gymothelloenv_b200.tournament import rollout
    ranks = common.Ranks(torch)
    dev = ranks.dev
    games = (1 << 21) if args.num_envs == (1 << 20) else args.num_envs
    K, W = max(1, min(args.steps, 50)), max(1, min(args.warmup, 5))
    first = ranks.rank * games
    int_peak, sms = common.int32_peak(torch, ranks)
    counts = common.int_counts() or {}
    nodes = {}
    try:
        import json
        import os
        nodes = json.load(open(os.path.join(common.ROOT, "profiles", "search_nodes.json")))
    except (OSError, ValueError):
        pass

    def launch(black, white, tally):
        rollout(games, _label(black), _label(white), board_size=8, initial_rand_steps=10, seed=0, env_id_base=first,
                device=dev, tally=tally)

    def measure(black, white, reps, warm):
        tally = torch.zeros((4,), dtype=torch.int64, device=dev)
        for _ in range(warm):
            launch(black, white, tally)
        runs = []
        for _ in range(reps):
            tally.zero_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            ranks.barrier()
            a.record()
            launch(black, white, tally)
            b.record()
            ranks.barrier()
            runs.append(ranks.max_over_ranks(a.elapsed_time(b)))
        tot = ranks.sum_over_ranks(tally.tolist())                 # the tournament's only collective
        ms = sorted(runs)[len(runs) // 2]
        n = ranks.world * games
        return {"black": _label(black), "white": _label(white), "games": n, "ms_per_tournament": ms,
                "games_per_s": n / (ms * 1e-3), "outer_plies_per_s": tot[3] / (ms * 1e-3),
                "black_wdl_pct": [round(100.0 * x / n, 3) for x in tot[:3]], "plies_per_game": tot[3] / n,
                "timed_runs_ms": runs}

    sampler = common.ClockSampler(ranks.local).start()
    sampler.wait_for(1, timeout=5.0)
    t_h0 = time.perf_counter()
    main_name = args.pairing if args.pairing in PAIRINGS else "greedy_vs_maximin2"
    rows = {}
    for name, (bk, wh) in PAIRINGS.items():
        reps, warm = (K, W) if name == main_name else (1, 1)
        if "maximin3" in name and name != main_name:
            reps, warm = 1, 0
        rows[name] = measure(bk, wh, reps, warm)
        key = "k_rollout_" + name
        if counts.get(key):
            c = counts[key]
            per_gpu_plies = rows[name]["outer_plies_per_s"] / ranks.world
            rows[name].update(common.int_fraction({"int_thread_inst_per_unit": c["int_thread_inst_per_unit"],
                                                   "source_digest": c.get("source_digest"), "source": c.get("source")},
                                                  per_gpu_plies, int_peak))
        sn = nodes.get(name)
        if sn:                                                      # profiles/search_nodes.json (counted on the GPU)
            rows[name]["search_nodes_per_ply"] = sn["nodes_per_ply"]
            rows[name]["search_nodes_per_s"] = rows[name]["outer_plies_per_s"] * sn["nodes_per_ply"]
            if rows[name].get("int_thread_inst_per_unit"):
                rows[name]["int_thread_inst_per_search_node"] = rows[name]["int_thread_inst_per_unit"] / sn["nodes_per_ply"]
    clocks = sampler.stop(t_h0 + 0.05, time.perf_counter())
    r = rows[main_name]

    # e2e: host call -> fused tournament -> D2H of the tally (32 bytes); nothing else ever crosses PCIe
    tally = torch.zeros((4,), dtype=torch.int64, device=dev)
    h_tally = torch.empty((4,), dtype=torch.int64, pin_memory=True)
    bk, wh = PAIRINGS[main_name]
    ranks.barrier()
    t0 = time.perf_counter()
    tally.zero_()
    launch(bk, wh, tally)
    h_tally.copy_(tally, non_blocking=True)
    torch.cuda.current_stream(dev).synchronize()
    ranks.barrier()
    dt = ranks.max_over_ranks(time.perf_counter() - t0)
    e2e = {"value": ranks.world * games / dt, "unit": "games/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 32, "steps": 1}

    def cpu_fn(arms):
        secs = max(4.0, args.cpu_seconds / 2.0)
        out = {"rand_vs_greedy": arms.two_agent(8, ("rand", 1), ("greedy", 1), 10, secs),
               "greedy_vs_maximin2": arms.two_agent(8, ("greedy", 1), ("maximin", 2), 10, secs)}
        if args.cpu_maximin3:
            out["greedy_vs_maximin3"] = arms.two_agent(8, ("greedy", 1), ("maximin", 3), 10, 1.0, min_games=1)
        return out
    cpu = _cpu(args, ranks, cpu_fn)
    if ranks.rank == 0:
        roof = {"bound": "int32", "achieved": None, "peak": int_peak / 1e12, "unit": "Tinst/s", "frac": r.get("int_frac"),
                "traffic": 0, "kernel": "k_rollout_coop<8>" if "maximin" in main_name else "k_rollout<8>",
                "note": "fused rollouts keep boards in registers: ~0 HBM bytes per ply (4 tally words per warp), so the "
                        "binding ceiling is the INT32 pipe; frac = counted integer thread-instructions per ply (ncu, "
                        "profiles/int_count.json) x plies/s / the INT32 stream peak measured in this run",
                "int32_peak_measured_thread_inst_per_s": int_peak}
        if r.get("int_thread_inst_per_unit"):
            roof["achieved"] = r["int_thread_inst_per_unit"] * r["outer_plies_per_s"] / ranks.world / 1e12
            roof.update({k: r[k] for k in ("int_thread_inst_per_unit", "int_count_stale", "int_count_source") if k in r})
        common.emit({
            "metric": "8x8 tournament games/sec (%s, fused rollout)" % main_name, "value": r["games_per_s"],
            "unit": "games/s", "n_gpus": ranks.world, "steps": K, "warmup": W, "ms_per_step": r["ms_per_tournament"],
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
            "config": {"workload": "8x8 greedy vs maximin-2/3 tournament, 2^21 games per GPU sharded by global game id, "
                                   "first <= 10 plies random, configs[2] (+ rand vs greedy = configs[0] protocol)",
                       "games_per_gpu": games, "pairing": main_name, "l2": "no HBM working set: games live in registers"},
            "pairings": rows, "roofline": roof,
            "cpu_baseline": (cpu or {}).get(main_name if main_name in (cpu or {}) else "greedy_vs_maximin2"),
            "cpu_baseline_pairings": cpu, "e2e": e2e, "gpu_launches": K, "clocks": clocks,
        })
    ranks.finish()
    return 0


# ----------------------------------------------------------------------------------------- configs[0]: scalar drop-in
def run_scalar(args):
    """run.py --no-render protocol through the SCALAR drop-in API (gymothelloenv_b200.othello.SimpleOthelloEnv +
    simple_policies, one env, one kernel launch per call): rand(black) vs greedy(white), 100 games, first 10 moves
    random, the run_2agent.py:119-167 loop -- next to the reference's own wall time for the same loop, and next to
    the batched answer to the same question (one fused rollout launch)."""
    import numpy as np
    import torch
    from gymothelloenv_b200 import othello as oth, simple_policies as sp, util as u
    from gymothelloenv_b200.tournament import rollout
    ranks = common.Ranks(torch)
    games = 100

    def play(n):
        env = oth.SimpleOthelloEnv(board_size=8, seed=0, initial_rand_steps=10, device=ranks.dev)
        pol = {-1: sp.RandomPolicy(seed=0), 1: sp.GreedyPolicy()}
        tally, plies = [0, 0, 0], 0
        for _ in range(n):
            obs = env.reset()
            for p in pol.values():
                p.reset(env)
            done = False
            while not done:
                turn = int(env.player_turn)
                state = u.make_state(np.array(obs, copy=True), env)
                obs, reward, done, _ = env.step(int(pol[turn].get_action(state)))
                plies += 1
            w = int(env.env.winner)
            tally[0 if w < 0 else (1 if w == 0 else 2)] += 1
        return tally, plies
    play(3)                                                      # warm-up: module load, allocations
    torch.cuda.synchronize(ranks.dev)
    t0 = time.perf_counter()
    tally, plies = play(games)
    dt = time.perf_counter() - t0
    # the batched form of the same protocol: 2^20 games in one launch
    rollout(4096, "rand", "greedy", seed=1, device=ranks.dev)
    torch.cuda.synchronize(ranks.dev)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    t, _ = rollout(1 << 20, "rand", "greedy", board_size=8, initial_rand_steps=10, seed=0, device=ranks.dev)
    b.record()
    torch.cuda.synchronize(ranks.dev)
    ms = a.elapsed_time(b)
    tl = [int(x) for x in t.tolist()]
    cpu = _cpu(args, ranks, lambda arms: arms.two_agent(8, ("rand", 1), ("greedy", 1), 10, 0.0,
                                                        min_games=max(1, games // arms.procs)))
    if ranks.rank == 0:
        common.emit({
            "metric": "8x8 rand vs greedy games/sec through the scalar drop-in API (run.py protocol)", "value": games / dt,
            "unit": "games/s", "n_gpus": 1, "steps": games, "warmup": 3, "ms_per_step": dt / games * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
            "config": {"workload": "8x8 random vs greedy, 100 games, first 10 moves random (run.py --no-render), "
                                   "configs[0], ONE env stepped through the Gym API",
                       "note": "a B=1 env is a launch-latency workload (one or two launches that write pinned host memory + one stream sync per call); "
                               "the batched API is the product, this line reports the drop-in honestly"},
            "scalar_api": {"games": games, "seconds": dt, "plies": plies, "plies_per_s": plies / dt, "black_wdl": tally},
            "batched_same_protocol": {"games": 1 << 20, "ms": ms, "games_per_s": (1 << 20) / (ms * 1e-3),
                                      "black_wdl_pct": [round(100.0 * x / (1 << 20), 2) for x in tl[:3]]},
            "roofline": {"bound": "latency", "achieved": None, "peak": None, "unit": None, "frac": None, "traffic": None,
                         "note": "one env per launch: bounded by launch + D2H sync latency, not by a throughput roofline"},
            "cpu_baseline": cpu,
            "e2e": {"value": games / dt, "unit": "games/s", "h2d_bytes_per_step": 8, "d2h_bytes_per_step": 8 * 8 * 2 + 16 + 256 + 8,
                    "note": "the scalar API IS end to end: every step() reads its action from and writes obs / mask / record / planes to pinned host memory"},
            "gpu_launches": None, "clocks": None,
        })
    ranks.finish()
    return 0
