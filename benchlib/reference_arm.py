"""bench.py --impl reference: the reference's own CPU implementation of the path (oracle/_ref, built by
oracle/make_ref.py from /root/reference; falls back to the port oracle/rules_py.py when that build is absent, and
says which) on all host cores, on the config / metric / unit of our arm.  No GPU, no product code on this path."""
import os
import time

from . import common, headline
from .configs2 import PAIRINGS


def run(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0                                   # under torchrun rank 0 alone runs the CPU arm
    import numpy  # noqa: F401  (imported in the parent, before the fork)
    from oracle import cpu_arms
    t_all = time.perf_counter()
    arms = cpu_arms.CpuArms()
    per = max(1.0, min(60.0, args.cpu_seconds))    # default 12 s per worker and sample (>= 10 s)
    K = max(1, min(args.steps, 5, int(120 // per)))   # bounded: a "step" = one sample; the arm ends within minutes
    cfg = args.config
    samples, extra = [], {}
    for _ in range(min(1, args.warmup)):           # one short untimed pass
        arms.playout(8, 1.0)
    if cfg == "headline":
        metric, unit, workload = headline.METRIC, headline.UNIT, headline.WORKLOAD
        for _ in range(K):
            samples.append(arms.playout(8, per))
    elif cfg == "ppo_feed":
        metric, unit = "8x8 env steps/sec (PPO feed: planes4 float32 obs + mask)", "env_steps/s"
        workload = "PPO self-play rollout feed, 65,536 on-GPU envs, torch obs tensors, configs[3]"
        for _ in range(K):
            samples.append(arms.playout(8, per, with_state=True))
    elif cfg == "sizes":
        S = args.board_size if args.board_size in (6, 10, 12) else 10
        metric, unit = "%dx%d env steps/sec (lockstep batch)" % (S, S), "env_steps/s"
        workload = "variable board_size 6/10/12 random-policy batch, 2^20 envs per GPU, configs[4]"
        for _ in range(K):
            samples.append(arms.playout(S, per))
        extra["sizes"] = {str(s): arms.playout(s, max(5.0, per / 2)) for s in (6, 10, 12) if s != S}
    elif cfg == "tournament":
        name = args.pairing if args.pairing in PAIRINGS else "greedy_vs_maximin2"
        bk, wh = PAIRINGS[name]
        metric, unit = "8x8 tournament games/sec (%s, fused rollout)" % name, "games/s"
        workload = ("8x8 greedy vs maximin-2/3 tournament, 2^21 games per GPU sharded by global game id, "
                    "first <= 10 plies random, configs[2] (+ rand vs greedy = configs[0] protocol)")
        for _ in range(K):
            samples.append(arms.two_agent(8, bk, wh, 10, per if "maximin3" not in name else 1.0, min_games=1))
    else:                                          # scalar: configs[0], 100 games rand vs greedy
        metric = "8x8 rand vs greedy games/sec through the scalar drop-in API (run.py protocol)"
        unit = "games/s"
        workload = ("8x8 random vs greedy, 100 games, first 10 moves random (run.py --no-render), configs[0], "
                    "ONE env stepped through the Gym API")
        # like for like with the scalar API: ONE process, 100 games
        one = cpu_arms.CpuArms(procs=1)
        for _ in range(K):
            samples.append(one.two_agent(8, ("rand", 1), ("greedy", 1), 10, 0.0, min_games=100))
        one.close()
        extra["all_cores"] = arms.two_agent(8, ("rand", 1), ("greedy", 1), 10, per)
    arms.close()
    samples = [s for s in samples if s]
    if not samples:
        common.emit({"impl": "reference", "unavailable": "oracle/_ref missing: this config needs the reference's policies"})
        return 0
    vals = sorted(s["value"] for s in samples)
    value = vals[len(vals) // 2]
    base = dict(samples[0])
    base["value"] = value
    base["sample"] = "median of %d samples; each: %s" % (len(samples), samples[0]["sample"])
    line = {
        "impl": "reference", "metric": metric, "value": value, "unit": unit, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": None, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "int64", "data": "synthetic", "config": {"workload": workload},
        "cpu_baseline": base, "samples": [s["value"] for s in samples],
        "e2e": {"value": value, "unit": unit, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "wall_s": time.perf_counter() - t_all,
    }
    line.update(extra)
    common.emit(line)
    return 0
