"""bench.py --config sizes | ppo_feed | tournament | scalar: the BASELINE.json configs beside the headline, each with
the contract's JSON line (value, roofline, cpu_baseline, e2e) so the driver can re-measure them."""
import time

from . import common


def survey_bytes(S, obs_bytes_per_cell=1, planes=1):
    """SURVEY 8(d) 'general S': the 197 B contract with 64 -> S*S and 8 -> 8*ceil(S*S/64)."""
    N, W8 = S * S, 8 * ((S * S + 63) // 64)
    state = 2 * (2 * W8 + W8 + 4)                       # bitboards, cached legal set, flags: read + written
    return state + 8 + 5 + N + planes * N * obs_bytes_per_cell


def _stagger(env, plies):
    env.reset()
    for _ in range(plies):
        env.step(policy="random")


def _cpu(args, ranks, fn):
    if ranks.rank != 0 or ranks.world != 1 or args.no_cpu_baseline:
        return None
    from oracle import cpu_arms
    arms = cpu_arms.CpuArms()
    try:
        return fn(arms)
    finally:
        arms.close()


# ------------------------------------------------------------------------------------------ configs[4]: board sizes
def run_sizes(args):
    """Random-policy lockstep batch of 2^20 envs per GPU at board size 6 / 10 / 12 (multi-word bitboards above 8x8).
    `value` is the --board-size row (default 10); every size is measured and reported under `sizes`."""
    import torch
    from gymothelloenv_b200 import VecOthelloEnv
    ranks = common.Ranks(torch)
    B, K, W = args.num_envs, args.steps, max(3, args.warmup)
    peak, peak_src = common.hbm_peak()
    int_peak, _ = common.int32_peak(torch, ranks)
    counts = common.int_counts() or {}
    rows, clocks_all = {}, {}
    for S in (6, 10, 12):
        env = VecOthelloEnv(B, board_size=S, device=ranks.dev, seed=0, auto_reset=True, obs_layout="board",
                            obs_dtype=torch.int8, env_id_base=ranks.rank * B)
        _stagger(env, 2 * S)
        ms_run, runs, clocks = common.steady_state_runs(torch, ranks, lambda: env.step(policy="random"), K, W,
                                                        repeats=args.repeats)
        rate = B / (ms_run / K * 1e-3)
        bps = survey_bytes(S)
        row = {"board_size": S, "words_per_colour": env.words, "ms_per_step": ms_run / K, "env_steps_per_s": ranks.world * rate,
               "bytes_per_env_step": bps, "achieved_gbs": bps * rate / 1e9, "hbm_frac": bps * rate / 1e9 / peak,
               "timed_runs_ms": runs}
        row.update(common.int_fraction(counts.get("k_step_%d" % S), rate, int_peak))
        rows[S] = row
        clocks_all[S] = clocks
        del env
        torch.cuda.empty_cache()
    S = args.board_size if args.board_size in rows else 10
    r = rows[S]
    # e2e: the chosen size through host buffers
    env = VecOthelloEnv(B, board_size=S, device=ranks.dev, seed=0, auto_reset=True, obs_layout="board",
                        obs_dtype=torch.int8, env_id_base=ranks.rank * B)
    _stagger(env, 2 * S)
    Ke = max(3, args.e2e_steps)
    acts = torch.empty((Ke + 2, B), dtype=torch.int16, pin_memory=True)
    saved = env.get_state()
    for s in range(Ke + 2):
        env.step(policy="random")
        acts[s].copy_(env.actions_taken.to(torch.int16), non_blocking=True)
    torch.cuda.synchronize(ranks.dev)
    env.set_state(saved)
    for s in range(2):
        env.step_host(acts[s])
    ranks.barrier()
    t0 = time.perf_counter()
    for s in range(2, Ke + 2):
        env.step_host(acts[s])
    ranks.barrier()
    dt = ranks.max_over_ranks(time.perf_counter() - t0)
    h2d, d2h = env.host_bytes_per_step()
    e2e = {"value": ranks.world * B * Ke / dt, "unit": "env_steps/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
           "steps": Ke, "board_size": S}
    cpu = _cpu(args, ranks, lambda a: {str(s): a.playout(s, max(3.0, args.cpu_seconds / 3.0)) for s in (6, 10, 12)})
    if ranks.rank == 0:
        common.emit({
            "metric": "%dx%d env steps/sec (lockstep batch)" % (S, S), "value": r["env_steps_per_s"], "unit": "env_steps/s",
            "n_gpus": ranks.world, "steps": K, "warmup": W, "ms_per_step": r["ms_per_step"], "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
            "config": {"workload": "variable board_size 6/10/12 random-policy batch, 2^20 envs per GPU, configs[4]",
                       "envs_per_gpu": B, "value_is_board_size": S, "obs": "int8 (B,S,S) + bool mask (B,S*S)",
                       "l2": "working set per step > 126 MB L2 at every size",
                       "timing": "median of %d runs of K steps after >= 0.3 s of untimed steps" % args.repeats},
            "sizes": rows,
            "roofline": {"bound": "hbm", "achieved": r["achieved_gbs"], "peak": peak, "unit": "GB/s", "frac": r["hbm_frac"],
                         "traffic": None, "kernel": "k_step<%d,RANDOM,FAST>" % S, "bytes_per_env_step": r["bytes_per_env_step"],
                         "hbm_frac": r["hbm_frac"], "int_frac": r.get("int_frac"), "peak_source": peak_src,
                         "int32_peak_measured_thread_inst_per_s": int_peak},
            "cpu_baseline": (cpu or {}).get(str(S)), "cpu_baseline_sizes": cpu, "e2e": e2e, "gpu_launches": K,
            "clocks": clocks_all[S],
        })
    ranks.finish()
    return 0


# --------------------------------------------------------------------------------------------- configs[3]: PPO feed
def run_ppo_feed(args):
    """65,536 on-GPU envs feeding torch obs tensors: planes4 float32 (B,4,8,8) + bool mask + reward + done per step.
    value = env-only steps/s (actions already on the device); e2e = the rollout step a learner pays: a network of the
    reference's CNNBase shape -> oth_masked_categorical -> oth_step, with the per-step D2H of reward/done."""
    import torch
    from gymothelloenv_b200 import LegalMoveSampler, VecOthelloEnv
    ranks = common.Ranks(torch)
    B = 65536 if args.num_envs == (1 << 20) else args.num_envs
    K, W = args.steps, max(3, args.warmup)
    env = VecOthelloEnv(B, board_size=8, device=ranks.dev, seed=0, auto_reset=True, obs_layout="planes4",
                        obs_dtype=torch.float32, env_id_base=ranks.rank * B)
    _stagger(env, 20)
    # L2 flush between timed steps: the working set (B * 1.1 KB = 76 MB) fits the 126 MB L2
    flush = torch.empty(192 << 20, dtype=torch.uint8, device=ranks.dev)

    def step():
        flush.fill_(1)
        env.step(policy="random")
    ms_both, _, _ = common.steady_state_runs(torch, ranks, step, K, W, repeats=args.repeats)
    ms_flush, _, _ = common.steady_state_runs(torch, ranks, lambda: flush.fill_(1), K, W, repeats=args.repeats, spin_s=0.05)
    ms_run_nf, runs_nf, clocks = common.steady_state_runs(torch, ranks, lambda: env.step(policy="random"), K, W,
                                                          repeats=args.repeats)
    ms_step = max(1e-6, (ms_both - ms_flush) / K)                 # the env step alone, cold L2
    rate = B / (ms_step * 1e-3)
    bps = survey_bytes(8, obs_bytes_per_cell=4, planes=4)          # 1157
    peak, peak_src = common.hbm_peak()

    # rollout step with the caller's network (random-init, reference CNNBase shape, fp32 torch = the caller's code)
    import torch.nn as nn
    torch.manual_seed(0)
    body = nn.Sequential(nn.Conv2d(4, 32, 3, stride=2, padding=1), nn.ReLU(), nn.Conv2d(32, 64, 2), nn.ReLU(),
                         nn.Conv2d(64, 64, 2), nn.ReLU(), nn.Flatten())
    with torch.no_grad():
        feat = body(torch.zeros(1, 4, 8, 8)).shape[1]
    trunk = nn.Sequential(body, nn.Linear(feat, 512), nn.ReLU()).to(ranks.dev)
    actor, critic = nn.Linear(512, 64).to(ranks.dev), nn.Linear(512, 1).to(ranks.dev)
    sampler = LegalMoveSampler(env)
    h_rd = torch.empty((B, 5), dtype=torch.uint8, pin_memory=True)
    d_rd = torch.empty((B, 5), dtype=torch.uint8, device=ranks.dev)

    @torch.no_grad()
    def rollout_step():
        h = trunk(env.obs)
        a, lp = sampler.sample(actor(h))
        critic(h)
        env.step(a)
        d_rd.copy_(env._packed[:, :5])                              # reward f32 + done: what the rollout loop reads
        h_rd.copy_(d_rd, non_blocking=True)
        torch.cuda.current_stream(ranks.dev).synchronize()
    for _ in range(5):
        rollout_step()
    Ke = max(10, args.e2e_steps)
    ranks.barrier()
    t0 = time.perf_counter()
    for _ in range(Ke):
        rollout_step()
    ranks.barrier()
    dt = ranks.max_over_ranks(time.perf_counter() - t0)
    e2e = {"value": ranks.world * B * Ke / dt, "unit": "env_steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 5 * B,
           "steps": Ke, "what": "network forward (CNNBase shape, fp32 torch) + masked categorical + env step + D2H of "
                                "reward/done; observations stay on the GPU by design (north_star)"}
    cpu = _cpu(args, ranks, lambda a: a.playout(8, args.cpu_seconds, with_state=True))
    if ranks.rank == 0:
        common.emit({
            "metric": "8x8 env steps/sec (PPO feed: planes4 float32 obs + mask)", "value": ranks.world * rate,
            "unit": "env_steps/s", "n_gpus": ranks.world, "steps": K, "warmup": W, "ms_per_step": ms_step,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
            "config": {"workload": "PPO self-play rollout feed, 65,536 on-GPU envs, torch obs tensors, configs[3]",
                       "envs_per_gpu": B, "obs": "float32 (B,4,8,8) + bool mask (B,64)",
                       "l2": "L2 flushed between timed steps (192 MB fill; its own time, measured separately, is subtracted)",
                       "ms_per_step_warm_l2": ms_run_nf / K},
            "roofline": {"bound": "hbm", "achieved": bps * rate / 1e9, "peak": peak, "unit": "GB/s",
                         "frac": bps * rate / 1e9 / peak, "traffic": None, "kernel": "k_step<8,RANDOM,general> planes4 f32",
                         "bytes_per_env_step": bps, "peak_source": peak_src},
            "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": K, "clocks": clocks,
        })
    ranks.finish()
    return 0
