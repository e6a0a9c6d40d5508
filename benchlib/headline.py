"""bench.py --config headline: BASELINE.json configs[1], 8x8 lockstep batch of 2^20 random-policy envs per GPU."""
import os
import time

from . import common

METRIC = "8x8 env steps/sec (lockstep batch)"
UNIT = "env_steps/s"
WORKLOAD = "8x8 lockstep batch, 2^20 random-policy envs per GPU (step + legal-move mask + obs), configs[1]"

# the ALGORITHMIC bytes of one env step of this workload -- state r+w 56 (bitboards 2x8 B, cached legal
# set 8 B, flags/counters 4 B; each read and written), action 8, reward f32 + done 5, dense legal mask 64, int8 obs 64.
BYTES_PER_STEP = 56 + 8 + 5 + 64 + 64          # = 197, what roofline.achieved is computed from
# what the kernel actually moves per env step: state 32 B read + 32 B written (two u32 words, not 4 B: meta and the
# Philox counter), obs 64, mask 64, one 16-byte scalar record (reward | done | turn | winner | action)
BYTES_MOVED_PER_STEP = 32 + 32 + 64 + 64 + 16  # = 208


def staggered_start(env, torch, phases=60):
    """env i is advanced by (i mod phases) random plies: every launch sees all game phases."""
    env.reset()
    idx = torch.arange(env.num_envs, device=env.device) % phases
    final = env.get_state()
    for t in range(1, phases):
        env.step(policy="random")
        sel = idx == t
        cur = env.get_state()
        for k in ("black", "white", "legal"):
            final[k][:, sel] = cur[k][:, sel]
        for k in ("meta", "ctr"):
            final[k][sel] = cur[k][sel]
    env.set_state(final)


def e2e_host_buffers(torch, ranks, env, B, Ke, packed=False):
    """The same action stream through HOST buffers (VecOthelloEnv.step_host): pinned actions in, obs / mask / reward /
    done out, copies inside the timed region.  Wall clock around Ke synchronous calls, max over ranks."""
    from gymothelloenv_b200 import VecOthelloEnv
    dev = ranks.dev
    twin = VecOthelloEnv(B, board_size=8, device=dev, seed=0, auto_reset=True, obs_layout="board",
                         obs_dtype=torch.int8, env_id_base=ranks.rank * B)
    host_actions = torch.empty((Ke + 3, B), dtype=torch.int16, pin_memory=True)   # an action is 0..63: 2 bytes on PCIe
    saved = env.get_state()
    for s in range(Ke + 3):                        # the action stream the device policy would play
        env.step(policy="random")
        host_actions[s].copy_(env.actions_taken.to(torch.int16), non_blocking=True)
    torch.cuda.synchronize(dev)
    twin.set_state(saved)
    kw = {"packed": True} if packed else {}
    for s in range(3):
        twin.step_host(host_actions[s], **kw)
    ranks.barrier()
    t0 = time.perf_counter()
    for s in range(3, Ke + 3):
        twin.step_host(host_actions[s], **kw)
    ranks.barrier()
    e2e_s = ranks.max_over_ranks(time.perf_counter() - t0)
    same = bool(torch.equal(twin._black, env._black) and torch.equal(twin._white, env._white))
    h2d, d2h = twin.host_bytes_per_step()
    return {"value": ranks.world * B * Ke / e2e_s, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
            "steps": Ke, "replay_matches_device_policy": same,
            "host_link_gbs_per_gpu": (h2d + d2h) * Ke / e2e_s / 1e9}


def run(args):
    import torch
    from gymothelloenv_b200 import VecOthelloEnv
    ranks = common.Ranks(torch)
    dev, rank, world = ranks.dev, ranks.rank, ranks.world
    all_cpus = os.sched_getaffinity(0)
    numa = common.bind_to_gpu_numa_node(torch, ranks.local)     # before any pinned allocation
    B, K, W = args.num_envs, args.steps, max(3, args.warmup)

    env = VecOthelloEnv(B, board_size=8, device=dev, seed=0, auto_reset=True, obs_layout="board",
                        obs_dtype=torch.int8, env_id_base=rank * B)
    staggered_start(env, torch)
    step = (lambda: env.step(policy="random"))
    if args.graph:                                   # one-step CUDA graph replayed per step (same kernel, same args)
        g = env.capture_steps("random", 1)
        step = g.replay
    ms_run, runs, clocks = common.steady_state_runs(torch, ranks, step, K, W, repeats=args.repeats,
                                                    profile=args.profile_timed_region)
    ms_per_step = ms_run / K
    value = world * B * K / (ms_run * 1e-3)

    # sanity of the timed work: games really finish and restart (auto-reset) at the expected rate
    env.step(policy="random")
    frac_done = float(env.done.sum().item()) / B

    # ---- per-launch kernel time in isolation (events around single launches; includes launch latency) ----
    per = []
    for _ in range(20):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); step(); b.record()
        torch.cuda.synchronize(dev)
        per.append(a.elapsed_time(b))
    per.sort()

    # ---- the two ceilings: HBM on the algorithmic bytes, INT32 ALU on the counted instructions ----
    peak, peak_src = common.hbm_peak()
    per_gpu_rate = B / (ms_per_step * 1e-3)
    achieved = BYTES_PER_STEP * per_gpu_rate / 1e9
    hbm_frac = achieved / peak
    int_peak, sms = common.int32_peak(torch, ranks)
    counts = common.int_counts() or {}
    intf = common.int_fraction(counts.get("k_step_8"), per_gpu_rate, int_peak)
    hbm_ceiling = peak * 1e9 / BYTES_PER_STEP
    int_ceiling = (int_peak / intf["int_thread_inst_per_unit"]) if intf.get("int_thread_inst_per_unit") else None
    bound = "hbm" if (int_ceiling is None or hbm_ceiling <= int_ceiling) else "int32"
    roofline = {"bound": bound, "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": hbm_frac,
                "traffic": (counts.get("k_step_8") or {}).get("dram_bytes_per_launch"),
                "kernel": "k_step<8,RANDOM,FAST>", "hbm_frac": hbm_frac, "int_frac": intf.get("int_frac"),
                "bytes_per_env_step": BYTES_PER_STEP, "bytes_moved_per_env_step": BYTES_MOVED_PER_STEP,
                "moved_frac": BYTES_MOVED_PER_STEP * per_gpu_rate / 1e9 / peak,
                "kernel_ms_avg": ms_per_step, "kernel_ms_isolated_median": per[len(per) // 2],
                "peak_source": peak_src, "int32_peak_measured_thread_inst_per_s": int_peak,
                "hbm_ceiling_env_steps_per_s": hbm_ceiling, "int_ceiling_env_steps_per_s": int_ceiling,
                "note": "bound = the lower of the two ceilings; frac = achieved / peak of that ceiling's resource "
                        "(HBM here: algorithmic 197 B per env step over the measured copy bandwidth)"}
    roofline.update({k: v for k, v in intf.items() if k not in ("int_frac",)})
    if bound == "int32" and intf.get("int_frac") is not None:
        roofline.update({"achieved": intf["int_thread_inst_per_unit"] * per_gpu_rate / 1e12, "peak": int_peak / 1e12,
                         "unit": "Tinst/s", "frac": intf["int_frac"]})

    # ---- e2e through host buffers: dense (the headline e2e) and the packed variant beside it ----
    Ke = max(3, args.e2e_steps)
    e2e = e2e_host_buffers(torch, ranks, env, B, Ke)
    e2e_packed = e2e_host_buffers(torch, ranks, env, B, Ke, packed=True)
    # the box's ceiling for a step that only moves those bytes (all ranks at once), and e2e as a fraction of it
    for r in (e2e, e2e_packed):
        ceil_gbs = common.host_copy_ceiling(torch, ranks, r["d2h_bytes_per_step"], r["h2d_bytes_per_step"], Ke)
        r["host_copy_ceiling_gbs_per_gpu"] = ceil_gbs
        r["frac_of_host_copy_ceiling"] = r["host_link_gbs_per_gpu"] / ceil_gbs
    e2e["numa_node_of_rank0"] = numa
    os.sched_setaffinity(0, all_cpus)                           # the CPU baseline below uses every core again

    # ---- CPU baseline (rank 0, N = 1 only) ----
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        from oracle import cpu_arms
        arms = cpu_arms.CpuArms()
        cpu = arms.playout(8, args.cpu_seconds)
        arms.close()

    total = ranks.sum_over_ranks([B * K * args.repeats, K * args.repeats])
    if rank == 0:
        common.emit({
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u64", "data": "synthetic",
            "config": {"workload": WORKLOAD, "envs_per_gpu": B, "board_size": 8, "policy": "random (in-kernel Philox)",
                       "auto_reset": True, "obs": "int8 (B,8,8) + bool mask (B,64)",
                       "l2": "working set per step %.0f MB > 126 MB L2 (state 32 B + outputs 144 B per env)"
                             % (B * (32 + 144) / 1e6),
                       "finished_games_per_step_frac": frac_done,
                       "timing": "median of %d runs of K steps after >= 0.3 s of untimed steps" % args.repeats,
                       "launch": ("one-step CUDA graph replayed per step" if args.graph
                                  else "python -> ctypes launch per step")},
            "timed_runs_ms": runs, "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "e2e_packed": e2e_packed,
            "gpu_launches": K, "clocks": clocks, "total_env_steps": total[0],
        })
    ranks.finish()
    return 0
