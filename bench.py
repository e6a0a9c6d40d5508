#!/usr/bin/env python
"""bench.py -- 8x8 env steps/s of the lockstep batch (BASELINE.json configs[1]).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

One "step" = ONE lockstep ply of a batch of 2^20 8x8 envs per GPU: in-kernel random
policy (Philox) + rules step + auto-reset + the declared outputs (state, int8 obs
(B,8,8), dense legal mask (B,64), reward f32, done, player_turn, winner, action) -- a
single kernel launch (oth_step).  Envs are staggered over game phases (env i starts
i mod 60 plies into a game) so every launch sees the natural mix of positions.
`value` = env steps/s with everything resident in HBM, all ranks (weak scaling: 2^20
envs per GPU, no data-path collective; one 4-word all-reduce of counters at the end).
`e2e` = the same step through HOST buffers (pinned actions in, obs/mask/reward/done out).
`--impl reference` = the reference's CPU algorithm (oracle/rules_py.py, a Python port with
the reference's loop structure -- the reference itself is Python and cannot travel to the
GPU box) on all host cores.  Prints ONE JSON line.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "8x8 env steps/sec (lockstep batch)"
UNIT = "env_steps/s"
WORKLOAD = "8x8 lockstep batch, 2^20 random-policy envs per GPU (step + legal-move mask + obs), configs[1]"

# Algorithmic bytes per env step for this workload (DESIGN.md "bytes per unit"):
#   state read+write: black, white, cached legal u64 (3*8) + meta u32 + ctr u32 = 32 B, x2
#   outputs: obs int8 64 + legal mask 64 + reward 4 + done 1 + player_turn 1 + winner 1 + action int64 8
BYTES_PER_STEP = 2 * 32 + 64 + 64 + 4 + 1 + 1 + 1 + 8   # = 207


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--warmup", type=int, default=100)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--num-envs", type=int, default=1 << 20, help="envs per GPU")
    ap.add_argument("--e2e-steps", type=int, default=30)
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="bounded CPU-baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--graph", action="store_true",
                    help="replay a one-step CUDA graph per timed step instead of a Python -> ctypes launch "
                         "(measured identical: 45.05 vs 45.06 us, the loop is not host-paced)")
    ap.add_argument("--profile-timed-region", action="store_true",
                    help="cudaProfilerStart/Stop around the K timed steps (ncu --profile-from-start off then lists "
                         "exactly the launches of the timed region)")
    return ap.parse_args()


# ----------------------------------------------------------------------------- CPU arms
def _py_worker(args):
    seed, seconds = args
    from oracle.rules_py import random_playout_steps
    return random_playout_steps(8, seed, seconds)


def cpu_python_port(seconds, procs=None):
    """All host cores, one process each, uniformly random 8x8 games on the Python port
    (BASELINE.md section 3 recipe).  -> (steps/s, cores)"""
    import multiprocessing as mp
    procs = procs or os.cpu_count() or 1
    ctx = mp.get_context("fork")
    with ctx.Pool(procs) as pool:
        t0 = time.perf_counter()
        res = pool.map(_py_worker, [(1000 + i, seconds) for i in range(procs)])
        wall = time.perf_counter() - t0
    return sum(r[0] for r in res) / wall, procs


def cpu_c_oracle(seconds, threads=None):
    """C oracle (cell-array ray walk), one thread per core; ctypes releases the GIL."""
    from concurrent.futures import ThreadPoolExecutor
    from oracle import oracle
    threads = threads or os.cpu_count() or 1
    oracle.lib()
    s, _ = oracle.random_playouts(8, 1, 200)
    t0 = time.perf_counter()
    oracle.random_playouts(8, 2, 2000)
    per_game = (time.perf_counter() - t0) / 2000
    games = max(100, int(seconds / per_game))
    with ThreadPoolExecutor(threads) as ex:
        t0 = time.perf_counter()
        res = list(ex.map(lambda i: oracle.random_playouts(8, 100 + i, games)[0], range(threads)))
        wall = time.perf_counter() - t0
    return sum(res) / wall, threads


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    sample_s = max(1.0, min(args.cpu_seconds, 60.0))
    vals = []
    for _ in range(min(3, max(1, args.warmup))):   # warm-up passes: short and bounded
        cpu_python_port(0.5)
    t_all = time.perf_counter()
    per_step_seconds = max(1.0, sample_s / max(1, args.steps)) if args.steps <= 20 else 1.0
    k = max(1, min(args.steps, 20))                # bounded: at most 20 timed samples
    for _ in range(k):
        v, cores = cpu_python_port(per_step_seconds)
        vals.append(v)
    value = sum(vals) / len(vals)
    c_val, c_threads = cpu_c_oracle(2.0)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": None, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "int64", "data": "synthetic",
        "config": {"workload": WORKLOAD, "note": "CPU arm: uniformly random 8x8 games, one env per core"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": "%d x %.1fs of random 8x8 playouts on %d processes (oracle/rules_py.py, Python port "
                                   "with the reference's loop structure)" % (k, per_step_seconds, cores)},
        "cpu_baseline_c": {"value": c_val, "unit": UNIT, "cores": c_threads, "kind": "port",
                           "sample": "2s of random playouts per thread, C oracle (oracle/othello_oracle.c)"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "wall_s": time.perf_counter() - t_all,
    }
    print(json.dumps(line))
    return 0


# ----------------------------------------------------------------------------- clocks
class ClockSampler(object):
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), [x.strip() for x in line.split(",")]))

    def stop(self, t0, t1):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        rows = [r for (t, r) in self.rows if t0 <= t <= t1 + 0.11] or [r for (_, r) in self.rows[-3:]]
        sm, mx, reasons = [], None, set()
        for r in rows:
            try:
                sm.append(float(r[0]))
                mx = float(r[1])
            except (ValueError, IndexError):
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


# ----------------------------------------------------------------------------- GPU arm
def staggered_start(env, torch, phases=60):
    """env i is advanced by (i mod phases) random plies."""
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


def run_ours(args):
    import torch
    import torch.distributed as dist
    from gymothelloenv_b200 import VecOthelloEnv, _cabi
    import ctypes as C

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    B, K, W = args.num_envs, args.steps, max(3, args.warmup)

    env = VecOthelloEnv(B, board_size=8, device=dev, seed=0, auto_reset=True, obs_layout="board",
                        obs_dtype=torch.int8, env_id_base=rank * B)
    staggered_start(env, torch)
    lib = _cabi.load()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    # ---- warm-up, then EXACTLY K timed steps (one launch each) ----
    for _ in range(W):
        env.step(policy="random")
    # --graph: the timed loop replays a ONE-step CUDA graph (same kernel, same arguments) -- a check that the
    # Python -> ctypes launch path does not pace the back-to-back launches (it does not: profiles/r02i_graph.md)
    step_graph = env.capture_steps("random", 1) if args.graph else None
    if step_graph is not None:
        for _ in range(3):
            step_graph.replay()
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    time.sleep(0.25)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    t_host0 = time.perf_counter()
    done_total = torch.zeros((), dtype=torch.int64, device=dev)     # allocated before the timed region
    torch.cuda.synchronize(dev)
    if args.profile_timed_region:
        torch.cuda.profiler.start()
    ev0.record()
    if step_graph is None:
        for _ in range(K):
            env.step(policy="random")
    else:
        for _ in range(K):
            step_graph.replay()                  # one k_step<8> launch per replay
    ev1.record()
    barrier()
    if args.profile_timed_region:
        torch.cuda.profiler.stop()
    t_host1 = time.perf_counter()
    clocks = sampler.stop(t_host0, t_host1)
    ms_total = ev0.elapsed_time(ev1)
    # max over ranks of the device time
    t = torch.tensor([ms_total], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total = float(t.item())
    ms_per_step = ms_total / K
    value = world * B * K / (ms_total * 1e-3)

    # sanity of the timed work: games really finish and restart (auto-reset) at the expected rate
    env.step(policy="random")
    done_total += env.done.sum()
    frac_done = float(done_total.item()) / B

    # ---- per-launch kernel time (events around single launches) ----
    per = []
    for _ in range(20):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); env.step(policy="random"); b.record()
        torch.cuda.synchronize(dev)
        per.append(a.elapsed_time(b))
    per.sort()
    kernel_ms_single = per[len(per) // 2]
    kernel_ms = ms_per_step                       # back-to-back launches: the step IS the kernel
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.isfile(peaks_path):
        peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    else:
        peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    achieved = BYTES_PER_STEP * B / (kernel_ms * 1e-3) / 1e9
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": None, "kernel": "k_step<8>", "bytes_per_env_step": BYTES_PER_STEP,
                "kernel_ms_avg": kernel_ms, "kernel_ms_isolated_median": kernel_ms_single, "peak_source": peak_src}
    traffic_path = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.isfile(traffic_path):
        try:
            roofline["traffic"] = json.load(open(traffic_path)).get("k_step_8_bytes_per_launch")
        except Exception:
            pass

    # ---- INT32 ALU probe (the other candidate bound) ----
    sink = torch.zeros(4, dtype=torch.int32, device=dev)
    n_instr = C.c_uint64(0)
    sms = torch.cuda.get_device_properties(dev).multi_processor_count
    stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    for _ in range(2):
        lib.oth_int_alu_probe(_cabi.ptr(sink), sms * 8, 256, 2000, C.byref(n_instr), stream)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); lib.oth_int_alu_probe(_cabi.ptr(sink), sms * 8, 256, 2000, C.byref(n_instr), stream); b.record()
    torch.cuda.synchronize(dev)
    alu_peak = n_instr.value / (a.elapsed_time(b) * 1e-3)
    alu = {"int32_alu_peak_measured_inst_per_s": alu_peak, "sms": sms,
           "note": "LOP3/SHF/IADD stream, 8x256 threads per SM; see profiles/ for the instruction count per env step"}
    # the other roofline north_star names: ALU-pipe instructions k_step<8> executes per launch (an ncu count committed
    # under profiles/, NOT measured in this run) over this run's measured step time, against the pipe's issue limit
    # (16 lanes per SM sub-partition per clock = one warp instruction every 2 clocks) at the clock sampled under load
    alu_path = os.path.join(ROOT, "profiles", "alu_pipe.json")
    if os.path.isfile(alu_path):
        try:
            aj = json.load(open(alu_path))
            scale = B / float(aj["envs_per_launch"])
            inst = aj["k_step_8_alu_pipe_warp_inst_per_launch"] * scale
            mhz = (clocks or {}).get("sm_mhz") or 1965.0
            pipe_peak = sms * 4 * 0.5 * mhz * 1e6                      # warp instructions / s
            alu.update({"alu_pipe_warp_inst_per_launch": inst, "alu_pipe_peak_warp_inst_per_s": pipe_peak,
                        "alu_pipe_frac": inst / (kernel_ms * 1e-3) / pipe_peak,
                        "alu_pipe_active_pct_ncu": aj["k_step_8_alu_pipe_active_pct"], "count_source": aj["source"]})
        except Exception:
            pass

    # ---- e2e through host buffers ----
    Ke = max(3, args.e2e_steps)
    twin = VecOthelloEnv(B, board_size=8, device=dev, seed=0, auto_reset=True, obs_layout="board",
                         obs_dtype=torch.int8, env_id_base=rank * B)
    twin.set_state(env.get_state())
    host_actions = torch.empty((Ke + 3, B), dtype=torch.int16, pin_memory=True)   # an action is 0..63: 2 bytes on PCIe
    saved = env.get_state()
    for s in range(Ke + 3):                        # the action stream the device policy would play
        env.step(policy="random")
        host_actions[s].copy_(env.actions_taken.to(torch.int16), non_blocking=True)
    torch.cuda.synchronize(dev)
    twin.set_state(saved)
    for s in range(3):
        twin.step_host(host_actions[s])
    barrier()
    t0 = time.perf_counter()
    for s in range(3, Ke + 3):
        twin.step_host(host_actions[s])
    barrier()
    e2e_s = time.perf_counter() - t0
    te = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_s = float(te.item())
    same = bool(torch.equal(twin._black, env._black) and torch.equal(twin._white, env._white))
    h2d, d2h = twin.host_bytes_per_step()
    e2e = {"value": world * B * Ke / e2e_s, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
           "steps": Ke, "replay_matches_device_policy": same}

    # ---- CPU baseline (rank 0, N=1 only) ----
    cpu = None
    cpu_c = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        v, cores = cpu_python_port(args.cpu_seconds)
        cpu = {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
               "sample": "%.0fs of uniformly random 8x8 playouts on %d processes, oracle/rules_py.py (Python port "
                         "with the reference's loop structure)" % (args.cpu_seconds, cores)}
        vc, th = cpu_c_oracle(3.0)
        cpu_c = {"value": vc, "unit": UNIT, "cores": th, "kind": "port",
                 "sample": "3s of random playouts per thread, C oracle (oracle/othello_oracle.c)"}

    counters = torch.tensor([B * K, K, 0, 0], dtype=torch.int64, device=dev)
    if world > 1:
        dist.all_reduce(counters)                 # the only collective: 4 counters, after the timed region
    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u64", "data": "synthetic",
            "config": {"workload": WORKLOAD, "envs_per_gpu": B, "board_size": 8, "policy": "random (in-kernel Philox)",
                       "auto_reset": True, "obs": "int8 (B,8,8) + bool mask (B,64)",
                       "l2": "working set per step %.0f MB > 126 MB L2 (state 32 B + outputs 143 B per env)"
                             % (B * (32 + 143) / 1e6),
                       "finished_games_per_step_frac": frac_done,
                       "launch": ("python -> ctypes launch per step" if step_graph is None
                                  else "one-step CUDA graph replayed per step (VecOthelloEnv.capture_steps)")},
            "roofline": roofline, "alu": alu, "cpu_baseline": cpu, "cpu_baseline_c": cpu_c, "e2e": e2e,
            "gpu_launches": K, "clocks": clocks, "total_env_steps": int(counters[0].item()),
        }
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    a = parse()
    sys.exit(run_reference(a) if a.impl == "reference" else run_ours(a))
