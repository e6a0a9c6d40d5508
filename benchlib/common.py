"""Shared pieces of bench.py: process-group setup, clock sampling, steady-state timing, peaks, JSON line."""
import json
import os
import subprocess
import threading
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class Ranks(object):
    """RANK / LOCAL_RANK / WORLD_SIZE from the environment (torchrun), one process per GPU."""

    def __init__(self, torch, need_cuda=True):
        self.torch = torch
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        self.dist = None
        if need_cuda:
            if not torch.cuda.is_available():
                raise SystemExit("bench.py needs a CUDA device (no CPU fallback)")
            torch.cuda.set_device(self.local)
            self.dev = torch.device("cuda", self.local)
            if self.world > 1:
                import torch.distributed as dist
                dist.init_process_group("nccl", device_id=self.dev)
                self.dist = dist

    def barrier(self):
        if self.dist is not None:
            self.dist.barrier()
        self.torch.cuda.synchronize(self.dev)

    def max_over_ranks(self, x):
        t = self.torch.tensor([float(x)], dtype=self.torch.float64, device=self.dev)
        if self.dist is not None:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(self, values):
        t = self.torch.tensor([int(v) for v in values], dtype=self.torch.int64, device=self.dev)
        if self.dist is not None:
            self.dist.all_reduce(t)                      # NCCL sum: the only collective of any config
        return [int(v) for v in t.tolist()]

    def finish(self):
        if self.dist is not None:
            self.dist.destroy_process_group()


class ClockSampler(object):
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "50"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), [x.strip() for x in line.split(",")]))

    def wait_for(self, n=1, timeout=5.0, since=0.0):
        """Blocks until `n` samples newer than `since` have arrived (nvidia-smi can take more than a second to print
        its first line on a fresh box: a short bench used to end before it did and reported no clocks at all)."""
        if self.proc is None:
            return False
        t0 = time.perf_counter()
        while time.perf_counter() - t0 < timeout:
            if sum(1 for (t, _) in self.rows if t >= since) >= n:
                return True
            if self.proc.poll() is not None:
                return False
            time.sleep(0.01)
        return False

    def stop(self, t0, t1):
        """Samples taken in [t0, t1] (the GPU is under the bench's own load for all of it)."""
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.12)
        self.proc.terminate()
        rows = [r for (t, r) in self.rows if t0 <= t <= t1 + 0.06] or [r for (_, r) in self.rows[-3:]]
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


def steady_state_runs(torch, ranks, step, K, W, repeats=5, spin_s=0.3, profile=False):
    """W warm-up steps, then at least `spin_s` seconds of UNTIMED back-to-back steps whatever W says (clocks, L2 and
    the allocator reach the state a long run has: with --steps 20 --warmup 5 alone the driver's run used to read
    4.5 % slower than a 2000-step run), then `repeats` timed runs of EXACTLY K steps each, every run bracketed by
    barrier + synchronize and timed with CUDA events on the launch stream; each run's time is the max over ranks.
    -> (median ms per run, [ms per run], clocks dict)."""
    for _ in range(max(3, W)):
        step()
    ranks.barrier()
    sampler = ClockSampler(ranks.local).start()
    sampler.wait_for(1, timeout=5.0)                 # the sampler is up before the load starts
    t_spin = time.perf_counter()
    while True:
        for _ in range(64):
            step()
        torch.cuda.synchronize(ranks.dev)
        dt = time.perf_counter() - t_spin
        # at least spin_s, and until three clock samples were taken UNDER this load (bounded: 3 s)
        if dt >= spin_s and (dt >= 3.0 or sampler.proc is None or
                             sum(1 for (t, _) in sampler.rows if t >= t_spin + 0.1) >= 3):
            break
    t_host0 = t_spin + 0.1
    runs = []
    for r in range(repeats):
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ranks.barrier()
        step()                        # untimed pre-roll: the GPU is busy when ev0 fires, so the K timed steps run back
        if profile and r == 0:        # to back from the first one (an idle GPU adds its launch latency to step 1)
            torch.cuda.profiler.start()
        ev0.record()
        for _ in range(K):
            step()
        ev1.record()
        ranks.barrier()
        if profile and r == 0:
            torch.cuda.profiler.stop()
        runs.append(ranks.max_over_ranks(ev0.elapsed_time(ev1)))
    clocks = sampler.stop(t_host0, time.perf_counter())
    return sorted(runs)[len(runs) // 2], runs, clocks


def hbm_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.isfile(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def int32_peak(torch, ranks):
    """Measured INT32 ALU stream peak (thread-level LOP3/SHF/IADD3 per second): the denominator of int_frac."""
    import ctypes as C
    from gymothelloenv_b200 import _cabi
    lib = _cabi.load()
    dev = ranks.dev
    sink = torch.zeros(4, dtype=torch.int32, device=dev)
    n = C.c_uint64(0)
    sms = torch.cuda.get_device_properties(dev).multi_processor_count
    stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    for _ in range(2):
        lib.oth_int_alu_probe(_cabi.ptr(sink), sms * 8, 256, 2000, C.byref(n), stream)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    lib.oth_int_alu_probe(_cabi.ptr(sink), sms * 8, 256, 2000, C.byref(n), stream)
    b.record()
    torch.cuda.synchronize(dev)
    return n.value / (a.elapsed_time(b) * 1e-3), sms


def int_counts():
    """Integer thread-instructions per unit of each kernel, from the ncu counter
    smsp__sass_thread_inst_executed_op_integer_pred_on.sum committed under profiles/ (scripts/update_int_count.py)
    together with the digest of the sources the profiled library was built from."""
    p = os.path.join(ROOT, "profiles", "int_count.json")
    if not os.path.isfile(p):
        return None
    try:
        return json.load(open(p))
    except ValueError:
        return None


def int_fraction(entry, units_per_s, peak_inst_per_s):
    """-> dict(int_frac | int_count_stale)"""
    from gymothelloenv_b200 import build
    if entry is None:
        return {"int_frac": None, "int_count": "no ncu count committed for this kernel"}
    out = {"int_thread_inst_per_unit": entry["int_thread_inst_per_unit"], "int_count_source": entry.get("source")}
    if entry.get("source_digest") != build.source_digest():
        out["int_count_stale"] = True
        out["int_frac"] = None
    else:
        out["int_frac"] = entry["int_thread_inst_per_unit"] * units_per_s / peak_inst_per_s
    return out


def emit(line):
    print(json.dumps(line), flush=True)


def bind_to_gpu_numa_node(torch, local_rank):
    """Pinned host buffers are placed by first touch: run this process on the cores of the NUMA node its GPU hangs
    off BEFORE allocating them (the e2e path is a PCIe + host-memory workload).  -> node id or None."""
    try:
        bus = torch.cuda.get_device_properties(local_rank).pci_bus_id
        dom = getattr(torch.cuda.get_device_properties(local_rank), "pci_domain_id", 0)
        dev = torch.cuda.get_device_properties(local_rank).pci_device_id
        path = "/sys/bus/pci/devices/%04x:%02x:%02x.0/numa_node" % (dom, bus, dev)
        node = int(open(path).read().strip())
        if node < 0:
            return None
        cpus = []
        for part in open("/sys/devices/system/node/node%d/cpulist" % node).read().strip().split(","):
            a, _, b = part.partition("-")
            cpus.extend(range(int(a), int(b or a) + 1))
        allowed = sorted(set(cpus) & set(os.sched_getaffinity(0)))
        if allowed:
            os.sched_setaffinity(0, allowed)
        return node
    except Exception:
        return None


def host_copy_ceiling(torch, ranks, d2h_bytes, h2d_bytes, iters):
    """What the box gives a step that ONLY moves the e2e path's bytes: every rank copies h2d_bytes in and d2h_bytes
    out through pinned buffers `iters` times, all ranks at once (the 8 ranks of a node share one host memory system).
    -> GB/s per GPU (max-over-ranks time)."""
    dev = ranks.dev
    d_out = torch.empty(d2h_bytes, dtype=torch.uint8, device=dev)
    h_out = torch.empty(d2h_bytes, dtype=torch.uint8, pin_memory=True)
    d_in = torch.empty(max(1, h2d_bytes), dtype=torch.uint8, device=dev)
    h_in = torch.empty(max(1, h2d_bytes), dtype=torch.uint8, pin_memory=True)
    for _ in range(2):
        d_in.copy_(h_in, non_blocking=True)
        h_out.copy_(d_out, non_blocking=True)
    ranks.barrier()
    t0 = time.perf_counter()
    for _ in range(iters):
        d_in.copy_(h_in, non_blocking=True)
        h_out.copy_(d_out, non_blocking=True)
        torch.cuda.current_stream(dev).synchronize()
    ranks.barrier()
    dt = ranks.max_over_ranks(time.perf_counter() - t0)
    return (d2h_bytes + h2d_bytes) * iters / dt / 1e9
