#!/usr/bin/env python
"""Can the SM overlap the rules phase and the writer phase of k_step<8> when they come from DIFFERENT warps?
Runs, over 2^20 envs in total and inside one CUDA graph (no host launch gaps):
  serial      8 x [rules-only step of 2^17 envs] then 8 x [writer (oth_observe) of 2^17 envs], one stream
  concurrent  the same 16 launches on two streams (rules on one, writer on the other) so that CTAs of both kinds
              are co-resident on the SMs
  fused       8 x [fused step of 2^17 envs] (today's kernel)
concurrent << serial means a warp-specialised k_step (rules warps + writer warps in one CTA) would pay.
usage: python scripts/bench_phase_overlap.py [parts]"""
import json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gymothelloenv_b200 import VecOthelloEnv

PARTS = int(sys.argv[1]) if len(sys.argv) > 1 else 8
SUB = (1 << 20) // PARTS


def mk(layout, mask):
    envs = []
    for k in range(PARTS):
        e = VecOthelloEnv(SUB, board_size=8, seed=0, auto_reset=True, obs_layout=layout, obs_dtype=torch.int8,
                          with_legal_mask=mask, env_id_base=k * SUB)
        e.reset()
        for _ in range(30):
            e.step(policy="random")
        envs.append(e)
    return envs


rules, writer, fused = mk("none", False), mk("board", True), mk("board", True)
torch.cuda.synchronize()


def capture(fn):
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        fn()
    return g


def serial():
    for e in rules:
        e.step(policy="random")
    for e in writer:
        e.observe()


side = torch.cuda.Stream()


def concurrent():
    cur = torch.cuda.current_stream()
    side.wait_stream(cur)
    with torch.cuda.stream(side):
        for e in writer:
            e.observe()
    for e in rules:
        e.step(policy="random")
    cur.wait_stream(side)


def fused_fn():
    for e in fused:
        e.step(policy="random")


def timed(g, reps=200):
    for _ in range(10):
        g.replay()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        g.replay()
    b.record()
    torch.cuda.synchronize()
    return round(a.elapsed_time(b) * 1e3 / reps, 2)


row = {"envs": PARTS * SUB, "parts": PARTS}
for name, fn in (("serial_us", serial), ("concurrent_us", concurrent), ("fused_us", fused_fn)):
    row[name] = timed(capture(fn))
print(json.dumps(row))
