#!/usr/bin/env python
"""Side measurements for the BASELINE.json configs that are NOT the bench.py headline line
(configs[0], [2], [3], [4]): parity-test workloads timed once for DESIGN.md / profiles/.

    python scripts/bench_configs.py [--quick] > gpurun_out/<tag>/configs.jsonl

Prints one JSON object per measurement.  CUDA events on the launch stream, after warm-up.
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

from gymothelloenv_b200 import VecOthelloEnv  # noqa: E402
from gymothelloenv_b200.tournament import rollout  # noqa: E402


def timed(fn, reps, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) * 1e-3 / reps


def emit(**kw):
    print(json.dumps(kw), flush=True)


def stagger(env, phases):
    env.reset()
    for _ in range(phases):
        env.step(policy="random")


def config_board_sizes(quick):
    """configs[4]: variable board size random-policy batch (multi-word bitboards for S > 8)."""
    B = 1 << 20
    for S in (6, 8, 10, 12):
        env = VecOthelloEnv(B, board_size=S, seed=0, auto_reset=True, obs_layout="board", obs_dtype=torch.int8)
        stagger(env, 20)
        dt = timed(lambda: env.step(policy="random"), 50 if quick else 300)
        N, W = S * S, env.words
        bytes_per = 2 * (3 * 8 * W + 8) + 2 * N + 4 + 1 + 1 + 1 + 8
        emit(config="board_size", S=S, words=W, envs=B, ms_per_step=dt * 1e3, env_steps_per_s=B / dt,
             bytes_per_env_step=bytes_per, achieved_gbs=bytes_per * B / dt / 1e9)
        del env
        torch.cuda.empty_cache()


def config_ppo_feed(quick):
    """configs[3]: 65,536 envs, planes4 float32 obs + bool mask, external actions (uniform legal
    sampling done with torch ops on the mask = a stand-in for the policy net's masked categorical)."""
    B = 65536
    env = VecOthelloEnv(B, board_size=8, seed=0, auto_reset=True, obs_layout="planes4", obs_dtype=torch.float32)
    stagger(env, 20)
    acts = env.policy_actions("rand")
    dt_env = timed(lambda: env.step(acts), 100 if quick else 500)   # invalid actions allowed: same cost

    def rollout_step():
        a = env.policy_actions("rand")
        env.step(a)
    dt_pol = timed(rollout_step, 100 if quick else 500)
    emit(config="ppo_feed", envs=B, obs="planes4 f32 (B,4,8,8) + bool mask (B,64)", ms_env_step=dt_env * 1e3,
         env_steps_per_s=B / dt_env, bytes_per_env_step=1167, achieved_gbs=1167 * B / dt_env / 1e9,
         ms_policy_plus_step=dt_pol * 1e3, rollout_steps_per_s=B / dt_pol)
    B = 1 << 20
    env = VecOthelloEnv(B, board_size=8, seed=0, auto_reset=True, obs_layout="planes4", obs_dtype=torch.float32)
    stagger(env, 20)
    dt = timed(lambda: env.step(policy="random"), 30 if quick else 200)
    emit(config="ppo_feed_1M", envs=B, ms_env_step=dt * 1e3, env_steps_per_s=B / dt, bytes_per_env_step=1167,
         achieved_gbs=1167 * B / dt / 1e9)


def reference_shaped_net(S):
    """Random-init network with the layer shapes of the reference's CNNBase + actor / critic heads
    (a2c_ppo_acktr/model.py:288-305) for a (4,S,S) input; the caller's dense NN, plain torch, fp32."""
    import torch.nn as nn
    body = nn.Sequential(nn.Conv2d(4, 32, 3, stride=2, padding=1), nn.ReLU(), nn.Conv2d(32, 64, 2), nn.ReLU(),
                         nn.Conv2d(64, 64, 2), nn.ReLU(), nn.Flatten())
    with torch.no_grad():
        feat = body(torch.zeros(1, 4, S, S)).shape[1]
    trunk = nn.Sequential(body, nn.Linear(feat, 512), nn.ReLU())
    return trunk.cuda(), nn.Linear(512, S * S).cuda(), nn.Linear(512, 1).cuda()


def config_ppo_rollout(quick):
    """configs[3] end to end: network forward -> legal-move-restricted sampling (oth_masked_categorical) -> env step,
    (a) as a plain lockstep rollout on VecOthelloEnv, (b) through the self-play protocol PPOEnvs.step."""
    from gymothelloenv_b200 import LegalMoveSampler, PPOEnvs
    B, S = 65536, 8
    torch.manual_seed(0)
    trunk, actor, critic = reference_shaped_net(S)
    env = VecOthelloEnv(B, board_size=S, seed=0, auto_reset=True, obs_layout="planes4", obs_dtype=torch.float32)
    stagger(env, 20)
    sampler = LegalMoveSampler(env)

    @torch.no_grad()
    def net_only():
        h = trunk(env.obs)
        return actor(h), critic(h)

    @torch.no_grad()
    def rollout_step():
        logits, value = net_only()
        a, lp = sampler.sample(logits)
        env.step(a)
    reps = 50 if quick else 300
    dt_net = timed(net_only, reps)
    dt_all = timed(rollout_step, reps)
    emit(config="ppo_rollout_lockstep", envs=B, net="CNNBase-shaped (conv 4-32-64-64, fc 512) + actor/critic heads, fp32 torch",
         ms_net_forward=dt_net * 1e3, ms_net_sample_step=dt_all * 1e3, ms_sampler_plus_env=(dt_all - dt_net) * 1e3,
         rollout_steps_per_s=B / dt_all)

    class Pol(object):
        sampler = None

        @torch.no_grad()
        def act(self, obs, hxs, masks, choices):
            h = trunk(obs)
            a, lp = self.sampler.sample(actor(h))
            return critic(h).squeeze(1), a.unsqueeze(1), lp, torch.zeros_like(lp)
    pol = Pol()
    envs = PPOEnvs(B, None, pol, "cuda:0", board_size=S, num_disk_as_reward=True, seed=3)
    pol.sampler = LegalMoveSampler(envs.vec)
    games = 1 if quick else 3
    tuples = calls = acts = 0
    torch.cuda.synchronize()
    t0 = time.time()
    for g in range(games):
        envs.reset()
        while True:
            obs, actions, rewards, dones, infos, outputs, masks, bad = envs.step(None)
            calls += 1
            acts += envs.policy_calls
            tuples += int((infos.type == 0).sum())
            if bool(dones.all()):
                break
    torch.cuda.synchronize()
    dt = time.time() - t0
    emit(config="ppo_selfplay_protocol", envs=B, games_per_env=games, step_calls=calls, seconds=dt,
         learning_tuples_per_s=tuples / dt, tuples=tuples, ms_per_step_call=dt / calls * 1e3)


def config_tournaments(quick):
    """configs[0] and [2]: fused rollouts, README protocol (<= 10 random opening plies, fixed colours)."""
    cases = [("rand", "rand", 1 << 22), ("rand", "greedy", 1 << 22), ("greedy", "rand", 1 << 22),
             ("greedy", "greedy", 1 << 22), ("greedy", "maximin2", 1 << 21), ("maximin2", "greedy", 1 << 21),
             ("greedy", "maximin3", 1 << (17 if quick else 21)), ("maximin3", "greedy", 1 << (17 if quick else 21))]
    for black, white, games in cases:
        rollout(4096, black, white, seed=1)          # warm-up / module load
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        tally, _ = rollout(games, black, white, board_size=8, initial_rand_steps=10, seed=0)
        b.record()
        torch.cuda.synchronize()
        dt = a.elapsed_time(b) * 1e-3
        w, d, l, plies = [int(x) for x in tally.tolist()]
        emit(config="tournament", black=black, white=white, games=games, seconds=dt, games_per_s=games / dt,
             env_steps_per_s=plies / dt, black_wdl_pct=[round(100.0 * x / games, 2) for x in (w, d, l)],
             plies_per_game=plies / games)


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--quick", action="store_true")
    ap.add_argument("--only", default="")
    args = ap.parse_args()
    t0 = time.time()
    emit(config="device", name=torch.cuda.get_device_name(0))
    if args.only in ("", "sizes"):
        config_board_sizes(args.quick)
    if args.only in ("", "ppo"):
        config_ppo_feed(args.quick)
    if args.only in ("", "ppo", "ppo_rollout"):
        config_ppo_rollout(args.quick)
    if args.only in ("", "tournament"):
        config_tournaments(args.quick)
    emit(config="done", wall_s=time.time() - t0)
