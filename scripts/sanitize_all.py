"""Drives every kernel of libothello_b200.so once or twice at small sizes: a 7-second pre-flight that every
entry point launches and completes for every board size / layout / dtype (2 138 launches on a B200, r01u), and
the workload to put under compute-sanitizer where that tool is available:

    compute-sanitizer --tool memcheck  --error-exitcode 1 python scripts/sanitize_all.py
    compute-sanitizer --tool racecheck --error-exitcode 1 python scripts/sanitize_all.py --quick

(compute-sanitizer is closed on the GPU pool this repo is developed on -- the wrapper refuses to start it -- so
memory safety rests on the bit-exact parity tests over full-warp AND ragged batches for every size x layout x
dtype, and on the C ABI's argument checks, tests/test_cabi_errors.py.)

What it touches: oth_reset / oth_step / oth_observe for every board size 4..12 x obs layout x obs dtype with
a FULL-warp batch (the warp-cooperative shared-memory writers) and a RAGGED one (the generic per-cell writer),
with and without auto-reset; every action source; oth_policy (random / greedy / maximin 1..4, per-thread and
lane-cooperative); oth_rollout (per-thread and cooperative search, small-list batching); set_board_state /
set_player_turn / update paths; the self-play and teacher/student protocols; the masked categorical sampler
and the epsilon-greedy chooser.  The sizes are tiny on purpose: the sanitizer slows kernels 10-100x.
Results are not checked here (the parity tests do that); the sanitizer's report is the result.
"""
import argparse
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np   # noqa: E402
import torch         # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--quick", action="store_true", help="fewer sizes / steps (racecheck and synccheck are slower)")
    a = ap.parse_args()
    from gymothelloenv_b200 import VecOthelloEnv, PPOEnvs, PPOTeacherStudentEnvs
    from gymothelloenv_b200 import sampling, tournament
    from tests.selfplay_fake import FakePolicy

    dev = "cuda:0"
    t0 = time.time()
    launches = 0
    sizes = (5, 6, 8, 9, 12, 16) if a.quick else tuple(range(4, 17))
    steps = 3 if a.quick else 6
    layouts = ("none", "board", "board+mask", "planes4", "planes3", "absolute")
    for S in sizes:
        for layout in layouts:
            for dt in (torch.int8, torch.float32):
                if layout == "none" and dt is torch.float32:
                    continue
                for B, auto in ((128, True), (77, False)):
                    try:
                        env = VecOthelloEnv(B, S, dev, seed=S, initial_rand_steps=4, obs_layout=layout, obs_dtype=dt,
                                            auto_reset=auto, num_disk_as_reward=bool(B & 1))
                    except ValueError:
                        continue       # layout name not offered by this build
                    env.reset()
                    for k in range(steps):
                        env.step(policy="random" if k % 3 else "greedy")
                        launches += 1
                    acts = env.policy_actions("random")
                    acts[::7] = S * S + 3                    # out of range -> sudden death
                    acts[1::7] = 0                           # mostly illegal
                    env.step(acts)
                    env.observe()
                    launches += 4
        # policies on mid-game positions
        env = VecOthelloEnv(96 if a.quick else 160, S, dev, seed=11, obs_layout="none", with_legal_mask=False)
        env.reset()
        for _ in range(S * S // 3):
            env.step(policy="random")
        depths = (1, 2, 3) if (a.quick or S > 8) else (1, 2, 3, 4)
        env.policy_actions("random"); env.policy_actions("greedy")
        for d in depths:
            env.policy_actions("maximin", d)
            launches += 1
        # state injection
        st = env.get_state()
        env.set_state(st)
        cells = torch.zeros((env.num_envs, S, S), dtype=torch.int8, device=dev)
        c = S // 2
        cells[:, c - 1, c - 1] = 1; cells[:, c, c] = 1; cells[:, c - 1, c] = -1; cells[:, c, c - 1] = -1
        env.set_board_state(cells, perspective=1, turn=-1, recompute_legal=True, clear_terminal=True)
        env.set_player_turn(-1)
        env.count_disks(); env.possible_moves(0)
        # rollouts
        pairs = [("rand", "greedy"), ("greedy", "maximin2"), ("maximin3", "rand")]
        if a.quick:
            pairs = pairs[:2]
        for bk, wh in pairs:
            for kw in ({}, {"search_small_lists": True}, {"search_per_thread": True}):
                tournament.rollout(64, bk, wh, board_size=S, initial_rand_steps=4, games_per_env=2, seed=3, device=dev,
                                   record=True, trace_len=2 * S * S, **kw)
                launches += 1
        # self-play protocol
        B = 64
        sp = PPOEnvs(B, None, FakePolicy(S), dev, board_size=S, num_disk_as_reward=True, initial_rand_steps=2)
        rng = np.random.RandomState(S)
        sp.reset(protagonists=rng.choice([-1, 1], size=B))
        for _ in range(4 if a.quick else S * S):
            ret = sp.step(None)
            if bool(np.asarray(ret[3].cpu()).all()):
                break
        sp.test("greedy", B, None)
        sp.over()
        ts = PPOTeacherStudentEnvs(B, None, FakePolicy(S, 0), FakePolicy(S, 4321), dev, board_size=S,
                                   num_disk_as_reward=True)
        ts.reset(1, {"rand": 0.625, "greedy": 0.25}, {"rand": 0.5, "greedy": 0.125})
        for _ in range(4 if a.quick else 2 * S * S):
            ret = ts.step(None, None)
            if bool((ret[5].type == 2).all()):
                break
        ts.over()
        # samplers
        env = VecOthelloEnv(200, S, dev, seed=5, obs_layout="planes4", obs_dtype=torch.float32)
        env.reset()
        for _ in range(5):
            env.step(policy="random")
        logits = torch.randn((200, S * S), device=dev)
        smp = sampling.LegalMoveSampler(env, seed=9)
        act, logp = smp.sample(logits)[:2]
        smp.mode(logits)
        smp.log_probs(logits, act)
        eg = sampling.EpsilonGreedy(env, anealing_steps=10, initial_replay_size=0, seed=2)
        eg.get_action(logits); eg.get_action_at_test(logits)
        launches += 5
        # scalar drop-in API (B = 1 launches: reset / step / update_board / set_board_state / greedy + maximin)
        from gymothelloenv_b200 import othello as oth
        base = oth.OthelloBaseEnv(S, mute=True, device=dev)
        base.reset()
        base.update_board(base.possible_moves[0])
        base.set_player_turn(1)
        base.step(base.possible_moves[0])
        two = oth.SimpleOthelloEnv(S, initial_rand_steps=2, seed=1, device=dev)
        obs = two.reset()
        for _ in range(6):
            if two.env.terminated or not two.possible_moves:
                break
            obs, _, _, _ = two.step(two.possible_moves[0])
        torch.cuda.synchronize()
        print("S=%d ok (%.1fs)" % (S, time.time() - t0), flush=True)
    print("sanitize_all: done, >= %d launches, %.1fs" % (launches, time.time() - t0))


if __name__ == "__main__":
    main()
