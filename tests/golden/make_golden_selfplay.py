#!/usr/bin/env python
"""Golden vectors for the PPO self-play vec-env protocol, from the LIVE reference
(build container only; nothing here is imported by the product).

    python tests/golden/make_golden_selfplay.py      # rewrites tests/golden/selfplay.npz

What runs, unmodified: the reference's ``envs.PPOEnvs`` (envs.py:7-41,188-257) with real
forked worker processes executing ``subproc_worker`` / ``rule_base_game``
(ppo_run_self_play.py:244-441; the two functions are compiled from the reference file at
run time -- the module itself cannot be imported because of its learner dependencies,
SURVEY 3.4) over ``othello.SimpleOthelloEnv`` + ``util.make_state``.  The policy network is
replaced by a deterministic stand-in, ``tests/selfplay_fake.py::FakePolicy``, whose action /
value / log-prob / hidden outputs are a pure function of (env index, observation planes,
choices): the device implementation is fed by the same function, so every tensor
``PPOEnvs.step`` returns can be compared bit for bit.

Recorded per ``PPOEnvs.step`` call: obs (B,4,S,S), actions, rewards, dones, info type and
choices, outputs (v, logp, hidden), masks, bad_masks.  Per ``PPOEnvs.test`` call:
(games, wins) and the protagonist colours of the test games.
"""
import ast
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle import ref_shim  # noqa: E402
from tests.selfplay_fake import FakePolicy  # noqa: E402

othello, simple_policies, util = ref_shim.load_reference()
import envs as ref_envs  # noqa: E402  (reference envs.py)


def reference_worker_functions():
    """subproc_worker and rule_base_game compiled from the reference source file."""
    path = os.path.join(ref_shim.REFERENCE_ROOT, "ppo_run_self_play.py")
    tree = ast.parse(open(path).read(), filename=path)
    keep = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name in ("subproc_worker", "rule_base_game")]
    assert len(keep) == 2
    mod = ast.Module(body=keep, type_ignores=[])
    ns = {"np": np, "make_state": util.make_state, "create_policy": util.create_policy}
    exec(compile(mod, path, "exec"), ns)
    return ns["subproc_worker"], ns["rule_base_game"]


def run_case(S, num_envs, games, disk_reward, np_seed, test_games):
    subproc_worker, _ = reference_worker_functions()
    env_list = []
    for i in range(num_envs):
        with ref_shim.muted():
            env = othello.SimpleOthelloEnv(board_size=S, seed=i, initial_rand_steps=0, num_disk_as_reward=disk_reward,
                                           render_in_step=False)
        env.rand_steps_holder = 0
        env.test_rand_steps_holder = 0
        env_list.append(env)
    policy = FakePolicy(S)
    np.random.seed(np_seed)                       # the workers inherit this state at fork (protagonist draws)
    fork_state = np.random.get_state()
    devnull = os.open(os.devnull, os.O_WRONLY)
    saved_stdout = os.dup(1)
    os.dup2(devnull, 1)                           # workers print on every reset / terminal
    try:
        vec = ref_envs.PPOEnvs(env_list, subproc_worker, policy, "cpu")
        calls = []
        for g in range(games):
            vec.reset()
            over = False
            while not over:
                obs, actions, rewards, dones, infos, outputs, masks, bad_masks = vec.step(None)
                N = S * S
                ch = np.zeros((num_envs, N), dtype=np.uint8)
                for i, info in enumerate(infos):
                    if len(info["choices"]):
                        ch[i, np.asarray(info["choices"], dtype=np.int64)] = 1
                calls.append(dict(
                    game=g,
                    obs=obs.numpy().astype(np.int8), actions=actions.numpy().reshape(-1).astype(np.int64),
                    rewards=rewards.numpy().reshape(-1).astype(np.int32), dones=np.array([bool(d) for d in dones], dtype=np.uint8),
                    types=np.array([0 if info["type"] is None else 1 for info in infos], dtype=np.uint8),
                    choices=ch, outputs=outputs.numpy().astype(np.float32),
                    masks=masks.numpy().reshape(-1).astype(np.float32), bad_masks=bad_masks.numpy().reshape(-1).astype(np.float32)))
                over = all(dones)
        test_result = None
        if test_games:
            test_result = vec.test("greedy", test_games, None)
        vec.over()
        for p in vec.procs:
            p.join(timeout=10)
    finally:
        os.dup2(saved_stdout, 1)
        os.close(devnull)
    # protagonist colours: every worker draws np.random.randint(2) once per game from the state it
    # inherited at fork (ppo_run_self_play.py:265-266,389-390), so all workers share one sequence
    rs = np.random.RandomState()
    rs.set_state(fork_state)
    per_worker_test = test_games // num_envs if test_games else 0
    prot = [(-1 if rs.randint(2) == 0 else 1) for _ in range(games + per_worker_test)]
    return calls, prot, test_result


def main():
    cases = [
        # S, envs, games, disk_reward, numpy seed, test games
        (8, 8, 4, True, 11, 16),
        (8, 6, 3, False, 5, 0),
        (6, 8, 4, True, 3, 16),
        (4, 8, 6, False, 2, 8),
        (5, 4, 3, True, 9, 0),
        (10, 4, 2, True, 4, 0),
    ]
    out = {"num_cases": np.int64(len(cases))}
    for c, (S, B, G, disk, seed, tg) in enumerate(cases):
        calls, prot, test_result = run_case(S, B, G, disk, seed, tg)
        pre = "c%d_" % c
        out[pre + "cfg"] = np.array([S, B, G, int(disk), seed, tg], dtype=np.int64)
        out[pre + "protagonist"] = np.array(prot, dtype=np.int8)
        out[pre + "test"] = np.array(test_result if test_result else (0, 0), dtype=np.int64)
        for k in ("game", "obs", "actions", "rewards", "dones", "types", "choices", "outputs", "masks", "bad_masks"):
            out[pre + k] = np.stack([np.asarray(x[k]) for x in calls])
        n_tr = int(sum((x["types"] == 0).sum() for x in calls))
        print("case %d: S=%d envs=%d games=%d disk=%d -> %d step() calls, %d learning transitions, test=%s"
              % (c, S, B, G, disk, len(calls), n_tr, test_result))
    np.savez_compressed(os.path.join(HERE, "selfplay.npz"), **out)
    print("wrote", os.path.join(HERE, "selfplay.npz"))


if __name__ == "__main__":
    main()
