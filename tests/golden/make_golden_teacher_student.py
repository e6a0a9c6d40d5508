#!/usr/bin/env python
"""Golden vectors for the teacher/student vec-env protocol, from the LIVE reference
(build container only; nothing here is imported by the product).

    python tests/golden/make_golden_teacher_student.py   # rewrites tests/golden/teacher_student.npz

What runs, unmodified: the reference's ``envs.PPOTeacherStudentEnvs`` (envs.py:44-184) with real
forked worker processes executing ``othello_teacher_vs_student`` (ppo_run_teacher_vs_student.py:425-573,
compiled from the reference file at run time -- the module itself cannot be imported because of its
learner dependencies) over ``othello.SimpleOthelloEnv`` + ``util.make_state``.  The two networks are
replaced by ``tests/selfplay_fake.py::FakePolicy`` with different salts.  The driver below issues the
call sequence of the reference main loop (ppo_run_teacher_vs_student.py:216-246): ``reset(teacher,
win_avg, last_win_avg)``, then ``step`` until -- checked at the top of an iteration, with one more call
after it turns true -- every env has reported done in its teacher or its student slot.

Every ``step`` call runs under a SIGALRM timeout: a call that blocks is a finding about the reference,
recorded in the output (``cN_blocked``), not a hang of this script.
"""
import ast
import os
import signal
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle import ref_shim  # noqa: E402
from tests.selfplay_fake import FakePolicy  # noqa: E402

othello, simple_policies, util = ref_shim.load_reference()
import envs as ref_envs  # noqa: E402  (reference envs.py)

STUDENT_SALT = 4321
TYPE_CODE = {"black_transition": 0, "white_transition": 1, "over": 2}


class StepBlocked(Exception):
    pass


def _alarm(signum, frame):
    raise StepBlocked()


def reference_worker():
    path = os.path.join(ref_shim.REFERENCE_ROOT, "ppo_run_teacher_vs_student.py")
    tree = ast.parse(open(path).read(), filename=path)
    keep = [n for n in tree.body if isinstance(n, ast.FunctionDef)
            and n.name in ("othello_teacher_vs_student", "rule_base_game")]
    assert len(keep) == 2
    ns = {"np": np, "make_state": util.make_state, "create_policy": util.create_policy}
    exec(compile(ast.Module(body=keep, type_ignores=[]), path, "exec"), ns)
    return ns["othello_teacher_vs_student"]


def run_case(S, num_envs, teachers, disk_reward, win_pairs, step_timeout=20):
    worker = reference_worker()
    env_list = []
    for i in range(num_envs):
        with ref_shim.muted():
            env = othello.SimpleOthelloEnv(board_size=S, seed=i, initial_rand_steps=0, num_disk_as_reward=disk_reward,
                                           render_in_step=False)
        env.rand_steps_holder = 0
        env.test_rand_steps_holder = 0
        env_list.append(env)
    teacher_pol, student_pol = FakePolicy(S, 0), FakePolicy(S, STUDENT_SALT)
    devnull = os.open(os.devnull, os.O_WRONLY)
    saved_stdout = os.dup(1)
    os.dup2(devnull, 1)                           # the worker prints on every command it receives
    calls, blocked = [], None
    signal.signal(signal.SIGALRM, _alarm)
    vec = None
    try:
        vec = ref_envs.PPOTeacherStudentEnvs(env_list, worker, teacher_pol, student_pol, "cpu")
        N = S * S
        for ep, teacher in enumerate(teachers):
            win_avg, last_win_avg = win_pairs[ep]
            vec.reset(teacher, win_avg, last_win_avg)
            done_ts = [[0] * num_envs, [0] * num_envs]
            over, k = False, 0
            while not over:
                over = all(np.array(done_ts[0]) + np.array(done_ts[1]))        # :230
                signal.alarm(step_timeout)
                try:
                    t_or_s, obs, actions, rewards, dones, infos, outputs, masks, bad_masks = vec.step(None, None)
                except StepBlocked:
                    blocked = (ep, k, bool(over))
                    break
                finally:
                    signal.alarm(0)
                ch = np.zeros((num_envs, N), dtype=np.uint8)
                for i, info in enumerate(infos):
                    if len(info["choices"]):
                        ch[i, np.asarray(info["choices"], dtype=np.int64)] = 1
                calls.append(dict(
                    episode=ep, teacher=teacher, who=np.array([0 if w == "teacher" else 1 for w in t_or_s], dtype=np.uint8),
                    obs=obs.numpy().astype(np.int8), actions=actions.numpy().reshape(-1).astype(np.int64),
                    rewards=rewards.numpy().reshape(-1).astype(np.float32),
                    dones=np.array([bool(d) for d in dones], dtype=np.uint8),
                    types=np.array([TYPE_CODE[info["type"]] for info in infos], dtype=np.uint8),
                    choices=ch, outputs=outputs.numpy().astype(np.float32),
                    masks=masks.numpy().reshape(-1).astype(np.float32),
                    bad_masks=bad_masks.numpy().reshape(-1).astype(np.float32)))
                for i, who in enumerate(t_or_s):                                # save_to_buffer, :194-199
                    done_ts[0 if who == "teacher" else 1][i] = dones[i]
                k += 1
            if blocked:
                break
    finally:
        os.dup2(saved_stdout, 1)
        os.close(devnull)
        if vec is not None:
            for p in vec.procs:
                p.terminate()
            for p in vec.procs:
                p.join(timeout=5)
    return calls, blocked


def main():
    wp = [({"rand": 0.625, "greedy": 0.25}, {"rand": 0.5, "greedy": 0.125}),
          ({"rand": 0.5, "greedy": 0.5}, {"rand": 0.75, "greedy": 0.125}),
          ({"rand": 0.0, "greedy": 0.0}, {"rand": 0.0, "greedy": 0.0})]
    cases = [
        # S, envs, teacher colour per episode, disk_reward
        (8, 6, [1, -1, 1], True),
        (6, 8, [-1, 1, -1], False),
        (4, 8, [1, -1, -1], True),
        (5, 4, [-1, 1], True),
        (10, 3, [1, -1], False),
    ]
    only = int(sys.argv[1]) if len(sys.argv) > 1 else None
    out = {"num_cases": np.int64(len(cases))}
    for c, (S, B, teachers, disk) in enumerate(cases):
        if only is not None and c != only:
            continue
        calls, blocked = run_case(S, B, teachers, disk, wp)
        pre = "c%d_" % c
        out[pre + "cfg"] = np.array([S, B, len(teachers), int(disk)], dtype=np.int64)
        out[pre + "teachers"] = np.array(teachers, dtype=np.int8)
        out[pre + "blocked"] = np.array(blocked if blocked else (-1, -1, 0), dtype=np.int64)
        for k in ("episode", "teacher", "who", "obs", "actions", "rewards", "dones", "types", "choices", "outputs",
                  "masks", "bad_masks"):
            out[pre + k] = np.stack([np.asarray(x[k]) for x in calls]) if calls else np.zeros(0)
        per_ep = [sum(1 for x in calls if x["episode"] == e) for e in range(len(teachers))]
        print("case %d: S=%d envs=%d teachers=%s disk=%d -> step() calls per episode %s, blocked=%s"
              % (c, S, B, teachers, disk, per_ep, blocked))
    if only is None:
        np.savez_compressed(os.path.join(HERE, "teacher_student.npz"), **out)
        print("wrote", os.path.join(HERE, "teacher_student.npz"))
    return out


if __name__ == "__main__":
    main()
