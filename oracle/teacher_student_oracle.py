"""TEST INFRASTRUCTURE -- CPU restatement of the teacher/student vec-env protocol.

Restates, over the C oracle env (oracle/othello_oracle.c via oracle.OracleVec), what the reference's
``PPOTeacherStudentEnvs`` + ``othello_teacher_vs_student`` compute (envs.py:44-140,
ppo_run_teacher_vs_student.py:425-560): which (who, state, action, reward, done, type, choices,
outputs) tuple each env hands the learner per ``PPOTeacherStudentEnvs.step`` call.
Pinned by tests/golden/teacher_student.npz = outputs of the live reference (real forked workers)
under tests/selfplay_fake.FakePolicy.  Importers allowed: tests/ only.

The worker, seen per env, is this machine ("pending tuple of colour c" = state, action, choices,
outputs of c's latest decision; black moves first, so white is the only colour that can lack one):
  * c needs an action (:480-484 / :523-527): when c's run of consecutive plies STARTS the saved state
    of c becomes the current make_state (:474 `state_b`, :522 `state_w = next_state_b`, :557
    `state_b = next_state_w`); inside a forced-pass run it is NOT refreshed (:486-496 / :529-539 send
    `next_state_*` to the policy but keep `state_*` in the tuple -- reference quirk, kept);
    action / choices / outputs are always refreshed.
  * c moves again because the other side must pass (:486-488 / :529-531): c's pending tuple is emitted
    with c's own reward, done False.
  * c's run ends (turn passes, or the game is over): the other colour's pending tuple, if any, is
    emitted with MINUS c's last reward and c's done flag (:498-503 / :541-545).  If the game is over,
    c's own tuple follows in the NEXT step() call with c's reward, done True (:505-509 / :546-551); with
    no pending tuple on the other side (white before its first move) it is emitted at once.
  * the teacher colour's reward is replaced (:457-466): sum_k(win_avg[k] - last_win_avg[k]) when the
    tuple's done flag is set, else 0.
  * a finished env answers every further call with zeros, done True, type 'over', choices [] (:470-471).
PPOTeacherStudentEnvs.step (envs.py:63-140) serves black's requests first (`turn = -1 if
'need_black_action' in types else 1`, :78) with policy[turn] on the whole batch, and labels each row
teacher / student from the tuple's colour; 'over' rows count as white (:124-130).
"""
import numpy as np

from oracle.oracle import OracleVec
from oracle.selfplay_oracle import SelfPlayOracle

TYPE_BLACK, TYPE_WHITE, TYPE_OVER = 0, 1, 2
WHO_TEACHER, WHO_STUDENT = 0, 1


def _ci(colour):          # -1 (black) -> 0, +1 (white) -> 1
    return 0 if colour < 0 else 1


class TeacherStudentOracle(object):
    def __init__(self, num_envs, board_size=8, disk_reward=True, legacy_stale_state=True):
        self.B, self.S = int(num_envs), int(board_size)
        self.N = self.S * self.S
        self.disk_reward = bool(disk_reward)
        self.legacy = bool(legacy_stale_state)
        self.orc = None

    def reset(self, teacher, win_avg, last_win_avg):
        B, S, N = self.B, self.S, self.N
        self.orc = OracleVec(B, S, sudden_death=True, disk_reward=self.disk_reward)
        self.teacher = int(teacher)
        self.teacher_done_reward = float(sum(win_avg[k] - last_win_avg[k] for k in win_avg.keys()))   # :460-461
        self.finished = np.zeros(B, dtype=bool)
        self.queued = np.zeros(B, dtype=bool)            # own final tuple owed at the next call
        self.queued_colour = np.zeros(B, dtype=np.int64)
        self.queued_reward = np.zeros(B, dtype=np.int64)
        self.pending = np.zeros((2, B), dtype=bool)
        self.prev_mover = np.zeros(B, dtype=np.int64)
        self.sav_state = np.zeros((2, B, 4, S, S), dtype=np.int8)
        self.sav_action = np.zeros((2, B), dtype=np.int64)
        self.sav_choices = np.zeros((2, B, N), dtype=np.uint8)
        self.sav_outputs = np.zeros((2, B, 3), dtype=np.float32)

    def _emit(self, out, i, colour, raw_reward, done):
        c = _ci(colour)
        out["obs"][i] = self.sav_state[c, i]
        out["actions"][i] = self.sav_action[c, i]
        if colour == self.teacher:
            out["rewards"][i] = np.float32(self.teacher_done_reward if done else 0.0)
        else:
            out["rewards"][i] = np.float32(raw_reward)
        out["dones"][i] = 1 if done else 0
        out["types"][i] = TYPE_BLACK if colour < 0 else TYPE_WHITE
        out["choices"][i] = self.sav_choices[c, i]
        out["outputs"][i] = self.sav_outputs[c, i]
        out["who"][i] = WHO_TEACHER if colour == self.teacher else WHO_STUDENT

    def step(self, policies):
        """One PPOTeacherStudentEnvs.step.  policies = {-1: black's network, +1: white's network}."""
        B, N, S = self.B, self.N, self.S
        out = dict(obs=np.zeros((B, 4, S, S), dtype=np.int8), actions=np.zeros(B, dtype=np.int64),
                   rewards=np.zeros(B, dtype=np.float32), dones=np.zeros(B, dtype=np.uint8),
                   types=np.zeros(B, dtype=np.uint8), choices=np.zeros((B, N), dtype=np.uint8),
                   outputs=np.zeros((B, 3), dtype=np.float32), who=np.zeros(B, dtype=np.uint8))
        emitted = np.zeros(B, dtype=bool)
        for i in np.nonzero(self.finished)[0]:
            if self.queued[i]:
                self._emit(out, i, int(self.queued_colour[i]), int(self.queued_reward[i]), True)
                self.queued[i] = False
            else:                                                       # 'over' (:470-471, envs.py:124-130)
                out["dones"][i], out["types"][i] = 1, TYPE_OVER
                out["who"][i] = WHO_TEACHER if self.teacher == 1 else WHO_STUDENT
            emitted[i] = True
        self.policy_calls = []
        while not emitted.all():                                        # envs.py:76
            board, legal, turn, term, winner, _ = self.orc.export()
            need = ~emitted
            colour = -1 if (need & (turn == -1)).any() else 1           # envs.py:78
            sel = need & (turn == colour)
            st = self.orc.make_state()
            obs_batch = out["obs"].copy()
            obs_batch[need] = st[need]
            lists = [np.nonzero(legal[i])[0].tolist() if need[i] else np.nonzero(out["choices"][i])[0].tolist()
                     for i in range(B)]
            v, act, lp, hid = policies[colour].act(obs_batch.astype(np.float32), None, None, lists)
            self.policy_calls.append(colour)
            act = np.asarray(act).reshape(B)
            c, o = _ci(colour), _ci(-colour)
            for i in np.nonzero(sel)[0]:
                if not (self.legacy and self.prev_mover[i] == colour):
                    self.sav_state[c, i] = st[i]
                self.sav_action[c, i] = act[i]
                self.sav_choices[c, i] = legal[i]
                self.sav_outputs[c, i] = (v[i], lp[i], hid[i])
                self.pending[c, i] = True
            applied = np.where(sel, act, -1)
            reward, done = SelfPlayOracle._step_subset(self, applied, sel)
            _, _, turn2, _, _, _ = self.orc.export()
            for i in np.nonzero(sel)[0]:
                r, d = int(reward[i]), bool(done[i])
                if not d and turn2[i] == colour:                        # the other side passes (:486-488 / :529-531)
                    self._emit(out, i, colour, r, False)
                    emitted[i] = True
                else:                                                   # c's run ends
                    if self.pending[o, i]:
                        self._emit(out, i, -colour, -r, d)              # :498-503 / :541-545
                        emitted[i] = True
                        if d:
                            self.queued[i], self.queued_colour[i], self.queued_reward[i] = True, colour, r
                    elif d:
                        self._emit(out, i, colour, r, True)             # :505-509 with init still True
                        emitted[i] = True
                self.prev_mover[i] = colour
                if d:
                    self.finished[i] = True
        return out
