"""TEST INFRASTRUCTURE -- CPU restatement of the PPO self-play vec-env protocol.

Restates, over the C oracle env (oracle/othello_oracle.c via oracle.OracleVec), what the
reference's ``PPOEnvs`` + ``subproc_worker`` + ``rule_base_game`` compute
(envs.py:188-257, ppo_run_self_play.py:244-441): which (state, action, reward, done,
choices, outputs) tuple each env hands the learner per ``PPOEnvs.step`` call.
Pinned by tests/golden/selfplay.npz = outputs of the live reference (real forked workers)
under tests/selfplay_fake.FakePolicy.  Importers allowed: tests/ only.

The worker (ppo_run_self_play.py:282-355), seen from the protagonist colour P chosen at
reset (:265-267), reduces to this per-env machine; "the pending tuple" is
(state, action, choices, outputs) of P's latest decision:
  * a ply of P (:285-290 / :325-330): when P's run of consecutive plies STARTS the saved
    state becomes the current make_state (:360 `state_b = next_state_w`, :323
    `state_w = next_state_b`); action/choices/outputs are always refreshed.  If the ply ends
    the game, the pending tuple is emitted with P's own reward (:313-317 / :350-354).  If P
    moves again because the opponent must pass, it is emitted with reward 0, done False
    (:292-295 / :332-335) -- and the saved state is NOT refreshed inside that loop, so the
    next tuple carries the state from the start of the run (reference quirk, kept);
  * a ply of the other colour (:323-341 for P=black): when its run ends (turn back to P, or
    game over) the pending tuple is emitted with MINUS the last reward (:305-309, :344-347);
    with nothing pending (P=white, first round) nothing is emitted (:305 `if not init`), or,
    if the game is already over, an all-zero tuple with done=True, choices [] (:318-321).
Finished envs answer every further call with zeros, done=True, type 'over' and the stale
possible_moves list (:273-274).
"""
import numpy as np

from oracle.oracle import OracleVec

TYPE_TRANSITION, TYPE_OVER = 0, 1


def _make_state(orc):
    return orc.make_state()          # int8 (B,4,S,S), legal plane empty when <2 moves (util.py:55)


class SelfPlayOracle(object):
    def __init__(self, num_envs, board_size=8, disk_reward=True, legacy_stale_state=True):
        self.B, self.S = int(num_envs), int(board_size)
        self.N = self.S * self.S
        self.disk_reward = bool(disk_reward)
        self.legacy = bool(legacy_stale_state)
        self.orc = None

    def reset(self, protagonists):
        B = self.B
        self.orc = OracleVec(B, self.S, sudden_death=True, disk_reward=self.disk_reward)
        self.P = np.asarray(protagonists, dtype=np.int64).reshape(B)
        self.finished = np.zeros(B, dtype=bool)
        self.pending = np.zeros(B, dtype=bool)
        self.prev_mover = np.zeros(B, dtype=np.int64)              # 0 = no ply yet in this game
        self.sav_state = np.zeros((B, 4, self.S, self.S), dtype=np.int8)
        self.sav_action = np.zeros(B, dtype=np.int64)
        self.sav_choices = np.zeros((B, self.N), dtype=np.uint8)
        self.sav_outputs = np.zeros((B, 3), dtype=np.float32)

    def step(self, policy, applied_actions_hook=None):
        """One PPOEnvs.step (envs.py:189-227).  -> dict of arrays, one entry per env."""
        B, N = self.B, self.N
        out = dict(obs=np.zeros((B, 4, self.S, self.S), dtype=np.int8), actions=np.zeros(B, dtype=np.int64),
                   rewards=np.zeros(B, dtype=np.int32), dones=np.zeros(B, dtype=np.uint8),
                   types=np.zeros(B, dtype=np.uint8), choices=np.zeros((B, N), dtype=np.uint8),
                   outputs=np.zeros((B, 3), dtype=np.float32))
        board, legal, turn, term, winner, _ = self.orc.export()
        emitted = np.zeros(B, dtype=bool)
        for i in np.nonzero(self.finished)[0]:                      # 'over' (:273-274)
            out["dones"][i], out["types"][i], out["choices"][i] = 1, TYPE_OVER, legal[i]
            emitted[i] = True
        self.policy_calls = 0
        while not emitted.all():                                    # envs.py:200
            board, legal, turn, term, winner, _ = self.orc.export()
            need = ~emitted
            obs_batch = out["obs"].copy()
            st = _make_state(self.orc)
            obs_batch[need] = st[need]
            choice_lists = [np.nonzero(legal[i])[0].tolist() if need[i] else np.nonzero(out["choices"][i])[0].tolist()
                            for i in range(B)]
            v, act, lp, hid = policy.act(obs_batch.astype(np.float32), None, None, choice_lists)   # envs.py:203
            self.policy_calls += 1
            act = np.asarray(act).reshape(B)
            mover = turn.astype(np.int64)
            is_p = need & (mover == self.P)
            for i in np.nonzero(is_p)[0]:
                if not (self.legacy and self.prev_mover[i] == self.P[i]):
                    self.sav_state[i] = st[i]
                self.sav_action[i] = act[i]
                self.sav_choices[i] = legal[i]
                self.sav_outputs[i] = (v[i], lp[i], hid[i])
                self.pending[i] = True
            applied = act.copy()
            if applied_actions_hook is not None:                   # random-opening override (othello.py:70-73)
                applied = applied_actions_hook(applied, need)
            applied[~need] = -1
            # only envs that need an action step; the others are frozen by an invalid no-op
            reward, done = self._step_subset(applied, need)
            _, _, turn2, _, _, _ = self.orc.export()
            for i in np.nonzero(need)[0]:
                P = self.P[i]
                run_ends = bool(done[i]) or turn2[i] != mover[i]
                if mover[i] == P:
                    if done[i] or turn2[i] == P:
                        self._emit(out, i, int(reward[i]), bool(done[i]))
                        emitted[i] = True
                elif run_ends:
                    if self.pending[i]:
                        self._emit(out, i, -int(reward[i]), bool(done[i]))
                        emitted[i] = True
                    elif done[i]:                                   # :318-321
                        out["dones"][i] = 1
                        emitted[i] = True
                self.prev_mover[i] = mover[i]
                if done[i]:
                    self.finished[i] = True
        return out

    def _emit(self, out, i, reward, done):
        out["obs"][i] = self.sav_state[i]
        out["actions"][i] = self.sav_action[i]
        out["rewards"][i] = reward
        out["dones"][i] = 1 if done else 0
        out["choices"][i] = self.sav_choices[i]
        out["outputs"][i] = self.sav_outputs[i]

    def _step_subset(self, actions, need):
        """Steps only the envs in `need` (the C oracle steps whole batches: run a twin and copy back)."""
        B = self.B
        board, legal, turn, term, winner, _ = self.orc.export()
        reward = np.zeros(B, dtype=np.int32)
        done = np.zeros(B, dtype=np.uint8)
        if need.all():
            r, d, _ = self.orc.step(actions)
            return r, d
        # snapshot, step all, restore the envs that must not move
        saved = self.orc._buf.copy()
        r, d, _ = self.orc.step(np.where(need, actions, 0))
        L = self.orc._buf.size // B
        keep = ~need
        buf = self.orc._buf.reshape(B, L)
        buf[keep] = saved.reshape(B, L)[keep]
        reward[need], done[need] = r[need], d[need]
        return reward, done


def rule_base_games(policy, opponent, num_games_per_env, board_size, protagonists, disk_reward=True):
    """rule_base_game (ppo_run_self_play.py:371-441) for B envs in lockstep, opening budget 0:
    protagonists[g][i] = colour of the net in game g of env i; the other colour is played by
    `opponent` ('greedy').  -> wins int64[B] (reward from the net's side > 0, :432-439)."""
    prot = np.asarray(protagonists, dtype=np.int64)
    G, B = prot.shape
    assert G == num_games_per_env
    wins = np.zeros(B, dtype=np.int64)
    for g in range(G):
        orc = OracleVec(B, board_size, sudden_death=True, disk_reward=disk_reward)
        P = prot[g]
        live = np.ones(B, dtype=bool)
        while live.any():
            board, legal, turn, term, winner, _ = orc.export()
            st = orc.make_state()
            lists = [np.nonzero(legal[i])[0].tolist() for i in range(B)]
            v, act, lp, hid = policy.act(st.astype(np.float32), None, None, lists)
            act = np.asarray(act).reshape(B)
            scripted = orc.policy(opponent)
            a = np.where(turn == P, act, scripted)
            reward, done, _ = orc.step(np.where(live, a, -1))
            for i in np.nonzero(live & (done != 0))[0]:
                r_from_p = int(reward[i]) * (1 if turn[i] == P[i] else -1)
                wins[i] += r_from_p > 0
                live[i] = False
    return wins
