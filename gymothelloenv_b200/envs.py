"""Envs / PPOEnvs -- the PPO self-play vec-env of the reference, on the device.

Mirrors the reference's ``envs.py`` (``Envs`` :7-41, ``PPOEnvs`` :188-257) and the worker
side of the protocol (``subproc_worker`` / ``rule_base_game``,
``ppo_run_self_play.py:244-441``).  The reference forks one process per env and moves
every ply through pickled pipes; here B envs live on one B200 and a "round"
(``PPOEnvs.step``) is: one ``oth_selfplay_begin`` launch, then per ``policy.act`` call one
``oth_selfplay_advance`` launch (bookkeeping + rules step + next observation fused), then
two ``oth_observe`` launches that expand the finished tuples.  What each env hands the
learner per round is bit-identical to the reference worker's message (see
tests/test_gpu_selfplay.py): the protagonist colour's (state, action, reward, done,
choices, (value, log-prob, hidden)), reward = its own terminal reward or MINUS the reward
of the opponent's reply, finished envs answering zeros / done / 'over'.

Differences a caller sees (all tensors stay on the GPU):
  * ``dones`` is a bool CUDA tensor [B] (reference: python list) -- use ``dones.all()``;
  * ``infos`` is a ``SelfPlayInfos`` (``.type`` uint8 [B]: 0 = learning tuple, 1 = 'over';
    ``.choices_mask`` bool [B, S*S]); ``infos[i]`` builds the reference's dict lazily;
  * ``policy.act(obs, rnn_hxs, masks, choices)`` receives ``choices`` as the bool mask
    [B, S*S] (``choices_as='lists'`` gives the reference's list of lists, with a sync);
  * protagonist colours come from Philox (reference: the worker's global numpy RNG), or
    from ``reset(protagonists=...)``.
No CPU fallback: needs libothello_b200.so and a CUDA device.
"""
import ctypes as C

import torch

from . import _cabi
from .vec_env import VecOthelloEnv


class SelfPlayInfos(object):
    """infos of one PPOEnvs.step: tensors + lazy reference-style dict access."""

    def __init__(self, type_u8, choices_mask):
        self.type = type_u8
        self.choices_mask = choices_mask

    def __len__(self):
        return self.type.numel()

    def __getitem__(self, i):
        t = int(self.type[i].item())
        return {"type": None if t == 0 else "over",
                "choices": torch.nonzero(self.choices_mask[i]).flatten().tolist()}

    def __iter__(self):
        types = self.type.cpu().tolist()
        masks = self.choices_mask.cpu()
        for t, m in zip(types, masks):
            yield {"type": None if t == 0 else "over", "choices": torch.nonzero(m).flatten().tolist()}

    def choices(self):
        """list of lists, like [info['choices'] for info in infos] (ppo_run_self_play.py:172)"""
        return [d["choices"] for d in self]


class _RuleBaseGames(object):
    """rule_base_game (ppo_run_self_play.py:371-441 == ppo_run_teacher_vs_student.py:575-645, identical text)
    for all envs in lockstep; mixed into Envs."""

    def _rule_base_games(self, policy, name, num_games, ex_hiddens, protagonists=None):
        if name not in ("rand", "greedy"):
            raise ValueError("test opponent must be 'rand' or 'greedy'")
        v = self.vec
        B, N = self.num_processes, v.num_cells
        per = int(num_games) // B
        wins = torch.zeros((B,), dtype=torch.int64, device=self.device)
        for g in range(per):
            pin = None
            if protagonists is not None:
                pin = torch.as_tensor(protagonists[g]).to(device=self.device, dtype=torch.int8).reshape(-1)
                pin = pin.expand(B).contiguous() if pin.numel() == 1 else pin.contiguous()
            with v._on_device():
                _cabi.check(self._lib.oth_selfplay_new_game(v._state(), self._c_sp, B, v.board_size, v.seed_value,
                                                            v.env_id_base, self.test_rand_steps, self._game_index,
                                                            _cabi.ptr(pin), v._stream()), "oth_selfplay_new_game")
            self._game_index += 1
            saved_rand = v.initial_rand_steps
            v.initial_rand_steps = self.test_rand_steps                 # :372
            try:
                v.refresh_outputs()
                P = self.protagonist.clone()
                hold = torch.full((B,), _cabi.ACTION_HOLD, dtype=torch.int64, device=self.device)
                ones = torch.ones((B, 1), dtype=torch.float32, device=self.device)
                # The reference's workers run at their own pace: each plays the scripted side's plies on its own and
                # only then asks the learner (ppo_run_self_play.py:381-387), and Envs.test (envs.py:233-250) serves
                # one request of EVERY worker per policy.act.  In lockstep: while some game waits for the scripted
                # side, those games move and the others are held (OTH_ACTION_HOLD); once every running game waits
                # for the learner, one act on the whole batch (row i = env i, as in the reference) moves them all.
                # About half as many forward passes as one per ply.
                for rounds in range(N * N + 1):                          # every round places a disc in >= 1 running game
                    if rounds == N * N:
                        raise RuntimeError("test games did not finish")
                    live = ~v.terminated()
                    theirs = live & (v.player_turn != P)
                    n_theirs, n_live = (int(x) for x in torch.stack((theirs.sum(), live.sum())).tolist())   # one sync
                    if n_live == 0:
                        break
                    if n_theirs > 0:
                        moved, sign = theirs, -1.0
                        v.step(torch.where(theirs, v.policy_actions(name), hold))
                    else:
                        with torch.no_grad():
                            _, actions, _, _ = policy.act(v.obs, ex_hiddens, ones, self._choices_arg(v.legal_mask))
                        self.test_policy_calls += 1
                        moved, sign = live, 1.0
                        v.step(torch.where(live, self._as_dev(actions, torch.int64), hold))
                    wins += (moved & v.done & (v.reward * sign > 0)).to(torch.int64)   # :432-439
            finally:
                v.initial_rand_steps = saved_rand                        # :440
        return per * B, int(wins.sum().item())


class Envs(_RuleBaseGames):
    """envs.py:7-41.  `envs`: an int (number of lockstep envs) or, drop-in style, a list of
    env objects whose first element supplies board_size / initial_rand_steps /
    num_disk_as_reward / rand_seed / test_rand_steps_holder (ppo_run_self_play.py:99-109);
    `subproc_worker` is accepted and ignored (there are no worker processes)."""

    def __init__(self, envs, subproc_worker=None, policy=None, device=None, board_size=8, initial_rand_steps=0,
                 num_disk_as_reward=True, seed=0, test_rand_steps=10, choices_as="mask", legacy_stale_state=True,
                 env_id_base=0):
        if isinstance(envs, int):
            num = envs
        else:
            envs = list(envs)
            num = len(envs)
            e0 = envs[0]
            board_size = getattr(e0, "board_size", board_size)
            initial_rand_steps = getattr(e0, "initial_rand_steps", initial_rand_steps)
            inner = getattr(e0, "env", e0)
            num_disk_as_reward = getattr(inner, "num_disk_as_reward", num_disk_as_reward)
            seed = getattr(e0, "rand_seed", seed)
            test_rand_steps = getattr(e0, "test_rand_steps_holder", test_rand_steps)
        if choices_as not in ("mask", "lists"):
            raise ValueError("choices_as must be 'mask' or 'lists'")
        self.num_processes = int(num)
        self.policy = policy
        self.choices_as = choices_as
        self.test_rand_steps = int(test_rand_steps)
        self.legacy_stale_state = bool(legacy_stale_state)
        self.vec = VecOthelloEnv(self.num_processes, board_size=board_size, device=device, seed=seed,
                                 initial_rand_steps=initial_rand_steps, sudden_death_on_invalid_move=True,
                                 num_disk_as_reward=num_disk_as_reward, obs_layout="planes4",
                                 obs_dtype=torch.float32, auto_reset=False, env_id_base=env_id_base)
        self.device = self.vec.device
        self._lib = self.vec._lib
        S, N, B, W, dev = self.vec.board_size, self.vec.num_cells, self.num_processes, self.vec.words, self.device
        self.board_size = S
        self.observation_shape = (4, S, S)        # spaces.Box(zeros((4,S,S)), ones(...)), envs.py:16
        self.num_actions = N                      # spaces.Discrete(S**2), envs.py:17
        z = lambda shape, dt: torch.zeros(shape, dtype=dt, device=dev)   # noqa: E731
        self.protagonist = z((B,), torch.int8)
        self._phase = z((B,), torch.uint8)
        self._sav = {k: z((W, B), torch.int64) for k in ("black", "white", "legal", "choices")}
        self._sav_meta = z((B,), torch.int32)
        self._tr_action = z((B,), torch.int64)
        self._tr_reward = z((B,), torch.float32)
        self._tr_done = z((B,), torch.uint8)
        self._tr_type = z((B,), torch.uint8)
        self._tr_outputs = z((B, 3), torch.float32)
        self._pend_outputs = z((B, 3), torch.float32)
        self._pend_action = z((B,), torch.int64)
        self._need = z((1,), torch.int32)
        self._tr_obs = z((B, 4, S, S), torch.float32)
        self._tr_choices_u8 = z((B, N), torch.uint8)
        self._dummy_ctr = z((B,), torch.int32)
        self._game_index = 0
        self._c_sp = _cabi.OthSelfPlay(*[_cabi.ptr(t) for t in (
            self.protagonist, self._phase, self._sav["black"], self._sav["white"], self._sav["legal"], self._sav_meta,
            self._sav["choices"], self._tr_action, self._tr_reward, self._tr_done, self._tr_type, self._tr_outputs,
            self._pend_outputs, self._pend_action)])
        # the saved decision states seen as env states, for oth_observe
        self._c_sav_state = _cabi.OthState(_cabi.ptr(self._sav["black"]), _cabi.ptr(self._sav["white"]),
                                           _cabi.ptr(self._sav["legal"]), _cabi.ptr(self._sav_meta),
                                           _cabi.ptr(self._dummy_ctr))
        self._c_choice_state = _cabi.OthState(_cabi.ptr(self._sav["black"]), _cabi.ptr(self._sav["white"]),
                                              _cabi.ptr(self._sav["choices"]), _cabi.ptr(self._sav_meta),
                                              _cabi.ptr(self._dummy_ctr))
        self._c_tr_obs_out = _cabi.OthOutputs(_cabi.ptr(self._tr_obs), None, None, None, None, None, None,
                                              _cabi.OBS_PLANES4, _cabi.DTYPE_F32)
        self._c_tr_choice_out = _cabi.OthOutputs(None, _cabi.ptr(self._tr_choices_u8), None, None, None, None, None,
                                                 _cabi.OBS_NONE, _cabi.DTYPE_I8)
        self.policy_calls = 0
        self.test_policy_calls = 0            # policy.act calls made by test(): one per learner ply, not one per ply

    # ------------------------------------------------------------------ envs.py:27-35
    def over(self):
        """envs.py:27-29 tells the workers to exit; nothing to tear down here."""

    def reset(self, protagonists=None):
        """envs.py:31-35 + the worker's 'reset' (ppo_run_self_play.py:261-270): new game in
        every env, protagonist colour drawn per env (or given: int8/long tensor of -1/+1)."""
        v = self.vec
        pin = None
        if protagonists is not None:
            pin = torch.as_tensor(protagonists).to(device=self.device, dtype=torch.int8).reshape(-1).contiguous()
            if pin.numel() != self.num_processes:
                raise ValueError("protagonists must have one entry per env")
        with v._on_device():
            _cabi.check(self._lib.oth_selfplay_new_game(v._state(), self._c_sp, self.num_processes, v.board_size,
                                                        v.seed_value, v.env_id_base, v.initial_rand_steps,
                                                        self._game_index, _cabi.ptr(pin), v._stream()),
                        "oth_selfplay_new_game")
        self._game_index += 1

    def step(self, *args):
        raise NotImplementedError()

    def test(self, *args):
        raise NotImplementedError()

    # ------------------------------------------------------------------ helpers
    def _flags(self):
        return self.vec._flags() | (0 if self.legacy_stale_state else _cabi.F_SP_FRESH_STATE)

    def _choices_arg(self, mask):
        if self.choices_as == "mask":
            return mask
        m = mask.cpu()
        return [torch.nonzero(r).flatten().tolist() for r in m]

    def _as_dev(self, x, dtype):
        t = torch.as_tensor(x)
        return t.to(device=self.device, dtype=dtype).reshape(-1).contiguous()


class PPOEnvs(Envs):
    """envs.py:188-257 (self-play)."""

    def step(self, ex_hiddens):
        v = self.vec
        B = self.num_processes
        flags = self._flags()
        with v._on_device():
            _cabi.check(self._lib.oth_selfplay_begin(v._state(), self._c_sp, B, v.board_size, flags,
                                                     v._outputs(with_scalars=False), _cabi.ptr(self._need),
                                                     v._stream()), "oth_selfplay_begin")
            need = int(self._need.item())
            while need > 0:                                              # envs.py:200
                emitted = (self._phase & _cabi.SP_EMITTED) != 0
                masks = torch.where(emitted, 1.0 - self._tr_done.float(), torch.ones((), device=self.device)).unsqueeze(1)
                with torch.no_grad():
                    value, actions, logp, hidden = self.policy.act(v.obs, ex_hiddens, masks,
                                                                   self._choices_arg(v.legal_mask))   # envs.py:203
                self.policy_calls += 1
                a = self._as_dev(actions, torch.int64)
                val, lp, hid = (self._as_dev(x, torch.float32) for x in (value, logp, hidden))
                if a.numel() != B or val.numel() != B or lp.numel() != B or hid.numel() != B:
                    raise ValueError("policy.act must return one value / action / log-prob / hidden per env")
                _cabi.check(self._lib.oth_selfplay_advance(v._state(), self._c_sp, B, v.board_size, flags, _cabi.ptr(a),
                                                           _cabi.ptr(val), _cabi.ptr(lp), _cabi.ptr(hid), v.seed_value,
                                                           v.env_id_base, v.initial_rand_steps,
                                                           v._outputs(with_scalars=False), _cabi.ptr(self._need),
                                                           v._stream()), "oth_selfplay_advance")
                need = int(self._need.item())
            # expand the tuples: observation planes from the saved decision state, choices from its bitset
            zf = (flags | _cabi.F_ZERO_IF_TERMINATED) & ~_cabi.F_AUTO_RESET
            _cabi.check(self._lib.oth_observe(self._c_sav_state, B, v.board_size, zf, self._c_tr_obs_out, v._stream()),
                        "oth_observe")
            _cabi.check(self._lib.oth_observe(self._c_choice_state, B, v.board_size, zf, self._c_tr_choice_out,
                                              v._stream()), "oth_observe")
        dones = self._tr_done.view(torch.bool)
        masks = (1.0 - self._tr_done.float()).unsqueeze(1)               # envs.py:198
        bad_masks = torch.ones((B, 1), dtype=torch.float32, device=self.device)   # envs.py:199: never 'bad_transition'
        infos = SelfPlayInfos(self._tr_type, self._tr_choices_u8.view(torch.bool))
        return (self._tr_obs, self._tr_action.float().unsqueeze(1), self._tr_reward.unsqueeze(1), dones, infos,
                self._tr_outputs, masks, bad_masks)

    def test(self, name, num_games, ex_hiddens, protagonists=None):
        """envs.py:229-257 + rule_base_game (ppo_run_self_play.py:371-441): the learner plays
        `num_games // num_envs` games per env against 'rand' | 'greedy' with the test opening
        budget; a win = final reward from the learner's side > 0.  -> (games, wins)."""
        return self._rule_base_games(self.policy, name, num_games, ex_hiddens, protagonists)


class WhoList(object):
    """``t_or_s`` of PPOTeacherStudentEnvs.step: uint8 CUDA tensor (0 = teacher, 1 = student) that iterates
    as the reference's list of 'teacher' / 'student' strings (envs.py:122-134)."""

    def __init__(self, who_u8):
        self.who = who_u8

    @property
    def is_teacher(self):
        return self.who == 0

    def __len__(self):
        return self.who.numel()

    def __getitem__(self, i):
        return "teacher" if int(self.who[i].item()) == 0 else "student"

    def __iter__(self):
        for w in self.who.cpu().tolist():
            yield "teacher" if w == 0 else "student"


class TeacherStudentInfos(SelfPlayInfos):
    """infos of one PPOTeacherStudentEnvs.step: ``.type`` uint8 [B] (0 = 'black_transition', 1 =
    'white_transition', 2 = 'over'), ``.choices_mask`` bool [B, S*S]; dict access builds the reference's
    {'type', 'choices'} lazily."""
    _NAMES = ("black_transition", "white_transition", "over")

    def __getitem__(self, i):
        return {"type": self._NAMES[int(self.type[i].item())],
                "choices": torch.nonzero(self.choices_mask[i]).flatten().tolist()}

    def __iter__(self):
        types = self.type.cpu().tolist()
        masks = self.choices_mask.cpu()
        for t, m in zip(types, masks):
            yield {"type": self._NAMES[t], "choices": torch.nonzero(m).flatten().tolist()}


class PPOTeacherStudentEnvs(Envs):
    """envs.py:44-184 (teacher vs student) + the worker othello_teacher_vs_student
    (ppo_run_teacher_vs_student.py:425-573), on the device.

    Both colours learn: every ``step`` returns, per env, ONE tuple -- a black or a white transition (or
    'over') -- labelled teacher / student by colour.  Inside a call the reference serves all pending
    ``need_black_action`` requests before the white ones (envs.py:78) with ``policy[turn].act`` on the whole
    batch; here each such act is followed by one ``oth_ts_advance`` launch for that colour.  The tensors
    returned are bit-identical to the live reference's (tests/test_gpu_teacher_student.py, golden from real
    forked workers).  Tensor-typed returns as in PPOEnvs: ``t_or_s`` is a WhoList, ``dones`` a bool tensor,
    ``infos`` a TeacherStudentInfos."""

    def __init__(self, envs, subproc_worker=None, teacher_policy=None, student_policy=None, device=None, **kw):
        super(PPOTeacherStudentEnvs, self).__init__(envs, subproc_worker, teacher_policy, device, **kw)
        self.teacher_policy = teacher_policy
        self.student_policy = student_policy
        self.teacher = None
        self.student = None
        self.policy = {}
        self.int_to_str = {-1: "black", 1: "white"}
        self._teacher_done_reward = 0.0
        S, N, B, W, dev = self.board_size, self.num_actions, self.num_processes, self.vec.words, self.device
        z = lambda shape, dt: torch.zeros(shape, dtype=dt, device=dev)   # noqa: E731
        self._ts_phase = z((B,), torch.uint8)
        self._ts_sav = {k: z((2, W, B), torch.int64) for k in ("black", "white", "legal", "choices")}
        self._ts_sav_meta = z((2, B), torch.int32)
        self._ts_sav_action = z((2, B), torch.int64)
        self._ts_sav_outputs = z((2, B, 3), torch.float32)
        self._ts_queued_reward = z((B,), torch.int32)
        self._ts_tr = {k: z((W, B), torch.int64) for k in ("black", "white", "legal", "choices")}
        self._ts_tr_meta = z((B,), torch.int32)
        self._ts_tr_action = z((B,), torch.int64)
        self._ts_tr_reward = z((B,), torch.float32)
        self._ts_tr_done = z((B,), torch.uint8)
        self._ts_tr_type = z((B,), torch.uint8)
        self._ts_tr_who = z((B,), torch.uint8)
        self._ts_tr_outputs = z((B, 3), torch.float32)
        self._ts_need = z((2,), torch.int32)
        self._c_ts = _cabi.OthTeacherStudent(*[_cabi.ptr(t) for t in (
            self._ts_phase, self._ts_sav["black"], self._ts_sav["white"], self._ts_sav["legal"], self._ts_sav_meta,
            self._ts_sav["choices"], self._ts_sav_action, self._ts_sav_outputs, self._ts_queued_reward,
            self._ts_tr["black"], self._ts_tr["white"], self._ts_tr["legal"], self._ts_tr["choices"], self._ts_tr_meta,
            self._ts_tr_action, self._ts_tr_reward, self._ts_tr_done, self._ts_tr_type, self._ts_tr_who,
            self._ts_tr_outputs)])
        self._c_ts_state = _cabi.OthState(_cabi.ptr(self._ts_tr["black"]), _cabi.ptr(self._ts_tr["white"]),
                                          _cabi.ptr(self._ts_tr["legal"]), _cabi.ptr(self._ts_tr_meta),
                                          _cabi.ptr(self._dummy_ctr))
        self._c_ts_choice_state = _cabi.OthState(_cabi.ptr(self._ts_tr["black"]), _cabi.ptr(self._ts_tr["white"]),
                                                 _cabi.ptr(self._ts_tr["choices"]), _cabi.ptr(self._ts_tr_meta),
                                                 _cabi.ptr(self._dummy_ctr))
        self.policy_calls = []

    def reset(self, teacher, win_avg, last_win_avg):
        """envs.py:54-61 + the worker's 'reset' (:444-455).  teacher = -1 | +1."""
        teacher = int(teacher)
        if teacher not in (-1, 1):
            raise ValueError("teacher must be -1 (black) or +1 (white)")
        self.teacher = teacher
        self.student = -teacher
        self.policy = {self.teacher: self.teacher_policy, self.student: self.student_policy}
        r = 0
        for k in win_avg.keys():                                         # :459-461, same accumulation order
            r += win_avg[k] - last_win_avg[k]
        self._teacher_done_reward = float(r)
        v = self.vec
        with v._on_device():
            _cabi.check(self._lib.oth_ts_new_game(v._state(), self._c_ts, self.num_processes, v.board_size,
                                                  v.seed_value, v.env_id_base, v.initial_rand_steps, v._stream()),
                        "oth_ts_new_game")

    def step(self, ex_hiddens_student, ex_hiddens_teacher):
        if self.teacher is None:
            raise RuntimeError("call reset(teacher, win_avg, last_win_avg) first")
        v = self.vec
        B = self.num_processes
        flags = self._flags()
        tdr = self._teacher_done_reward
        self.policy_calls = []
        with v._on_device():
            _cabi.check(self._lib.oth_ts_begin(v._state(), self._c_ts, B, v.board_size, flags, self.teacher, tdr,
                                               v._outputs(with_scalars=False), _cabi.ptr(self._ts_need), v._stream()),
                        "oth_ts_begin")
            need_b, need_w = (int(x) for x in self._ts_need.tolist())
            while need_b > 0 or need_w > 0:                              # envs.py:76
                turn = -1 if need_b > 0 else 1                           # envs.py:78
                ex_hiddens = ex_hiddens_teacher if turn == self.teacher else ex_hiddens_student
                emitted = (self._ts_phase & _cabi.TS_EMITTED) != 0
                masks = torch.where(emitted, 1.0 - self._ts_tr_done.float(),
                                    torch.ones((), device=self.device)).unsqueeze(1)
                with torch.no_grad():
                    value, actions, logp, hidden = self.policy[turn].act(v.obs, ex_hiddens, masks,
                                                                         self._choices_arg(v.legal_mask))   # :83
                self.policy_calls.append(turn)
                a = self._as_dev(actions, torch.int64)
                val, lp, hid = (self._as_dev(x, torch.float32) for x in (value, logp, hidden))
                if a.numel() != B or val.numel() != B or lp.numel() != B or hid.numel() != B:
                    raise ValueError("policy.act must return one value / action / log-prob / hidden per env")
                _cabi.check(self._lib.oth_ts_advance(v._state(), self._c_ts, B, v.board_size, flags, self.teacher, tdr,
                                                     turn, _cabi.ptr(a), _cabi.ptr(val), _cabi.ptr(lp), _cabi.ptr(hid),
                                                     v.seed_value, v.env_id_base, v.initial_rand_steps,
                                                     v._outputs(with_scalars=False), _cabi.ptr(self._ts_need),
                                                     v._stream()), "oth_ts_advance")
                need_b, need_w = (int(x) for x in self._ts_need.tolist())
            zf = (flags | _cabi.F_ZERO_IF_TERMINATED) & ~_cabi.F_AUTO_RESET
            _cabi.check(self._lib.oth_observe(self._c_ts_state, B, v.board_size, zf, self._c_tr_obs_out, v._stream()),
                        "oth_observe")
            _cabi.check(self._lib.oth_observe(self._c_ts_choice_state, B, v.board_size, zf, self._c_tr_choice_out,
                                              v._stream()), "oth_observe")
        dones = self._ts_tr_done.view(torch.bool)
        masks = (1.0 - self._ts_tr_done.float()).unsqueeze(1)            # envs.py:99
        bad_masks = torch.ones((B, 1), dtype=torch.float32, device=self.device)   # envs.py:100
        infos = TeacherStudentInfos(self._ts_tr_type, self._tr_choices_u8.view(torch.bool))
        return (WhoList(self._ts_tr_who), self._tr_obs, self._ts_tr_action.float().unsqueeze(1),
                self._ts_tr_reward.unsqueeze(1), dones, infos, self._ts_tr_outputs, masks, bad_masks)

    def test(self, name, num_games, ex_hiddens, teacher=False, protagonists=None):
        """envs.py:142-184: rule_base_game with the teacher's or the student's network.  -> (games, wins)"""
        policy = self.teacher_policy if teacher else self.student_policy
        return self._rule_base_games(policy, name, num_games, ex_hiddens, protagonists)
