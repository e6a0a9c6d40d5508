"""Legal-move-restricted action sampling on the device.

Replaces the per-env Python loops of the reference learner's ``Policy.act`` /
``Policy.evaluate_actions`` (``pytorch_a2c_ppo_acktr_gail/a2c_ppo_acktr/model.py:60-150,
156-198``): ``FixedCategorical(logits=logits[i][choices[i]])`` -> sample / mode / log_probs,
one env at a time on the host.  Here one kernel (``oth_masked_categorical``, a warp per
env) reads the learner's logits [B, S*S] and the env's legal-move BITSET (no dense mask,
no lists) and returns the action and its log-probability under softmax(logits[legal]).
Edge cases follow the reference: no legal move -> action 0, log-prob 0 (:71-73); in
evaluate mode an action that is not a legal choice gets log-prob 0 (:172-173).

``masked_categorical`` (sample / mode / evaluate) is the ROLLOUT side and carries no gradient.  The PPO update's
``evaluate_actions`` (differentiated at ``algo/ppo.py:51-66``) is ``evaluate_actions`` below: an autograd function
with a hand-written backward kernel plus the full-distribution entropy; ``ChoiceStore`` keeps the rollout's
`choices` as device bitsets (``storage.py:19,48-57,133-145``).
"""
import ctypes as C

import torch

from . import _cabi

_MODES = {"sample": _cabi.CAT_SAMPLE, "mode": _cabi.CAT_MODE, "evaluate": _cabi.CAT_EVALUATE}


def masked_categorical(logits, legal_words, board_size, mode="sample", actions=None, seed=0, counter=0, env_id_base=0):
    """logits: float32 CUDA [B, S*S]; legal_words: int64 CUDA [W, B] bitsets (VecOthelloEnv._legal layout:
    bit a of word a//64 = action a).  -> (actions int64 [B], log_probs float32 [B])."""
    if mode not in _MODES:
        raise ValueError("mode must be sample | mode | evaluate")
    if not (isinstance(logits, torch.Tensor) and logits.is_cuda):
        raise _cabi.OthelloB200Error("masked_categorical needs CUDA tensors (no CPU path)")
    lib = _cabi.load()
    S = int(board_size)
    B = logits.shape[0]
    if logits.dim() != 2 or logits.shape[1] != S * S:
        raise ValueError("logits must be [B, board_size**2]")
    logits = logits.detach().to(torch.float32).contiguous()
    legal_words = legal_words.contiguous()
    if legal_words.dtype != torch.int64 or tuple(legal_words.shape) != (_cabi.words_per_board(S), B):
        raise ValueError("legal_words must be int64 [W, B]")
    a_in = None
    if mode == "evaluate":
        if actions is None:
            raise ValueError("evaluate mode needs actions")
        a_in = actions.to(device=logits.device, dtype=torch.int64).reshape(-1).contiguous()
        if a_in.numel() != B:
            raise ValueError("actions must have one entry per env")
    out_a = torch.empty((B,), dtype=torch.int64, device=logits.device)
    out_lp = torch.empty((B,), dtype=torch.float32, device=logits.device)
    with torch.cuda.device(logits.device):
        stream = C.c_void_p(torch.cuda.current_stream(logits.device).cuda_stream)
        _cabi.check(lib.oth_masked_categorical(_cabi.ptr(logits), _cabi.ptr(legal_words), B, S, _MODES[mode],
                                               _cabi.ptr(a_in), int(seed) & 0xFFFFFFFFFFFFFFFF, int(counter),
                                               int(env_id_base), _cabi.ptr(out_a), _cabi.ptr(out_lp), stream),
                    "oth_masked_categorical")
    return out_a, out_lp


def _stream(dev):
    return C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)


class _MaskedLogProb(torch.autograd.Function):
    """(log-prob of `actions` under softmax over the legal subset, entropy of the full distribution) with a
    hand-written backward kernel (oth_masked_logprob_forward / _backward)."""

    @staticmethod
    def forward(ctx, logits, legal_words, actions, board_size):
        lib = _cabi.load()
        B = logits.shape[0]
        x = logits.detach().to(torch.float32).contiguous()
        logp = torch.empty((B,), dtype=torch.float32, device=x.device)
        ent = torch.empty((B,), dtype=torch.float32, device=x.device)
        with torch.cuda.device(x.device):
            _cabi.check(lib.oth_masked_logprob_forward(_cabi.ptr(x), _cabi.ptr(legal_words), B, int(board_size),
                                                       _cabi.ptr(actions), _cabi.ptr(logp), _cabi.ptr(ent),
                                                       _stream(x.device)), "oth_masked_logprob_forward")
        ctx.save_for_backward(x, legal_words, actions)
        ctx.board_size = int(board_size)
        ctx.in_dtype = logits.dtype
        return logp, ent

    @staticmethod
    def backward(ctx, g_logp, g_ent):
        x, legal_words, actions = ctx.saved_tensors
        lib = _cabi.load()
        B = x.shape[0]
        grad = torch.empty_like(x)
        gl = None if g_logp is None else g_logp.to(torch.float32).contiguous()
        ge = None if g_ent is None else g_ent.to(torch.float32).contiguous()
        with torch.cuda.device(x.device):
            _cabi.check(lib.oth_masked_logprob_backward(_cabi.ptr(x), _cabi.ptr(legal_words), B, ctx.board_size,
                                                        _cabi.ptr(actions), _cabi.ptr(gl), _cabi.ptr(ge),
                                                        _cabi.ptr(grad), _stream(x.device)),
                        "oth_masked_logprob_backward")
        return grad.to(ctx.in_dtype), None, None, None


def evaluate_actions(logits, legal_words, actions, board_size):
    """The distribution part of ``Policy.evaluate_actions`` (model.py:156-198) for a whole minibatch, differentiable
    w.r.t. ``logits``: -> (action_log_probs float32 [B, 1], dist_entropy scalar).

    logits float32 CUDA [B, S*S] (the actor head's output, requires_grad); legal_words int64 CUDA [W, B]: the
    `choices` of every sample as bitsets (``ChoiceStore.gather``); actions int64 [B] or [B, 1].
    action_log_probs[i] is the log-probability of actions[i] under softmax(logits[i][choices_i]); it is the
    constant 0 -- with zero gradient -- when actions[i] is not among the choices or there are none (:167-168).
    dist_entropy is the mean entropy of the FULL S*S-way distribution, as the reference computes it (:159,178)."""
    if not (isinstance(logits, torch.Tensor) and logits.is_cuda):
        raise _cabi.OthelloB200Error("evaluate_actions needs CUDA tensors (no CPU path)")
    S = int(board_size)
    B = logits.shape[0]
    if logits.dim() != 2 or logits.shape[1] != S * S:
        raise ValueError("logits must be [B, board_size**2]")
    legal_words = legal_words.contiguous()
    if legal_words.dtype != torch.int64 or tuple(legal_words.shape) != (_cabi.words_per_board(S), B):
        raise ValueError("legal_words must be int64 [W, B]")
    a = actions.to(device=logits.device, dtype=torch.int64).reshape(-1).contiguous()
    if a.numel() != B:
        raise ValueError("actions must have one entry per sample")
    logp, ent = _MaskedLogProb.apply(logits, legal_words, a, S)
    return logp.unsqueeze(1), ent.mean()


class ChoiceStore(object):
    """The `choices` of a PPO rollout on the device: int64 bitsets [T, W, B] instead of RolloutStorage's list of lists
    (storage.py:19 `self.choices`, :48-57 `insert`), and the minibatch gather of :133-145 as ONE indexed copy."""

    def __init__(self, num_steps, num_envs, board_size, device):
        self.num_steps, self.num_envs, self.board_size = int(num_steps), int(num_envs), int(board_size)
        self.words = _cabi.words_per_board(self.board_size)
        self.data = torch.zeros((self.num_steps, self.words, self.num_envs), dtype=torch.int64, device=device)
        self.step = 0

    def insert(self, legal_words):
        """storage.py:56: the legal set [W, B] of the state the step's action was chosen in (None = no choices)."""
        if legal_words is None:
            self.data[self.step].zero_()
        else:
            self.data[self.step].copy_(legal_words)
        self.step = (self.step + 1) % self.num_steps                     # storage.py:64

    def gather(self, indices):
        """choices_batch for flat sample indices i = t * num_envs + b (the order of `.view(-1, ...)` in
        feed_forward_generator, storage.py:127-145) -> int64 [W, len(indices)], ready for evaluate_actions."""
        idx = torch.as_tensor(indices, dtype=torch.int64, device=self.data.device)
        t = torch.div(idx, self.num_envs, rounding_mode="floor")
        b = idx - t * self.num_envs
        return self.data[t, :, b].t().contiguous()               # [M, W] gathered rows -> [W, M]; the store is not copied

    def as_lists(self, indices):
        """The reference's representation (lists of ascending action ids), for tests and debugging."""
        words = self.gather(indices).cpu().tolist()
        out = []
        for j in range(len(words[0])):
            moves = []
            for w in range(self.words):
                v = words[w][j] & 0xFFFFFFFFFFFFFFFF
                while v:
                    low = v & -v
                    moves.append(w * 64 + low.bit_length() - 1)
                    v ^= low
            out.append(moves)
        return out


class LegalMoveSampler(object):
    """Stateful convenience over masked_categorical for one VecOthelloEnv: every sample() call
    uses the next Philox counter, so a rollout is reproducible from (seed, env ids)."""

    def __init__(self, env, seed=None):
        self.env = env
        self.seed = env.seed_value if seed is None else int(seed)
        self.counter = 0

    def sample(self, logits):
        a, lp = masked_categorical(logits, self.env._legal, self.env.board_size, "sample", seed=self.seed,
                                   counter=self.counter, env_id_base=self.env.env_id_base)
        self.counter += 1
        return a, lp

    def mode(self, logits):
        return masked_categorical(logits, self.env._legal, self.env.board_size, "mode")

    def log_probs(self, logits, actions, legal_words=None):
        """Log-probabilities of `actions` WITHOUT a gradient path (rollout bookkeeping).  The PPO update uses
        `evaluate_actions`, which is differentiable."""
        lw = self.env._legal if legal_words is None else legal_words
        return masked_categorical(logits, lw, self.env.board_size, "evaluate", actions=actions)[1]

    def evaluate_actions(self, logits, actions, legal_words=None):
        lw = self.env._legal if legal_words is None else legal_words
        return evaluate_actions(logits, lw, actions, self.env.board_size)


def epsilon_greedy(q_values, legal_words, board_size, epsilon, seed=0, counter=0, env_id_base=0):
    """The DQN agent's action choice (reference ``dqn.py:264-275`` / ``:478-488``) for a whole batch: with
    probability ``epsilon`` a uniformly random legal move, else the legal move with the largest q (first
    occurrence on ties).  q_values: float32 CUDA [B, S*S]; legal_words: int64 CUDA [W, B].
    -> (actions int64 [B], q_chosen float32 [B], explored bool [B])."""
    if not (isinstance(q_values, torch.Tensor) and q_values.is_cuda):
        raise _cabi.OthelloB200Error("epsilon_greedy needs CUDA tensors (no CPU path)")
    lib = _cabi.load()
    S = int(board_size)
    B = q_values.shape[0]
    if q_values.dim() != 2 or q_values.shape[1] != S * S:
        raise ValueError("q_values must be [B, board_size**2]")
    if not 0.0 <= float(epsilon) <= 1.0:
        raise ValueError("epsilon must be in [0, 1]")
    q_values = q_values.detach().to(torch.float32).contiguous()
    legal_words = legal_words.contiguous()
    if legal_words.dtype != torch.int64 or tuple(legal_words.shape) != (_cabi.words_per_board(S), B):
        raise ValueError("legal_words must be int64 [W, B]")
    out_a = torch.empty((B,), dtype=torch.int64, device=q_values.device)
    out_q = torch.empty((B,), dtype=torch.float32, device=q_values.device)
    out_e = torch.empty((B,), dtype=torch.uint8, device=q_values.device)
    with torch.cuda.device(q_values.device):
        stream = C.c_void_p(torch.cuda.current_stream(q_values.device).cuda_stream)
        _cabi.check(lib.oth_epsilon_greedy(_cabi.ptr(q_values), _cabi.ptr(legal_words), B, S, float(epsilon),
                                           int(seed) & 0xFFFFFFFFFFFFFFFF, int(counter), int(env_id_base),
                                           _cabi.ptr(out_a), _cabi.ptr(out_q), _cabi.ptr(out_e), stream),
                    "oth_epsilon_greedy")
    return out_a, out_q, out_e.bool()


class EpsilonGreedy(object):
    """The exploration schedule of the reference DQN agent (``dqn.py:144-146,175,196-201,264-286,403``) over
    one VecOthelloEnv: epsilon starts at ``initial_epsilon``; while ``t < initial_replay_size`` every action is
    random (``:266``); afterwards epsilon drops by ``(initial - final) / anealing_steps`` per get_action call
    until it reaches ``final_epsilon`` (``:283-284``).  The guard is the reference's ``epsilon > final_epsilon``
    on the accumulated float, so epsilon can end one decrement BELOW ``final_epsilon`` (1.0 -> 0.1 over 10 steps
    settles at 0.01); that is kept, not clamped.  ``t`` counts agent steps (``:403``): one per lockstep
    call here, as the reference counts one per env step of its single env."""

    def __init__(self, env, anealing_steps=1000000, initial_epsilon=1.0, final_epsilon=0.1,
                 initial_replay_size=20000, test_epsilon=0.05, seed=None):
        self.env = env
        self.epsilon = float(initial_epsilon)
        self.final_epsilon = float(final_epsilon)
        self.anealing_by_step = (float(initial_epsilon) - float(final_epsilon)) / anealing_steps
        self.initial_replay_size = int(initial_replay_size)
        self.test_epsilon = float(test_epsilon)
        self.seed = env.seed_value if seed is None else int(seed)
        self.t = 0
        self.counter = 0

    def _choose(self, q_values, eps):
        out = epsilon_greedy(q_values, self.env._legal, self.env.board_size, eps, seed=self.seed,
                             counter=self.counter, env_id_base=self.env.env_id_base)
        self.counter += 1
        return out

    def get_action(self, q_values):
        """dqn.py:264-286.  -> (actions, q_chosen, explored); call ``tick()`` once per stored transition."""
        eps = 1.0 if self.t < self.initial_replay_size else min(1.0, max(0.0, self.epsilon))
        out = self._choose(q_values, eps)
        if self.epsilon > self.final_epsilon and self.t >= self.initial_replay_size:
            self.epsilon -= self.anealing_by_step
        return out

    def get_action_at_test(self, q_values):
        """dqn.py:478-488: epsilon fixed at 0.05."""
        return self._choose(q_values, self.test_epsilon)

    def tick(self, n=1):
        """dqn.py:403 ``self.t += 1`` (the reference advances t in ``run``, once per transition)."""
        self.t += int(n)
