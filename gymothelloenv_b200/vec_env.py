"""VecOthelloEnv -- lockstep batch of Othello envs resident on one B200.

Replaces the reference's subprocess vec-env (``envs.py:7-41,188-257``: N forked workers,
one ``SimpleOthelloEnv`` each, pickled pipes) by B envs stepped by ONE kernel launch; obs,
reward, done and the legal-action mask are torch CUDA tensors that never leave the GPU.
Per-env semantics are those of ``SimpleOthelloEnv`` over ``OthelloBaseEnv``
(``othello.py:21-93,217-501``): optional random opening plies, sudden death / forfeited
turn on invalid moves, automatic passes, both reward modes, the cached (possibly stale)
``possible_moves`` set.

All compute is in libothello_b200.so (hand-written sm_100a CUDA, see csrc/); there is no
CPU fallback -- constructing this class without a CUDA device raises.
"""
import ctypes as C

import torch

from . import _cabi

_LAYOUTS = {
    "none": _cabi.OBS_NONE,
    "board": _cabi.OBS_BOARD,             # (B,S,S)   mover=+1, opponent=-1      othello.py:363-369
    "board+mask": _cabi.OBS_BOARD_MASK,   # (B,2,S,S) + cached legal grid         othello.py:370-376
    "planes4": _cabi.OBS_PLANES4,         # (B,4,S,S) black, white, turn, legal   util.py:48-74
    "planes3": _cabi.OBS_PLANES3,         # (B,3,S,S) black, white, turn          run_2agent.py:29-46
    "absolute": _cabi.OBS_ABSOLUTE,       # (B,S,S)   board_state: WHITE=+1       othello.py:10-12
}
_DTYPES = {torch.int8: _cabi.DTYPE_I8, torch.float32: _cabi.DTYPE_F32}
_POLICY_KINDS = {"rand": _cabi.POLICY_RANDOM, "random": _cabi.POLICY_RANDOM, "greedy": _cabi.POLICY_GREEDY,
                 "maximin": _cabi.POLICY_MAXIMIN, "alphabeta": _cabi.POLICY_ALPHABETA}
_ACT_SOURCES = {"random": _cabi.ACT_RANDOM, "rand": _cabi.ACT_RANDOM, "greedy": _cabi.ACT_GREEDY}


class _NullCtx(object):
    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False


_NULL_CTX = _NullCtx()


def policy_spec(kind, depth=1):
    if kind not in _POLICY_KINDS:
        raise ValueError("unknown policy kind %r (rand | greedy | maximin | alphabeta)" % (kind,))
    depth = int(depth)
    if _POLICY_KINDS[kind] >= _cabi.POLICY_MAXIMIN and not (1 <= depth <= _cabi.MAX_SEARCH_DEPTH):
        raise ValueError("search depth must be in 1..%d" % _cabi.MAX_SEARCH_DEPTH)
    return _cabi.OthPolicySpec(_POLICY_KINDS[kind], max(1, depth))


def require_cuda(device=None):
    if not torch.cuda.is_available():
        raise _cabi.OthelloB200Error("gymothelloenv_b200 needs a CUDA device (B200 / sm_100a); there is no CPU path")
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    if dev.type != "cuda":
        raise _cabi.OthelloB200Error("device must be a CUDA device, got %s" % dev)
    if dev.index is None:
        dev = torch.device("cuda", torch.cuda.current_device())
    return dev


def shard_range(total, rank, world_size):
    """Contiguous block of envs owned by `rank`: -> (first global id, count)."""
    total, rank, world_size = int(total), int(rank), int(world_size)
    if not (0 <= rank < world_size):
        raise ValueError("rank out of range")
    base, rem = divmod(total, world_size)
    count = base + (1 if rank < rem else 0)
    first = rank * base + min(rank, rem)
    return first, count


class VecOthelloEnv(object):
    def __init__(self, num_envs, board_size=8, device=None, seed=0, initial_rand_steps=0,
                 sudden_death_on_invalid_move=True, num_disk_as_reward=False, obs_layout="board",
                 obs_dtype=torch.int8, auto_reset=False, legacy_single_move_plane=True, env_id_base=0,
                 with_legal_mask=True):
        self.device = require_cuda(device)
        self._lib = _cabi.load()
        self.num_envs = int(num_envs)
        if self.num_envs < 1:
            raise ValueError("num_envs must be >= 1")
        self.board_size = max(_cabi.MIN_BOARD_SIZE, int(board_size))          # othello.py:230
        if self.board_size > _cabi.MAX_BOARD_SIZE:
            raise ValueError("board_size %d not supported by this build (max %d)" % (board_size, _cabi.MAX_BOARD_SIZE))
        self.num_cells = self.board_size ** 2
        self.words = _cabi.words_per_board(self.board_size)
        self.seed_value = int(seed) & 0xFFFFFFFFFFFFFFFF
        self.initial_rand_steps = int(initial_rand_steps)
        if not (0 <= self.initial_rand_steps <= 254):
            raise ValueError("initial_rand_steps must be in 0..254")
        self.env_id_base = int(env_id_base)
        self.sudden_death_on_invalid_move = bool(sudden_death_on_invalid_move)
        self.num_disk_as_reward = bool(num_disk_as_reward)
        self.auto_reset = bool(auto_reset)
        self.legacy_single_move_plane = bool(legacy_single_move_plane)
        if obs_layout not in _LAYOUTS:
            raise ValueError("obs_layout must be one of %s" % sorted(_LAYOUTS))
        if obs_dtype not in _DTYPES:
            raise ValueError("obs_dtype must be torch.int8 or torch.float32")
        self.obs_layout, self.obs_dtype = obs_layout, obs_dtype

        B, W, S, N, dev = self.num_envs, self.words, self.board_size, self.num_cells, self.device
        # ---- state (SoA) ----
        self._black = torch.zeros((W, B), dtype=torch.int64, device=dev)
        self._white = torch.zeros((W, B), dtype=torch.int64, device=dev)
        self._legal = torch.zeros((W, B), dtype=torch.int64, device=dev)
        self._meta = torch.zeros((B,), dtype=torch.int32, device=dev)
        self._ctr = torch.zeros((B,), dtype=torch.int32, device=dev)
        # ---- outputs, allocated once and overwritten by every step (like the reference's
        # aliased obs, othello.py:366) ----
        P = _cabi.OBS_PLANES[_LAYOUTS[obs_layout]]
        self.obs = None
        if P:
            shape = (B, S, S) if P == 1 else (B, P, S, S)
            self.obs = torch.zeros(shape, dtype=obs_dtype, device=dev)
        self._mask_u8 = torch.zeros((B, N), dtype=torch.uint8, device=dev) if with_legal_mask else None
        self.legal_mask = self._mask_u8.view(torch.bool) if with_legal_mask else None
        # the five per-env scalars live in ONE 16-byte record per env (oth_outputs.packed: one store per env instead
        # of five); reward / done / player_turn / winner / actions_taken are strided views of it
        self._packed = torch.zeros((B, 16), dtype=torch.uint8, device=dev)
        self._packed[:, 4] = 1                                                # done until the first reset()
        self.reward = self._packed.view(torch.float32)[:, 0]
        self._done_u8 = self._packed[:, 4]
        self.done = self._done_u8.view(torch.bool)
        self.player_turn = self._packed.view(torch.int8)[:, 5]
        self.winner = self._packed.view(torch.int8)[:, 6]
        self.actions_taken = self._packed.view(torch.int64)[:, 1]
        self._policy_counter = 0
        self._host = None
        self._c_state = None
        self._c_out = {}
        self.observation_shape = None if self.obs is None else tuple(self.obs.shape[1:])
        self.num_actions = N                                                  # Discrete(S*S), othello.py:245

    # ------------------------------------------------------------------ plumbing
    def _state(self):
        if self._c_state is None:   # tensors are allocated once: the struct can be cached
            self._c_state = _cabi.OthState(_cabi.ptr(self._black), _cabi.ptr(self._white), _cabi.ptr(self._legal),
                                           _cabi.ptr(self._meta), _cabi.ptr(self._ctr))
        return self._c_state

    def _flags(self):
        return ((_cabi.F_SUDDEN_DEATH if self.sudden_death_on_invalid_move else 0)
                | (_cabi.F_DISK_REWARD if self.num_disk_as_reward else 0)
                | (_cabi.F_AUTO_RESET if self.auto_reset else 0)
                | (_cabi.F_LEGACY_SINGLE_MOVE_PLANE if self.legacy_single_move_plane else 0))

    def _outputs(self, with_scalars=True):
        """oth_outputs of this env.  The scalars always travel through the packed record: oth_step writes all of it,
        oth_observe (with_scalars=False at the call sites) only its status bytes."""
        if "out" not in self._c_out:
            self._c_out["out"] = _cabi.OthOutputs(
                _cabi.ptr(self.obs), _cabi.ptr(self._mask_u8), None, None, None, None, None,
                _LAYOUTS[self.obs_layout], _DTYPES[self.obs_dtype], _cabi.ptr(self._packed))
        return self._c_out["out"]

    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def _on_device(self):
        """Context that makes self.device current (no-op object when it already is)."""
        if torch.cuda.current_device() == self.device.index:
            return _NULL_CTX
        return torch.cuda.device(self.device)

    # ------------------------------------------------------------------ Gym-like API
    def seed(self, seed):
        """SimpleOthelloEnv.seed (othello.py:53-58): new stream key and rewound stream position, effective AT ONCE:
        the key is an argument of every later step() (opening plies, fused random policy, auto-reset budgets), so
        games in progress continue on the new stream.  Call reset() next for a fresh, reproducible batch (the
        reference swaps its RandomState object the same way, mid-game included)."""
        self.seed_value = int(seed) & 0xFFFFFFFFFFFFFFFF
        self._ctr.zero_()

    HOLD = _cabi.ACTION_HOLD          # step(actions): the env sits this ply out

    def reset(self, where=None):
        """Resets every env (or those with where[i] True) and returns the observation tensor."""
        only = None
        if where is not None:
            only = where.to(device=self.device).view(-1).to(torch.uint8).contiguous()
            if only.numel() != self.num_envs:
                raise ValueError("where must have num_envs entries")
        with self._on_device():
            _cabi.check(self._lib.oth_reset(self._state(), self.num_envs, self.board_size, self.seed_value,
                                            self.env_id_base, self.initial_rand_steps, 0, _cabi.ptr(only),
                                            self._stream()), "oth_reset")
            self.reward.zero_()
            _cabi.check(self._lib.oth_observe(self._state(), self.num_envs, self.board_size, self._flags(),
                                              self._outputs(with_scalars=False), self._stream()), "oth_observe")
        return self.obs

    def step(self, actions=None, policy=None):
        """One ply in every env.

        actions: int64 CUDA tensor [B] (``env.step(a)``), or None with policy='random' |
        'greedy' to let the kernel play the side to move (fused policy + step).  An entry equal to
        ``VecOthelloEnv.HOLD`` (OTH_ACTION_HOLD) leaves that env alone for this ply: state kept, outputs
        re-emitted, reward 0, done False.
        Returns (obs, reward f32[B], done bool[B], info) with info = {'legal_mask',
        'player_turn', 'winner', 'actions'}; all CUDA tensors owned by the env and
        overwritten by the next call."""
        if actions is None:
            if policy not in _ACT_SOURCES:
                raise ValueError("step() needs actions or policy='random'|'greedy'")
            src, aptr = _ACT_SOURCES[policy], None
        else:
            if not (isinstance(actions, torch.Tensor) and actions.is_cuda):
                raise TypeError("actions must be a CUDA tensor (use step_host() for host buffers)")
            actions = actions.to(device=self.device, dtype=torch.int64).contiguous().view(-1)
            if actions.numel() != self.num_envs:
                raise ValueError("actions must have num_envs entries")
            src, aptr = _cabi.ACT_EXTERNAL, _cabi.ptr(actions)
        with self._on_device():
            _cabi.check(self._lib.oth_step(self._state(), self.num_envs, self.board_size, self._flags(), src, aptr,
                                           self.seed_value, self.env_id_base, self.initial_rand_steps,
                                           self._outputs(), self._stream()), "oth_step")
        return self.obs, self.reward, self.done, self.info

    def capture_steps(self, policy="random", steps=1):
        """CUDA graph of `steps` consecutive fused policy + step launches on this env (the in-kernel 'random' or
        'greedy' player).  `graph.replay()` then costs one graph launch instead of `steps` Python -> ctypes calls:
        useful when launches are short (small batches); at 2^20 envs the 45 us launch hides the host path either
        way (bench.py --graph measures both: 45.05 vs 45.06 us).  The graph
        holds raw pointers into this env's tensors, so it is valid for the lifetime of the env; outputs of a replay
        are those of its last step."""
        if policy not in _ACT_SOURCES:
            raise ValueError("capture_steps() needs policy='random'|'greedy'")
        torch.cuda.synchronize(self.device)
        graph = torch.cuda.CUDAGraph()
        with self._on_device():
            with torch.cuda.graph(graph):
                for _ in range(int(steps)):
                    self.step(policy=policy)
        return graph

    @property
    def info(self):
        return {"legal_mask": self.legal_mask, "player_turn": self.player_turn, "winner": self.winner,
                "actions": self.actions_taken}

    # ------------------------------------------------------------------ host-buffer path
    def step_host(self, actions_host, packed=False):
        """The same step through HOST buffers: pinned H2D of the actions, kernel, pinned D2H
        of obs / legal mask / reward / done.  `actions_host`: pinned integer CPU tensor [B]; int64 like the
        device API, or a narrower signed type (int16 holds every action of a 16x16 board and -1): it then
        crosses PCIe at its own width and is widened on the device.
        Returns CPU tensors (pinned, reused across calls).

        packed=True: instead of the dense obs + mask (2*S*S bytes per env) the call returns the step's result in its
        packed form -- a dict of pinned CPU tensors {black, white, legal: int64 (W,B) bitboards; reward f32 (B,);
        done u8 (B,); player_turn i8 (B,)} = 24*W + 6 bytes per env (30 instead of 133 at 8x8) -- for host consumers
        that expand lazily (`expand_packed_host` does it for any subset of envs).  The dense path stays the default."""
        if self._host is None:
            pin = lambda t: torch.empty(t.shape, dtype=t.dtype, pin_memory=True)   # noqa: E731
            self._host = {
                "actions_dev": torch.empty((self.num_envs,), dtype=torch.int64, device=self.device),
                "obs": pin(self.obs) if self.obs is not None else None,
                "mask": pin(self._mask_u8) if self._mask_u8 is not None else None,
                "reward": torch.empty((self.num_envs,), dtype=torch.float32, pin_memory=True),
                "done": torch.empty((self.num_envs,), dtype=torch.uint8, pin_memory=True),
                "reward_dev": torch.empty((self.num_envs,), dtype=torch.float32, device=self.device),
                "done_dev": torch.empty((self.num_envs,), dtype=torch.uint8, device=self.device),
            }
        h = self._host
        if actions_host.dtype == torch.int64:
            h["actions_dev"].copy_(actions_host, non_blocking=True)
        else:
            if actions_host.dtype not in (torch.int8, torch.int16, torch.int32):
                raise TypeError("actions_host must be a signed integer tensor")
            key = "actions_" + str(actions_host.dtype)
            if key not in h:
                h[key] = torch.empty((self.num_envs,), dtype=actions_host.dtype, device=self.device)
            h[key].copy_(actions_host, non_blocking=True)       # H2D at the narrow width
            h["actions_dev"].copy_(h[key])                       # widened on the device
        self._host_action_bytes = actions_host.element_size()
        self.step(h["actions_dev"])
        self._host_packed = bool(packed)
        if packed:
            if "black" not in h:
                for name in ("black", "white", "legal"):
                    h[name] = torch.empty((self.words, self.num_envs), dtype=torch.int64, pin_memory=True)
                h["turn"] = torch.empty((self.num_envs,), dtype=torch.int8, pin_memory=True)
                h["turn_dev"] = torch.empty((self.num_envs,), dtype=torch.int8, device=self.device)
            h["reward_dev"].copy_(self.reward)
            h["done_dev"].copy_(self._done_u8)
            h["turn_dev"].copy_(self.player_turn)
            for name in ("black", "white", "legal"):
                h[name].copy_(getattr(self, "_" + name), non_blocking=True)
            h["reward"].copy_(h["reward_dev"], non_blocking=True)
            h["done"].copy_(h["done_dev"], non_blocking=True)
            h["turn"].copy_(h["turn_dev"], non_blocking=True)
            torch.cuda.current_stream(self.device).synchronize()
            return {"black": h["black"], "white": h["white"], "legal": h["legal"], "reward": h["reward"],
                    "done": h["done"], "player_turn": h["turn"]}
        if h["obs"] is not None:
            h["obs"].copy_(self.obs, non_blocking=True)
        if h["mask"] is not None:
            h["mask"].copy_(self._mask_u8, non_blocking=True)
        h["reward_dev"].copy_(self.reward)                  # gather the strided views on the device: 5 B per env
        h["done_dev"].copy_(self._done_u8)                  # cross PCIe, not the 16-byte records
        h["reward"].copy_(h["reward_dev"], non_blocking=True)
        h["done"].copy_(h["done_dev"], non_blocking=True)
        torch.cuda.current_stream(self.device).synchronize()
        return h["obs"], h["reward"], h["done"], h["mask"]

    def host_bytes_per_step(self):
        """(h2d, d2h) bytes moved by one step_host() call (h2d at the width of the last actions_host passed)."""
        d2h = self.reward.numel() * 4 + self._done_u8.numel()
        if getattr(self, "_host_packed", False):
            return (self.num_envs * getattr(self, "_host_action_bytes", 8),
                    d2h + self.num_envs * (24 * self.words + 1))
        if self.obs is not None:
            d2h += self.obs.numel() * self.obs.element_size()
        if self._mask_u8 is not None:
            d2h += self._mask_u8.numel()
        return self.num_envs * getattr(self, "_host_action_bytes", 8), d2h

    def expand_packed_host(self, packed, index=None):
        """Host-side (numpy) expansion of step_host(packed=True) results for the envs in `index` (all when None):
        -> (obs int8 (n,S,S): mover +1 / opponent -1, othello.py:363-369; legal_mask bool (n,S*S))."""
        import numpy as np
        S, N = self.board_size, self.num_cells
        idx = slice(None) if index is None else np.asarray(index)
        cells = np.arange(N)
        word, bit = cells >> 6, (cells & 63).astype(np.uint64)

        def bits(t):
            w = t.numpy().view(np.uint64)[:, idx]                       # (W, n)
            return ((w[word, :] >> bit[:, None]) & np.uint64(1)).T.astype(np.int8)   # (n, N)
        black, white, legal = bits(packed["black"]), bits(packed["white"]), bits(packed["legal"])
        turn = packed["player_turn"].numpy()[idx].astype(np.int8)[:, None]
        obs = (white - black) * turn
        return obs.reshape(-1, S, S), legal.astype(bool)

    # ------------------------------------------------------------------ accessors
    def observe(self, layout=None, dtype=None, out=None):
        """Observation of the current state in any layout, without stepping."""
        layout = self.obs_layout if layout is None else layout
        dtype = self.obs_dtype if dtype is None else dtype
        P = _cabi.OBS_PLANES[_LAYOUTS[layout]]
        S, B = self.board_size, self.num_envs
        shape = (B, S, S) if P == 1 else (B, P, S, S)
        if out is None:
            out = torch.empty(shape, dtype=dtype, device=self.device)
        o = _cabi.OthOutputs(_cabi.ptr(out), None, None, None, None, None, None, _LAYOUTS[layout], _DTYPES[dtype])
        with self._on_device():
            _cabi.check(self._lib.oth_observe(self._state(), B, S, self._flags(), o, self._stream()), "oth_observe")
        return out

    def board_state(self):
        """board_state of every env: int8 (B,S,S), WHITE=+1, BLACK=-1 (othello.py:232)."""
        return self.observe("absolute", torch.int8)

    def refresh_outputs(self):
        """Re-derives obs / legal_mask / done / player_turn / winner from the state."""
        with self._on_device():
            _cabi.check(self._lib.oth_observe(self._state(), self.num_envs, self.board_size, self._flags(),
                                              self._outputs(with_scalars=False), self._stream()), "oth_observe")

    def possible_moves(self, i):
        """The cached possible_moves list of env i (ascending ints) -- synchronises."""
        words = self._legal[:, i].cpu().tolist()
        out = []
        for w, v in enumerate(words):
            v &= 0xFFFFFFFFFFFFFFFF
            while v:
                low = v & -v
                out.append(w * 64 + low.bit_length() - 1)
                v ^= low
        return out

    def terminated(self):
        return (self._meta & _cabi.META_TERMINATED) != 0

    def rand_steps(self):
        """-> (rand_step_cnt, max_rand_steps) int32 tensors (othello.py:50-51)"""
        return (self._meta >> _cabi.META_RANDCNT_SHIFT) & 0xFF, (self._meta >> _cabi.META_RANDMAX_SHIFT) & 0xFF

    def count_disks(self):
        """(white_cnt, black_cnt) like OthelloBaseEnv.count_disks (othello.py:468-471)."""
        b = self.board_state()
        return (b == 1).flatten(1).sum(1), (b == -1).flatten(1).sum(1)

    # ------------------------------------------------------------------ state injection
    def get_state(self):
        return {"black": self._black.clone(), "white": self._white.clone(), "legal": self._legal.clone(),
                "meta": self._meta.clone(), "ctr": self._ctr.clone(), "policy_counter": self._policy_counter}

    def set_state(self, state):
        for name in ("black", "white", "legal", "meta", "ctr"):
            getattr(self, "_" + name).copy_(state[name])
        self._policy_counter = int(state.get("policy_counter", 0))
        self.refresh_outputs()

    def set_board_state(self, cells, perspective=1, turn=None, recompute_legal=None, clear_terminal=False):
        """set_board_state (+ set_player_turn when `turn` is given) for every env
        (othello.py:380-389,464-466).  cells: (B,S,S) or (B,S*S) integer tensor seen from
        `perspective` (+1 WHITE / -1 BLACK; int or per-env tensor); turn: None, int or (B,)."""
        B, N = self.num_envs, self.num_cells
        cells = torch.as_tensor(cells).to(device=self.device, dtype=torch.int8).reshape(B, N).contiguous()

        def per_env(v):
            if v is None:
                return None
            if isinstance(v, torch.Tensor):
                return v.to(device=self.device, dtype=torch.int8).reshape(B).contiguous()
            return torch.full((B,), int(v), dtype=torch.int8, device=self.device)
        persp, turn_t = per_env(perspective), per_env(turn)
        if recompute_legal is None:
            recompute_legal = turn_t is not None
        with self._on_device():
            _cabi.check(self._lib.oth_set_board(self._state(), B, self.board_size, _cabi.ptr(cells), _cabi.ptr(persp),
                                                _cabi.ptr(turn_t), int(bool(recompute_legal)),
                                                int(bool(clear_terminal)), self._stream()), "oth_set_board")
        self.refresh_outputs()

    def set_player_turn(self, turn):
        """set_player_turn (othello.py:464-466): sets the side to move and recomputes possible_moves."""
        B = self.num_envs
        t = (turn.to(device=self.device, dtype=torch.int8).reshape(B).contiguous() if isinstance(turn, torch.Tensor)
             else torch.full((B,), int(turn), dtype=torch.int8, device=self.device))
        with self._on_device():
            _cabi.check(self._lib.oth_set_board(self._state(), B, self.board_size, None, None, _cabi.ptr(t), 1, 0,
                                                self._stream()), "oth_set_board")
        self.refresh_outputs()

    # ------------------------------------------------------------------ batched policies
    def policy_actions(self, kind, depth=1, out=None):
        """policy.get_action for the side to move of every env -> int64 [B] (-1 where terminated)."""
        spec = policy_spec(kind, depth)
        if out is None:
            out = torch.empty((self.num_envs,), dtype=torch.int64, device=self.device)
        with self._on_device():
            _cabi.check(self._lib.oth_policy(self._state(), self.num_envs, self.board_size, spec, self.seed_value,
                                             self._policy_counter, self.env_id_base, _cabi.ptr(out), self._stream()),
                        "oth_policy")
        self._policy_counter += 1
        return out

    def search_actions(self, depth, prune=True, cooperative=True, with_nodes=False):
        """MaxiMinPolicy.get_action at any depth 1..8 for every env (oth_search): -> actions int64 [B], or
        (actions, positions scored per env int64 [B]) with with_nodes.  prune: alpha-beta (same moves, fewer
        positions); cooperative: 8 lanes per position on boards up to 8x8."""
        flags = (_cabi.SEARCH_PRUNE if prune else 0) | (_cabi.SEARCH_COOPERATIVE if cooperative else 0)
        out = torch.empty((self.num_envs,), dtype=torch.int64, device=self.device)
        nodes = torch.zeros((self.num_envs,), dtype=torch.int64, device=self.device) if with_nodes else None
        with self._on_device():
            _cabi.check(self._lib.oth_search(self._state(), self.num_envs, self.board_size, int(depth), flags,
                                             _cabi.ptr(out), _cabi.ptr(nodes), self._stream()), "oth_search")
        return (out, nodes) if with_nodes else out
