"""Scalar Gym-style API, same surface as the reference's ``othello.py`` -- one env.

``OthelloBaseEnv`` / ``SimpleOthelloEnv`` / ``OthelloEnv`` keep the constructor
arguments, methods, attributes and error behaviour of the reference classes
(``othello.py:21-93``, ``:96-214``, ``:217-501``) so that its drivers and scripted
policies run unchanged, but the rules are NOT computed here: every call drives the
B=1 instance of the CUDA engine (``VecOthelloEnv``) and mirrors the resulting state to
the host.  The random-opening plies of the two wrappers use ``numpy.random.RandomState``
exactly like the reference, so seeded games reproduce the reference's action streams.

Differences (documented): ``mute`` defaults to True on the wrappers (the reference prints
on every reset/terminal); ``render`` only supports the text mode (no graphics stack);
``gym`` is not required -- ``action_space`` / ``observation_space`` are light stand-ins
with the same ``n`` / ``shape`` / ``low`` / ``high`` metadata.
"""
import numpy as np
import torch

from . import _cabi
from .vec_env import VecOthelloEnv

BLACK_DISK = -1
NO_DISK = 0
WHITE_DISK = 1


class Discrete(object):
    """Shape metadata of gym.spaces.Discrete (othello.py:245)."""

    def __init__(self, n):
        self.n = int(n)
        self.shape = ()

    def contains(self, x):
        return 0 <= int(x) < self.n

    def __repr__(self):
        return "Discrete(%d)" % self.n


class Box(object):
    """Shape metadata of gym.spaces.Box (othello.py:248-254)."""

    def __init__(self, low, high):
        self.low, self.high = np.asarray(low), np.asarray(high)
        self.shape = self.low.shape

    def __repr__(self):
        return "Box%s" % (self.shape,)


class OthelloBaseEnv(object):
    """Othello base environment (reference: othello.py:217-501) on the CUDA engine."""

    metadata = {'render.modes': ['np_array', 'human']}

    def __init__(self, board_size=8, sudden_death_on_invalid_move=True, num_disk_as_reward=False,
                 possible_actions_in_obs=False, mute=False, device=None):
        self.board_size = max(4, board_size)                                   # othello.py:230
        self.sudden_death_on_invalid_move = sudden_death_on_invalid_move
        self.num_disk_as_reward = num_disk_as_reward
        self.mute = mute
        self.possible_actions_in_obs = possible_actions_in_obs
        self.viewer = None
        self._vec = VecOthelloEnv(1, board_size=self.board_size, device=device,
                                  sudden_death_on_invalid_move=bool(sudden_death_on_invalid_move),
                                  num_disk_as_reward=bool(num_disk_as_reward),
                                  obs_layout="board+mask" if possible_actions_in_obs else "board",
                                  obs_dtype=torch.int8)
        S, N = self.board_size, self.board_size ** 2
        # Host mirrors in PINNED memory, which the device addresses directly (unified addressing): the kernels of a
        # scalar call write obs / legal mask / the 16-byte result record / the make_state planes straight into them and
        # read the action from one, so a call is kernel launches + ONE stream synchronisation -- no memcpy, no
        # per-call device tensor.  (A B = 1 env is a launch-latency workload; this is what its latency is made of.)
        v = self._vec
        self._h_obs = torch.zeros(tuple(v.obs.shape[1:]), dtype=torch.int8).pin_memory()
        self._h_mask = torch.zeros((N,), dtype=torch.uint8).pin_memory()
        self._h_packed = torch.zeros((16,), dtype=torch.uint8).pin_memory()
        self._h_planes = torch.zeros((4, S, S), dtype=torch.int8).pin_memory()
        self._h_action = torch.zeros((1,), dtype=torch.int64).pin_memory()
        self._h_pick = torch.zeros((1,), dtype=torch.int64).pin_memory()
        self._np_obs, self._np_mask, self._np_planes = self._h_obs.numpy(), self._h_mask.numpy(), self._h_planes.numpy()
        self._np_reward = self._h_packed.view(torch.float32).numpy()          # record: f32 reward | done | turn | winner
        self._np_status = self._h_packed.view(torch.int8).numpy()
        none5 = (None,) * 5
        self._c_host_out = _cabi.OthOutputs(_cabi.ptr(self._h_obs), _cabi.ptr(self._h_mask), *none5,
                                            _cabi.OBS_BOARD_MASK if possible_actions_in_obs else _cabi.OBS_BOARD,
                                            _cabi.DTYPE_I8, _cabi.ptr(self._h_packed))
        self._c_host_planes = _cabi.OthOutputs(_cabi.ptr(self._h_planes), None, *none5, _cabi.OBS_PLANES4, _cabi.DTYPE_I8)
        self._action_dev = torch.zeros((1,), dtype=torch.int64, device=v.device)

        # internal state as in the reference right after __init__ (othello.py:232,239-242)
        self._vec.reset()
        self._pull()
        self.possible_moves = []

        self.action_space = Discrete(N)
        shape = [2, S, S] if possible_actions_in_obs else [S, S]
        self.observation_space = Box(np.zeros(shape), np.ones(shape))

    # ------------------------------------------------------------------ device <-> host
    def _sync_host(self):
        """Waits for the kernels that wrote the host mirrors and decodes them: board_state, player_turn, winner,
        terminated, possible_moves, obs."""
        torch.cuda.current_stream(self._vec.device).synchronize()
        st = self._np_status
        self.terminated = bool(st[4])
        self.player_turn = WHITE_DISK if st[5] > 0 else BLACK_DISK
        self.winner = int(st[6])
        obs = self._np_obs.astype(np.int64)
        self._obs_host = obs
        mover_view = obs[0] if obs.ndim == 3 else obs                         # the mover's discs are +1 (othello.py:363-369)
        self.board_state = mover_view * self.player_turn                      # int board like np.zeros(dtype=int)
        self.possible_moves = np.flatnonzero(self._np_mask).tolist()

    def _emit_planes(self):
        v = self._vec
        _cabi.check(v._lib.oth_observe(v._state(), 1, self.board_size, v._flags(), self._c_host_planes, v._stream()),
                    "oth_observe")

    def _pull(self):
        """Mirror the device state (after anything but step() changed it)."""
        v = self._vec
        with v._on_device():
            _cabi.check(v._lib.oth_observe(v._state(), 1, self.board_size, v._flags(), self._c_host_out, v._stream()),
                        "oth_observe")
            self._emit_planes()
        self._sync_host()

    def current_planes(self):
        """util.make_state's four planes of the CURRENT state (int8 (4,S,S), written by the device with every call)."""
        return self._np_planes

    def device_pick(self, kind, depth):
        """policy.get_action on the device state: one oth_policy launch writing the action into pinned memory."""
        from .vec_env import policy_spec
        v = self._vec
        with v._on_device():
            _cabi.check(v._lib.oth_policy(v._state(), 1, self.board_size, policy_spec(kind, depth), v.seed_value,
                                          v._policy_counter, v.env_id_base, _cabi.ptr(self._h_pick), v._stream()),
                        "oth_policy")
            v._policy_counter += 1
            torch.cuda.current_stream(v.device).synchronize()
        return int(self._h_pick[0])

    def _push_board(self):
        """board_state may have been assigned by the caller: send it to the device."""
        self._vec.set_board_state(np.asarray(self.board_state, dtype=np.int8).reshape(1, -1), perspective=1,
                                  turn=None, recompute_legal=False)

    # ------------------------------------------------------------------ reference API
    def reset(self):
        self._vec.reset()
        self._pull()
        return self.get_observation()

    def get_possible_actions(self, board=None):
        """othello.py:313-343.  `board` (optional) is seen from the mover: own discs = +1."""
        scratch = self._scratch()
        if board is None:
            cells, persp = self.board_state, 1
            turn = self.player_turn
        else:
            cells, persp, turn = np.asarray(board), 1, WHITE_DISK
        scratch.set_board_state(np.asarray(cells, dtype=np.int8).reshape(1, -1), perspective=persp, turn=turn,
                                recompute_legal=True)
        return scratch.possible_moves(0)

    def _scratch(self):
        if getattr(self, "_scratch_env", None) is None:
            self._scratch_env = VecOthelloEnv(1, board_size=self.board_size, device=self._vec.device)
            self._scratch_env.reset()
        return self._scratch_env

    def print_board(self, print_valid_moves=True):
        valid_actions = self.get_possible_actions()
        board = self.board_state.copy().ravel()
        if print_valid_moves:
            for p in valid_actions:
                board[p] = 2
        board = board.reshape(*self.board_state.shape)
        print('Turn: {}'.format('WHITE' if self.player_turn == WHITE_DISK else 'BLACK'))
        print('Valid actions: {}'.format(valid_actions))
        for row in board:
            print(' '.join(map(lambda x: ['B', 'O', 'W', 'V'][x + 1], row)))
        print('-' * 10)

    def get_observation(self):
        """othello.py:363-378: mover's discs +1; (2,S,S) with the cached legal grid when
        possible_actions_in_obs.  A fresh int64 array (the reference returns an alias of
        board_state on WHITE's turn -- callers had to copy it anyway)."""
        return self._obs_host.copy()

    def set_board_state(self, board_state, perspective=WHITE_DISK):
        """othello.py:380-389: forces the board; turn and possible_moves are left alone."""
        state = board_state[0] if np.ndim(board_state) > 2 else board_state
        cells = np.asarray(state)
        self._vec.set_board_state(cells.astype(np.int8).reshape(1, -1), perspective=int(perspective), turn=None,
                                  recompute_legal=False)
        moves = self.possible_moves
        self._pull()
        self.possible_moves = moves

    def update_board(self, action):
        """othello.py:391-410"""
        self._action_dev.fill_(int(action))
        v = self._vec
        with v._on_device():
            _cabi.check(v._lib.oth_update_board(v._state(), 1, self.board_size, _cabi.ptr(self._action_dev),
                                                v._stream()), "oth_update_board")
        v.refresh_outputs()
        moves = self.possible_moves
        self._pull()
        self.possible_moves = moves

    def step(self, action):
        """othello.py:412-462 -> (obs, reward, done, None)"""
        if self.terminated:
            raise ValueError('Game has terminated!')
        try:
            a = int(action)
        except (TypeError, ValueError):
            a = -1
        if not (-(1 << 62) < a < (1 << 62)):
            a = -1
        v = self._vec
        self._h_action[0] = a
        was_terminated = self.terminated
        with v._on_device():
            _cabi.check(v._lib.oth_step(v._state(), 1, self.board_size, v._flags(), _cabi.ACT_EXTERNAL,
                                        _cabi.ptr(self._h_action), v.seed_value, v.env_id_base, v.initial_rand_steps,
                                        self._c_host_out, v._stream()), "oth_step")
            self._emit_planes()
        self._sync_host()
        reward = int(self._np_reward[0])
        if self.terminated and not was_terminated and not self.mute:
            self._announce()
        return self.get_observation(), reward, self.terminated, None

    def _announce(self):
        white_cnt, black_cnt = self.count_disks()
        print('white: {}, black: {}'.format(white_cnt, black_cnt))
        print({WHITE_DISK: 'WHITE wins', BLACK_DISK: 'BLACK wins', NO_DISK: 'DRAW'}[self.winner])

    def set_player_turn(self, turn):
        """othello.py:464-466: sets the turn and recomputes possible_moves."""
        self._vec.set_player_turn(int(turn))
        self._pull()

    def count_disks(self):
        """othello.py:468-471 -> (white_cnt, black_cnt)"""
        white_cnt = (self.board_state == WHITE_DISK).sum()
        black_cnt = (self.board_state == BLACK_DISK).sum()
        return white_cnt, black_cnt

    def determine_winner(self, sudden_death=False):
        """othello.py:473-501: marks the game terminated and returns the winner."""
        if sudden_death:
            winner = BLACK_DISK if self.player_turn == WHITE_DISK else WHITE_DISK
        else:
            white_cnt, black_cnt = self.count_disks()
            winner = WHITE_DISK if white_cnt > black_cnt else (BLACK_DISK if black_cnt > white_cnt else NO_DISK)
        code = 1 if winner == WHITE_DISK else (2 if winner == BLACK_DISK else 0)
        v = self._vec
        v._meta.bitwise_and_(~(3 << _cabi.META_WINNER_SHIFT)).bitwise_or_(
            _cabi.META_TERMINATED | (code << _cabi.META_WINNER_SHIFT))
        self.terminated = True
        self.winner = winner
        return winner

    def render(self, mode='human', close=False):
        if close:
            return
        if mode == 'np_array':
            self.print_board()
        else:
            # the reference's pyglet/OpenGL viewer (othello.py:511-597) needs a graphics
            # stack, which this box does not have; the text board is the only renderer
            self.print_board()

    def close(self):
        self.viewer = None

    def seed(self, seed=None):
        return [seed]


class SimpleOthelloEnv(object):
    """Two-sided wrapper (reference: othello.py:21-93)."""

    metadata = {'render.modes': ['np_array', 'human']}

    def __init__(self, board_size=8, initial_rand_steps=0, seed=0, sudden_death_on_invalid_move=True,
                 render_in_step=False, num_disk_as_reward=False, possible_actions_in_obs=False, mute=True,
                 device=None):
        self.board_size = board_size
        self.num_disk_as_reward = num_disk_as_reward
        self.mute = mute
        self.env = OthelloBaseEnv(board_size=board_size, num_disk_as_reward=num_disk_as_reward,
                                  sudden_death_on_invalid_move=sudden_death_on_invalid_move,
                                  possible_actions_in_obs=possible_actions_in_obs, mute=mute, device=device)
        self.observation_space = self.env.observation_space
        self.action_space = self.env.action_space
        self.render_in_step = render_in_step
        self.initial_rand_steps = initial_rand_steps
        self.rand_seed = seed
        self.rnd = np.random.RandomState(seed=self.rand_seed)
        self.max_rand_steps = 0
        self.rand_step_cnt = 0

    def seed(self, seed=None):
        if seed is not None:
            self.rand_seed = seed
            self.rnd = np.random.RandomState(seed=self.rand_seed)

    def reset(self):
        obs = self.env.reset()
        self.max_rand_steps = self.rnd.randint(low=0, high=self.initial_rand_steps // 2 + 1) * 2   # othello.py:62-63
        self.rand_step_cnt = 0
        if not self.mute:
            print('The initial {} steps will be random'.format(self.max_rand_steps))
        return obs

    def step(self, action):
        if self.rand_step_cnt < self.max_rand_steps:                            # othello.py:70-73
            ix = self.rnd.randint(0, len(self.possible_moves))
            action = self.possible_moves[ix]
            self.rand_step_cnt += 1
        obs, reward, done, _ = self.env.step(action)
        if self.render_in_step:
            self.render()
        return obs, reward, done, None

    def render(self, mode='human', close=False):
        self.env.render(mode=mode, close=close)

    def close(self):
        self.env.close()

    @property
    def player_turn(self):
        return self.env.player_turn

    @property
    def possible_moves(self):
        return self.env.possible_moves


class OthelloEnv(object):
    """Protagonist-vs-embedded-opponent wrapper (reference: othello.py:96-214)."""

    metadata = {'render.modes': ['np_array', 'human']}

    def __init__(self, white_policy=None, black_policy=None, protagonist=WHITE_DISK, board_size=8,
                 initial_rand_steps=0, seed=0, sudden_death_on_invalid_move=True, render_in_step=False,
                 num_disk_as_reward=False, possible_actions_in_obs=False, mute=True, device=None):
        self.board_size = board_size
        self.num_disk_as_reward = num_disk_as_reward
        self.mute = mute
        self.env = OthelloBaseEnv(board_size=board_size, num_disk_as_reward=num_disk_as_reward,
                                  sudden_death_on_invalid_move=sudden_death_on_invalid_move,
                                  possible_actions_in_obs=possible_actions_in_obs, mute=mute, device=device)
        self.observation_space = self.env.observation_space
        self.action_space = self.env.action_space
        self.render_in_step = render_in_step
        self.initial_rand_steps = initial_rand_steps
        self.rand_seed = seed
        self.rnd = np.random.RandomState(seed=self.rand_seed)
        self.max_rand_steps = 0
        self.rand_step_cnt = 0
        self.protagonist = protagonist
        self.opponent = white_policy if self.protagonist == BLACK_DISK else black_policy   # othello.py:133-136

    def switch_color(self):
        # like the reference (othello.py:138-142) this flips the protagonist colour only;
        # the opponent policy object is kept
        self.protagonist = WHITE_DISK if self.protagonist == BLACK_DISK else BLACK_DISK

    def seed(self, seed=None):
        if seed is not None:
            self.rand_seed = seed
            self.rnd = np.random.RandomState(seed=self.rand_seed)
            if self.opponent is not None and hasattr(self.opponent, 'seed'):
                self.opponent.seed(self.rand_seed)

    def reset(self):
        obs = self.env.reset()
        self.max_rand_steps = self.rnd.randint(low=0, high=self.initial_rand_steps // 2 + 1) * 2
        self.rand_step_cnt = 0
        if not self.mute:
            print('The initial {} steps will be random'.format(self.max_rand_steps))
        if hasattr(self.opponent, 'reset'):
            try:
                self.opponent.reset(self)
            except TypeError:
                pass
        if self.env.player_turn == self.protagonist:
            return obs
        action = self.opponent.get_action(obs)                                   # othello.py:165-174
        obs, _, done, _ = self.env.step(action)
        if done:
            if not self.mute:
                print('done==True in reset(), do it again.')
            return self.reset()
        return obs

    def step(self, action):
        assert self.env.player_turn == self.protagonist                         # othello.py:177
        if self.rand_step_cnt < self.max_rand_steps:
            ix = self.rnd.randint(0, len(self.possible_moves))
            action = self.possible_moves[ix]
            self.rand_step_cnt += 1
        obs, reward, done, _ = self.env.step(action)
        if self.render_in_step:
            self.render()
        if done:
            return obs, reward, done, None
        while not done and self.env.player_turn != self.protagonist:            # othello.py:190-199
            if self.rand_step_cnt < self.max_rand_steps:
                ix = self.rnd.randint(0, len(self.possible_moves))
                opponent_move = self.possible_moves[ix]
                self.rand_step_cnt += 1
            else:
                opponent_move = self.opponent.get_action(obs)
            obs, reward, done, _ = self.env.step(opponent_move)
            if self.render_in_step:
                self.render()
        return obs, -reward, done, None

    def render(self, mode='human', close=False):
        self.env.render(mode=mode, close=close)

    def close(self):
        self.env.close()

    @property
    def player_turn(self):
        return self.env.player_turn

    @property
    def possible_moves(self):
        return self.env.possible_moves
