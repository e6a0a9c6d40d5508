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
        self._abs_dev = torch.zeros((1, S, S), dtype=torch.int8, device=self._vec.device)
        self._action_dev = torch.zeros((1,), dtype=torch.int64, device=self._vec.device)
        # pinned host mirrors, refreshed by ONE synchronisation per call
        self._h_abs = torch.zeros((S, S), dtype=torch.int8).pin_memory()
        self._h_obs = torch.zeros(tuple(self._vec.obs.shape[1:]), dtype=torch.int8).pin_memory()
        self._h_mask = torch.zeros((N,), dtype=torch.uint8).pin_memory()
        self._h_meta = torch.zeros((1,), dtype=torch.int32).pin_memory()
        self._h_reward = torch.zeros((1,), dtype=torch.float32).pin_memory()

        # internal state as in the reference right after __init__ (othello.py:232,239-242)
        self._vec.reset()
        self._pull()
        self.possible_moves = []

        self.action_space = Discrete(N)
        shape = [2, S, S] if possible_actions_in_obs else [S, S]
        self.observation_space = Box(np.zeros(shape), np.ones(shape))

    # ------------------------------------------------------------------ device <-> host
    def _pull(self):
        """Mirror the device state: board_state, player_turn, winner, terminated, possible_moves, obs."""
        v = self._vec
        v.observe("absolute", torch.int8, out=self._abs_dev)
        self._h_abs.copy_(self._abs_dev[0], non_blocking=True)
        self._h_obs.copy_(v.obs[0], non_blocking=True)
        self._h_mask.copy_(v._mask_u8[0], non_blocking=True)
        self._h_meta.copy_(v._meta, non_blocking=True)
        self._h_reward.copy_(v.reward, non_blocking=True)
        torch.cuda.current_stream(v.device).synchronize()
        meta = int(self._h_meta[0])
        self.board_state = self._h_abs.numpy().astype(np.int64)                # int board like np.zeros(dtype=int)
        self.player_turn = WHITE_DISK if meta & _cabi.META_TURN_WHITE else BLACK_DISK
        self.terminated = bool(meta & _cabi.META_TERMINATED)
        code = (meta >> _cabi.META_WINNER_SHIFT) & 3
        self.winner = WHITE_DISK if code == 1 else (BLACK_DISK if code == 2 else NO_DISK)
        self.possible_moves = np.nonzero(self._h_mask.numpy())[0].tolist()
        self._obs_host = self._h_obs.numpy().astype(np.int64)

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
        self._action_dev.fill_(a)
        self._vec.step(self._action_dev)
        was_terminated = self.terminated
        self._pull()
        reward = int(self._h_reward[0])
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
