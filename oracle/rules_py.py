"""TEST / BASELINE INFRASTRUCTURE -- pure-Python (numpy array) port of the reference env.

Why this exists next to the C oracle: the reference is interpreted Python over a numpy
board (``othello.py:217-501``) and cannot travel to the GPU box.  ``bench.py``'s
``cpu_baseline`` / ``--impl reference`` legs need a CPU arm whose cost model is the
reference's own -- a Python double loop over cells calling a Python ray walk ~200 times
per step -- not a C rewrite that would be ~100x faster than what users of the reference
actually run.  This port keeps that structure (same asymptotics, same interpreter, same
numpy element indexing) while being written from scratch; the build container measured
it within a few percent of the live reference's steps/s.

Pinned by tests/test_oracle_golden.py::test_python_port_replays_golden.  Never imported
by the product package.
"""
import numpy as np

WHITE, BLACK, EMPTY = 1, -1, 0
_RAYS = ((1, 1), (1, 0), (1, -1), (0, 1), (0, -1), (-1, 1), (-1, 0), (-1, -1))


class PyOthello(object):
    """Semantics of OthelloBaseEnv (othello.py:217-501), mute, no rendering."""

    def __init__(self, board_size=8, sudden_death_on_invalid_move=True, num_disk_as_reward=False):
        self.size = max(4, board_size)                      # othello.py:230
        self.sudden_death = sudden_death_on_invalid_move
        self.disk_reward = num_disk_as_reward
        self.reset()

    # othello.py:256-271
    def reset(self):
        n = self.size
        self.grid = np.zeros((n, n), dtype=int)
        h = n // 2
        self.grid[h - 1][h - 1] = self.grid[h][h] = WHITE
        self.grid[h][h - 1] = self.grid[h - 1][h] = BLACK
        self.turn = BLACK
        self.winner = EMPTY
        self.over = False
        self.moves = self._scan_moves()
        return self.observation()

    # othello.py:273-311 -- "ours" is +1 after the perspective flip done by the callers
    def _captures(self, view, r, c, dr, dc):
        n = self.size
        r += dr
        c += dc
        if r < 0 or r >= n or c < 0 or c >= n or view[r][c] != -1:
            return 0
        k = 0
        while 0 <= r < n and 0 <= c < n and view[r][c] == -1:
            r += dr
            c += dc
            k += 1
        if r < 0 or r >= n or c < 0 or c >= n or view[r][c] != 1:
            return 0
        return k

    # othello.py:313-343
    def _scan_moves(self):
        view = self.grid if self.turn == WHITE else -self.grid
        n = self.size
        hit = self._captures
        found = []
        for r in range(n):
            for c in range(n):
                if view[r][c] == EMPTY and (
                        hit(view, r, c, 1, 1) or hit(view, r, c, 1, 0) or hit(view, r, c, 1, -1) or
                        hit(view, r, c, 0, 1) or hit(view, r, c, 0, -1) or hit(view, r, c, -1, 1) or
                        hit(view, r, c, -1, 0) or hit(view, r, c, -1, -1)):
                    found.append(r * n + c)
        return found

    # othello.py:363-369
    def observation(self):
        return self.grid if self.turn == WHITE else -self.grid

    # othello.py:391-410
    def _place(self, action):
        n = self.size
        r, c = divmod(action, n)
        if self.turn == BLACK:
            self.grid = -self.grid
        for dr in (-1, 0, 1):
            for dc in (-1, 0, 1):
                if dr or dc:
                    for i in range(1, self._captures(self.grid, r, c, dr, dc) + 1):
                        self.grid[r + i * dr][c + i * dc] = 1
        self.grid[r][c] = 1
        if self.turn == BLACK:
            self.grid = -self.grid

    # othello.py:468-501
    def _counts(self):
        return int((self.grid == WHITE).sum()), int((self.grid == BLACK).sum())

    def _finish(self, sudden):
        self.over = True
        if sudden:
            return -self.turn
        w, b = self._counts()
        return WHITE if w > b else (BLACK if b > w else EMPTY)

    def _give_turn(self, turn):            # othello.py:464-466
        self.turn = turn
        self.moves = self._scan_moves()

    # othello.py:412-462
    def step(self, action):
        if self.over:
            raise ValueError("Game has terminated!")
        bad = action not in self.moves
        if not bad:
            self._place(action)
        empties = (self.grid == EMPTY).sum()
        sudden = bad and self.sudden_death
        mover = self.turn
        if sudden or empties == 0:
            self.winner = self._finish(sudden)
        else:
            self._give_turn(-self.turn)
            if not self.moves:
                self._give_turn(-self.turn)
                if not self.moves:
                    self.winner = self._finish(False)
        reward = 0
        if self.over:
            if self.disk_reward:
                if sudden:
                    reward = -(self.size ** 2)
                else:
                    w, b = self._counts()
                    mine, theirs = (w, b) if mover == WHITE else (b, w)
                    reward = self.size ** 2 if theirs == 0 else mine - theirs
            else:
                reward = self.winner * mover
        return self.observation(), reward, self.over, None


def random_playout_steps(board_size, seed, seconds):
    """Uniformly random games for ~`seconds`; -> (env steps, wall seconds).  One process."""
    import time
    env = PyOthello(board_size)
    rnd = np.random.RandomState(seed)
    steps = 0
    t0 = time.perf_counter()
    while True:
        env.reset()
        while not env.over:
            env.step(env.moves[rnd.randint(0, len(env.moves))])
            steps += 1
        dt = time.perf_counter() - t0
        if dt >= seconds:
            return steps, dt


def make_state(env):
    """util.py:48-74 on the port's env: planes [black, white, turn, legal]; the legal plane stays empty unless the
    mover has >= 2 moves (util.py:55-56).  Only the CPU-baseline fallback uses it (cost model of configs[3])."""
    n = env.size
    obs = np.array(env.observation(), copy=True)
    legal = np.zeros((n, n))
    if len(env.moves) > 1:
        idx = np.array(env.moves)
        legal[idx // n, idx % n] = 1
    mine, theirs = (obs == 1).astype(float), (obs == -1).astype(float)
    if env.turn == BLACK:
        return np.stack([mine, theirs, np.zeros((n, n)), legal])
    return np.stack([theirs, mine, np.ones((n, n)), legal])
