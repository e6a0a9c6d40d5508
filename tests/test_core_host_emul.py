"""The DEVICE rules engine (csrc/oth_core.cuh) compiled for the host and checked against the C oracle.

tests/host_emul/emul_core.cpp includes the header the kernels are built from, with the CUDA intrinsics shimmed,
so the bitboard arithmetic itself (carry-propagation moves and flips, Kogge-Stone fills, multi-word carry chains,
the stop-times-pattern flip reconstruction, greedy / maximin picks) is exercised on CPU for every board size --
a wrong mask or shift shows up here, before any GPU time is spent.  The GPU parity tests stay the proof for the
kernels; nothing in the product loads this library."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from oracle import oracle

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "host_emul", "emul_core.cpp")
CORE = os.path.join(os.path.dirname(HERE), "gymothelloenv_b200", "csrc", "oth_core.cuh")
SEARCH = os.path.join(os.path.dirname(HERE), "gymothelloenv_b200", "csrc", "oth_search.cuh")
LIB = os.path.join(os.path.dirname(HERE), "oracle", "_build", "libcore_host_emul.so")

F_SUDDEN, F_DISK = 1, 2


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


@pytest.fixture(scope="module")
def emul():
    newest = max(os.path.getmtime(SRC), os.path.getmtime(CORE), os.path.getmtime(SEARCH))
    if not (os.path.isfile(LIB) and os.path.getmtime(LIB) >= newest):
        os.makedirs(os.path.dirname(LIB), exist_ok=True)
        tmp = LIB + ".%d.tmp" % os.getpid()
        subprocess.check_call(["g++", "-O1", "-std=c++17", "-fPIC", "-shared", "-o", tmp, SRC])
        os.replace(tmp, LIB)
    return C.CDLL(LIB)


class EmulVec(object):
    def __init__(self, lib, B, S):
        self.lib, self.B, self.S = lib, B, S
        self.W = lib.emul_words(S)
        self.black = np.zeros((self.W, B), dtype=np.uint64)
        self.white = np.zeros((self.W, B), dtype=np.uint64)
        self.legal = np.zeros((self.W, B), dtype=np.uint64)
        self.meta = np.zeros(B, dtype=np.uint32)
        assert lib.emul_reset(S, C.c_int64(B), _p(self.black), _p(self.white), _p(self.legal), _p(self.meta)) == 0

    def step(self, actions, flags):
        a = np.ascontiguousarray(actions, dtype=np.int64)
        reward = np.zeros(self.B, dtype=np.int32)
        done = np.zeros(self.B, dtype=np.uint8)
        winner = np.zeros(self.B, dtype=np.int8)
        assert self.lib.emul_step(self.S, C.c_int64(self.B), _p(self.black), _p(self.white), _p(self.legal),
                                  _p(self.meta), _p(a), flags, _p(reward), _p(done), _p(winner)) == 0
        return reward, done, winner

    def policy(self, kind, depth):
        out = np.zeros(self.B, dtype=np.int64)
        assert self.lib.emul_policy(self.S, C.c_int64(self.B), _p(self.black), _p(self.white), _p(self.legal),
                                    _p(self.meta), kind, depth, _p(out)) == 0
        return out

    def search(self, depth, prune):
        out = np.zeros(self.B, dtype=np.int64)
        nodes = np.zeros(self.B, dtype=np.int64)
        assert self.lib.emul_search(self.S, C.c_int64(self.B), _p(self.black), _p(self.white), _p(self.legal),
                                    _p(self.meta), depth, int(prune), _p(out), _p(nodes)) == 0
        return out, nodes

    def _bits(self, words):
        N = self.S * self.S
        out = np.zeros((self.B, N), dtype=np.uint8)
        for a in range(N):
            out[:, a] = (words[a >> 6] >> np.uint64(a & 63)) & np.uint64(1)
        return out

    def board(self):
        return (self._bits(self.white).astype(np.int8) - self._bits(self.black).astype(np.int8)).reshape(
            self.B, self.S, self.S)

    def legal_dense(self):
        return self._bits(self.legal)


@pytest.mark.parametrize("S", range(4, 17))
@pytest.mark.parametrize("sudden,disk", [(True, False), (False, True)])
def test_random_games_match_oracle(emul, S, sudden, disk):
    """Whole games with ~6 % invalid actions: board, cached legal set, turn, reward, done after every ply."""
    B = 48 if S <= 10 else 24
    rng = np.random.RandomState(100 * S + sudden)
    ev, orc = EmulVec(emul, B, S), oracle.OracleVec(B, S, sudden_death=sudden, disk_reward=disk)
    flags = (F_SUDDEN if sudden else 0) | (F_DISK if disk else 0)
    for t in range(2 * S * S):
        board, legal, turn, term, winner, n_moves = orc.export()
        assert np.array_equal(ev.board(), board), "board before ply %d" % t
        live = term == 0
        assert np.array_equal(ev.legal_dense()[live], legal[live]), "legal set before ply %d" % t
        assert np.array_equal(((ev.meta & 1).astype(np.int32) * 2 - 1)[live], turn[live])
        if not live.any():
            break
        acts = np.zeros(B, dtype=np.int64)
        for i in range(B):
            moves = np.flatnonzero(legal[i])
            if len(moves) and rng.rand() > 0.06:
                acts[i] = moves[rng.randint(len(moves))]
            else:
                acts[i] = rng.randint(-2, S * S + 2)
        r0, d0, _ = orc.step(acts)
        r1, d1, w1 = ev.step(acts, flags)
        assert np.array_equal(r0[live], r1[live]) and np.array_equal(d0[live], d1[live])
        _, _, _, _, winner, _ = orc.export()
        fin = live & (d0 == 1)
        assert np.array_equal(winner[fin], w1[fin].astype(np.int32))
    else:
        raise AssertionError("games did not finish")


@pytest.mark.parametrize("S,depths", [(4, (1, 2, 3, 4)), (5, (1, 2, 3)), (6, (1, 2, 3)), (7, (1, 2, 3)), (8, (1, 2, 3)), (9, (1, 2)), (10, (1, 2)),
                                      (12, (1, 2)), (13, (1,)), (16, (1,))])
def test_policy_picks_match_oracle(emul, S, depths):
    """greedy and maximin-k choices on positions reached by random play (first-occurrence ties included)."""
    B = 24
    rng = np.random.RandomState(7 * S)
    ev, orc = EmulVec(emul, B, S), oracle.OracleVec(B, S)
    for t in range(S * S):
        board, legal, turn, term, _, _ = orc.export()
        if term.all():
            break
        if t % 3 == 0:
            live = term == 0
            assert np.array_equal(ev.policy(1, 1)[live], orc.policy("greedy")[live])
            for d in depths:
                assert np.array_equal(ev.policy(2, d)[live], orc.policy("maximin", d)[live]), (S, t, d)
        acts = np.zeros(B, dtype=np.int64)
        for i in range(B):
            moves = np.flatnonzero(legal[i])
            acts[i] = moves[rng.randint(len(moves))] if len(moves) else 0
        orc.step(acts)
        ev.step(acts, F_SUDDEN)


@pytest.mark.parametrize("S,depths", [(4, (1, 2, 3, 4, 5, 6, 7, 8)), (5, (1, 3, 5, 6)), (6, (1, 2, 3, 4, 5)), (8, (1, 2, 3, 4)),
                                      (10, (1, 2, 3)), (13, (1, 2))])
def test_alpha_beta_search_matches_oracle_maximin(emul, S, depths):
    """search_pick (oth_search.cuh): the pruned and the unpruned generic-depth search choose the move the oracle's
    plain maximin chooses (first-occurrence ties), at every depth; pruning never scores more positions."""
    B = 12
    rng = np.random.RandomState(11 * S)
    ev, orc = EmulVec(emul, B, S), oracle.OracleVec(B, S)
    checked = 0
    for t in range(S * S):
        board, legal, turn, term, _, _ = orc.export()
        if term.all():
            break
        if t % 5 == 2:
            live = term == 0
            for d in depths:
                want = orc.policy("maximin", d)
                a0, n0 = ev.search(d, False)
                a1, n1 = ev.search(d, True)
                assert np.array_equal(a0[live], want[live]), (S, t, d, "unpruned")
                assert np.array_equal(a1[live], want[live]), (S, t, d, "pruned")
                assert (n1 <= n0).all() and (d > 2 or True)
                if d == 1:
                    assert np.array_equal(n0, n1)
                checked += int(live.sum())
        acts = np.zeros(B, dtype=np.int64)
        for i in range(B):
            moves = np.flatnonzero(legal[i])
            acts[i] = moves[rng.randint(len(moves))] if len(moves) else 0
        orc.step(acts)
        ev.step(acts, F_SUDDEN)
    assert checked > 20


@pytest.mark.parametrize("S", [4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 16])
def test_bit_sliced_flip_counts_equal_per_move_flips(emul, S):
    """flip_planes1 / flip_planes / max_flips (every move scored at once: greedy and the last ply of the searches)
    against one flips_at + popcount per cell and legal<S>, on ARBITRARY disc patterns -- sparse, dense, reachable or
    not -- including long runs (rows / columns / diagonals filled with one colour)."""
    rng = np.random.RandomState(100 + S)
    N, W = S * S, (S * S + 63) // 64
    B = 6000 if W == 1 else 1500
    fill = rng.choice([0.15, 0.4, 0.7, 0.95], size=B)
    cells = rng.rand(B, N)
    colour = rng.rand(B, N) < rng.choice([0.1, 0.2, 0.5, 0.8], size=(B, 1))
    occupied = cells < fill[:, None]
    ownc, oppc = occupied & colour, occupied & ~colour
    dense = np.zeros((2, B, 64 * W), dtype=np.uint64)
    dense[0, :, :N], dense[1, :, :N] = ownc, oppc
    weights = (np.uint64(1) << np.arange(64, dtype=np.uint64))
    words = (dense.reshape(2, B, W, 64) * weights).sum(3).astype(np.uint64)          # (2, B, W)
    own, opp = np.ascontiguousarray(words[0].T), np.ascontiguousarray(words[1].T)    # [W][B]
    sliced, per_move = np.zeros((B, 64 * W), dtype=np.int8), np.zeros((B, 64 * W), dtype=np.int8)
    mv_s, mv_l = np.zeros((W, B), dtype=np.uint64), np.zeros((W, B), dtype=np.uint64)
    mx = np.zeros(B, dtype=np.int32)
    assert emul.emul_flipcount(S, C.c_int64(B), _p(own), _p(opp), _p(sliced), _p(per_move), _p(mv_s), _p(mv_l), _p(mx)) == 0
    assert np.array_equal(sliced, per_move)
    assert np.array_equal(mv_s, mv_l)
    assert np.array_equal(mx, per_move.max(1).clip(min=0))
    assert per_move.max() >= min(2 * (S - 2), 8) and (mv_s != 0).any(0).mean() > 0.3      # the cases are not trivial


@pytest.mark.parametrize("S", [6, 8, 10, 12, 13, 16])
def test_bit_sliced_flip_counts_on_star_patterns(emul, S):
    """One empty cell with a bracketed opponent run of a chosen length in each of the 8 directions (up to the longest
    the board allows): the large sums and every carry of the bit-sliced adders, which random boards rarely reach."""
    rng = np.random.RandomState(7 + S)
    N, W, B = S * S, (S * S + 63) // 64, 800
    own_c, opp_c = np.zeros((B, N), dtype=bool), np.zeros((B, N), dtype=bool)
    want = np.zeros(B, dtype=np.int64)
    centre = np.zeros(B, dtype=np.int64)
    for i in range(B):
        r0, c0 = rng.randint(S), rng.randint(S)
        centre[i] = r0 * S + c0
        total = 0
        for dr, dc in ((0, 1), (0, -1), (1, 0), (-1, 0), (1, 1), (-1, -1), (1, -1), (-1, 1)):
            room = 0                                          # cells available behind the centre in this direction
            r, c = r0 + dr, c0 + dc
            while 0 <= r < S and 0 <= c < S:
                room += 1
                r, c = r + dr, c + dc
            if room < 2:
                continue
            k = room - 1 if rng.rand() < 0.6 else rng.randint(0, room)      # mostly the longest run
            closed = rng.rand() < 0.85
            for j in range(1, k + 1):
                opp_c[i, (r0 + j * dr) * S + c0 + j * dc] = True
            if closed and k > 0:
                own_c[i, (r0 + (k + 1) * dr) * S + c0 + (k + 1) * dc] = True
                total += k
        want[i] = total
    dense = np.zeros((2, B, 64 * W), dtype=np.uint64)
    dense[0, :, :N], dense[1, :, :N] = own_c, opp_c
    weights = (np.uint64(1) << np.arange(64, dtype=np.uint64))
    words = (dense.reshape(2, B, W, 64) * weights).sum(3).astype(np.uint64)
    own, opp = np.ascontiguousarray(words[0].T), np.ascontiguousarray(words[1].T)
    sliced, per_move = np.zeros((B, 64 * W), dtype=np.int8), np.zeros((B, 64 * W), dtype=np.int8)
    mv_s, mv_l = np.zeros((W, B), dtype=np.uint64), np.zeros((W, B), dtype=np.uint64)
    mx = np.zeros(B, dtype=np.int32)
    assert emul.emul_flipcount(S, C.c_int64(B), _p(own), _p(opp), _p(sliced), _p(per_move), _p(mv_s), _p(mv_l), _p(mx)) == 0
    assert np.array_equal(sliced, per_move) and np.array_equal(mv_s, mv_l)
    assert np.array_equal(sliced[np.arange(B), centre], want)             # the construction itself
    assert want.max() >= 3 * (S - 3)                                       # sums well into the upper planes


def test_nth_set_bit(emul):
    rng = np.random.RandomState(3)
    for S in (8, 10, 13, 16):
        W, B = emul.emul_words(S), 512
        legal = rng.randint(0, 2 ** 62, size=(W, B)).astype(np.uint64) * np.uint64(3)
        if S * S % 64:
            legal[W - 1] &= np.uint64((1 << (S * S % 64)) - 1)
        dense = np.zeros((B, 64 * W), dtype=np.uint8)
        for a in range(64 * W):
            dense[:, a] = (legal[a >> 6] >> np.uint64(a & 63)) & np.uint64(1)
        cnt = dense.sum(1)
        n = (rng.rand(B) * cnt).astype(np.int32)
        out = np.zeros(B, dtype=np.int64)
        assert emul.emul_nth(S, C.c_int64(B), _p(legal), _p(n), _p(out)) == 0
        for i in range(B):
            assert out[i] == np.flatnonzero(dense[i])[n[i]]


def test_philox_block_matches_numpy_reference(emul):
    from tests import philox_ref
    out = np.zeros(4, dtype=np.uint32)
    for c in [(0, 0, 0, 0), (1, 2, 3, 4), (0xFFFFFFFF, 0, 12345, 1)]:
        emul.emul_philox(*[C.c_uint32(x) for x in c], C.c_uint32(0xDEADBEEF), C.c_uint32(42), _p(out))
        ref = philox_ref.philox4x32_10(c[0], c[1], c[2], c[3], 0xDEADBEEF, 42)
        assert [int(x) for x in ref] == out.tolist()
