/*
 * TEST INFRASTRUCTURE -- CPU oracle for the Othello env hot path.
 *
 * A plain-C restatement of the reference's *algorithm* (cell array + ray walking, the
 * same structure as the Python, NOT bitboards -- so it is independent of the CUDA
 * kernels it checks).  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may load this library; the product package
 * never does.  Parity status: PINNED -- tests/test_oracle_golden.py checks every
 * function below against the npz/json files under tests/golden, which are outputs of the live
 * reference (tests/golden/make_golden.py), and oracle/validate_against_reference.py
 * fuzzes it against the imported reference in the build container.
 *
 * Each function cites the reference lines it follows (paths relative to the
 * reference root, omurammm/GymOthelloEnv).
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define ORC_MAX_S 16
#define ORC_MAX_N (ORC_MAX_S * ORC_MAX_S)

#define ORC_WHITE 1   /* othello.py:12 */
#define ORC_BLACK (-1) /* othello.py:10 */
#define ORC_EMPTY 0   /* othello.py:11 */

typedef struct orc_env {
    int32_t S;               /* board_size, clamped to >=4 (othello.py:230) */
    int32_t sudden_death;    /* sudden_death_on_invalid_move (othello.py:231) */
    int32_t disk_reward;     /* num_disk_as_reward (othello.py:234) */
    int32_t turn;            /* player_turn (othello.py:239) */
    int32_t winner;          /* othello.py:240 */
    int32_t terminated;      /* othello.py:241 */
    int32_t n_moves;         /* len(possible_moves) (othello.py:242) */
    int16_t moves[ORC_MAX_N];/* possible_moves, ascending row*S+col */
    int8_t board[ORC_MAX_N]; /* board_state, absolute colours, row-major */
} orc_env;

int orc_sizeof_env(void) { return (int)sizeof(orc_env); }
int orc_max_board_size(void) { return ORC_MAX_S; }

/* othello.py:256-263 */
static void place_start(orc_env *e) {
    int S = e->S, c = S / 2;
    memset(e->board, 0, sizeof(e->board));
    e->board[(c - 1) * S + (c - 1)] = ORC_WHITE;
    e->board[c * S + c] = ORC_WHITE;
    e->board[c * S + (c - 1)] = ORC_BLACK;
    e->board[(c - 1) * S + c] = ORC_BLACK;
}

/* othello.py:273-311.  `sign` plays the role of the perspective flip at
 * othello.py:316-319 / 395-396: a cell holds "our" disc when board*sign == +1. */
static int run_length(const int8_t *board, int S, int sign, int r, int c, int dr, int dc) {
    int rr = r + dr, cc = c + dc, n = 0;
    while (rr >= 0 && rr < S && cc >= 0 && cc < S && board[rr * S + cc] * sign == -1) {
        rr += dr;
        cc += dc;
        n++;
    }
    if (n == 0) return 0;
    if (rr < 0 || rr >= S || cc < 0 || cc >= S || board[rr * S + cc] * sign != 1) return 0;
    return n;
}

/* othello.py:313-343 (scan order: row-major over empty cells, any of 8 rays) */
static int list_moves(const int8_t *board, int S, int sign, int16_t *out) {
    int n = 0;
    for (int r = 0; r < S; r++)
        for (int c = 0; c < S; c++) {
            if (board[r * S + c] != ORC_EMPTY) continue;
            int ok = 0;
            for (int dr = 1; dr >= -1 && !ok; dr--)
                for (int dc = 1; dc >= -1 && !ok; dc--)
                    if ((dr || dc) && run_length(board, S, sign, r, c, dr, dc)) ok = 1;
            if (ok) out[n++] = (int16_t)(r * S + c);
        }
    return n;
}

/* othello.py:464-466 */
void orc_set_player_turn(orc_env *e, int turn) {
    e->turn = turn;
    e->n_moves = list_moves(e->board, e->S, turn, e->moves);
}

/* othello.py:222-243 */
void orc_init(orc_env *e, int board_size, int sudden_death, int disk_reward) {
    memset(e, 0, sizeof(*e));
    e->S = board_size < 4 ? 4 : board_size;
    if (e->S > ORC_MAX_S) e->S = ORC_MAX_S;
    e->sudden_death = sudden_death;
    e->disk_reward = disk_reward;
    place_start(e);
    e->turn = ORC_BLACK;
    e->winner = ORC_EMPTY;
    e->terminated = 0;
    e->n_moves = 0;
}

/* othello.py:265-271 */
void orc_reset(orc_env *e) {
    place_start(e);
    e->turn = ORC_BLACK;
    e->winner = ORC_EMPTY;
    e->terminated = 0;
    orc_set_player_turn(e, ORC_BLACK);
}

/* othello.py:468-471 (note the order: white first) */
void orc_count_disks(const orc_env *e, int *white, int *black) {
    int w = 0, b = 0, N = e->S * e->S;
    for (int i = 0; i < N; i++) {
        w += e->board[i] == ORC_WHITE;
        b += e->board[i] == ORC_BLACK;
    }
    *white = w;
    *black = b;
}

/* othello.py:380-389.  `cells` is an S*S grid seen from `perspective`. */
void orc_set_board_state(orc_env *e, const int8_t *cells, int perspective) {
    int N = e->S * e->S;
    for (int i = 0; i < N; i++) e->board[i] = (int8_t)(cells[i] * perspective);
}

/* othello.py:391-410 */
static void apply_move(orc_env *e, int action) {
    int S = e->S, r = action / S, c = action % S, me = e->turn;
    for (int dr = -1; dr <= 1; dr++)
        for (int dc = -1; dc <= 1; dc++) {
            if (!dr && !dc) continue;
            int k = run_length(e->board, S, me, r, c, dr, dc);
            for (int i = 1; i <= k; i++) e->board[(r + i * dr) * S + (c + i * dc)] = (int8_t)me;
        }
    e->board[r * S + c] = (int8_t)me;
}

/* othello.py:473-501 */
static int decide_winner(orc_env *e, int sudden) {
    e->terminated = 1;
    if (sudden) return -e->turn;
    int w, b;
    orc_count_disks(e, &w, &b);
    return w > b ? ORC_WHITE : (b > w ? ORC_BLACK : ORC_EMPTY);
}

/* othello.py:412-462.  Returns 0, or -1 for ValueError('Game has terminated!'). */
int orc_step(orc_env *e, int action, int *reward_out, int *done_out) {
    if (e->terminated) return -1;
    int invalid = 1;
    for (int i = 0; i < e->n_moves; i++)   /* `action not in self.possible_moves`: cached list */
        if (e->moves[i] == action) { invalid = 0; break; }
    if (!invalid) apply_move(e, action);

    int N = e->S * e->S, vacant = 0;
    for (int i = 0; i < N; i++) vacant += e->board[i] == ORC_EMPTY;
    int sudden = invalid && e->sudden_death;
    int done = sudden || vacant == 0;
    int mover = e->turn;
    if (done) {
        e->winner = decide_winner(e, sudden);      /* turn and moves stay untouched (stale) */
    } else {
        orc_set_player_turn(e, -e->turn);
        if (e->n_moves == 0) {
            orc_set_player_turn(e, -e->turn);      /* pass */
            if (e->n_moves == 0) e->winner = decide_winner(e, 0);
        }
    }
    int reward = 0;
    if (e->terminated) {
        if (e->disk_reward) {
            if (sudden) {
                reward = -N;
            } else {
                int w, b;
                orc_count_disks(e, &w, &b);
                if (mover == ORC_WHITE) { reward = w - b; if (b == 0) reward = N; }
                else                    { reward = b - w; if (w == 0) reward = N; }
            }
        } else {
            reward = e->winner * mover;
        }
    }
    *reward_out = reward;
    *done_out = e->terminated;
    return 0;
}

/* othello.py:363-378.  planes = 1, or 2 with possible_actions_in_obs. */
void orc_get_observation(const orc_env *e, int with_mask, int8_t *out) {
    int N = e->S * e->S;
    for (int i = 0; i < N; i++) out[i] = (int8_t)(e->board[i] * e->turn);
    if (with_mask) {
        memset(out + N, 0, (size_t)N);
        for (int i = 0; i < e->n_moves; i++) out[N + e->moves[i]] = 1;
    }
}

/* util.py:48-74: [black, white, turn, legal]; legal plane empty unless >=2 moves */
void orc_make_state(const orc_env *e, int8_t *out) {
    int N = e->S * e->S;
    for (int i = 0; i < N; i++) {
        out[i] = e->board[i] == ORC_BLACK;
        out[N + i] = e->board[i] == ORC_WHITE;
        out[2 * N + i] = e->turn == ORC_WHITE;
        out[3 * N + i] = 0;
    }
    if (e->n_moves > 1)
        for (int i = 0; i < e->n_moves; i++) out[3 * N + e->moves[i]] = 1;
}

/* simple_policies.py:12-18 + the inject sequence used at :78-81 / :135-137 */
static void clone_position(const orc_env *src, int turn, orc_env *dst) {
    orc_init(dst, src->S, src->sudden_death, 0);
    memcpy(dst->board, src->board, sizeof(dst->board));
    orc_set_player_turn(dst, turn);
}

static int own_count(const orc_env *e, int me) {
    int w, b;
    orc_count_disks(e, &w, &b);
    return me == ORC_WHITE ? w : b;
}

/* simple_policies.py:69-92: own disc count after each cached legal move; first max.
 * Returns -1 when there is no legal move. */
int orc_greedy(const orc_env *e) {
    int best = -1, best_cnt = -1, me = e->turn;
    for (int i = 0; i < e->n_moves; i++) {
        orc_env t;
        clone_position(e, me, &t);
        int r, d;
        orc_step(&t, e->moves[i], &r, &d);
        int cnt = own_count(&t, me);
        if (cnt > best_cnt) { best_cnt = cnt; best = e->moves[i]; }
    }
    return best;
}

/* simple_policies.py:111-155 */
static int mm_search(const orc_env *e, int depth, int max_depth, int perspective, int me, int *move_out) {
    if (e->terminated || depth >= max_depth || e->n_moves == 0) {
        if (move_out) *move_out = -1;
        return own_count(e, me);
    }
    int best_val = 0, best_mv = -1;
    for (int i = 0; i < e->n_moves; i++) {
        orc_env t;
        clone_position(e, perspective, &t);
        int r, d;
        orc_step(&t, e->moves[i], &r, &d);
        if (!t.terminated && t.turn == perspective) orc_set_player_turn(&t, -perspective); /* :139-144 */
        int v = mm_search(&t, depth + 1, max_depth, -perspective, me, 0);
        int better = (i == 0) || (perspective == me ? v > best_val : v < best_val); /* first arg{max,min} */
        if (better) { best_val = v; best_mv = e->moves[i]; }
    }
    if (move_out) *move_out = best_mv;
    return best_val;
}

/* simple_policies.py:157-163 */
int orc_maximin(const orc_env *e, int max_depth) {
    int mv = -1;
    mm_search(e, 0, max_depth, e->turn, e->turn, &mv);
    return mv;
}

/* ---- perft through set_player_turn + step ---- */
int64_t orc_perft(const orc_env *e, int depth) {
    if (depth == 0 || e->terminated) return 1;
    int64_t total = 0;
    for (int i = 0; i < e->n_moves; i++) {
        orc_env t = *e;
        int r, d;
        orc_step(&t, e->moves[i], &r, &d);
        total += orc_perft(&t, depth - 1);
    }
    return total;
}

/* ---- batch helpers for the parity tests (array of envs, lockstep) ---- */

void orc_vec_init(orc_env *envs, int64_t B, int S, int sudden, int disk) {
    for (int64_t i = 0; i < B; i++) { orc_init(&envs[i], S, sudden, disk); orc_reset(&envs[i]); }
}

/* Steps every env that is not terminated; terminated envs get reward 0 / done 1 and,
 * if auto_reset, are reset *after* the terminal transition was reported one call
 * earlier (i.e. the caller resets them explicitly with orc_vec_reset_done). */
void orc_vec_step(orc_env *envs, int64_t B, const int64_t *actions, int32_t *reward, uint8_t *done, uint8_t *was_valid) {
    for (int64_t i = 0; i < B; i++) {
        orc_env *e = &envs[i];
        if (e->terminated) { reward[i] = 0; done[i] = 1; if (was_valid) was_valid[i] = 0; continue; }
        if (was_valid) {
            int v = 0;
            for (int k = 0; k < e->n_moves; k++) v |= e->moves[k] == actions[i];
            was_valid[i] = (uint8_t)v;
        }
        int r, d;
        orc_step(e, (int)actions[i], &r, &d);
        reward[i] = r;
        done[i] = (uint8_t)d;
    }
}

void orc_vec_reset_done(orc_env *envs, int64_t B) {
    for (int64_t i = 0; i < B; i++) if (envs[i].terminated) orc_reset(&envs[i]);
}

/* Dense export: board[B,N] absolute, legal[B,N], meta[B,4] = turn, terminated, winner, n_moves */
void orc_vec_export(const orc_env *envs, int64_t B, int8_t *board, uint8_t *legal, int32_t *meta) {
    for (int64_t i = 0; i < B; i++) {
        const orc_env *e = &envs[i];
        int N = e->S * e->S;
        memcpy(board + i * N, e->board, (size_t)N);
        memset(legal + i * N, 0, (size_t)N);
        for (int k = 0; k < e->n_moves; k++) legal[i * N + e->moves[k]] = 1;
        meta[i * 4 + 0] = e->turn;
        meta[i * 4 + 1] = e->terminated;
        meta[i * 4 + 2] = e->winner;
        meta[i * 4 + 3] = e->n_moves;
    }
}

void orc_vec_observe(const orc_env *envs, int64_t B, int with_mask, int8_t *obs, int8_t *state4) {
    for (int64_t i = 0; i < B; i++) {
        int N = envs[i].S * envs[i].S;
        if (obs) orc_get_observation(&envs[i], with_mask, obs + i * N * (with_mask ? 2 : 1));
        if (state4) orc_make_state(&envs[i], state4 + i * 4 * N);
    }
}

void orc_vec_policy(const orc_env *envs, int64_t B, int kind, int depth, int64_t *actions) {
    for (int64_t i = 0; i < B; i++) {
        if (envs[i].terminated || envs[i].n_moves == 0) { actions[i] = -1; continue; }
        actions[i] = kind == 1 ? orc_greedy(&envs[i]) : orc_maximin(&envs[i], depth);
    }
}

/* Inject dense absolute boards + turns (recomputes the cached list, like
 * set_board_state + set_player_turn, othello.py:380-389,464-466). */
void orc_vec_inject(orc_env *envs, int64_t B, const int8_t *board, const int8_t *turn) {
    for (int64_t i = 0; i < B; i++) {
        orc_env *e = &envs[i];
        int N = e->S * e->S;
        memset(e->board, 0, sizeof(e->board));
        memcpy(e->board, board + i * N, (size_t)N);
        e->terminated = 0;
        e->winner = 0;
        orc_set_player_turn(e, turn[i]);
    }
}

/* ---- CPU baseline: uniformly random playouts, one thread (callers fork/pthread) ----
 * xorshift64* picks among the cached legal list; returns env steps executed. */
static inline uint64_t xs64(uint64_t *s) {
    uint64_t x = *s;
    x ^= x >> 12; x ^= x << 25; x ^= x >> 27;
    *s = x;
    return x * 2685821657736338717ULL;
}

int64_t orc_random_playouts(int S, uint64_t seed, int64_t games, int with_obs, int64_t *tally /*[3] black W/D/L*/) {
    orc_env e;
    orc_init(&e, S, 1, 0);
    uint64_t st = seed * 0x9E3779B97F4A7C15ULL + 0x1234567ULL;
    int64_t steps = 0;
    int8_t obs[ORC_MAX_N * 2];
    volatile int8_t sink = 0;
    for (int64_t g = 0; g < games; g++) {
        orc_reset(&e);
        while (!e.terminated) {
            int a = e.moves[(xs64(&st) >> 33) % (uint64_t)e.n_moves];
            int r, d;
            orc_step(&e, a, &r, &d);
            if (with_obs) { orc_get_observation(&e, 1, obs); sink ^= obs[a]; }
            steps++;
        }
        if (tally) tally[e.winner == ORC_BLACK ? 0 : (e.winner == 0 ? 1 : 2)]++;
    }
    (void)sink;
    return steps;
}

/* Greedy/maximin tournament on the CPU: black policy vs white policy, kinds: 0 random,
 * 1 greedy, 2 maximin(depth).  First `rand_plies(g)` plies random like SimpleOthelloEnv
 * (othello.py:62-73): rand budget = 2*U{0..init/2}.  Returns plies. */
int64_t orc_tournament(int S, uint64_t seed, int64_t games, int kind_b, int depth_b, int kind_w, int depth_w,
                       int init_rand, int64_t *tally /*[3] black W/D/L*/) {
    orc_env e;
    orc_init(&e, S, 1, 0);
    uint64_t st = seed * 0x9E3779B97F4A7C15ULL + 0x7654321ULL;
    int64_t steps = 0;
    for (int64_t g = 0; g < games; g++) {
        orc_reset(&e);
        int budget = 2 * (int)((xs64(&st) >> 33) % (uint64_t)(init_rand / 2 + 1)), ply = 0;
        while (!e.terminated) {
            int kind = e.turn == ORC_BLACK ? kind_b : kind_w, depth = e.turn == ORC_BLACK ? depth_b : depth_w;
            int a;
            if (ply < budget || kind == 0) a = e.moves[(xs64(&st) >> 33) % (uint64_t)e.n_moves];
            else if (kind == 1) a = orc_greedy(&e);
            else a = orc_maximin(&e, depth);
            int r, d;
            orc_step(&e, a, &r, &d);
            ply++;
            steps++;
        }
        tally[e.winner == ORC_BLACK ? 0 : (e.winner == 0 ? 1 : 2)]++;
    }
    return steps;
}
