/*
 * othello_b200.h -- C ABI of libothello_b200.so, the B200 (sm_100a) Othello env engine.
 *
 * This is the drop-in boundary for the env hot path of omurammm/GymOthelloEnv.  The
 * reference has no FFI of its own (it is pure Python); each entry point below names
 * the reference interface it replaces (paths relative to the reference root).  The
 * Python host package (gymothelloenv_b200/) binds these with ctypes and mirrors the
 * reference's class API on top; INTEGRATION.md shows the stub a maintainer of the
 * reference would add.
 *
 * Conventions
 *   - plain pointers and sizes only; every pointer is a DEVICE pointer unless its name
 *     ends in _host; buffers are owned by the caller (PyTorch's caching allocator) and
 *     borrowed for the duration of one call;
 *   - every call is asynchronous on `stream` (a cudaStream_t passed as void*; NULL =
 *     the legacy default stream) and never synchronises the host;
 *   - return value: 0 on success, otherwise an OTH_ERR_* / cudaError_t-derived code;
 *     oth_last_error() gives a thread-local message.  Nothing throws across the ABI;
 *   - no global mutable state: safe from several host threads on distinct buffers.
 *
 * State layout (SoA, one slot per env, W = ceil(S*S/64) 64-bit words per colour):
 *   black[w*B + i], white[w*B + i]   bit (r*S + c) of the S*S-bit board = cell (r, c),
 *                                    i.e. bit a <-> action a (othello.py:342,392-393)
 *   legal[w*B + i]                   the CACHED possible_moves set (othello.py:242,466),
 *                                    including its stale value after a full-board /
 *                                    sudden-death terminal
 *   meta[i]  bit0 turn (0 = BLACK to move, 1 = WHITE)        (othello.py:239)
 *            bit1 terminated                                  (othello.py:241)
 *            bits2-3 winner: 0 none/draw, 1 WHITE, 2 BLACK    (othello.py:240)
 *            bits8-15 rand_step_cnt, bits16-23 max_rand_steps (othello.py:50-51)
 *   ctr[i]   number of Philox draws this env has consumed (RNG stream position)
 */
#ifndef OTHELLO_B200_H
#define OTHELLO_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define OTH_ABI_VERSION 2
#define OTH_MIN_BOARD_SIZE 4   /* othello.py:230 clamps to >= 4 */
#define OTH_MAX_BOARD_SIZE 16   /* 256 cells = 4 words per colour; actions fit the int16 traces */

typedef struct oth_state {
    uint64_t *black;   /* [W][B] */
    uint64_t *white;   /* [W][B] */
    uint64_t *legal;   /* [W][B] */
    uint32_t *meta;    /* [B] */
    uint32_t *ctr;     /* [B] */
} oth_state;

/* meta word helpers */
#define OTH_META_TURN_WHITE 0x1u
#define OTH_META_TERMINATED 0x2u
#define OTH_META_WINNER_SHIFT 2
#define OTH_META_RANDCNT_SHIFT 8
#define OTH_META_RANDMAX_SHIFT 16

/* flags for oth_step / oth_rollout */
#define OTH_F_SUDDEN_DEATH   0x01  /* sudden_death_on_invalid_move (othello.py:224,427) */
#define OTH_F_DISK_REWARD    0x02  /* num_disk_as_reward (othello.py:225,446-459) */
#define OTH_F_AUTO_RESET     0x04  /* finished envs restart inside the same step call */
#define OTH_F_LEGACY_SINGLE_MOVE_PLANE 0x08 /* util.py:55: make_state legal plane is empty when exactly one move */
#define OTH_F_ZERO_IF_TERMINATED 0x10 /* oth_observe: envs whose meta has the terminated bit get an all-zero obs
                                         (the worker's `o`, ppo_run_self_play.py:246,274); legal_mask unaffected */
#define OTH_F_SP_FRESH_STATE 0x20     /* oth_selfplay_advance: refresh the saved state on EVERY protagonist ply
                                         (default: keep ppo_run_self_play.py:291-301's stale state in pass runs) */
#define OTH_F_SEARCH_PER_THREAD 0x80   /* oth_rollout: every maximin decision by the per-thread DFS (no lane-cooperative
                                           search); same results, for A/B timing and cross-checks */
#define OTH_F_SEARCH_SMALL_LISTS 0x40  /* oth_rollout, tests only: the lane-cooperative maximin search (depth >= 3) keeps at
                                           most 3 root moves, so its per-thread fallback runs constantly */

/* observation layouts (plane count P, each plane S*S cells, env-major [B][P][S][S]) */
#define OTH_OBS_NONE        0
#define OTH_OBS_BOARD       1  /* P=1: mover's discs +1, opponent -1 (othello.py:363-369) */
#define OTH_OBS_BOARD_MASK  2  /* P=2: + cached legal grid (possible_actions_in_obs, othello.py:370-376) */
#define OTH_OBS_PLANES4     3  /* P=4: black, white, turn, legal (util.py:48-74) */
#define OTH_OBS_PLANES3     4  /* P=3: black, white, turn (run_2agent.py:29-46) */
#define OTH_OBS_ABSOLUTE    5  /* P=1: board_state itself, WHITE=+1 / BLACK=-1 (othello.py:10-12) */

#define OTH_DTYPE_I8  0
#define OTH_DTYPE_F32 1

/* where oth_step takes its actions from */
#define OTH_ACT_EXTERNAL 0   /* `actions` array, like env.step(a) */
/* actions[i] == OTH_ACTION_HOLD: env i sits this ply out -- state untouched, outputs re-emitted, reward 0, done 0.
 * The reference's workers each step at their own pace (ppo_run_self_play.py:371-441: a worker plays the scripted
 * opponent's plies on its own and only then asks for an action); a lockstep batch needs a way to leave an env alone. */
#define OTH_ACTION_HOLD (-0x7FFFFFFFFFFFFFFFLL - 1)
#define OTH_ACT_RANDOM   1   /* in-kernel RandomPolicy (simple_policies.py:37-41), Philox stream */
#define OTH_ACT_GREEDY   2   /* in-kernel GreedyPolicy (simple_policies.py:69-92) */

/* policy kinds for oth_policy / oth_rollout (util.py:28-45 create_policy) */
#define OTH_POLICY_RANDOM  0
#define OTH_POLICY_GREEDY  1
#define OTH_POLICY_MAXIMIN 2
#define OTH_POLICY_ALPHABETA 3  /* maximin with alpha-beta pruning (README.md:55-56 "can be improved by pruning"): the
                                   SAME move as OTH_POLICY_MAXIMIN at the same depth, fewer positions searched */
#define OTH_MAX_SEARCH_DEPTH 8  /* MAXIMIN depth 2..4: full-width kernels (lane-cooperative on boards <= 8x8); depth
                                   5..8 runs the pruned search -- same choice, never a silently shallower one */

#define OTH_ERR_BAD_ARG        1001
#define OTH_ERR_BAD_BOARD_SIZE 1002
#define OTH_ERR_NULL_POINTER   1003
#define OTH_ERR_CUDA           2000  /* + cudaError_t */

/* outputs of one lockstep step / observe; any pointer may be NULL (not written) */
typedef struct oth_outputs {
    void    *obs;          /* [B][P][S*S] int8 or float32, see obs_layout / obs_dtype */
    uint8_t *legal_mask;   /* [B][S*S] 0/1 bytes (torch.bool compatible): the cached possible_moves */
    float   *reward;       /* [B] (othello.py:444-461), mover's perspective */
    uint8_t *done;         /* [B] (othello.py:462 `self.terminated`) */
    int8_t  *player_turn;  /* [B] -1 BLACK / +1 WHITE to move AFTER the call (othello.py:239) */
    int8_t  *winner;       /* [B] result of a transition that ended a game: -1/0/+1, else 0 */
    int64_t *actions_out;  /* [B] the action actually applied (after policy / random-opening override) */
    int32_t obs_layout;    /* OTH_OBS_* */
    int32_t obs_dtype;     /* OTH_DTYPE_* */
    void    *packed;       /* [B] 16-byte records (16-byte aligned), or NULL: the five per-env scalars in ONE store --
                              bytes 0-3 float reward | 4 uint8 done | 5 int8 player_turn | 6 int8 winner | 7 zero |
                              8-15 int64 action applied.  oth_step writes the whole record; oth_observe and the protocol
                              calls, which do not produce a reward / action, write bytes 4-7 only (the protocol calls
                              bytes 4-5).  Written in addition to whichever of the separate arrays are not NULL. */
} oth_outputs;

typedef struct oth_policy_spec {
    int32_t kind;    /* OTH_POLICY_* */
    int32_t depth;   /* maximin / alpha-beta search depth 1..OTH_MAX_SEARCH_DEPTH (simple_policies.py:101-103) */
} oth_policy_spec;

int         oth_abi_version(void);
const char *oth_last_error(void);
/* number of 64-bit words per colour for a board size, 0 if unsupported */
int         oth_words_per_board(int board_size);

/* OthelloBaseEnv.reset (othello.py:265-271) + SimpleOthelloEnv.reset's opening budget
 * max_rand_steps = 2*U{0..init_rand_steps/2} (othello.py:62-64), for every env, or only
 * for envs with only_where[i] != 0 when only_where is not NULL.  env_id_base + i is the
 * global env id that keys the Philox stream (identical trajectories for any sharding).
 * reset_rng != 0 rewinds ctr[i] to 0 (SimpleOthelloEnv.seed, othello.py:53-58). */
int oth_reset(oth_state st, int64_t B, int board_size, uint64_t seed, int64_t env_id_base,
              int init_rand_steps, int reset_rng, const uint8_t *only_where, void *stream);

/* Writes obs / legal_mask / player_turn / done / winner for the CURRENT state without
 * stepping: get_observation (othello.py:363-378), util.make_state (util.py:48-74),
 * the possible_moves property (othello.py:91-93). reward/actions_out are ignored. */
int oth_observe(oth_state st, int64_t B, int board_size, int flags, oth_outputs out, void *stream);

/* One lockstep ply for every env: OthelloBaseEnv.step (othello.py:412-462) preceded by
 * SimpleOthelloEnv.step's random-opening override (othello.py:68-73).
 * action_source = OTH_ACT_EXTERNAL reads actions[i] (any int64: out-of-range or not in
 * the cached list = invalid move => sudden death or forfeited turn, othello.py:417-428;
 * OTH_ACTION_HOLD = leave this env alone).
 * Envs that are already terminated are a no-op reporting done=1, reward=0 and a zero
 * observation (the reference raises ValueError, othello.py:415-416; the batched
 * contract mirrors the worker's 'over' message, ppo_run_self_play.py:274). */
int oth_step(oth_state st, int64_t B, int board_size, int flags, int action_source,
             const int64_t *actions, uint64_t seed, int64_t env_id_base, int init_rand_steps,
             oth_outputs out, void *stream);

/* Batched policy.get_action for the side to move in every env
 * (simple_policies.py:37-41 random, :69-92 greedy, :111-163 maximin).
 * Terminated envs get -1.  `counter` selects the Philox block of the random policy's
 * own stream (callers pass their step index). */
int oth_policy(oth_state st, int64_t B, int board_size, oth_policy_spec spec, uint64_t seed,
               uint64_t counter, int64_t env_id_base, int64_t *actions_out, void *stream);

/* MaxiMinPolicy.get_action (simple_policies.py:111-163) at any depth 1..OTH_MAX_SEARCH_DEPTH for the side to move in
 * every env, with the search's cost exposed: flags & OTH_SEARCH_PRUNE = alpha-beta pruning (same moves as without:
 * full window at the root, first-occurrence argmax kept), flags & OTH_SEARCH_COOPERATIVE = 8 lanes per position
 * (boards <= 8x8; the replies to the first root move, then the other root moves, are searched side by side).
 * actions_out[i] = the move (-1 for finished envs / no move); nodes_out[i] (may be NULL) = positions scored. */
#define OTH_SEARCH_PRUNE       0x1
#define OTH_SEARCH_COOPERATIVE 0x2
int oth_search(oth_state st, int64_t B, int board_size, int depth, int flags, int64_t *actions_out, int64_t *nodes_out,
               void *stream);

/* Fused tournament (run_2agent.py:92-203 with fixed colours; rule_base_game,
 * ppo_run_self_play.py:371-441): every env plays games_per_env complete games, black
 * policy vs white policy, first max_rand_steps plies uniformly random; no state is read
 * or written, only tally[0..3] += {black wins, draws, white wins, plies} (atomics; the
 * buffer is the NCCL all-reduce send buffer).  Optional per-game records for tests:
 * game_result[g] (int8 winner), game_plies[g] (int16), game_counts[2g..] (int16 black, white discs; 16x16 holds up to 256),
 * trace[g*trace_stride + ply] (int16 action applied), with g = i*games_per_env + k. */
int oth_rollout(int64_t B, int board_size, int flags, oth_policy_spec black, oth_policy_spec white,
                int init_rand_steps, int games_per_env, uint64_t seed, int64_t env_id_base,
                unsigned long long *tally, int8_t *game_result, int16_t *game_plies, int16_t *game_counts,
                int16_t *trace, int trace_stride, void *stream);

/* oth_rollout for callers that cannot allocate device memory (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Same games, same arguments and the same tally += {black wins, draws, white wins, plies} as oth_rollout (without
 * its optional per-game records), but the per-game records the reduction needs live in a file-scope device array of
 * the library -- OTH_NOALLOC_MAX_GAMES * 3 bytes (6 MiB) held by every GPU that loads the library until exit -- and
 * the call performs no cudaMalloc / cudaFree or any allocating routine.  B * games_per_env above
 * OTH_NOALLOC_MAX_GAMES is refused with OTH_ERR_BAD_ARG (split the call, or use oth_rollout).  All calls on a device
 * share the array: calls whose kernels overlap on the GPU (different streams) race; calls on one stream are safe.
 * The tally is accumulated by a deterministic-input reduction (block partial sums, 4 atomics per block). */
#define OTH_NOALLOC_MAX_GAMES (1 << 21)
int oth_rollout_noalloc(int64_t B, int board_size, int flags, oth_policy_spec black, oth_policy_spec white,
                        int init_rand_steps, int games_per_env, uint64_t seed, int64_t env_id_base,
                        unsigned long long *tally, void *stream);

/* set_board_state + set_player_turn (othello.py:380-389,464-466): dense cells [B][S*S]
 * int8 seen from perspective[i] (+1: WHITE=+1, -1: negated), new turn[i] (-1/+1).
 * recompute_legal != 0 refreshes the cached list (set_player_turn); 0 keeps it
 * (set_board_state alone does not touch possible_moves). Clears terminated/winner only
 * when clear_terminal != 0 (the reference never clears them outside reset). */
int oth_set_board(oth_state st, int64_t B, int board_size, const int8_t *cells, const int8_t *perspective,
                  const int8_t *turn, int recompute_legal, int clear_terminal, void *stream);

/* OthelloBaseEnv.update_board (othello.py:391-410) alone: the side to move places a disc on
 * actions[i] and turns the bracketed runs; no validity check (a cell without brackets just
 * receives the disc), turn / cached list / terminal flags untouched; out-of-range = no-op. */
int oth_update_board(oth_state st, int64_t B, int board_size, const int64_t *actions, void *stream);

/* ---------------------------------------------------------------------------------------
 * PPO self-play protocol (envs.py:188-227 PPOEnvs.step + ppo_run_self_play.py:244-368
 * subproc_worker), batched: every env plays BOTH colours with the same learner; the tuple
 * handed to the learner per round is the protagonist colour's (state, action, reward, done,
 * choices, (v, logp, hidden)) -- reward = its own terminal reward, or MINUS the reward of
 * the opponent's reply (:306,345).  One "round" = one PPOEnvs.step call:
 *   oth_selfplay_begin    clears the per-round flags; finished envs answer 'over' (:273-274)
 *   loop while *need_count > 0:  policy net on (obs, legal_mask) -> actions, v, logp, hidden
 *       oth_selfplay_advance     applies them to the envs that still owe a tuple
 *   oth_observe on {sav_black, sav_white, sav_legal, sav_meta} with OTH_F_ZERO_IF_TERMINATED
 *       gives the tuple's observation planes; sav_choices is its `choices` list as a bitset.
 */
typedef struct oth_selfplay {
    int8_t   *protagonist;  /* [B] colour the learner is credited for in this game: -1 BLACK / +1 WHITE (:265-267) */
    uint8_t  *phase;        /* [B] OTH_SP_* bits */
    uint64_t *sav_black;    /* [W][B] state_b / state_w of the pending decision (:283,323,360) */
    uint64_t *sav_white;    /* [W][B] */
    uint64_t *sav_legal;    /* [W][B] its possible_moves at that time (plane 3 of make_state) */
    uint32_t *sav_meta;     /* [B] bit0 turn plane; bit1 set = the tuple's observation is all zeros */
    uint64_t *sav_choices;  /* [W][B] info['choices'] of the tuple (:286-287 choice_b / choice_w) */
    int64_t  *tr_action;    /* [B] action_b / action_w (the learner's choice, even when the opening override replaced it) */
    float    *tr_reward;    /* [B] */
    uint8_t  *tr_done;      /* [B] */
    uint8_t  *tr_type;      /* [B] 0 = learning tuple (info['type'] None), 1 = 'over' */
    float    *tr_outputs;   /* [B][3] value, log-prob, hidden of the decision (output_b / output_w) */
    float    *pend_outputs; /* [B][3] same, for the decision still waiting for its reward */
    int64_t  *pend_action;  /* [B] */
} oth_selfplay;

#define OTH_SP_PENDING  0x01  /* a protagonist decision waits for its reward */
#define OTH_SP_EMITTED  0x02  /* this env has produced its tuple for the current round */
#define OTH_SP_FINISHED 0x04  /* game over: answers 'over' until the next oth_selfplay_new_game */
#define OTH_SP_PREV_P   0x08  /* the previous ply of this game was the protagonist's (pass run) */

/* `cmd == 'reset'` of the worker (:261-270) for every env: oth_reset + phase cleared + protagonist colour drawn
 * from Philox stream 2 (counter = game_index) unless protagonist_in (device, [B], -1/+1) is given. */
int oth_selfplay_new_game(oth_state st, oth_selfplay sp, int64_t B, int board_size, uint64_t seed, int64_t env_id_base,
                          int init_rand_steps, uint64_t game_index, const int8_t *protagonist_in, void *stream);

/* Start of a round: clears OTH_SP_EMITTED, emits the 'over' tuple of finished envs, writes out.obs / out.legal_mask /
 * out.player_turn for the CURRENT state (the first `need_action` message, :285) and sets *need_count (device int32)
 * to the number of envs that still owe a tuple. */
int oth_selfplay_begin(oth_state st, oth_selfplay sp, int64_t B, int board_size, int flags, oth_outputs out,
                       int32_t *need_count, void *stream);

/* One `policy.act` worth of progress (envs.py:200-217): every env that still owes a tuple applies actions[i]
 * (subject to the random-opening override, othello.py:70-73), with value[i] / logp[i] / hidden[i] recorded when
 * the mover is the protagonist.  Envs whose tuple is now complete get OTH_SP_EMITTED and their tr_* / sav_* rows
 * filled; the others stay in `need_action`.  out.* is rewritten for the new current state; *need_count is set to
 * the number of envs still owing a tuple. */
int oth_selfplay_advance(oth_state st, oth_selfplay sp, int64_t B, int board_size, int flags, const int64_t *actions,
                         const float *value, const float *logp, const float *hidden, uint64_t seed,
                         int64_t env_id_base, int init_rand_steps, oth_outputs out, int32_t *need_count, void *stream);

/* ---- teacher / student protocol (envs.py:44-140 PPOTeacherStudentEnvs.step, ppo_run_teacher_vs_student.py:
 * 425-560 othello_teacher_vs_student) --------------------------------------------------------------------------
 * Both colours are learners.  Per env and colour c (index 0 = black, 1 = white) the engine keeps c's latest
 * decision; arrays marked [2] are colour-major: element (c, k, i) at ((c * W) + k) * B + i for bitboards,
 * (c * B + i) for scalars.  The tuple an env owes per step() call is assembled in tr_*:
 *   - c moves again because the other side must pass: c's tuple, c's own reward, done 0      (:486-488 / :529-531)
 *   - c's run ends: the OTHER colour's tuple with minus c's last reward and c's done flag      (:498-503 / :541-545);
 *     when the game is over c's own tuple follows in the NEXT call (OTH_TS_QUEUED)            (:505-509 / :546-551);
 *     with nothing pending on the other side (white before its first move) it is emitted at once
 *   - finished env: zeros, done 1, type OTH_TS_TYPE_OVER, no choices                          (:470-471)
 * The saved state of c is taken when c's run of plies STARTS and kept through forced-pass runs (reference quirk;
 * OTH_F_SP_FRESH_STATE refreshes it on every ply).  The teacher colour's reward is replaced by
 * teacher_done_reward when the tuple's done flag is set, else 0 (:457-466).
 * Observation planes of the tuple: oth_observe on {tr_black, tr_white, tr_legal, tr_meta} with
 * OTH_F_ZERO_IF_TERMINATED; its choices: the same call on {.., tr_choices, ..}. */
typedef struct oth_teacher_student {
    uint8_t  *phase;          /* [B] OTH_TS_* bits */
    uint64_t *sav_black;      /* [2][W][B] board of c's saved state */
    uint64_t *sav_white;      /* [2][W][B] */
    uint64_t *sav_legal;      /* [2][W][B] legal set of that state (the planes4 legal plane) */
    uint32_t *sav_meta;       /* [2][B]    turn bit of that state */
    uint64_t *sav_choices;    /* [2][W][B] possible_moves at c's latest decision */
    int64_t  *sav_action;     /* [2][B] */
    float    *sav_outputs;    /* [2][B][3] value, log-prob, hidden of c's latest decision */
    int32_t  *queued_reward;  /* [B] raw reward of the queued final tuple */
    uint64_t *tr_black, *tr_white, *tr_legal, *tr_choices;   /* [W][B] the emitted tuple's state / choices */
    uint32_t *tr_meta;        /* [B] turn bit, or OTH_META_TERMINATED for an all-zero tuple */
    int64_t  *tr_action;      /* [B] */
    float    *tr_reward;      /* [B] */
    uint8_t  *tr_done;        /* [B] */
    uint8_t  *tr_type;        /* [B] OTH_TS_TYPE_* */
    uint8_t  *tr_who;         /* [B] 0 = teacher, 1 = student (envs.py:122-134; 'over' rows count as white) */
    float    *tr_outputs;     /* [B][3] */
} oth_teacher_student;

#define OTH_TS_EMITTED      0x01  /* this env has produced its tuple for the current call */
#define OTH_TS_FINISHED     0x02  /* game over */
#define OTH_TS_PEND_BLACK   0x04  /* black has a decision waiting for its reward */
#define OTH_TS_PEND_WHITE   0x08
#define OTH_TS_PREV_BLACK   0x10  /* colour of the previous ply of this game (neither bit: no ply yet) */
#define OTH_TS_PREV_WHITE   0x20
#define OTH_TS_QUEUED       0x40  /* the last mover's own final tuple is owed at the next call */
#define OTH_TS_QUEUED_WHITE 0x80  /* ... and that mover was white */
#define OTH_TS_TYPE_BLACK 0   /* 'black_transition' */
#define OTH_TS_TYPE_WHITE 1   /* 'white_transition' */
#define OTH_TS_TYPE_OVER  2   /* 'over' */

/* The worker's 'reset' (:444-455): new game in every env, all protocol state cleared. */
int oth_ts_new_game(oth_state st, oth_teacher_student ts, int64_t B, int board_size, uint64_t seed,
                    int64_t env_id_base, int init_rand_steps, void *stream);
/* Start of a step() call: clears OTH_TS_EMITTED; finished envs emit their queued final tuple or an 'over' row.
 * teacher = -1 | +1 (colour of the teacher this episode).  Writes out.obs / out.legal_mask for the current states
 * and need_count[0] / need_count[1] = number of envs waiting for a BLACK / WHITE action. */
int oth_ts_begin(oth_state st, oth_teacher_student ts, int64_t B, int board_size, int flags, int teacher,
                 float teacher_done_reward, oth_outputs out, int32_t *need_count, void *stream);
/* One policy[colour].act worth of progress (envs.py:83-98): every env still owing a tuple whose side to move is
 * `colour` records (actions[i], value[i], logp[i], hidden[i]) as that colour's decision and plays it (subject to
 * the random-opening override, othello.py:70-73).  Same outputs as oth_ts_begin. */
int oth_ts_advance(oth_state st, oth_teacher_student ts, int64_t B, int board_size, int flags, int teacher,
                   float teacher_done_reward, int colour, const int64_t *actions, const float *value,
                   const float *logp, const float *hidden, uint64_t seed, int64_t env_id_base, int init_rand_steps,
                   oth_outputs out, int32_t *need_count, void *stream);

/* Legal-move-restricted categorical over the learner's logits (model.py:60-150 Policy.act, :156-198
 * evaluate_actions): the distribution is softmax(logits[i][choices_i]), choices_i = set bits of legal[.][i].
 *   mode OTH_CAT_SAMPLE    action ~ that distribution (inverse CDF in ascending action order, one Philox draw,
 *                          stream 3, counter = `counter`); logp_out = its log-probability
 *   mode OTH_CAT_MODE      action = first argmax (deterministic=True)
 *   mode OTH_CAT_EVALUATE  logp_out = log-probability of actions_in[i]; 0 when it is not a legal choice (:172-173)
 * No legal move: action 0, logp 0 (:71-73).  float32 arithmetic, max-subtracted log-sum-exp. */
#define OTH_CAT_SAMPLE   0
#define OTH_CAT_MODE     1
#define OTH_CAT_EVALUATE 2
int oth_masked_categorical(const float *logits, const uint64_t *legal, int64_t B, int board_size, int mode,
                           const int64_t *actions_in, uint64_t seed, uint64_t counter, int64_t env_id_base,
                           int64_t *actions_out, float *logp_out, void *stream);

/* Policy.evaluate_actions for the PPO update (a2c_ppo_acktr/model.py:156-198, differentiated by algo/ppo.py:51-66),
 * batched and DIFFERENTIABLE.  legal = the `choices` of every sample as bitsets [W][B] (storage.py:19,48-57 keeps
 * python lists).  forward:
 *   logp_out[i]    = log softmax(logits[i][choices_i])[actions[i]]; the constant 0 when actions[i] is not one of the
 *                    choices or there are none (model.py:167-168) -- such rows get no gradient
 *   entropy_out[i] = entropy of the FULL S*S-way softmax(logits[i]) (model.py:159,178 uses the unmasked dist);
 *                    the caller averages it (`dist.entropy().mean()`).  Either output may be NULL.
 * backward: grad_logits[i][a] = grad_logp[i] * (1[a == actions[i]] - p_choices_i[a]) on the choices, 0 elsewhere
 *                              + grad_entropy[i] * dH_i/dlogits[i][a]       (either incoming gradient may be NULL)
 * float32; the backward recomputes the forward (4*S*S bytes per sample), nothing is saved between the calls. */
int oth_masked_logprob_forward(const float *logits, const uint64_t *legal, int64_t B, int board_size,
                               const int64_t *actions, float *logp_out, float *entropy_out, void *stream);
int oth_masked_logprob_backward(const float *logits, const uint64_t *legal, int64_t B, int board_size,
                                const int64_t *actions, const float *grad_logp, const float *grad_entropy,
                                float *grad_logits, void *stream);

/* Epsilon-greedy over the legal moves, the DQN agent's action choice (dqn.py:264-275 get_action,
 * :478-488 get_action_at_test): with probability epsilon (`epsilon >= u`, u uniform in [0,1) as at dqn.py:266)
 * a uniformly random entry of the ascending legal list, else the legal move with the largest q value, FIRST
 * occurrence on ties (torch.argmax over q[0][possible_moves], :271-273).  q is float32 [B, S*S]; legal as above.
 * Randomness: Philox stream 4 of env (env_id_base + i); word 2*counter decides explore / exploit
 * (u = (word >> 8) * 2^-24), word 2*counter+1 picks the random entry (multiply-shift).  The reference draws
 * from Python's `random`, so only the DISTRIBUTION can match it, not the stream (stated in DESIGN.md).
 * q_chosen_out[i] = q[i][action] (may be null); explored_out[i] = 1 when the random branch was taken (may be
 * null).  No legal move: action 0, q_chosen 0, explored 0 (the reference never asks in that state). */
int oth_epsilon_greedy(const float *q, const uint64_t *legal, int64_t B, int board_size, float epsilon,
                       uint64_t seed, uint64_t counter, int64_t env_id_base, int64_t *actions_out,
                       float *q_chosen_out, uint8_t *explored_out, void *stream);

/* Measurement helper: a dependency-limited LOP3/SHF/IADD3 stream on every SM; writes
 * the number of thread-level integer instructions executed to *instr_out_host after the
 * launch is enqueued (the count is static). Used by bench.py to measure the INT32-ALU
 * roofline denominator that MEASURED_PEAKS.json does not contain. */
int oth_int_alu_probe(uint32_t *sink, int blocks, int threads, int iters, uint64_t *instr_out_host, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* OTHELLO_B200_H */
