// oth_cabi.cu -- the extern "C" boundary declared in include/othello_b200.h.
// Argument checking, board-size dispatch and error reporting only; all device work is
// in oth_kernels.cuh.  No torch types, no global mutable state, never throws.
#include <cstdio>
#include <cstring>
#include <pthread.h>
#include "oth_kernels.cuh"

#define OTH_API extern "C" __attribute__((visibility("default")))

#define OTH_DECL(S) extern "C" const OthSizeOps *oth_size_ops_##S();
OTH_DECL(4) OTH_DECL(5) OTH_DECL(6) OTH_DECL(7) OTH_DECL(8) OTH_DECL(9) OTH_DECL(10) OTH_DECL(11) OTH_DECL(12)
OTH_DECL(13) OTH_DECL(14) OTH_DECL(15) OTH_DECL(16)

static thread_local char g_err[256] = "";

static int fail(int code, const char *msg) {
    snprintf(g_err, sizeof(g_err), "%s", msg);
    return code;
}
static int cuda_result(cudaError_t e, const char *what) {
    if (e == cudaSuccess) return 0;
    snprintf(g_err, sizeof(g_err), "%s: %s", what, cudaGetErrorString(e));
    return OTH_ERR_CUDA + (int)e;
}

static const OthSizeOps *ops_for(int S) {
    switch (S) {
    case 4: return oth_size_ops_4();
    case 5: return oth_size_ops_5();
    case 6: return oth_size_ops_6();
    case 7: return oth_size_ops_7();
    case 8: return oth_size_ops_8();
    case 9: return oth_size_ops_9();
    case 10: return oth_size_ops_10();
    case 11: return oth_size_ops_11();
    case 12: return oth_size_ops_12();
    case 13: return oth_size_ops_13();
    case 14: return oth_size_ops_14();
    case 15: return oth_size_ops_15();
    case 16: return oth_size_ops_16();
    default: return nullptr;
    }
}

static int check_state(const oth_state &s, int64_t B, int S, const OthSizeOps **ops) {
    if (B < 0 || B > (int64_t)1 << 40) return fail(OTH_ERR_BAD_ARG, "num_envs out of range");
    *ops = ops_for(S);
    if (*ops == nullptr) {
        snprintf(g_err, sizeof(g_err), "board_size %d unsupported (this build: %d..%d)", S, OTH_MIN_BOARD_SIZE,
                 OTH_MAX_BOARD_SIZE);
        return OTH_ERR_BAD_BOARD_SIZE;
    }
    if (!s.black || !s.white || !s.legal || !s.meta || !s.ctr) return fail(OTH_ERR_NULL_POINTER, "null state pointer");
    return 0;
}

static int check_outputs(const oth_outputs &o) {
    if (o.obs != nullptr) {
        if (o.obs_layout < OTH_OBS_NONE || o.obs_layout > OTH_OBS_ABSOLUTE) return fail(OTH_ERR_BAD_ARG, "bad obs_layout");
        if (o.obs_dtype != OTH_DTYPE_I8 && o.obs_dtype != OTH_DTYPE_F32) return fail(OTH_ERR_BAD_ARG, "bad obs_dtype");
        if (((uintptr_t)o.obs & 15u) != 0) return fail(OTH_ERR_BAD_ARG, "obs must be 16-byte aligned");
    }
    if (o.legal_mask != nullptr && ((uintptr_t)o.legal_mask & 15u) != 0)
        return fail(OTH_ERR_BAD_ARG, "legal_mask must be 16-byte aligned");
    if (o.packed != nullptr && ((uintptr_t)o.packed & 15u) != 0)
        return fail(OTH_ERR_BAD_ARG, "packed must be 16-byte aligned");
    return 0;
}

OTH_API int oth_abi_version(void) { return OTH_ABI_VERSION; }
OTH_API const char *oth_last_error(void) { return g_err; }
OTH_API int oth_words_per_board(int board_size) {
    const OthSizeOps *o = ops_for(board_size);
    return o ? o->W : 0;
}

OTH_API int oth_reset(oth_state st, int64_t B, int board_size, uint64_t seed, int64_t env_id_base, int init_rand_steps,
                      int reset_rng, const uint8_t *only_where, void *stream) {
    const OthSizeOps *ops;
    if (int rc = check_state(st, B, board_size, &ops)) return rc;
    if (init_rand_steps < 0 || init_rand_steps > 254) return fail(OTH_ERR_BAD_ARG, "init_rand_steps must be in 0..254");
    if (B == 0) return 0;
    return cuda_result(ops->reset(st, B, seed, env_id_base, init_rand_steps, reset_rng, only_where, (cudaStream_t)stream),
                       "oth_reset");
}

OTH_API int oth_observe(oth_state st, int64_t B, int board_size, int flags, oth_outputs out, void *stream) {
    const OthSizeOps *ops;
    if (int rc = check_state(st, B, board_size, &ops)) return rc;
    if (int rc = check_outputs(out)) return rc;
    if (B == 0) return 0;
    return cuda_result(ops->observe(st, B, flags, out, (cudaStream_t)stream), "oth_observe");
}

OTH_API int oth_step(oth_state st, int64_t B, int board_size, int flags, int action_source, const int64_t *actions,
                     uint64_t seed, int64_t env_id_base, int init_rand_steps, oth_outputs out, void *stream) {
    const OthSizeOps *ops;
    if (int rc = check_state(st, B, board_size, &ops)) return rc;
    if (int rc = check_outputs(out)) return rc;
    if (action_source < OTH_ACT_EXTERNAL || action_source > OTH_ACT_GREEDY) return fail(OTH_ERR_BAD_ARG, "bad action_source");
    if (action_source == OTH_ACT_EXTERNAL && actions == nullptr) return fail(OTH_ERR_NULL_POINTER, "actions is null");
    if (init_rand_steps < 0 || init_rand_steps > 254) return fail(OTH_ERR_BAD_ARG, "init_rand_steps must be in 0..254");
    if (B == 0) return 0;
    StepParams p;
    p.st = st; p.B = B; p.flags = flags; p.action_source = action_source; p.actions = (const i64 *)actions;
    p.seed = seed; p.base = env_id_base; p.init_rand = init_rand_steps; p.out = out;
    return cuda_result(ops->step(p, (cudaStream_t)stream), "oth_step");
}

static int check_spec(const oth_policy_spec &s) {
    if (s.kind < OTH_POLICY_RANDOM || s.kind > OTH_POLICY_ALPHABETA) return fail(OTH_ERR_BAD_ARG, "bad policy kind");
    if (s.kind >= OTH_POLICY_MAXIMIN && (s.depth < 1 || s.depth > OTH_MAX_SEARCH_DEPTH))
        return fail(OTH_ERR_BAD_ARG, "search depth must be in 1..8");
    return 0;
}

OTH_API int oth_search(oth_state st, int64_t B, int board_size, int depth, int flags, int64_t *actions_out,
                       int64_t *nodes_out, void *stream) {
    const OthSizeOps *ops;
    if (int rc = check_state(st, B, board_size, &ops)) return rc;
    if (depth < 1 || depth > OTH_MAX_SEARCH_DEPTH) return fail(OTH_ERR_BAD_ARG, "search depth must be in 1..8");
    if (flags & ~(OTH_SEARCH_PRUNE | OTH_SEARCH_COOPERATIVE)) return fail(OTH_ERR_BAD_ARG, "bad search flags");
    if (actions_out == nullptr) return fail(OTH_ERR_NULL_POINTER, "actions_out is null");
    if (B == 0) return 0;
    return cuda_result(ops->search(st, B, depth, (flags & OTH_SEARCH_PRUNE) ? 1 : 0, (flags & OTH_SEARCH_COOPERATIVE) ? 1 : 0,
                                   (i64 *)actions_out, (i64 *)nodes_out, (cudaStream_t)stream), "oth_search");
}

OTH_API int oth_policy(oth_state st, int64_t B, int board_size, oth_policy_spec spec, uint64_t seed, uint64_t counter,
                       int64_t env_id_base, int64_t *actions_out, void *stream) {
    const OthSizeOps *ops;
    if (int rc = check_state(st, B, board_size, &ops)) return rc;
    if (int rc = check_spec(spec)) return rc;
    if (actions_out == nullptr) return fail(OTH_ERR_NULL_POINTER, "actions_out is null");
    if (B == 0) return 0;
    return cuda_result(ops->policy(st, B, spec.kind, spec.depth, seed, counter, env_id_base, (i64 *)actions_out,
                                   (cudaStream_t)stream), "oth_policy");
}

OTH_API int oth_rollout(int64_t B, int board_size, int flags, oth_policy_spec black, oth_policy_spec white,
                        int init_rand_steps, int games_per_env, uint64_t seed, int64_t env_id_base,
                        unsigned long long *tally, int8_t *game_result, int16_t *game_plies, int16_t *game_counts,
                        int16_t *trace, int trace_stride, void *stream) {
    const OthSizeOps *ops = ops_for(board_size);
    if (ops == nullptr) return fail(OTH_ERR_BAD_BOARD_SIZE, "board_size unsupported");
    if (B < 0 || games_per_env < 0) return fail(OTH_ERR_BAD_ARG, "negative count");
    if (int rc = check_spec(black)) return rc;
    if (int rc = check_spec(white)) return rc;
    if (init_rand_steps < 0 || init_rand_steps > 254) return fail(OTH_ERR_BAD_ARG, "init_rand_steps must be in 0..254");
    if (trace != nullptr && trace_stride <= 0) return fail(OTH_ERR_BAD_ARG, "trace_stride must be positive");
    if (B == 0 || games_per_env == 0) return 0;
    RolloutParams p;
    p.B = B; p.flags = flags | OTH_F_SUDDEN_DEATH; p.black = black; p.white = white; p.init_rand = init_rand_steps;
    p.games_per_env = games_per_env; p.seed = seed; p.base = env_id_base; p.tally = tally;
    p.game_result = game_result; p.game_plies = game_plies; p.game_counts = game_counts; p.trace = trace;
    p.trace_stride = trace_stride;
    return cuda_result(ops->rollout(p, (cudaStream_t)stream), "oth_rollout");
}

// ------------------------------------------------------------------------------------------------
// oth_rollout_noalloc: the fused tournament for callers that cannot allocate device memory (CUDA graph capture,
// tight serving loops, harnesses that forbid cudaMalloc).  The rollout kernels write their per-game records
// (winner int8, plies int16) into a FILE-SCOPE device array instead of caller buffers and a reduction kernel folds
// the records into the caller's 4-word tally.  Nothing here calls cudaMalloc / cudaFree or an allocating routine.
//   - every GPU that loads the library holds the array (OTH_NOALLOC_MAX_GAMES * 3 bytes = 6 MiB) until exit;
//   - all calls on one device share it: two calls whose kernels overlap on the GPU (different streams) race;
//   - nothing in it has to start at zero: the rollout writes records [0, games) and the reduction reads exactly
//     that range, so the contents left by the previous call are never observed.
// In host code `g_noalloc_scratch` is only a stand-in for the symbol: the device address is fetched with
// cudaGetSymbolAddress once per device (that device current) and handed to the kernels as a plain pointer.
static __device__ __align__(256) unsigned char g_noalloc_scratch[(size_t)OTH_NOALLOC_MAX_GAMES * 3];

constexpr int NOALLOC_MAX_DEVICES = 64;
static void *g_noalloc_addr[NOALLOC_MAX_DEVICES];      // written once per device; readers take the same lock
static pthread_mutex_t g_noalloc_lock = PTHREAD_MUTEX_INITIALIZER;

static int noalloc_scratch(unsigned char **out) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return cuda_result(e, "oth_rollout_noalloc: cudaGetDevice");
    if (dev < 0 || dev >= NOALLOC_MAX_DEVICES) return fail(OTH_ERR_BAD_ARG, "oth_rollout_noalloc: device index out of range");
    pthread_mutex_lock(&g_noalloc_lock);
    void *addr = g_noalloc_addr[dev];
    if (addr == nullptr) {
        e = cudaGetSymbolAddress(&addr, g_noalloc_scratch);       // `dev` is current: its copy of the array
        if (e == cudaSuccess) g_noalloc_addr[dev] = addr;
    }
    pthread_mutex_unlock(&g_noalloc_lock);
    if (e != cudaSuccess) return cuda_result(e, "oth_rollout_noalloc: cudaGetSymbolAddress");
    *out = (unsigned char *)addr;
    return 0;
}

// records -> tally[0..3] += {black wins, draws, white wins, plies}: block-reduced, 4 atomics per block
__global__ void __launch_bounds__(256) k_noalloc_reduce(const int16_t *plies, const int8_t *result, i64 games,
                                                        unsigned long long *tally) {
    __shared__ u32 part[4][8];
    u32 c[4] = {0u, 0u, 0u, 0u};
    for (i64 g = (i64)blockIdx.x * blockDim.x + threadIdx.x; g < games; g += (i64)gridDim.x * blockDim.x) {
        const int w = result[g];
        c[0] += w < 0; c[1] += w == 0; c[2] += w > 0; c[3] += (u32)plies[g];
    }
#pragma unroll
    for (int k = 0; k < 4; k++) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) c[k] += __shfl_down_sync(0xFFFFFFFFu, c[k], o);
        if ((threadIdx.x & 31) == 0) part[k][threadIdx.x >> 5] = c[k];
    }
    __syncthreads();
    if (threadIdx.x < 4) {
        unsigned long long s = 0;
        for (int w = 0; w < 8; w++) s += part[threadIdx.x][w];
        if (s) atomicAdd(&tally[threadIdx.x], s);
    }
}

OTH_API int oth_rollout_noalloc(int64_t B, int board_size, int flags, oth_policy_spec black, oth_policy_spec white,
                                int init_rand_steps, int games_per_env, uint64_t seed, int64_t env_id_base,
                                unsigned long long *tally, void *stream) {
    const OthSizeOps *ops = ops_for(board_size);
    if (ops == nullptr) return fail(OTH_ERR_BAD_BOARD_SIZE, "board_size unsupported");
    if (B < 0 || games_per_env < 0) return fail(OTH_ERR_BAD_ARG, "negative count");
    if (int rc = check_spec(black)) return rc;
    if (int rc = check_spec(white)) return rc;
    if (init_rand_steps < 0 || init_rand_steps > 254) return fail(OTH_ERR_BAD_ARG, "init_rand_steps must be in 0..254");
    if (tally == nullptr) return fail(OTH_ERR_NULL_POINTER, "tally is null");
    if (B > OTH_NOALLOC_MAX_GAMES || games_per_env > OTH_NOALLOC_MAX_GAMES ||
        B * (int64_t)games_per_env > OTH_NOALLOC_MAX_GAMES) {
        snprintf(g_err, sizeof(g_err), "oth_rollout_noalloc: %lld games exceed the static scratch (max %d games per call); "
                 "split the call or use oth_rollout", (long long)B * games_per_env, OTH_NOALLOC_MAX_GAMES);
        return OTH_ERR_BAD_ARG;
    }
    if (B == 0 || games_per_env == 0) return 0;
    unsigned char *scratch = nullptr;
    if (int rc = noalloc_scratch(&scratch)) return rc;
    const i64 games = B * (i64)games_per_env;
    int16_t *plies = reinterpret_cast<int16_t *>(scratch);                                   // [MAX] int16
    int8_t *result = reinterpret_cast<int8_t *>(scratch + (size_t)OTH_NOALLOC_MAX_GAMES * 2);   // [MAX] int8
    RolloutParams p;
    p.B = B; p.flags = flags | OTH_F_SUDDEN_DEATH; p.black = black; p.white = white; p.init_rand = init_rand_steps;
    p.games_per_env = games_per_env; p.seed = seed; p.base = env_id_base; p.tally = nullptr;
    p.game_result = result; p.game_plies = plies; p.game_counts = nullptr; p.trace = nullptr; p.trace_stride = 0;
    if (int rc = cuda_result(ops->rollout(p, (cudaStream_t)stream), "oth_rollout_noalloc (rollout)")) return rc;
    const unsigned blocks = (unsigned)((games + 255) / 256 < 592 ? (games + 255) / 256 : 592);   // <= 4 per SM
    k_noalloc_reduce<<<blocks, 256, 0, (cudaStream_t)stream>>>(plies, result, games, tally);
    return cuda_result(cudaGetLastError(), "oth_rollout_noalloc (reduce)");
}

OTH_API int oth_set_board(oth_state st, int64_t B, int board_size, const int8_t *cells, const int8_t *perspective,
                          const int8_t *turn, int recompute_legal, int clear_terminal, void *stream) {
    const OthSizeOps *ops;
    if (int rc = check_state(st, B, board_size, &ops)) return rc;
    if (B == 0) return 0;
    return cuda_result(ops->set_board(st, B, cells, perspective, turn, recompute_legal, clear_terminal,
                                      (cudaStream_t)stream), "oth_set_board");
}

OTH_API int oth_update_board(oth_state st, int64_t B, int board_size, const int64_t *actions, void *stream) {
    const OthSizeOps *ops;
    if (int rc = check_state(st, B, board_size, &ops)) return rc;
    if (actions == nullptr) return fail(OTH_ERR_NULL_POINTER, "actions is null");
    if (B == 0) return 0;
    return cuda_result(ops->update_board(st, B, (const i64 *)actions, (cudaStream_t)stream), "oth_update_board");
}

// ---- PPO self-play protocol ----------------------------------------------------------
static int check_selfplay(const oth_selfplay &sp) {
    if (!sp.protagonist || !sp.phase || !sp.sav_black || !sp.sav_white || !sp.sav_legal || !sp.sav_meta ||
        !sp.sav_choices || !sp.tr_action || !sp.tr_reward || !sp.tr_done || !sp.tr_type || !sp.tr_outputs ||
        !sp.pend_outputs || !sp.pend_action)
        return fail(OTH_ERR_NULL_POINTER, "null self-play pointer");
    return 0;
}

OTH_API int oth_selfplay_new_game(oth_state st, oth_selfplay sp, int64_t B, int board_size, uint64_t seed,
                                  int64_t env_id_base, int init_rand_steps, uint64_t game_index,
                                  const int8_t *protagonist_in, void *stream) {
    const OthSizeOps *ops;
    if (int rc = check_state(st, B, board_size, &ops)) return rc;
    if (int rc = check_selfplay(sp)) return rc;
    if (init_rand_steps < 0 || init_rand_steps > 254) return fail(OTH_ERR_BAD_ARG, "init_rand_steps must be in 0..254");
    if (B == 0) return 0;
    return cuda_result(ops->selfplay_new_game(st, sp, B, seed, env_id_base, init_rand_steps, (u32)game_index,
                                              protagonist_in, (cudaStream_t)stream), "oth_selfplay_new_game");
}

static int selfplay_launch(const OthSizeOps *ops, SelfPlayParams &p, void *stream, const char *what) {
    if (p.need_count == nullptr) return fail(OTH_ERR_NULL_POINTER, "need_count is null");
    cudaError_t e = cudaMemsetAsync(p.need_count, 0, sizeof(int), (cudaStream_t)stream);
    if (e != cudaSuccess) return cuda_result(e, what);
    return cuda_result(ops->selfplay(p, (cudaStream_t)stream), what);
}

OTH_API int oth_selfplay_begin(oth_state st, oth_selfplay sp, int64_t B, int board_size, int flags, oth_outputs out,
                               int32_t *need_count, void *stream) {
    const OthSizeOps *ops;
    if (int rc = check_state(st, B, board_size, &ops)) return rc;
    if (int rc = check_selfplay(sp)) return rc;
    if (int rc = check_outputs(out)) return rc;
    if (B == 0) return 0;
    SelfPlayParams p;
    memset(&p, 0, sizeof(p));
    p.st = st; p.sp = sp; p.B = B; p.flags = flags; p.mode = 0; p.out = out; p.need_count = need_count;
    return selfplay_launch(ops, p, stream, "oth_selfplay_begin");
}

OTH_API int oth_selfplay_advance(oth_state st, oth_selfplay sp, int64_t B, int board_size, int flags,
                                 const int64_t *actions, const float *value, const float *logp, const float *hidden,
                                 uint64_t seed, int64_t env_id_base, int init_rand_steps, oth_outputs out,
                                 int32_t *need_count, void *stream) {
    const OthSizeOps *ops;
    if (int rc = check_state(st, B, board_size, &ops)) return rc;
    if (int rc = check_selfplay(sp)) return rc;
    if (int rc = check_outputs(out)) return rc;
    if (actions == nullptr) return fail(OTH_ERR_NULL_POINTER, "actions is null");
    if (init_rand_steps < 0 || init_rand_steps > 254) return fail(OTH_ERR_BAD_ARG, "init_rand_steps must be in 0..254");
    if (B == 0) return 0;
    SelfPlayParams p;
    memset(&p, 0, sizeof(p));
    p.st = st; p.sp = sp; p.B = B; p.flags = flags; p.mode = 1; p.actions = (const i64 *)actions;
    p.value = value; p.logp = logp; p.hidden = hidden; p.seed = seed; p.base = env_id_base;
    p.init_rand = init_rand_steps; p.out = out; p.need_count = need_count;
    return selfplay_launch(ops, p, stream, "oth_selfplay_advance");
}

// ---- teacher / student protocol (envs.py:44-140, ppo_run_teacher_vs_student.py:425-560) ----
static int check_ts(const oth_teacher_student &t) {
    if (!t.phase || !t.sav_black || !t.sav_white || !t.sav_legal || !t.sav_meta || !t.sav_choices || !t.sav_action ||
        !t.sav_outputs || !t.queued_reward || !t.tr_black || !t.tr_white || !t.tr_legal || !t.tr_choices || !t.tr_meta ||
        !t.tr_action || !t.tr_reward || !t.tr_done || !t.tr_type || !t.tr_who || !t.tr_outputs)
        return fail(OTH_ERR_NULL_POINTER, "null teacher/student pointer");
    return 0;
}

OTH_API int oth_ts_new_game(oth_state st, oth_teacher_student ts, int64_t B, int board_size, uint64_t seed,
                            int64_t env_id_base, int init_rand_steps, void *stream) {
    const OthSizeOps *ops;
    if (int rc = check_state(st, B, board_size, &ops)) return rc;
    if (int rc = check_ts(ts)) return rc;
    if (init_rand_steps < 0 || init_rand_steps > 254) return fail(OTH_ERR_BAD_ARG, "init_rand_steps must be in 0..254");
    if (B == 0) return 0;
    return cuda_result(ops->ts_new_game(st, ts, B, seed, env_id_base, init_rand_steps, (cudaStream_t)stream),
                       "oth_ts_new_game");
}

static int ts_launch(const OthSizeOps *ops, TeacherStudentParams &p, void *stream, const char *what) {
    if (p.need_count == nullptr) return fail(OTH_ERR_NULL_POINTER, "need_count is null");
    if (p.teacher != 1 && p.teacher != -1) return fail(OTH_ERR_BAD_ARG, "teacher must be -1 or +1");
    cudaError_t e = cudaMemsetAsync(p.need_count, 0, 2 * sizeof(int), (cudaStream_t)stream);
    if (e != cudaSuccess) return cuda_result(e, what);
    return cuda_result(ops->teacher_student(p, (cudaStream_t)stream), what);
}

OTH_API int oth_ts_begin(oth_state st, oth_teacher_student ts, int64_t B, int board_size, int flags, int teacher,
                         float teacher_done_reward, oth_outputs out, int32_t *need_count, void *stream) {
    const OthSizeOps *ops;
    if (int rc = check_state(st, B, board_size, &ops)) return rc;
    if (int rc = check_ts(ts)) return rc;
    if (int rc = check_outputs(out)) return rc;
    if (B == 0) return 0;
    TeacherStudentParams p;
    memset(&p, 0, sizeof(p));
    p.st = st; p.ts = ts; p.B = B; p.flags = flags; p.mode = 0; p.teacher = teacher;
    p.teacher_done_reward = teacher_done_reward; p.out = out; p.need_count = need_count;
    return ts_launch(ops, p, stream, "oth_ts_begin");
}

OTH_API int oth_ts_advance(oth_state st, oth_teacher_student ts, int64_t B, int board_size, int flags, int teacher,
                           float teacher_done_reward, int colour, const int64_t *actions, const float *value,
                           const float *logp, const float *hidden, uint64_t seed, int64_t env_id_base,
                           int init_rand_steps, oth_outputs out, int32_t *need_count, void *stream) {
    const OthSizeOps *ops;
    if (int rc = check_state(st, B, board_size, &ops)) return rc;
    if (int rc = check_ts(ts)) return rc;
    if (int rc = check_outputs(out)) return rc;
    if (actions == nullptr) return fail(OTH_ERR_NULL_POINTER, "actions is null");
    if (colour != 1 && colour != -1) return fail(OTH_ERR_BAD_ARG, "colour must be -1 or +1");
    if (init_rand_steps < 0 || init_rand_steps > 254) return fail(OTH_ERR_BAD_ARG, "init_rand_steps must be in 0..254");
    if (B == 0) return 0;
    TeacherStudentParams p;
    memset(&p, 0, sizeof(p));
    p.st = st; p.ts = ts; p.B = B; p.flags = flags; p.mode = 1; p.teacher = teacher;
    p.teacher_done_reward = teacher_done_reward; p.colour = colour; p.actions = (const i64 *)actions;
    p.value = value; p.logp = logp; p.hidden = hidden; p.seed = seed; p.base = env_id_base;
    p.init_rand = init_rand_steps; p.out = out; p.need_count = need_count;
    return ts_launch(ops, p, stream, "oth_ts_advance");
}

// ---- legal-move-restricted categorical (model.py:60-150,156-198) ---------------------
// One warp per env; lane l holds actions l, l+32, ... (N <= 256).  float32, max-subtracted.
#define OTH_STREAM_SAMPLE 3u
constexpr int CAT_MAXK = 8;    // 32 lanes x 8 cells = 256 = the largest board (16 x 16)
__global__ void __launch_bounds__(128) k_masked_categorical(const float *logits, const u64 *legal, i64 B, int N, int mode,
                                                            const i64 *actions_in, u64 seed, u64 counter, i64 base,
                                                            i64 *actions_out, float *logp_out) {
    const i64 env = ((i64)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (env >= B) return;
    const int K = (N + 31) >> 5;
    const float NEG_INF = __int_as_float(0xFF800000u);
    float x[CAT_MAXK];
    bool ok[CAT_MAXK];
    float m = NEG_INF;
#pragma unroll
    for (int k = 0; k < CAT_MAXK; k++) {
        const int a = 32 * k + lane;
        ok[k] = false; x[k] = NEG_INF;
        if (k < K && a < N) {
            const u64 word = legal[(i64)(a >> 6) * B + env];
            ok[k] = (word >> (a & 63)) & 1ull;
            if (ok[k]) x[k] = logits[env * N + a];
        }
        m = fmaxf(m, x[k]);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xFFFFFFFFu, m, o));
    u32 any_legal = 0;
#pragma unroll
    for (int k = 0; k < CAT_MAXK; k++) any_legal |= __ballot_sync(0xFFFFFFFFu, ok[k]);
    if (any_legal == 0u) {                                   // model.py:71-73
        if (lane == 0) {
            if (actions_out) actions_out[env] = (mode == OTH_CAT_EVALUATE && actions_in) ? actions_in[env] : 0;
            if (logp_out) logp_out[env] = 0.f;
        }
        return;
    }
    float ex[CAT_MAXK], s = 0.f;
#pragma unroll
    for (int k = 0; k < CAT_MAXK; k++) { ex[k] = ok[k] ? expf(x[k] - m) : 0.f; s += ex[k]; }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xFFFFFFFFu, s, o);
    const float logZ = m + logf(s);
    int chosen = -1;
    if (mode == OTH_CAT_SAMPLE) {
        const u32 r = philox_draw(seed, base + env, OTH_STREAM_SAMPLE + (u32)(counter >> 32), (u32)counter);
        const float target = (float)(r >> 8) * (1.0f / 16777216.0f) * s;     // u in [0,1) times the mass
        float acc = 0.f;
        int last_legal = -1;
#pragma unroll
        for (int k = 0; k < CAT_MAXK; k++) {
            float scan = ex[k];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const float t = __shfl_up_sync(0xFFFFFFFFu, scan, o);
                if (lane >= o) scan += t;
            }
            const u32 hit = __ballot_sync(0xFFFFFFFFu, ok[k] && (acc + scan > target));
            const u32 lg = __ballot_sync(0xFFFFFFFFu, ok[k]);
            if (chosen < 0 && hit != 0u) chosen = 32 * k + __ffs(hit) - 1;
            if (lg != 0u) last_legal = 32 * k + 31 - __clz(lg);
            acc += __shfl_sync(0xFFFFFFFFu, scan, 31);
        }
        if (chosen < 0) chosen = last_legal;                 // rounding at the top of the CDF
    } else if (mode == OTH_CAT_MODE) {
        float best = NEG_INF;
        int bi = 0x7FFFFFFF;
#pragma unroll
        for (int k = 0; k < CAT_MAXK; k++)
            if (ok[k] && x[k] > best) { best = x[k]; bi = 32 * k + lane; }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float ob = __shfl_xor_sync(0xFFFFFFFFu, best, o);
            const int oi = __shfl_xor_sync(0xFFFFFFFFu, bi, o);
            if (ob > best || (ob == best && oi < bi)) { best = ob; bi = oi; }
        }
        chosen = bi;
    } else {
        const i64 a = actions_in[env];
        chosen = ((u64)a < (u64)N) ? (int)a : -1;
    }
    // log-probability of `chosen` (0 when it is not a legal choice, model.py:172-173)
    float xc = 0.f;
    bool legal_choice = false;
#pragma unroll
    for (int k = 0; k < CAT_MAXK; k++) {
        const float v = __shfl_sync(0xFFFFFFFFu, x[k], chosen & 31);
        const bool g = __shfl_sync(0xFFFFFFFFu, (int)ok[k], chosen & 31);
        if (chosen >= 0 && (chosen >> 5) == k) { xc = v; legal_choice = g; }
    }
    if (lane == 0) {
        if (actions_out) actions_out[env] = (mode == OTH_CAT_EVALUATE) ? actions_in[env] : (i64)chosen;
        if (logp_out) logp_out[env] = legal_choice ? (xc - logZ) : 0.f;
    }
}

OTH_API int oth_masked_categorical(const float *logits, const uint64_t *legal, int64_t B, int board_size, int mode,
                                   const int64_t *actions_in, uint64_t seed, uint64_t counter, int64_t env_id_base,
                                   int64_t *actions_out, float *logp_out, void *stream) {
    const OthSizeOps *ops = ops_for(board_size);
    if (ops == nullptr) return fail(OTH_ERR_BAD_BOARD_SIZE, "board_size unsupported");
    if (B < 0) return fail(OTH_ERR_BAD_ARG, "negative count");
    if (logits == nullptr || legal == nullptr) return fail(OTH_ERR_NULL_POINTER, "logits / legal is null");
    if (mode < OTH_CAT_SAMPLE || mode > OTH_CAT_EVALUATE) return fail(OTH_ERR_BAD_ARG, "bad mode");
    if (mode == OTH_CAT_EVALUATE && actions_in == nullptr) return fail(OTH_ERR_NULL_POINTER, "actions_in is null");
    if (board_size * board_size > 32 * CAT_MAXK) return fail(OTH_ERR_BAD_BOARD_SIZE, "board too large for the sampler");
    if (B == 0) return 0;
    const unsigned grid = (unsigned)((B * 32 + 127) / 128);
    k_masked_categorical<<<grid, 128, 0, (cudaStream_t)stream>>>(logits, (const u64 *)legal, B, board_size * board_size,
                                                                 mode, (const i64 *)actions_in, seed, counter,
                                                                 env_id_base, (i64 *)actions_out, logp_out);
    return cuda_result(cudaGetLastError(), "oth_masked_categorical");
}

// ---- evaluate_actions for the PPO update (model.py:156-198, algo/ppo.py:51-66), differentiable ------------------
// forward:  logp[i]    = log softmax(logits[i][choices_i])[action_i], 0 when the action is not a legal choice or there
//                        is no choice (model.py:167-168 returns the constant 0 there: no gradient either)
//           entropy[i] = entropy of the FULL S*S-way softmax(logits[i]) -- the reference takes dist.entropy() of the
//                        unmasked distribution (model.py:159,178)
// backward: grad_logits[i][a] = g_logp[i] * (1[a == action_i] - p_legal[i][a]) on legal a (0 elsewhere, all-zero row
//                               when logp was the constant 0)  +  g_ent[i] * dH/dx_a,  dH/dx_a = -p_a (log p_a + H)
// One warp per env, lane l holds cells l, l+32, ...; everything is recomputed in the backward (4*S*S bytes per env).
struct CatRow {
    float x[CAT_MAXK];
    bool ok[CAT_MAXK], in[CAT_MAXK];
    float m_all, s_all, m_leg, s_leg;
    bool any_legal;
};
OTH_DEV void cat_row_load(CatRow &r, const float *logits, const u64 *legal, i64 B, int N, i64 env, int lane) {
    const float NEG_INF = __int_as_float(0xFF800000u);
    const int K = (N + 31) >> 5;
    r.m_all = NEG_INF; r.m_leg = NEG_INF;
#pragma unroll
    for (int k = 0; k < CAT_MAXK; k++) {
        const int a = 32 * k + lane;
        r.in[k] = k < K && a < N;
        r.ok[k] = false; r.x[k] = NEG_INF;
        if (r.in[k]) {
            r.x[k] = logits[env * N + a];
            r.ok[k] = (legal[(i64)(a >> 6) * B + env] >> (a & 63)) & 1ull;
            r.m_all = fmaxf(r.m_all, r.x[k]);
            if (r.ok[k]) r.m_leg = fmaxf(r.m_leg, r.x[k]);
        }
    }
    u32 any = 0;
#pragma unroll
    for (int k = 0; k < CAT_MAXK; k++) any |= __ballot_sync(0xFFFFFFFFu, r.ok[k]);
    r.any_legal = any != 0u;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        r.m_all = fmaxf(r.m_all, __shfl_xor_sync(0xFFFFFFFFu, r.m_all, o));
        r.m_leg = fmaxf(r.m_leg, __shfl_xor_sync(0xFFFFFFFFu, r.m_leg, o));
    }
    r.s_all = 0.f; r.s_leg = 0.f;
#pragma unroll
    for (int k = 0; k < CAT_MAXK; k++) {
        if (r.in[k]) r.s_all += expf(r.x[k] - r.m_all);
        if (r.ok[k]) r.s_leg += expf(r.x[k] - r.m_leg);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        r.s_all += __shfl_xor_sync(0xFFFFFFFFu, r.s_all, o);
        r.s_leg += __shfl_xor_sync(0xFFFFFFFFu, r.s_leg, o);
    }
}
// is `a` one of the env's legal choices?  (warp-uniform result)
OTH_DEV bool cat_action_legal(const CatRow &r, i64 a, int N) {
    if ((u64)a >= (u64)N) return false;
    bool g = false;
#pragma unroll
    for (int k = 0; k < CAT_MAXK; k++) {
        const bool v = __shfl_sync(0xFFFFFFFFu, (int)r.ok[k], (int)a & 31);
        if (((int)a >> 5) == k) g = v;
    }
    return g;
}
// entropy of the full distribution: H = logZ - sum_a p_a x_a
OTH_DEV float cat_entropy(const CatRow &r) {
    float px = 0.f;
#pragma unroll
    for (int k = 0; k < CAT_MAXK; k++)
        if (r.in[k]) px += expf(r.x[k] - r.m_all) * r.x[k];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) px += __shfl_xor_sync(0xFFFFFFFFu, px, o);
    return r.m_all + logf(r.s_all) - px / r.s_all;
}

__global__ void __launch_bounds__(128) k_masked_logprob_fwd(const float *logits, const u64 *legal, i64 B, int N,
                                                            const i64 *actions, float *logp_out, float *entropy_out) {
    const i64 env = ((i64)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (env >= B) return;
    CatRow r;
    cat_row_load(r, logits, legal, B, N, env, lane);
    const i64 a = actions[env];
    const bool legal_choice = r.any_legal && cat_action_legal(r, a, N);
    float xa = 0.f;
    if (legal_choice) xa = logits[env * N + a];
    const float H = entropy_out != nullptr ? cat_entropy(r) : 0.f;
    if (lane == 0) {
        if (logp_out) logp_out[env] = legal_choice ? (xa - r.m_leg - logf(r.s_leg)) : 0.f;
        if (entropy_out) entropy_out[env] = H;
    }
}

__global__ void __launch_bounds__(128) k_masked_logprob_bwd(const float *logits, const u64 *legal, i64 B, int N,
                                                            const i64 *actions, const float *g_logp, const float *g_ent,
                                                            float *grad_logits) {
    const i64 env = ((i64)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (env >= B) return;
    CatRow r;
    cat_row_load(r, logits, legal, B, N, env, lane);
    const i64 a = actions[env];
    const bool legal_choice = r.any_legal && cat_action_legal(r, a, N);
    const float gl = (g_logp != nullptr && legal_choice) ? g_logp[env] : 0.f;
    const float ge = g_ent != nullptr ? g_ent[env] : 0.f;
    float H = 0.f, logZ = 0.f;
    if (g_ent != nullptr) { H = cat_entropy(r); logZ = r.m_all + logf(r.s_all); }
#pragma unroll
    for (int k = 0; k < CAT_MAXK; k++) {
        if (!r.in[k]) continue;
        const int c = 32 * k + lane;
        float g = 0.f;
        if (r.ok[k]) g = gl * ((c == (int)a ? 1.f : 0.f) - expf(r.x[k] - r.m_leg) / r.s_leg);
        if (g_ent != nullptr) {
            const float lp = r.x[k] - logZ;
            g += ge * (-expf(lp) * (lp + H));
        }
        grad_logits[env * N + c] = g;
    }
}

static int check_logprob_args(const float *logits, const uint64_t *legal, int64_t B, int board_size, const int64_t *actions) {
    if (ops_for(board_size) == nullptr) return fail(OTH_ERR_BAD_BOARD_SIZE, "board_size unsupported");
    if (B < 0) return fail(OTH_ERR_BAD_ARG, "negative count");
    if (logits == nullptr || legal == nullptr || actions == nullptr) return fail(OTH_ERR_NULL_POINTER, "logits / legal / actions is null");
    if (board_size * board_size > 32 * CAT_MAXK) return fail(OTH_ERR_BAD_BOARD_SIZE, "board too large for the sampler");
    return 0;
}

OTH_API int oth_masked_logprob_forward(const float *logits, const uint64_t *legal, int64_t B, int board_size,
                                       const int64_t *actions, float *logp_out, float *entropy_out, void *stream) {
    if (int rc = check_logprob_args(logits, legal, B, board_size, actions)) return rc;
    if (logp_out == nullptr && entropy_out == nullptr) return fail(OTH_ERR_NULL_POINTER, "no output requested");
    if (B == 0) return 0;
    const unsigned grid = (unsigned)((B * 32 + 127) / 128);
    k_masked_logprob_fwd<<<grid, 128, 0, (cudaStream_t)stream>>>(logits, (const u64 *)legal, B, board_size * board_size,
                                                                 (const i64 *)actions, logp_out, entropy_out);
    return cuda_result(cudaGetLastError(), "oth_masked_logprob_forward");
}

OTH_API int oth_masked_logprob_backward(const float *logits, const uint64_t *legal, int64_t B, int board_size,
                                        const int64_t *actions, const float *grad_logp, const float *grad_entropy,
                                        float *grad_logits, void *stream) {
    if (int rc = check_logprob_args(logits, legal, B, board_size, actions)) return rc;
    if (grad_logits == nullptr) return fail(OTH_ERR_NULL_POINTER, "grad_logits is null");
    if (B == 0) return 0;
    const unsigned grid = (unsigned)((B * 32 + 127) / 128);
    k_masked_logprob_bwd<<<grid, 128, 0, (cudaStream_t)stream>>>(logits, (const u64 *)legal, B, board_size * board_size,
                                                                 (const i64 *)actions, grad_logp, grad_entropy, grad_logits);
    return cuda_result(cudaGetLastError(), "oth_masked_logprob_backward");
}

// ---- epsilon-greedy over the legal moves (dqn.py:264-275,478-488) --------------------
// One warp per env, lane l holds q of actions l, l+32, ...  Explore: the n-th set bit of the
// legal bitset (= the n-th entry of the ascending possible_moves list).  Exploit: first argmax.
#define OTH_STREAM_EPSILON 4u
__global__ void __launch_bounds__(128) k_epsilon_greedy(const float *q, const u64 *legal, i64 B, int N, float epsilon,
                                                        u64 seed, u64 counter, i64 base, i64 *actions_out,
                                                        float *q_chosen_out, uint8_t *explored_out) {
    const i64 env = ((i64)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (env >= B) return;
    const int K = (N + 31) >> 5, W = (N + 63) >> 6;
    const float NEG_INF = __int_as_float(0xFF800000u);
    u64 words[4] = {0ull, 0ull, 0ull, 0ull};                       // W <= 4 (16 x 16)
    int n_legal = 0;
#pragma unroll
    for (int w = 0; w < 4; w++)
        if (w < W) { words[w] = legal[(i64)w * B + env]; n_legal += __popcll(words[w]); }
    float x[CAT_MAXK];
    float best = NEG_INF;
    int bi = 0x7FFFFFFF;
#pragma unroll
    for (int k = 0; k < CAT_MAXK; k++) {
        const int a = 32 * k + lane;
        x[k] = 0.f;
        if (k < K && a < N) {
            const u64 word = (a >> 6) == 0 ? words[0] : ((a >> 6) == 1 ? words[1] : ((a >> 6) == 2 ? words[2] : words[3]));
            const bool ok = (word >> (a & 63)) & 1ull;
            if (ok) {
                x[k] = q[env * N + a];
                // strict > over ascending actions inside the lane; NaN never wins, as in a max-scan
                if (bi == 0x7FFFFFFF || x[k] > best) { best = x[k]; bi = a; }
            }
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float ob = __shfl_xor_sync(0xFFFFFFFFu, best, o);
        const int oi = __shfl_xor_sync(0xFFFFFFFFu, bi, o);
        if (oi != 0x7FFFFFFF && (bi == 0x7FFFFFFF || ob > best || (ob == best && oi < bi))) { best = ob; bi = oi; }
    }
    int chosen = 0;
    bool explored = false;
    if (n_legal > 0) {
        const u32 ctr0 = (u32)(2ull * counter), hi = (u32)((2ull * counter) >> 32);
        const u32 r0 = philox_draw(seed, base + env, OTH_STREAM_EPSILON + 16u * hi, ctr0);
        const float u = (float)(r0 >> 8) * (1.0f / 16777216.0f);
        explored = epsilon >= u;                                   // dqn.py:266
        if (explored) {
            const u32 r1 = philox_draw(seed, base + env, OTH_STREAM_EPSILON + 16u * hi, ctr0 + 1u);
            int n = bounded(r1, n_legal);                          // random.randrange(len(possible_moves))
            int wsel = 0;
#pragma unroll
            for (int w = 0; w < 3; w++) {
                const int c = __popcll(words[w]);
                if (wsel == w && n >= c) { n -= c; wsel = w + 1; }
            }
            const u64 word = wsel == 0 ? words[0] : (wsel == 1 ? words[1] : (wsel == 2 ? words[2] : words[3]));
            chosen = 64 * wsel + nth_bit64(word, n);
        } else {
            chosen = bi;
        }
    }
    float qc = 0.f;
#pragma unroll
    for (int k = 0; k < CAT_MAXK; k++) {
        const float v = __shfl_sync(0xFFFFFFFFu, x[k], chosen & 31);
        if ((chosen >> 5) == k) qc = v;
    }
    if (lane == 0) {
        if (actions_out) actions_out[env] = (i64)chosen;
        if (q_chosen_out) q_chosen_out[env] = n_legal > 0 ? qc : 0.f;
        if (explored_out) explored_out[env] = explored ? 1 : 0;
    }
}

OTH_API int oth_epsilon_greedy(const float *q, const uint64_t *legal, int64_t B, int board_size, float epsilon,
                               uint64_t seed, uint64_t counter, int64_t env_id_base, int64_t *actions_out,
                               float *q_chosen_out, uint8_t *explored_out, void *stream) {
    const OthSizeOps *ops = ops_for(board_size);
    if (ops == nullptr) return fail(OTH_ERR_BAD_BOARD_SIZE, "board_size unsupported");
    if (B < 0) return fail(OTH_ERR_BAD_ARG, "negative count");
    if (q == nullptr || legal == nullptr) return fail(OTH_ERR_NULL_POINTER, "q / legal is null");
    if (actions_out == nullptr) return fail(OTH_ERR_NULL_POINTER, "actions_out is null");
    if (!(epsilon >= 0.f && epsilon <= 1.f)) return fail(OTH_ERR_BAD_ARG, "epsilon must be in [0, 1]");
    if (board_size * board_size > 32 * CAT_MAXK) return fail(OTH_ERR_BAD_BOARD_SIZE, "board too large for the sampler");
    if (B == 0) return 0;
    const unsigned grid = (unsigned)((B * 32 + 127) / 128);
    k_epsilon_greedy<<<grid, 128, 0, (cudaStream_t)stream>>>(q, (const u64 *)legal, B, board_size * board_size, epsilon,
                                                             seed, counter, env_id_base, (i64 *)actions_out,
                                                             q_chosen_out, explored_out);
    return cuda_result(cudaGetLastError(), "oth_epsilon_greedy");
}

// ---- INT32 ALU probe: 8 independent LOP3/SHF/IADD3 chains per thread -----------------
#define PROBE_UNROLL 16
__global__ void k_int_alu_probe(uint32_t *sink, int iters) {
    uint32_t a0 = threadIdx.x, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7;
    const uint32_t k = blockIdx.x | 0x9E3779B9u;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int u = 0; u < PROBE_UNROLL; u++) {
            // per chain: one LOP3 (xor-and-or), one funnel shift, one add = 3 ALU-pipe instructions
#define PROBE_CHAIN(x)                                                                   \
            asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(x) : "r"(k), "r"(a0 ^ 0)); \
            asm volatile("shf.l.wrap.b32 %0, %0, %0, 5;" : "+r"(x));                      \
            asm volatile("add.u32 %0, %0, %1;" : "+r"(x) : "r"(k));
            PROBE_CHAIN(a1) PROBE_CHAIN(a2) PROBE_CHAIN(a3) PROBE_CHAIN(a4)
            PROBE_CHAIN(a5) PROBE_CHAIN(a6) PROBE_CHAIN(a7)
        }
    }
    uint32_t r = a1 ^ a2 ^ a3 ^ a4 ^ a5 ^ a6 ^ a7;
    if (r == 0x12345u) sink[0] = r;   // practically never; keeps the chains live
}

OTH_API int oth_int_alu_probe(uint32_t *sink, int blocks, int threads, int iters, uint64_t *instr_out_host, void *stream) {
    if (sink == nullptr) return fail(OTH_ERR_NULL_POINTER, "sink is null");
    if (blocks <= 0 || threads <= 0 || threads > 1024 || iters <= 0) return fail(OTH_ERR_BAD_ARG, "bad launch shape");
    k_int_alu_probe<<<blocks, threads, 0, (cudaStream_t)stream>>>(sink, iters);
    if (instr_out_host) *instr_out_host = (uint64_t)blocks * threads * (uint64_t)iters * PROBE_UNROLL * 7 * 3;
    return cuda_result(cudaGetLastError(), "oth_int_alu_probe");
}
