// oth_cabi.cu -- the extern "C" boundary declared in include/othello_b200.h.
// Argument checking, board-size dispatch and error reporting only; all device work is
// in oth_kernels.cuh.  No torch types, no global mutable state, never throws.
#include <cstdio>
#include <cstring>
#include "oth_kernels.cuh"

#define OTH_API extern "C" __attribute__((visibility("default")))

#define OTH_DECL(S) extern "C" const OthSizeOps *oth_size_ops_##S();
OTH_DECL(4) OTH_DECL(5) OTH_DECL(6) OTH_DECL(7) OTH_DECL(8) OTH_DECL(9) OTH_DECL(10) OTH_DECL(11) OTH_DECL(12)

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
    if (s.kind < OTH_POLICY_RANDOM || s.kind > OTH_POLICY_MAXIMIN) return fail(OTH_ERR_BAD_ARG, "bad policy kind");
    if (s.kind == OTH_POLICY_MAXIMIN && (s.depth < 1 || s.depth > OTH_MAX_SEARCH_DEPTH))
        return fail(OTH_ERR_BAD_ARG, "maximin depth must be in 1..4");
    return 0;
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
                        unsigned long long *tally, int8_t *game_result, int16_t *game_plies, uint8_t *game_counts,
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
