// TEST INFRASTRUCTURE -- compiles gymothelloenv_b200/csrc/oth_core.cuh (the device rules engine) for the HOST with
// g++, CUDA intrinsics replaced by portable shims, so that the bit arithmetic of the kernels (carry-propagation
// moves / flips, Kogge-Stone fills, multi-word carry chains, n-th set bit, greedy / maximin picks) can be checked
// against the C oracle on a machine without a GPU (tests/test_core_host_emul.py, `-m "not gpu"`).
// This is NOT a CPU path of the product: nothing in gymothelloenv_b200/ builds, loads or calls it.
#include <algorithm>
#include <cstdint>
#include <cstring>

#define OTH_HOST_EMUL 1
#define __device__
#define __host__
#define __forceinline__ inline
#define __noinline__
static inline unsigned long long __brevll(unsigned long long x) {
    unsigned long long r = 0;
    for (int i = 0; i < 64; i++) r |= ((x >> i) & 1ull) << (63 - i);
    return r;
}
static inline int __popcll(unsigned long long x) { return __builtin_popcountll(x); }
static inline int __popc(unsigned x) { return __builtin_popcount(x); }
static inline int __ffsll(long long x) { return __builtin_ffsll(x); }
static inline int __ffs(int x) { return __builtin_ffs(x); }
static inline unsigned __umulhi(unsigned a, unsigned b) { return (unsigned)(((unsigned long long)a * b) >> 32); }
using std::max;
using std::min;

#include "../../gymothelloenv_b200/csrc/oth_core.cuh"
#include "../../gymothelloenv_b200/csrc/oth_search.cuh"

template <int S>
static void step_batch(int64_t B, uint64_t *black, uint64_t *white, uint64_t *legal_, uint32_t *meta, const int64_t *actions,
                       int flags, int32_t *reward, uint8_t *done, int8_t *winner) {
    constexpr int W = Geo<S>::W;
    for (int64_t i = 0; i < B; i++) {
        EnvRegs<S> e;
        for (int k = 0; k < W; k++) { e.black.w[k] = black[k * B + i]; e.white.w[k] = white[k * B + i]; e.legal.w[k] = legal_[k * B + i]; }
        e.meta = meta[i]; e.ctr = 0;
        if (e.meta & OTH_META_TERMINATED) { reward[i] = 0; done[i] = 1; winner[i] = (int8_t)meta_winner(e.meta); continue; }
        StepResult r = env_step<S>(e, actions[i], flags);
        for (int k = 0; k < W; k++) { black[k * B + i] = e.black.w[k]; white[k * B + i] = e.white.w[k]; legal_[k * B + i] = e.legal.w[k]; }
        meta[i] = e.meta;
        reward[i] = r.reward; done[i] = (uint8_t)r.done; winner[i] = (int8_t)r.winner;
    }
}
template <int S>
static void reset_batch(int64_t B, uint64_t *black, uint64_t *white, uint64_t *legal_, uint32_t *meta) {
    constexpr int W = Geo<S>::W;
    for (int64_t i = 0; i < B; i++) {
        EnvRegs<S> e;
        e.ctr = 0;
        env_reset<S>(e, (u64)0, i, 0);
        for (int k = 0; k < W; k++) { black[k * B + i] = e.black.w[k]; white[k * B + i] = e.white.w[k]; legal_[k * B + i] = e.legal.w[k]; }
        meta[i] = e.meta;
    }
}
template <int S>
static void policy_batch(int64_t B, const uint64_t *black, const uint64_t *white, const uint64_t *legal_, const uint32_t *meta,
                         int kind, int depth, int64_t *out) {
    constexpr int W = Geo<S>::W;
    for (int64_t i = 0; i < B; i++) {
        BB<W> b, w, l;
        for (int k = 0; k < W; k++) { b.w[k] = black[k * B + i]; w.w[k] = white[k * B + i]; l.w[k] = legal_[k * B + i]; }
        if ((meta[i] & OTH_META_TERMINATED) || !any(l)) { out[i] = -1; continue; }
        const bool wm = meta[i] & OTH_META_TURN_WHITE;
        out[i] = policy_pick<S, true>(kind, depth, wm ? w : b, wm ? b : w, l, 0u);
    }
}
// generic-depth maximin with / without alpha-beta pruning: action and positions scored per env
template <int S>
static void search_batch(int64_t B, const uint64_t *black, const uint64_t *white, const uint64_t *legal_, const uint32_t *meta,
                         int depth, int prune, int64_t *out, int64_t *nodes_out) {
    constexpr int W = Geo<S>::W;
    for (int64_t i = 0; i < B; i++) {
        BB<W> b, w, l;
        for (int k = 0; k < W; k++) { b.w[k] = black[k * B + i]; w.w[k] = white[k * B + i]; l.w[k] = legal_[k * B + i]; }
        unsigned long long nodes = 0;
        if ((meta[i] & OTH_META_TERMINATED) || !any(l)) { out[i] = -1; nodes_out[i] = 0; continue; }
        const bool wm = meta[i] & OTH_META_TURN_WHITE;
        out[i] = search_pick<S>(wm ? w : b, wm ? b : w, l, depth, prune != 0, nodes);
        nodes_out[i] = (int64_t)nodes;
    }
}
// n-th set bit of the cached legal set (the random policy's pick for a given index)
template <int S>
static void nth_batch(int64_t B, const uint64_t *legal_, const int32_t *n, int64_t *out) {
    constexpr int W = Geo<S>::W;
    for (int64_t i = 0; i < B; i++) {
        BB<W> l;
        for (int k = 0; k < W; k++) l.w[k] = legal_[k * B + i];
        out[i] = nth_bit<W>(l, n[i]);
    }
}

// bit-sliced flip counts (flip_planes1 / flip_planes / max_flips) next to the per-move flips_at + legal they replace;
// boards are [W][B] words, counts [B][64*W]
template <int S>
static void flipcount_batch(int64_t B, const uint64_t *own, const uint64_t *opp, int8_t *sliced, int8_t *per_move,
                            uint64_t *moves_sliced, uint64_t *moves_legal, int32_t *max_sliced) {
    constexpr int N = Geo<S>::N, W = Geo<S>::W;
    for (int64_t i = 0; i < B; i++) {
        BB<W> P, O;
        for (int k = 0; k < W; k++) { P.w[k] = own[k * B + i]; O.w[k] = opp[k * B + i]; }
        uint64_t plane[6][W];
        if constexpr (W == 1) {
            const FlipPlanes f = flip_planes1<S>(P.w[0], O.w[0]);
            for (int p = 0; p < 5; p++) plane[p][0] = f.c[p];
            plane[5][0] = 0;
        } else {
            const FlipPlanesW<W> f = flip_planes<S>(P, O);
            for (int p = 0; p < 6; p++) for (int k = 0; k < W; k++) plane[p][k] = f.c[p].w[k];
        }
        for (int a = 0; a < 64 * W; a++) {
            int c = 0;
            for (int p = 0; p < 6; p++) c |= (int)((plane[p][a >> 6] >> (a & 63)) & 1ull) << p;
            const bool empty = a < N && !(((P.w[a >> 6] | O.w[a >> 6]) >> (a & 63)) & 1ull);
            sliced[i * 64 * W + a] = (int8_t)(empty ? c : -1);
            per_move[i * 64 * W + a] = (int8_t)(empty ? popc(flips_at<S>(a, bb_bit<W>(a), P, O)) : -1);
        }
        BB<W> mv;
        max_sliced[i] = max_flips<S>(P, O, mv);
        const BB<W> lg = legal<S>(P, O);
        for (int k = 0; k < W; k++) { moves_sliced[k * B + i] = mv.w[k]; moves_legal[k * B + i] = lg.w[k]; }
    }
}

#define DISPATCH(S_, CALL)                                                                                   \
    switch (S_) {                                                                                            \
    case 4: CALL(4); break; case 5: CALL(5); break; case 6: CALL(6); break; case 7: CALL(7); break;          \
    case 8: CALL(8); break; case 9: CALL(9); break; case 10: CALL(10); break; case 11: CALL(11); break;      \
    case 12: CALL(12); break; case 13: CALL(13); break; case 14: CALL(14); break; case 15: CALL(15); break;  \
    case 16: CALL(16); break; default: return -1;                                                            \
    }

extern "C" {
int emul_words(int S) { return (S * S + 63) / 64; }
int emul_reset(int S, int64_t B, uint64_t *black, uint64_t *white, uint64_t *legal_, uint32_t *meta) {
#define CALL(N) reset_batch<N>(B, black, white, legal_, meta)
    DISPATCH(S, CALL)
#undef CALL
    return 0;
}
int emul_step(int S, int64_t B, uint64_t *black, uint64_t *white, uint64_t *legal_, uint32_t *meta, const int64_t *actions,
              int flags, int32_t *reward, uint8_t *done, int8_t *winner) {
#define CALL(N) step_batch<N>(B, black, white, legal_, meta, actions, flags, reward, done, winner)
    DISPATCH(S, CALL)
#undef CALL
    return 0;
}
int emul_policy(int S, int64_t B, const uint64_t *black, const uint64_t *white, const uint64_t *legal_, const uint32_t *meta,
                int kind, int depth, int64_t *out) {
#define CALL(N) policy_batch<N>(B, black, white, legal_, meta, kind, depth, out)
    DISPATCH(S, CALL)
#undef CALL
    return 0;
}
int emul_search(int S, int64_t B, const uint64_t *black, const uint64_t *white, const uint64_t *legal_, const uint32_t *meta,
                int depth, int prune, int64_t *out, int64_t *nodes_out) {
#define CALL(N) search_batch<N>(B, black, white, legal_, meta, depth, prune, out, nodes_out)
    DISPATCH(S, CALL)
#undef CALL
    return 0;
}
int emul_nth(int S, int64_t B, const uint64_t *legal_, const int32_t *n, int64_t *out) {
#define CALL(N) nth_batch<N>(B, legal_, n, out)
    DISPATCH(S, CALL)
#undef CALL
    return 0;
}
int emul_flipcount(int S, int64_t B, const uint64_t *own, const uint64_t *opp, int8_t *sliced, int8_t *per_move,
                   uint64_t *moves_sliced, uint64_t *moves_legal1, int32_t *max_sliced) {
#define CALL(N) flipcount_batch<N>(B, own, opp, sliced, per_move, moves_sliced, moves_legal1, max_sliced)
    DISPATCH(S, CALL)
#undef CALL
    return 0;
}
// Philox4x32-10 block, for tests/philox_ref.py cross-checks
void emul_philox(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1, uint32_t *out) {
    Philox4 b = philox4x32_10(c0, c1, c2, c3, k0, k1);
    for (int i = 0; i < 4; i++) out[i] = b.v[i];
}
}
