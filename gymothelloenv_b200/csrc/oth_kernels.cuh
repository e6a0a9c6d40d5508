// oth_kernels.cuh -- the sm_100a kernels of the Othello env engine for one board size S.
//
// Mapping: one thread = one env (bitboards in registers); 128-thread CTAs; state is SoA
// in HBM so every state access is a coalesced 8-byte (or 4-byte) load/store per lane.
// Dense outputs (obs planes, legal mask) are written WARP-COOPERATIVELY: each lane parks
// its env's bitboards in a per-warp shared-memory slab, then the 32 lanes walk the
// warp's contiguous output slab in 16-byte chunks so every store instruction covers
// 512 contiguous bytes (st.global.v4), instead of 32 strided 64-byte rows.
#pragma once
#include <cuda_runtime.h>
#include "oth_core.cuh"
#include "oth_search.cuh"
#include "oth_coop_search.cuh"

#ifndef OTH_STEPN_MIN_CTAS
#define OTH_STEPN_MIN_CTAS 0   // min resident CTAs per SM the fast step kernels of sizes other than 8 are compiled for
#endif
constexpr int OTH_BLOCK = 128;
constexpr int OTH_WARPS = OTH_BLOCK / 32;
constexpr int OTH_BLOCK_SMALL = 64;   // CTA size of the step kernels for batches that leave an SM only a few CTAs
#ifndef OTH_SMALL_BATCH_CTAS_PER_SM
#define OTH_SMALL_BATCH_CTAS_PER_SM 8  // half-size CTAs are considered below this many 128-env CTAs per SM (0: never)
#endif

enum { ROW_BLACK = 0, ROW_WHITE = 1, ROW_LEGAL = 2, ROW_ZERO = 3, ROW_ONES = 4, ROW_COUNT = 5 };
#define OTH_LAYOUT_MASK_INTERNAL 6   // the dense legal_mask buffer, written with the same machinery

// Even board sizes (S*S % 4 == 0) stage NIBBLES (see "nibble staging" below); odd sizes stage
// the raw bitboards and pay per-cell index arithmetic in warp_write_planes.
enum { NIB_POS0 = 0, NIB_POS1 = 1, NIB_POS2 = 2, NIB_POS3 = 3, NIB_NEG0 = 4, NIB_MASK = 5, NIB_ROWS = 6 };
template <int S, bool NIB = ((S * S) % 4 == 0)>
struct WarpStageT {
    union {
        u64 row[ROW_COUNT][Geo<S>::W][32];
        uint4 slab[(32 * S * S + 15) / 16];               // odd-size slab writer (int8 outputs): 32 envs x S*S bytes
    };
};
template <int S>
struct WarpStageT<S, true> {
    // [row][env lane][16 nibbles per 64 cells, one per byte]; the slab writers reuse the same bytes.  10x10 gets one
    // more row (1 KB per warp) so that TWO output slabs (obs + mask, 2 x 3200 B) fit: the bulk-store path needs both.
    uint4 nib[NIB_ROWS + (S == 10 ? 1 : 0)][32][Geo<S>::W];
};
template <int S>
using WarpStage = WarpStageT<S>;

__host__ __device__ constexpr int layout_planes(int layout) {
    return layout == OTH_OBS_BOARD_MASK ? 2 : (layout == OTH_OBS_PLANES4 ? 4 : (layout == OTH_OBS_PLANES3 ? 3 : 1));
}

// which staged bitboard gives the +1 cells / the -1 cells of plane p
template <int LAYOUT>
OTH_DEV void plane_rows(int p, bool turn_white, bool quirk_empty, bool zero, int &pos, int &neg) {
    neg = ROW_ZERO;
    if constexpr (LAYOUT == OTH_OBS_BOARD) {                 // othello.py:363-369
        pos = turn_white ? ROW_WHITE : ROW_BLACK;
        neg = turn_white ? ROW_BLACK : ROW_WHITE;
    } else if constexpr (LAYOUT == OTH_OBS_BOARD_MASK) {     // othello.py:370-376
        pos = p == 0 ? (turn_white ? ROW_WHITE : ROW_BLACK) : ROW_LEGAL;
        neg = p == 0 ? (turn_white ? ROW_BLACK : ROW_WHITE) : ROW_ZERO;
    } else if constexpr (LAYOUT == OTH_OBS_PLANES4 || LAYOUT == OTH_OBS_PLANES3) {   // util.py:48-74
        pos = p == 0 ? ROW_BLACK
            : p == 1 ? ROW_WHITE
            : p == 2 ? (turn_white ? ROW_ONES : ROW_ZERO)
                     : (quirk_empty ? ROW_ZERO : ROW_LEGAL);
    } else if constexpr (LAYOUT == OTH_OBS_ABSOLUTE) {       // board_state (othello.py:10-12)
        pos = ROW_WHITE;
        neg = ROW_BLACK;
    } else {                                                 // dense legal mask
        pos = ROW_LEGAL;
    }
    if (zero) { pos = ROW_ZERO; neg = ROW_ZERO; }
}

// 4 bits -> 4 bytes of 0/1 (bit i of n lands on bit 8*i: n*0x00204081 places it at
// i, i+7, i+14, i+21 without carries; the mask keeps position 8*i)
OTH_DEV u32 spread4(u32 n) { return (n * 0x00204081u) & 0x01010101u; }
// one staged nibble -> register (zero-extended byte load).  Spelled as PTX so that nvcc cannot
// merge four neighbouring byte loads into one LDS.32 + four PRMT extractions: the extraction
// would sit on the ALU pipe (the binding unit of k_step), the byte loads sit on the idle LSU.
OTH_DEV u32 lds_u8(u32 saddr) {
    u32 v;
    asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(saddr) : "memory");
    return v;
}
OTH_DEV u32 cells4_i8(u32 pn, u32 nn) { return spread4(nn) * 0xFFu + spread4(pn); }   // +1 / -1 bytes
OTH_DEV u32 cell_f32(u32 pbit, u32 nbit) { return (pbit ? 0x3F800000u : 0u) | (nbit ? 0xBF800000u : 0u); }

template <int S, int LAYOUT, int ES>
__device__ __noinline__ void warp_write_planes(void *base, i64 env0, int nenv, const WarpStage<S> &st,
                                               u32 turn_ballot, u32 quirk_ballot, u32 zero_ballot, int lane) {
    constexpr int N = Geo<S>::N;
    constexpr int P = layout_planes(LAYOUT);
    constexpr int EPC = 16 / ES;                                   // elements per 16-byte chunk
    constexpr int G = (ES == 1) ? (N % 16 == 0 ? 16 : (N % 4 == 0 ? 4 : 1)) : (N % 4 == 0 ? 4 : 1);
    constexpr int GPC = EPC / G;                                   // groups per chunk
    const int slab_bytes = nenv * P * N * ES;
    const int nchunks = (slab_bytes + 15) >> 4;
    char *slab = reinterpret_cast<char *>(base) + env0 * (i64)(P * N * ES);
    for (int c = lane; c < nchunks; c += 32) {
        u32 words[4] = {0u, 0u, 0u, 0u};
#pragma unroll
        for (int g = 0; g < GPC; g++) {
            const int fe = c * EPC + g * G;                        // first element of the group in the slab
            int e = fe / (P * N);
            const int rem = fe - e * (P * N);
            const int p = rem / N;
            const int sq = rem - p * N;
            e = min(e, 31);
            int pos, neg;
            plane_rows<LAYOUT>(p, (turn_ballot >> e) & 1u, (quirk_ballot >> e) & 1u, (zero_ballot >> e) & 1u, pos, neg);
            const int wi = (Geo<S>::W == 1) ? 0 : (sq >> 6);
            const u32 pb = (u32)(st.row[pos][wi][e] >> (sq & 63)) & ((1u << G) - 1u);
            const u32 nb = (u32)(st.row[neg][wi][e] >> (sq & 63)) & ((1u << G) - 1u);
            if constexpr (ES == 1) {
                if constexpr (G == 16) {
#pragma unroll
                    for (int k = 0; k < 4; k++) words[k] = cells4_i8((pb >> (4 * k)) & 0xFu, (nb >> (4 * k)) & 0xFu);
                } else if constexpr (G == 4) {
                    words[g] = cells4_i8(pb, nb);
                } else {
                    words[g >> 2] |= ((pb ? 0x01u : 0u) | (nb ? 0xFFu : 0u)) << (8 * (g & 3));
                }
            } else {
                if constexpr (G == 4) {
#pragma unroll
                    for (int k = 0; k < 4; k++) words[k] = cell_f32((pb >> k) & 1u, (nb >> k) & 1u);
                } else {
                    words[g] = cell_f32(pb, nb);
                }
            }
        }
        char *dst = slab + (i64)c * 16;
        if (c * 16 + 16 <= slab_bytes) {
            *reinterpret_cast<uint4 *>(dst) = make_uint4(words[0], words[1], words[2], words[3]);
        } else {   // ragged tail of the last warp
            for (int b = 0; c * 16 + b < slab_bytes; b++) dst[b] = (char)(words[b >> 2] >> (8 * (b & 3)));
        }
    }
}

template <int S, int ES>
OTH_DEV void warp_write_layout(int layout, void *base, i64 env0, int nenv, const WarpStage<S> &st, u32 turn_ballot,
                               u32 quirk_ballot, u32 zero_ballot, int lane) {
    switch (layout) {
    case OTH_OBS_BOARD: warp_write_planes<S, OTH_OBS_BOARD, ES>(base, env0, nenv, st, turn_ballot, quirk_ballot, zero_ballot, lane); break;
    case OTH_OBS_BOARD_MASK: warp_write_planes<S, OTH_OBS_BOARD_MASK, ES>(base, env0, nenv, st, turn_ballot, quirk_ballot, zero_ballot, lane); break;
    case OTH_OBS_PLANES4: warp_write_planes<S, OTH_OBS_PLANES4, ES>(base, env0, nenv, st, turn_ballot, quirk_ballot, zero_ballot, lane); break;
    case OTH_OBS_PLANES3: warp_write_planes<S, OTH_OBS_PLANES3, ES>(base, env0, nenv, st, turn_ballot, quirk_ballot, zero_ballot, lane); break;
    case OTH_OBS_ABSOLUTE: warp_write_planes<S, OTH_OBS_ABSOLUTE, ES>(base, env0, nenv, st, turn_ballot, quirk_ballot, zero_ballot, lane); break;
    default: break;
    }
}

// ---------------------------------------------------------------------------------
// 8x8 fast path ("nibble staging").  The producer lane splits each 64-bit plane source
// into 16 nibbles, one per byte (6 SIMD-in-register ops + 4 byte permutes per plane),
// and parks them as one 16-byte row in shared memory.  A consumer lane then needs, per
// 4 output cells, ONE LDS.U8 (the nibble), one IMAD (spread) and one LOP3 (mask): the
// ALU pipe -- the binding unit of k_step, see profiles/ -- does ~1 op per 4 cells
// instead of ~9, and the address of the 4 nibbles of int8 chunk c is simply 4*c.
// ---------------------------------------------------------------------------------

OTH_DEV uint4 nibble_bytes(u64 x) {
    const u32 lo = (u32)x, hi = (u32)(x >> 32);
    const u32 lo_e = lo & 0x0F0F0F0Fu, lo_o = (lo >> 4) & 0x0F0F0F0Fu;   // even / odd nibbles of cells 0..31
    const u32 hi_e = hi & 0x0F0F0F0Fu, hi_o = (hi >> 4) & 0x0F0F0F0Fu;
    uint4 r;
    r.x = __byte_perm(lo_e, lo_o, 0x5140);   // n0 n1 n2 n3
    r.y = __byte_perm(lo_e, lo_o, 0x7362);   // n4..n7
    r.z = __byte_perm(hi_e, hi_o, 0x5140);
    r.w = __byte_perm(hi_e, hi_o, 0x7362);
    return r;
}

// Full warp (32 envs), fully unrolled: lane l owns, for every plane p and k = 0..3, the
// (env, part) pair c' = 32k + l -> env c'>>2, 16-cell part c'&3.  Its nibbles sit at
// row_p + 4c' in the stage and its 16 output bytes at env*P*64 + p*64 + part*16, so every
// shared/global address is ONE per-lane base plus compile-time immediates: no index
// arithmetic, no loop, no branch.  A quad of lanes writes 64 contiguous bytes.
template <int LAYOUT, int ES>
OTH_DEV void fast8_full(void *base, i64 env0, const WarpStage<8> &st, int lane) {
    constexpr int P = layout_planes(LAYOUT);
    constexpr bool HAS_NEG = (LAYOUT == OTH_OBS_BOARD || LAYOUT == OTH_OBS_BOARD_MASK || LAYOUT == OTH_OBS_ABSOLUTE);
    constexpr int ROW0 = (LAYOUT == OTH_LAYOUT_MASK_INTERNAL) ? NIB_MASK : NIB_POS0;
    const unsigned char *rows = reinterpret_cast<const unsigned char *>(&st.nib[0][0][0]);
    if constexpr (ES == 1) {
        const u32 src = (u32)__cvta_generic_to_shared(rows) + ROW0 * 512 + lane * 4;
        const u32 nsrc = (u32)__cvta_generic_to_shared(rows) + NIB_NEG0 * 512 + lane * 4;
        char *dst = reinterpret_cast<char *>(base) + env0 * (i64)(P * 64) + (lane >> 2) * (P * 64) + (lane & 3) * 16;
#pragma unroll
        for (int p = 0; p < P; p++) {
#pragma unroll
            for (int k = 0; k < 4; k++) {
                u32 w[4];
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    w[j] = spread4(lds_u8(src + p * 512 + k * 128 + j));
                    if (HAS_NEG && p == 0) w[j] = spread4(lds_u8(nsrc + k * 128 + j)) * 0xFFu + w[j];
                }
                *reinterpret_cast<uint4 *>(dst + k * (8 * P * 64) + p * 64) = make_uint4(w[0], w[1], w[2], w[3]);
            }
        }
    } else {
        // f32: one nibble -> 4 floats = one 16-byte chunk; c' = 32k + l, k = 0..15 -> env c'>>4, nibble c'&15
        const unsigned char *src = rows + ROW0 * 512 + lane;
        const unsigned char *nsrc = rows + NIB_NEG0 * 512 + lane;
        char *dst = reinterpret_cast<char *>(base) + env0 * (i64)(P * 256) + (lane >> 4) * (P * 256) + (lane & 15) * 16;
#pragma unroll
        for (int p = 0; p < P; p++) {
#pragma unroll 4
            for (int k = 0; k < 16; k++) {
                const u32 n = src[p * 512 + k * 32];
                u32 w[4];
#pragma unroll
                for (int j = 0; j < 4; j++) w[j] = (n & (1u << j)) * (0x3F800000u >> j);
                if (HAS_NEG && p == 0) {
                    const u32 nn = nsrc[k * 32];
#pragma unroll
                    for (int j = 0; j < 4; j++) w[j] |= (nn & (1u << j)) * (0xBF800000u >> j);
                }
                *reinterpret_cast<uint4 *>(dst + k * (2 * P * 256) + p * 256) = make_uint4(w[0], w[1], w[2], w[3]);
            }
        }
    }
}

template <int LAYOUT, int ES>
__device__ __noinline__ void fast8_write(void *base, i64 env0, int nenv, const WarpStage<8> &st, int lane) {
    constexpr int P = layout_planes(LAYOUT);
    constexpr bool HAS_NEG = (LAYOUT == OTH_OBS_BOARD || LAYOUT == OTH_OBS_BOARD_MASK || LAYOUT == OTH_OBS_ABSOLUTE);
    constexpr int ROW0 = (LAYOUT == OTH_LAYOUT_MASK_INTERNAL) ? NIB_MASK : NIB_POS0;
    const unsigned char *rows = reinterpret_cast<const unsigned char *>(&st.nib[0][0][0]);
    char *slab = reinterpret_cast<char *>(base) + env0 * (i64)(P * 64 * ES);
    if (nenv == 32) {
        fast8_full<LAYOUT, ES>(base, env0, st, lane);
        return;
    }
    if constexpr (ES == 1) {
        const int nchunks = nenv * P * 4;                 // 16 cells per 16-byte chunk
        for (int c = lane; c < nchunks; c += 32) {
            int e, p, part;
            if constexpr (P == 1) { e = c >> 2; p = 0; part = c & 3; }
            else { part = c & 3; const int q = c >> 2; e = q / P; p = q - e * P; }
            const unsigned char *src = rows + ((ROW0 + p) * 32 + e) * 16 + part * 4;
            u32 w[4];
#pragma unroll
            for (int j = 0; j < 4; j++) w[j] = spread4(src[j]);
            if constexpr (HAS_NEG) {
                if (p == 0) {
                    const unsigned char *nsrc = rows + (NIB_NEG0 * 32 + e) * 16 + part * 4;
#pragma unroll
                    for (int j = 0; j < 4; j++) w[j] = spread4(nsrc[j]) * 0xFFu + w[j];
                }
            }
            *reinterpret_cast<uint4 *>(slab + (i64)c * 16) = make_uint4(w[0], w[1], w[2], w[3]);
        }
    } else {
        const int nchunks = nenv * P * 16;                // 4 cells (one nibble) per 16-byte chunk
        for (int c = lane; c < nchunks; c += 32) {
            const int k = c & 15, q = c >> 4;
            int e, p;
            if constexpr (P == 1) { e = q; p = 0; }
            else { e = q / P; p = q - e * P; }
            const u32 n = rows[((ROW0 + p) * 32 + e) * 16 + k];
            u32 w[4];
#pragma unroll
            for (int j = 0; j < 4; j++) w[j] = (n & (1u << j)) * (0x3F800000u >> j);
            if constexpr (HAS_NEG) {
                if (p == 0) {
                    const u32 nn = rows[(NIB_NEG0 * 32 + e) * 16 + k];
#pragma unroll
                    for (int j = 0; j < 4; j++) w[j] |= (nn & (1u << j)) * (0xBF800000u >> j);
                }
            }
            *reinterpret_cast<uint4 *>(slab + (i64)c * 16) = make_uint4(w[0], w[1], w[2], w[3]);
        }
    }
}

template <int ES>
OTH_DEV void fast8_write_layout(int layout, void *base, i64 env0, int nenv, const WarpStage<8> &st, int lane) {
    switch (layout) {
    case OTH_OBS_BOARD: fast8_write<OTH_OBS_BOARD, ES>(base, env0, nenv, st, lane); break;
    case OTH_OBS_BOARD_MASK: fast8_write<OTH_OBS_BOARD_MASK, ES>(base, env0, nenv, st, lane); break;
    case OTH_OBS_PLANES4: fast8_write<OTH_OBS_PLANES4, ES>(base, env0, nenv, st, lane); break;
    case OTH_OBS_PLANES3: fast8_write<OTH_OBS_PLANES3, ES>(base, env0, nenv, st, lane); break;
    case OTH_OBS_ABSOLUTE: fast8_write<OTH_OBS_ABSOLUTE, ES>(base, env0, nenv, st, lane); break;
    default: break;
    }
}

// Any even board size: group g (4 cells = one staged nibble) of the warp's output slab belongs to
// (env, plane) q = g / (N/4) and is nibble r = g % (N/4) of that plane; a 16-byte int8 chunk is four
// consecutive groups (it may straddle a plane or env boundary when N/4 is not a multiple of 4), a
// 16-byte float32 chunk is one group.  One division by a constant per chunk, then increments.
template <int S, int LAYOUT, int ES>
__device__ __noinline__ void nib_write(void *base, i64 env0, int nenv, const WarpStage<S> &st, int lane) {
    constexpr int N = Geo<S>::N, W = Geo<S>::W, P = layout_planes(LAYOUT);
    constexpr int G4 = N / 4, RB = 16 * W;
    constexpr bool HAS_NEG = (LAYOUT == OTH_OBS_BOARD || LAYOUT == OTH_OBS_BOARD_MASK || LAYOUT == OTH_OBS_ABSOLUTE);
    constexpr int ROW0 = (LAYOUT == OTH_LAYOUT_MASK_INTERNAL) ? NIB_MASK : NIB_POS0;
    const u32 rows = (u32)__cvta_generic_to_shared(&st.nib[0][0][0]);
    char *slab = reinterpret_cast<char *>(base) + env0 * (i64)(P * N * ES);
    const int ngroups = nenv * P * G4;
    if constexpr (ES == 1) {
        const int nchunks = (ngroups + 3) >> 2;
        for (int c = lane; c < nchunks; c += 32) {
            int q = (4 * c) / G4;
            int r = 4 * c - q * G4;
            int e = q / P, p = q - e * P;
            u32 w[4];
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const int ec = min(e, 31);                                  // ragged tail of the last warp
                w[j] = spread4(lds_u8(rows + ((ROW0 + p) * 32 + ec) * RB + r));
                if (HAS_NEG && p == 0) w[j] = spread4(lds_u8(rows + (NIB_NEG0 * 32 + ec) * RB + r)) * 0xFFu + w[j];
                if (++r == G4) { r = 0; if (++p == P) { p = 0; e++; } }
            }
            char *dst = slab + (i64)c * 16;
            if (4 * c + 4 <= ngroups) {
                *reinterpret_cast<uint4 *>(dst) = make_uint4(w[0], w[1], w[2], w[3]);
            } else {
                for (int j = 0; 4 * c + j < ngroups; j++) *reinterpret_cast<u32 *>(dst + 4 * j) = w[j];
            }
        }
    } else {
        for (int c = lane; c < ngroups; c += 32) {
            const int q = c / G4, r = c - q * G4;
            const int e = q / P, p = q - e * P;
            const u32 n = lds_u8(rows + ((ROW0 + p) * 32 + e) * RB + r);
            u32 w[4];
#pragma unroll
            for (int j = 0; j < 4; j++) w[j] = (n & (1u << j)) * (0x3F800000u >> j);
            if (HAS_NEG && p == 0) {
                const u32 nn = lds_u8(rows + (NIB_NEG0 * 32 + e) * RB + r);
#pragma unroll
                for (int j = 0; j < 4; j++) w[j] |= (nn & (1u << j)) * (0xBF800000u >> j);
            }
            *reinterpret_cast<uint4 *>(slab + (i64)c * 16) = make_uint4(w[0], w[1], w[2], w[3]);
        }
    }
}

template <int S, int ES>
OTH_DEV void nib_write_layout(int layout, void *base, i64 env0, int nenv, const WarpStage<S> &st, int lane) {
    if constexpr (S == 8) {
        fast8_write_layout<ES>(layout, base, env0, nenv, st, lane);
    } else {
        switch (layout) {
        case OTH_OBS_BOARD: nib_write<S, OTH_OBS_BOARD, ES>(base, env0, nenv, st, lane); break;
        case OTH_OBS_BOARD_MASK: nib_write<S, OTH_OBS_BOARD_MASK, ES>(base, env0, nenv, st, lane); break;
        case OTH_OBS_PLANES4: nib_write<S, OTH_OBS_PLANES4, ES>(base, env0, nenv, st, lane); break;
        case OTH_OBS_PLANES3: nib_write<S, OTH_OBS_PLANES3, ES>(base, env0, nenv, st, lane); break;
        case OTH_OBS_ABSOLUTE: nib_write<S, OTH_OBS_ABSOLUTE, ES>(base, env0, nenv, st, lane); break;
        default: break;
        }
    }
}

// Even sizes other than 8x8, int8 outputs, full warp: "slab" writer.  Phase 1: lane e expands its own
// env, 4 cells at a time (nibble -> 4 bytes), into the warp's shared-memory slab at byte e*N -- already the
// output order, so no consumer ever divides by N/4 (for N/4 odd, S = 6 and 10, the word stride between
// lanes is odd: conflict-free).  Phase 2: the 32 lanes copy the 32*N-byte slab to global memory as 16-byte
// vectors, 512 contiguous bytes per store instruction.
template <int S, bool HAS_NEG>
OTH_DEV void slab_expand(unsigned char *slab, int lane, const BB<Geo<S>::W> &pos, const BB<Geo<S>::W> &neg) {
    constexpr int N = Geo<S>::N, G4 = N / 4;
    static_assert(N % 4 == 0, "slab writer: even board sizes only");
    u32 *dst = reinterpret_cast<u32 *>(slab + lane * N);
    // S*S a multiple of 16 (S = 12): four words per 16-byte store -- a quarter-warp's 8 stores then fall into 8
    // different 16-byte bank groups (lane stride 9 x 16 B), where the 32-bit stores were 4-way conflicted (lane
    // stride 36 words), and there are 4x fewer store instructions.  S = 16 (stride 16 x 16 B) keeps 32-bit stores.
    constexpr bool VEC = (N % 16 == 0) && ((N / 16) % 2 == 1);
    u32 quad[4];
#pragma unroll
    for (int r = 0; r < G4; r++) {
        const int k = r >> 4, sh = 4 * (r & 7);
        const bool hi = (r & 8) != 0;
        const u32 ph = hi ? (u32)(pos.w[k] >> 32) : (u32)pos.w[k];
        u32 w = spread4((ph >> sh) & 0xFu);
        if constexpr (HAS_NEG) {
            const u32 nh = hi ? (u32)(neg.w[k] >> 32) : (u32)neg.w[k];
            w = spread4((nh >> sh) & 0xFu) * 0xFFu + w;
        }
        if constexpr (VEC) {
            quad[r & 3] = w;
            if ((r & 3) == 3) reinterpret_cast<uint4 *>(dst)[r >> 2] = make_uint4(quad[0], quad[1], quad[2], quad[3]);
        } else {
            dst[r] = w;
        }
    }
}
// Odd sizes (S*S % 4 != 0): lane e's S*S bytes start at byte e*S*S of the slab, `a` = that offset mod 4 past an
// aligned word.  The lane shifts its bitboards up by a, so aligned word j holds its cells 4j-a .. 4j-a+3 and the
// nibble extraction keeps compile-time shifts; words 1 .. N/4-1 are always wholly its own (one 32-bit store each),
// the first word and the last two may be shared with the neighbouring envs and are stored byte by byte.
template <int W> OTH_DEV BB<W> bb_shl_small(const BB<W> &x, int a) {     // 0 <= a <= 3
    BB<W> r;
    BB_FOR(i) {
        const u64 lo = i > 0 ? x.w[i > 0 ? i - 1 : 0] : 0ull;
        r.w[i] = (x.w[i] << a) | ((lo >> 1) >> (63 - a));
    }
    return r;
}
template <int S, bool HAS_NEG>
OTH_DEV void slab_expand_odd(unsigned char *slab, int lane, const BB<Geo<S>::W> &pos, const BB<Geo<S>::W> &neg) {
    constexpr int N = Geo<S>::N, W = Geo<S>::W, JF = N / 4;
    static_assert(N % 4 != 0 && N + 3 <= 64 * W, "odd board sizes whose shifted board still fits its words");
    const int start = lane * N, a = start & 3;
    const BB<W> ps = bb_shl_small<W>(pos, a), ns = bb_shl_small<W>(neg, a);
    unsigned char *base = slab + (start - a);
    u32 *dst = reinterpret_cast<u32 *>(base);
#pragma unroll
    for (int j = 0; j <= JF + 1; j++) {
        if (4 * j >= 64 * W) continue;
        const int k = j >> 4, sh = 4 * (j & 7);
        const bool hi = (j & 8) != 0;
        const u32 ph = hi ? (u32)(ps.w[k] >> 32) : (u32)ps.w[k];
        u32 w = spread4((ph >> sh) & 0xFu);
        if constexpr (HAS_NEG) {
            const u32 nh = hi ? (u32)(ns.w[k] >> 32) : (u32)ns.w[k];
            w = spread4((nh >> sh) & 0xFu) * 0xFFu + w;
        }
        if (j >= 1 && j < JF) {
            dst[j] = w;
        } else {
#pragma unroll
            for (int b = 0; b < 4; b++) {
                const int c = 4 * j + b - a;                         // cell index of byte b of word j
                if (c >= 0 && c < N) base[4 * j + b] = (unsigned char)(w >> (8 * b));
            }
        }
    }
}
// which bitboard gives the +1 cells of each plane of a layout, and the -1 cells of plane 0
template <int S, int LAYOUT>
OTH_DEV void layout_sources(BB<Geo<S>::W> *pos, BB<Geo<S>::W> &neg, const BB<Geo<S>::W> &black, const BB<Geo<S>::W> &white,
                            const BB<Geo<S>::W> &lg, bool tw, bool quirk_empty) {
    constexpr int W = Geo<S>::W;
    const BB<W> zero = bb_zero<W>();
    neg = zero;
    if constexpr (LAYOUT == OTH_OBS_BOARD || LAYOUT == OTH_OBS_BOARD_MASK) {   // othello.py:363-376
        pos[0] = bb_select(tw, white, black); neg = bb_select(tw, black, white);
        if constexpr (LAYOUT == OTH_OBS_BOARD_MASK) pos[1] = lg;
    } else if constexpr (LAYOUT == OTH_OBS_ABSOLUTE) {
        pos[0] = white; neg = black;
    } else {                                                                   // util.py:48-74, run_2agent.py:29-46
        pos[0] = black; pos[1] = white;
        pos[2] = bb_select(tw, geo_full<S>(), zero);
        if constexpr (LAYOUT == OTH_OBS_PLANES4) pos[3] = bb_select(quirk_empty, zero, lg);
    }
}

// float32 observations of odd sizes, full warp: no staging.  A cell is 4 bytes, so every (env, plane) row of S*S
// floats is 4-byte aligned; for env e = 0..31 and each plane, the owner lane's bitboard is broadcast by shuffle
// and lane l writes cells l, l+32, ...: one store instruction = 32 consecutive floats = 128 contiguous bytes.
// Cell l + 32*i is bit l of 32-bit half i of the board, so the half is picked at compile time.
template <int S, int LAYOUT>
OTH_DEV void rows_write_f32(void *base, i64 env0, int lane, const BB<Geo<S>::W> &black, const BB<Geo<S>::W> &white,
                            const BB<Geo<S>::W> &lg, bool tw, bool quirk_empty) {
    constexpr int N = Geo<S>::N, W = Geo<S>::W, P = layout_planes(LAYOUT);
    constexpr bool HAS_NEG = (LAYOUT == OTH_OBS_BOARD || LAYOUT == OTH_OBS_BOARD_MASK || LAYOUT == OTH_OBS_ABSOLUTE);
    BB<W> pos[P], neg;
    layout_sources<S, LAYOUT>(pos, neg, black, white, lg, tw, quirk_empty);
    u32 *out = reinterpret_cast<u32 *>(base) + env0 * (i64)(P * N);
#pragma unroll 2
    for (int e = 0; e < 32; e++) {
#pragma unroll
        for (int p = 0; p < P; p++) {
            u32 *row = out + (e * P + p) * N;
#pragma unroll
            for (int i = 0; i < (N + 31) / 32; i++) {
                const u64 pw = __shfl_sync(0xFFFFFFFFu, pos[p].w[i >> 1], e);
                const u32 ph = (i & 1) ? (u32)(pw >> 32) : (u32)pw;
                u32 v = ((ph >> lane) & 1u) * 0x3F800000u;
                if (HAS_NEG && p == 0) {
                    const u64 nw = __shfl_sync(0xFFFFFFFFu, neg.w[i >> 1], e);
                    const u32 nh = (i & 1) ? (u32)(nw >> 32) : (u32)nw;
                    v |= ((nh >> lane) & 1u) * 0xBF800000u;
                }
                const int c = lane + 32 * i;
                if (32 * i + 32 <= N || c < N) row[c] = v;
            }
        }
    }
}
template <int S>
OTH_DEV void rows_write_f32_layout(int layout, void *base, i64 env0, int lane, const BB<Geo<S>::W> &black,
                                   const BB<Geo<S>::W> &white, const BB<Geo<S>::W> &lg, bool tw, bool quirk_empty) {
    switch (layout) {
    case OTH_OBS_BOARD: rows_write_f32<S, OTH_OBS_BOARD>(base, env0, lane, black, white, lg, tw, quirk_empty); break;
    case OTH_OBS_BOARD_MASK: rows_write_f32<S, OTH_OBS_BOARD_MASK>(base, env0, lane, black, white, lg, tw, quirk_empty); break;
    case OTH_OBS_PLANES4: rows_write_f32<S, OTH_OBS_PLANES4>(base, env0, lane, black, white, lg, tw, quirk_empty); break;
    case OTH_OBS_PLANES3: rows_write_f32<S, OTH_OBS_PLANES3>(base, env0, lane, black, white, lg, tw, quirk_empty); break;
    case OTH_OBS_ABSOLUTE: rows_write_f32<S, OTH_OBS_ABSOLUTE>(base, env0, lane, black, white, lg, tw, quirk_empty); break;
    default: break;
    }
}

template <int S>
OTH_DEV void slab_copy(const unsigned char *slab, void *base, i64 env0, int lane) {
    constexpr int N = Geo<S>::N, CH = 2 * N;                       // 32*N bytes = 2N chunks of 16
    const uint4 *src = reinterpret_cast<const uint4 *>(slab);
    uint4 *dst = reinterpret_cast<uint4 *>(reinterpret_cast<char *>(base) + env0 * (i64)N);
#pragma unroll
    for (int k = 0; k < (CH + 31) / 32; k++) {
        const int c = lane + 32 * k;
        if (CH % 32 == 0 || c < CH) dst[c] = src[c];
    }
}

// int8 observations of any layout, any size but 8, full warp: the flat output [env][plane][cell] is a sequence of
// 32*P rows of S*S bytes; pass k handles rows 32k .. 32k+31: lane l fetches the bitboards of row 32k+l (env r/P,
// plane r%P) from their owner lane by shuffle, expands them into the slab at byte l*S*S (the alignment phase of the
// row's global offset equals that of l*S*S because 32*k*S*S is a multiple of 4), and the warp copies the slab out.
template <int S, int LAYOUT>
OTH_DEV void rows_i8_write(void *base, i64 env0, int lane, unsigned char *slab, const BB<Geo<S>::W> &black,
                           const BB<Geo<S>::W> &white, const BB<Geo<S>::W> &lg, bool tw, bool quirk_empty) {
    constexpr int N = Geo<S>::N, W = Geo<S>::W, P = layout_planes(LAYOUT);
    constexpr bool HAS_NEG = (LAYOUT == OTH_OBS_BOARD || LAYOUT == OTH_OBS_BOARD_MASK || LAYOUT == OTH_OBS_ABSOLUTE);
    BB<W> pos[P], neg;
    layout_sources<S, LAYOUT>(pos, neg, black, white, lg, tw, quirk_empty);
    char *out = reinterpret_cast<char *>(base) + env0 * (i64)(P * N);
#pragma unroll
    for (int k = 0; k < P; k++) {
        const int r = 32 * k + lane, e = r / P, p = r - e * P;
        BB<W> sp = bb_zero<W>(), sn = bb_zero<W>();
#pragma unroll
        for (int q = 0; q < P; q++) {
            BB<W> t;
            BB_FOR(i) t.w[i] = __shfl_sync(0xFFFFFFFFu, pos[q].w[i], e);
            if (P == 1 || p == q) sp = t;
        }
        if constexpr (HAS_NEG) {
            BB<W> t;
            BB_FOR(i) t.w[i] = __shfl_sync(0xFFFFFFFFu, neg.w[i], e);
            if (p == 0) sn = t;
        }
        if constexpr (N % 4 == 0) slab_expand<S, HAS_NEG>(slab, lane, sp, sn);
        else slab_expand_odd<S, HAS_NEG>(slab, lane, sp, sn);
        __syncwarp();
        slab_copy<S>(slab, out + (i64)k * 32 * N, 0, lane);
        __syncwarp();
    }
}
template <int S>
OTH_DEV void rows_i8_write_layout(int layout, void *base, i64 env0, int lane, unsigned char *slab, const BB<Geo<S>::W> &black,
                                  const BB<Geo<S>::W> &white, const BB<Geo<S>::W> &lg, bool tw, bool quirk_empty) {
    switch (layout) {
    case OTH_OBS_BOARD: rows_i8_write<S, OTH_OBS_BOARD>(base, env0, lane, slab, black, white, lg, tw, quirk_empty); break;
    case OTH_OBS_BOARD_MASK: rows_i8_write<S, OTH_OBS_BOARD_MASK>(base, env0, lane, slab, black, white, lg, tw, quirk_empty); break;
    case OTH_OBS_PLANES4: rows_i8_write<S, OTH_OBS_PLANES4>(base, env0, lane, slab, black, white, lg, tw, quirk_empty); break;
    case OTH_OBS_PLANES3: rows_i8_write<S, OTH_OBS_PLANES3>(base, env0, lane, slab, black, white, lg, tw, quirk_empty); break;
    case OTH_OBS_ABSOLUTE: rows_i8_write<S, OTH_OBS_ABSOLUTE>(base, env0, lane, slab, black, white, lg, tw, quirk_empty); break;
    default: break;
    }
}

template <int W>
OTH_DEV void stage_nibbles(uint4 *dst, const BB<W> &x) {
    BB_FOR(k) dst[k] = nibble_bytes(x.w[k]);
}

// even S: stage this lane's env as nibbles and write every requested dense output for the warp
template <int S>
OTH_DEV void emit_dense_nib(const EnvRegs<S> &e, bool active, bool zero_obs, bool zero_mask, int flags,
                            const oth_outputs &out, i64 warp_env0, i64 B, WarpStage<S> &st) {
    constexpr int W = Geo<S>::W;
    const int lane = threadIdx.x & 31;
    const bool live = active && !zero_obs;
    const bool tw = e.meta & OTH_META_TURN_WHITE;
    const BB<W> zero = bb_zero<W>();
    const BB<W> black = bb_select(live, e.black, zero), white = bb_select(live, e.white, zero);
    const BB<W> lg = bb_select(live, e.legal, zero);
    const BB<W> lg_mask = bb_select(active && !zero_mask, e.legal, zero);
    const int layout = (out.obs != nullptr) ? out.obs_layout : OTH_OBS_NONE;
    if constexpr (S != 8) {
        const bool obs_slab = layout == OTH_OBS_BOARD && out.obs_dtype == OTH_DTYPE_I8;
        const bool obs_rows8 = layout != OTH_OBS_NONE && layout != OTH_OBS_BOARD && out.obs_dtype == OTH_DTYPE_I8;
        if (warp_env0 + 32 <= B && (obs_slab || obs_rows8 || layout == OTH_OBS_NONE)) {   // warp-uniform
            static_assert(sizeof(st.nib) >= 32 * Geo<S>::N, "slab must fit the warp stage");
            unsigned char *slab = reinterpret_cast<unsigned char *>(&st.nib[0][0][0]);
            if (obs_rows8) {                                                 // multi-plane int8 layouts, row passes
                const bool quirk = (flags & OTH_F_LEGACY_SINGLE_MOVE_PLANE) && popc(lg) <= 1;   // util.py:55
                rows_i8_write_layout<S>(layout, out.obs, warp_env0, lane, slab, black, white, lg, live && tw, quirk);
            }
            if (obs_slab) {                                                  // othello.py:363-369
                slab_expand<S, true>(slab, lane, bb_select(tw, white, black), bb_select(tw, black, white));
                __syncwarp();
                slab_copy<S>(slab, out.obs, warp_env0, lane);
                __syncwarp();
            }
            if (out.legal_mask != nullptr) {
                slab_expand<S, false>(slab, lane, lg_mask, lg_mask);
                __syncwarp();
                slab_copy<S>(slab, out.legal_mask, warp_env0, lane);
                __syncwarp();
            }
            return;
        }
    }
    if (layout == OTH_OBS_BOARD || layout == OTH_OBS_BOARD_MASK) {           // othello.py:363-376
        stage_nibbles<W>(st.nib[NIB_POS0][lane], bb_select(tw, white, black));
        stage_nibbles<W>(st.nib[NIB_NEG0][lane], bb_select(tw, black, white));
        if (layout == OTH_OBS_BOARD_MASK) stage_nibbles<W>(st.nib[NIB_POS1][lane], lg);
    } else if (layout == OTH_OBS_ABSOLUTE) {
        stage_nibbles<W>(st.nib[NIB_POS0][lane], white);
        stage_nibbles<W>(st.nib[NIB_NEG0][lane], black);
    } else if (layout == OTH_OBS_PLANES4 || layout == OTH_OBS_PLANES3) {     // util.py:48-74
        stage_nibbles<W>(st.nib[NIB_POS0][lane], black);
        stage_nibbles<W>(st.nib[NIB_POS1][lane], white);
        const u32 t = (live && tw) ? 0x0F0F0F0Fu : 0u;
        BB_FOR(k) st.nib[NIB_POS2][lane][k] = make_uint4(t, t, t, t);
        if (layout == OTH_OBS_PLANES4) {
            const bool quirk = (flags & OTH_F_LEGACY_SINGLE_MOVE_PLANE) && popc(lg) <= 1;   // util.py:55
            stage_nibbles<W>(st.nib[NIB_POS3][lane], bb_select(quirk, zero, lg));
        }
    }
    if (out.legal_mask != nullptr) stage_nibbles<W>(st.nib[NIB_MASK][lane], lg_mask);
    __syncwarp();
    if (warp_env0 < B) {
        const int nenv = (int)min((i64)32, B - warp_env0);
        if (S == 8 && layout == OTH_OBS_BOARD && out.obs_dtype == OTH_DTYPE_I8 && nenv == 32) {
            if constexpr (S == 8) fast8_full<OTH_OBS_BOARD, 1>(out.obs, warp_env0, st, lane);   // the headline path, inlined
        } else if (layout != OTH_OBS_NONE) {
            if (out.obs_dtype == OTH_DTYPE_F32) nib_write_layout<S, 4>(layout, out.obs, warp_env0, nenv, st, lane);
            else nib_write_layout<S, 1>(layout, out.obs, warp_env0, nenv, st, lane);
        }
        if (out.legal_mask != nullptr) {
            if constexpr (S == 8) {
                if (nenv == 32) fast8_full<OTH_LAYOUT_MASK_INTERNAL, 1>(out.legal_mask, warp_env0, st, lane);
                else fast8_write<OTH_LAYOUT_MASK_INTERNAL, 1>(out.legal_mask, warp_env0, nenv, st, lane);
            } else {
                nib_write<S, OTH_LAYOUT_MASK_INTERNAL, 1>(out.legal_mask, warp_env0, nenv, st, lane);
            }
        }
    }
    __syncwarp();
}

// stage this lane's env and write every requested dense output for the warp
template <int S>
OTH_DEV void emit_dense(const EnvRegs<S> &e, bool active, bool zero_obs, bool zero_mask, int flags,
                        const oth_outputs &out, i64 warp_env0, i64 B, WarpStage<S> &st) {
    if constexpr ((S * S) % 4 == 0) {
        emit_dense_nib<S>(e, active, zero_obs, zero_mask, flags, out, warp_env0, B, st);
    } else {
        constexpr int W = Geo<S>::W;
        const int lane = threadIdx.x & 31;
        {
            const int layout = (out.obs != nullptr) ? out.obs_layout : OTH_OBS_NONE;
            const bool obs_slab = layout == OTH_OBS_BOARD && out.obs_dtype == OTH_DTYPE_I8;
            const bool obs_rows = layout != OTH_OBS_NONE && out.obs_dtype == OTH_DTYPE_F32;
            const bool obs_rows8 = layout != OTH_OBS_NONE && layout != OTH_OBS_BOARD && out.obs_dtype == OTH_DTYPE_I8;
            if (warp_env0 + 32 <= B) {                                           // warp-uniform: full warp
                const bool live = active && !zero_obs, tw = live && (e.meta & OTH_META_TURN_WHITE);
                const BB<W> zero = bb_zero<W>();
                const BB<W> black = bb_select(live, e.black, zero), white = bb_select(live, e.white, zero);
                const BB<W> lg_mask = bb_select(active && !zero_mask, e.legal, zero);
                unsigned char *slab = reinterpret_cast<unsigned char *>(&st.slab[0]);
                if (obs_rows || obs_rows8) {
                    const BB<W> lg = bb_select(live, e.legal, zero);
                    const bool quirk = (flags & OTH_F_LEGACY_SINGLE_MOVE_PLANE) && popc(lg) <= 1;   // util.py:55
                    if (obs_rows) rows_write_f32_layout<S>(layout, out.obs, warp_env0, lane, black, white, lg, tw, quirk);
                    else rows_i8_write_layout<S>(layout, out.obs, warp_env0, lane, slab, black, white, lg, tw, quirk);
                }
                if (obs_slab) {                                                  // othello.py:363-369
                    slab_expand_odd<S, true>(slab, lane, bb_select(tw, white, black), bb_select(tw, black, white));
                    __syncwarp();
                    slab_copy<S>(slab, out.obs, warp_env0, lane);
                    __syncwarp();
                }
                if (out.legal_mask != nullptr) {
                    slab_expand_odd<S, false>(slab, lane, lg_mask, lg_mask);
                    __syncwarp();
                    slab_copy<S>(slab, out.legal_mask, warp_env0, lane);
                    __syncwarp();
                }
                return;
            }
        }
        BB_FOR(i) {
            st.row[ROW_BLACK][i][lane] = active ? e.black.w[i] : 0ull;
            st.row[ROW_WHITE][i][lane] = active ? e.white.w[i] : 0ull;
            st.row[ROW_LEGAL][i][lane] = active ? e.legal.w[i] : 0ull;
            st.row[ROW_ZERO][i][lane] = 0ull;
            st.row[ROW_ONES][i][lane] = ~0ull;
        }
        const u32 turn_ballot = __ballot_sync(0xFFFFFFFFu, active && (e.meta & OTH_META_TURN_WHITE));
        const u32 quirk_ballot = __ballot_sync(0xFFFFFFFFu, (flags & OTH_F_LEGACY_SINGLE_MOVE_PLANE) && popc(e.legal) <= 1);
        const u32 zero_ballot = __ballot_sync(0xFFFFFFFFu, zero_obs || !active);
        const u32 zero_mask_ballot = __ballot_sync(0xFFFFFFFFu, zero_mask || !active);
        __syncwarp();
        if (warp_env0 < B) {
            const int nenv = (int)min((i64)32, B - warp_env0);
            if (out.obs != nullptr && out.obs_layout != OTH_OBS_NONE) {
                if (out.obs_dtype == OTH_DTYPE_F32)
                    warp_write_layout<S, 4>(out.obs_layout, out.obs, warp_env0, nenv, st, turn_ballot, quirk_ballot, zero_ballot, lane);
                else
                    warp_write_layout<S, 1>(out.obs_layout, out.obs, warp_env0, nenv, st, turn_ballot, quirk_ballot, zero_ballot, lane);
            }
            if (out.legal_mask != nullptr)
                warp_write_planes<S, OTH_LAYOUT_MASK_INTERNAL, 1>(out.legal_mask, warp_env0, nenv, st, turn_ballot, quirk_ballot, zero_mask_ballot, lane);
        }
        __syncwarp();
    }
}

template <int S>
OTH_DEV void load_env(EnvRegs<S> &e, const oth_state &s, i64 i, i64 B) {
    constexpr int W = Geo<S>::W;
    // ctr and meta first: they are needed first (the Philox draw depends on ctr alone, the terminated check on meta),
    // and a warp's first wait at the top of k_step was the LAST of five loads (profiles/r03a: 35 % of the
    // long-scoreboard samples sat on the first use of ctr)
    e.ctr = s.ctr[i];
    e.meta = s.meta[i];
    BB_FOR(k) {
        e.legal.w[k] = s.legal[k * B + i];
        e.black.w[k] = s.black[k * B + i];
        e.white.w[k] = s.white[k * B + i];
    }
}
template <int S>
OTH_DEV void store_env(const EnvRegs<S> &e, const oth_state &s, i64 i, i64 B) {
    constexpr int W = Geo<S>::W;
    BB_FOR(k) {
        s.black[k * B + i] = e.black.w[k];
        s.white[k * B + i] = e.white.w[k];
        s.legal[k * B + i] = e.legal.w[k];
    }
    s.meta[i] = e.meta;
    s.ctr[i] = e.ctr;
}
template <int S>
OTH_DEV EnvRegs<S> empty_env() {
    EnvRegs<S> e;
    e.black = bb_zero<Geo<S>::W>(); e.white = e.black; e.legal = e.black;
    e.meta = 0; e.ctr = 0;
    return e;
}

// ------------------------------------------------------------------- per-env scalars
// out.packed: one 16-byte record per env (include/othello_b200.h) = ONE address and one STG.128 instead of five
// stores with five 64-bit address computations; the separate arrays are still served when they are not null.
OTH_DEV u32 pack_status(int done, int turn, int winner) {
    return (u32)(done & 0xFF) | ((u32)(turn & 0xFF) << 8) | ((u32)(winner & 0xFF) << 16);
}
template <bool PACKED_ONLY>
OTH_DEV void put_step_scalars(const oth_outputs &o, i64 i, int reward, int done, int turn, int winner, i64 action) {
    if (PACKED_ONLY || o.packed != nullptr)
        reinterpret_cast<uint4 *>(o.packed)[i] = make_uint4(__float_as_uint((float)reward), pack_status(done, turn, winner),
                                                            (u32)(u64)action, (u32)((u64)action >> 32));
    if constexpr (!PACKED_ONLY) {
        if (o.reward) o.reward[i] = (float)reward;
        if (o.done) o.done[i] = (uint8_t)done;
        if (o.player_turn) o.player_turn[i] = (int8_t)turn;
        if (o.winner) o.winner[i] = (int8_t)winner;
        if (o.actions_out) o.actions_out[i] = action;
    }
}
// calls that produce no reward / action (oth_observe, the protocol kernels): status bytes only
OTH_DEV void put_status(const oth_outputs &o, i64 i, int done, int turn, int winner, bool with_winner) {
    if (o.packed != nullptr) {
        u32 *rec = reinterpret_cast<u32 *>(o.packed) + 4 * i + 1;
        if (with_winner) *rec = pack_status(done, turn, winner);
        else *reinterpret_cast<unsigned short *>(rec) = (unsigned short)pack_status(done, turn, 0);
    }
    if (o.done) o.done[i] = (uint8_t)done;
    if (o.player_turn) o.player_turn[i] = (int8_t)turn;
    if (with_winner && o.winner) o.winner[i] = (int8_t)winner;
}

// ---------------------------------------------------------------------------- reset
template <int S>
__global__ void __launch_bounds__(OTH_BLOCK) k_reset(oth_state s, i64 B, u64 seed, i64 base, int init_rand,
                                                     int reset_rng, const uint8_t *only_where) {
    const i64 i = (i64)blockIdx.x * OTH_BLOCK + threadIdx.x;
    if (i >= B) return;
    if (only_where != nullptr && only_where[i] == 0) return;
    EnvRegs<S> e;
    e.ctr = reset_rng ? 0u : s.ctr[i];
    env_reset<S>(e, seed, base + i, init_rand);
    store_env<S>(e, s, i, B);
}

// -------------------------------------------------------------------------- observe
template <int S>
__global__ void __launch_bounds__(OTH_BLOCK) k_observe(oth_state s, i64 B, int flags, oth_outputs out) {
    __shared__ WarpStage<S> stage[OTH_WARPS];
    const i64 i = (i64)blockIdx.x * OTH_BLOCK + threadIdx.x;
    const bool active = i < B;
    EnvRegs<S> e = empty_env<S>();
    bool zero_obs = false;
    if (active) {
        load_env<S>(e, s, i, B);
        put_status(out, i, (e.meta & OTH_META_TERMINATED) ? 1 : 0, (e.meta & OTH_META_TURN_WHITE) ? 1 : -1,
                   meta_winner(e.meta), true);
        zero_obs = (flags & OTH_F_ZERO_IF_TERMINATED) && (e.meta & OTH_META_TERMINATED);
    }
    emit_dense<S>(e, active, zero_obs, false, flags, out, i - (threadIdx.x & 31), B, stage[threadIdx.x >> 5]);
}

// ----------------------------------------------------------------------------- step
struct StepParams {
    oth_state st;
    i64 B;
    int flags;
    int action_source;
    const i64 *actions;
    u64 seed;
    i64 base;
    int init_rand;
    oth_outputs out;
    PhiloxKeys keys;     // philox_round_keys(seed), filled by the launcher
};

// 8x8 headline writer: the producer parks each 64-bit plane as FOUR words of nibbles-in-bytes in even / odd order
// (lo even, lo odd, hi even, hi odd: 4 LOP3 + 2 SHF, no byte permutes -- 12 ALU-pipe instructions fewer per env than
// nibble_bytes for the three planes); the consumer's byte address of nibble n = 4*part + j is then
// (part>>1)*8 + (part&1)*2 + (j&1)*4 + (j>>1): still one per-lane base plus compile-time immediates.
OTH_DEV uint4 nibble_bytes_eo(u64 x) {
    const u32 lo = (u32)x, hi = (u32)(x >> 32);
    return make_uint4(lo & 0x0F0F0F0Fu, (lo >> 4) & 0x0F0F0F0Fu, hi & 0x0F0F0F0Fu, (hi >> 4) & 0x0F0F0F0Fu);
}
template <bool HAS_NEG>
OTH_DEV void fast8_full_eo(void *base, i64 env0, const uint4 *pos_rows, const uint4 *neg_rows, int lane) {
    const u32 lane_off = (u32)((lane >> 2) * 16 + ((lane >> 1) & 1) * 8 + (lane & 1) * 2);
    const u32 src = (u32)__cvta_generic_to_shared(pos_rows) + lane_off;
    const u32 nsrc = (u32)__cvta_generic_to_shared(neg_rows) + lane_off;
    char *dst = reinterpret_cast<char *>(base) + env0 * (i64)64 + (lane >> 2) * 64 + (lane & 3) * 16;
#pragma unroll
    for (int k = 0; k < 4; k++) {
        u32 w[4];
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const int imm = k * 128 + (j & 1) * 4 + (j >> 1);
            w[j] = spread4(lds_u8(src + imm));
            if (HAS_NEG) w[j] = spread4(lds_u8(nsrc + imm)) * 0xFFu + w[j];
        }
        *reinterpret_cast<uint4 *>(dst + k * (8 * 64)) = make_uint4(w[0], w[1], w[2], w[3]);
    }
}

// Finished warp slab -> global memory as ONE bulk async copy (cp.async.bulk shared::cta -> global) issued by lane 0,
// instead of ceil(2N/32) LDS.128 + STG.128 pairs per lane.  Every lane fences its generic-proxy shared stores towards
// the async proxy, the warp synchronises, lane 0 issues the copy and commits the group.  The slab may be rewritten
// (and the warp may exit) only after slab_bulk_wait().
template <int S>
OTH_DEV void slab_bulk_store(const unsigned char *slab, void *base, i64 env0, int lane) {
    constexpr int BYTES = 32 * Geo<S>::N;
    static_assert(BYTES % 16 == 0, "bulk copies move multiples of 16 bytes");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncwarp();
    if (lane == 0) {
        char *dst = reinterpret_cast<char *>(base) + env0 * (i64)Geo<S>::N;
        asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
                     :: "l"(dst), "r"((u32)__cvta_generic_to_shared(slab)), "n"(BYTES) : "memory");
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    }
}
OTH_DEV void slab_bulk_wait(int lane) {
    if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
    __syncwarp();
}

// The headline output set with everything decided at compile time (oth_step dispatches here when it applies):
// int8 `board` observation + dense legal mask, full warps only; S == 8 uses the even/odd nibble stage, the other
// even sizes the slab writer.
template <int S>
OTH_DEV void emit_board_mask_fast(const EnvRegs<S> &e, bool zero, const oth_outputs &out, i64 warp_env0, WarpStage<S> &st) {
    constexpr int W = Geo<S>::W;
    static_assert((S * S) % 4 == 0, "fast path: even board sizes");
    const int lane = threadIdx.x & 31;
    const bool tw = e.meta & OTH_META_TURN_WHITE;
    const BB<W> z = bb_zero<W>();
    const BB<W> pos = bb_select(zero, z, bb_select(tw, e.white, e.black));     // othello.py:363-369
    const BB<W> neg = bb_select(zero, z, bb_select(tw, e.black, e.white));
    const BB<W> lg = bb_select(zero, z, e.legal);
    if constexpr (S == 8) {
        st.nib[NIB_POS0][lane][0] = nibble_bytes_eo(pos.w[0]);
        st.nib[NIB_NEG0][lane][0] = nibble_bytes_eo(neg.w[0]);
        st.nib[NIB_MASK][lane][0] = nibble_bytes_eo(lg.w[0]);
        __syncwarp();
        fast8_full_eo<true>(out.obs, warp_env0, &st.nib[NIB_POS0][0][0], &st.nib[NIB_NEG0][0][0], lane);
        fast8_full_eo<false>(out.legal_mask, warp_env0, &st.nib[NIB_MASK][0][0], &st.nib[NIB_MASK][0][0], lane);
        __syncwarp();
    } else {
        static_assert(sizeof(st.nib) >= 32 * Geo<S>::N, "slab must fit the warp stage");
        unsigned char *base = reinterpret_cast<unsigned char *>(&st.nib[0][0][0]);
        // (a nibble-staged expansion with LDS.U8 extraction, as in the 8x8 writer, measured 1-4 % slower here:
        // profiles/r03e_experiments.md)
        // bulk async stores where BOTH slabs (obs + mask) fit the stage and a slab is at least 3 KB (10x10, 12x12):
        // measured -3.3 % on 12x12; +2.7 % on 6x6 (1.1 KB copies do not pay for fence + wait) and neutral with a
        // single slab that must be drained between the two copies (profiles/r03q_bulk_store.md)
        constexpr int SLAB = 32 * Geo<S>::N;
        constexpr bool BULK = 2 * SLAB <= (int)sizeof(WarpStage<S>) && SLAB >= 3072;
        if constexpr (BULK) {
            slab_expand<S, true>(base, lane, pos, neg);
            slab_bulk_store<S>(base, out.obs, warp_env0, lane);
            slab_expand<S, false>(base + SLAB, lane, lg, lg);
            slab_bulk_store<S>(base + SLAB, out.legal_mask, warp_env0, lane);
            slab_bulk_wait(lane);
        } else {
            slab_expand<S, true>(base, lane, pos, neg);
            __syncwarp();
            slab_copy<S>(base, out.obs, warp_env0, lane);
            __syncwarp();
            slab_expand<S, false>(base, lane, lg, lg);
            __syncwarp();
            slab_copy<S>(base, out.legal_mask, warp_env0, lane);
            __syncwarp();
        }
    }
}

#ifndef OTH_STEP8_MIN_CTAS
#define OTH_STEP8_MIN_CTAS 0    // min resident CTAs per SM the 8x8 fast kernel is compiled for (0: compiler default; 8: register cap 64)
#endif
// One env's step (policy / opening override -> rules -> auto-reset -> state store -> scalars): shared by the
// one-tile general kernel and the multi-tile fast kernel.  `ext_action` = the caller's action for OTH_ACT_EXTERNAL.
template <int S, int SRC, bool PACKED_ONLY>
OTH_DEV void step_one_env(const StepParams &p, EnvRegs<S> &e, i64 i, i64 ext_action, bool &zero_obs) {
    StepResult r;
    r.reward = 0; r.done = 1; r.winner = 0; r.action = -1;
    // in-kernel random policy: the draw depends on ctr alone, so it is computed BEFORE anything waits for meta or the
    // bitboards (wasted only for finished envs, which do not advance ctr)
    u32 rnd = 0;
    if (SRC == OTH_ACT_RANDOM) rnd = philox_draw(p.keys, p.base + i, OTH_STREAM_ENV, e.ctr);
    if (e.meta & OTH_META_TERMINATED) {
        zero_obs = true;                                   // finished env: no-op ('over', ppo_run_self_play.py:274)
        r.winner = meta_winner(e.meta);                    // its result stays readable (env.winner, othello.py:240)
    } else if (SRC == OTH_ACT_EXTERNAL && ext_action == OTH_ACTION_HOLD) {
        r.done = 0;                                        // sits this ply out: state as it was, outputs re-emitted
    } else {
        i64 action = -1;
        const bool opening = meta_rand_cnt(e.meta) < meta_rand_max(e.meta);   // othello.py:70-73
        if (opening || SRC == OTH_ACT_RANDOM) {
            if (SRC != OTH_ACT_RANDOM) rnd = philox_draw(p.keys, p.base + i, OTH_STREAM_ENV, e.ctr);
            action = random_pick<S>(e.legal, rnd);
            e.ctr++;
            if (opening) e.meta += 1u << OTH_META_RANDCNT_SHIFT;
        } else if (SRC == OTH_ACT_GREEDY) {
            const bool wm = e.meta & OTH_META_TURN_WHITE;
            action = greedy_pick<S>(bb_select(wm, e.white, e.black), bb_select(wm, e.black, e.white), e.legal);
        } else {
            action = ext_action;
        }
        r = env_step<S>(e, action, p.flags);
        if (r.done && (p.flags & OTH_F_AUTO_RESET)) env_reset<S>(e, p.keys, p.base + i, p.init_rand);
        store_env<S>(e, p.st, i, p.B);
    }
    put_step_scalars<PACKED_ONLY>(p.out, i, r.reward, r.done, (e.meta & OTH_META_TURN_WHITE) ? 1 : -1, r.winner, r.action);
}

// SRC: OTH_ACT_* at compile time.  General kernel: one 128-env tile per CTA, any B, any output set.
template <int S, int SRC, int BLOCK = OTH_BLOCK>
__global__ void __launch_bounds__(BLOCK) k_step(StepParams p) {
    __shared__ WarpStage<S> stage[BLOCK / 32];
    const i64 i = (i64)blockIdx.x * BLOCK + threadIdx.x;
    const bool active = i < p.B;
    EnvRegs<S> e = empty_env<S>();
    bool zero_obs = false;
    if (active) {
        load_env<S>(e, p.st, i, p.B);
        step_one_env<S, SRC, false>(p, e, i, SRC == OTH_ACT_EXTERNAL ? p.actions[i] : (i64)-1, zero_obs);
    }
    emit_dense<S>(e, active, zero_obs, zero_obs, p.flags, p.out, i - (threadIdx.x & 31), p.B, stage[threadIdx.x >> 5]);
}

// Fast kernels: B is a multiple of the CTA size, the scalars go to out.packed only and the dense outputs are the int8
// `board` observation + legal mask (even S) -- no per-output null checks, no ragged-warp handling, no layout dispatch.
// One 128-env tile per CTA: the hardware's CTA scheduler staggers the CTAs' load / rules / store phases.  Multi-tile
// CTAs that prefetch the next tile's state (in registers, or staged in shared memory with cp.async) were built and
// measured: 39.7-43.8 us against 38.9 us, slower the more tiles a CTA owns (profiles/r03e_experiments.md).
template <int S, int SRC, int BLOCK = OTH_BLOCK>
__global__ void __launch_bounds__(BLOCK, S == 8 ? OTH_STEP8_MIN_CTAS : OTH_STEPN_MIN_CTAS) k_step_fast(StepParams p) {
    __shared__ WarpStage<S> stage[BLOCK / 32];
    const i64 i = (i64)blockIdx.x * BLOCK + threadIdx.x;
    EnvRegs<S> e;
    load_env<S>(e, p.st, i, p.B);
    bool zero_obs = false;
    step_one_env<S, SRC, true>(p, e, i, SRC == OTH_ACT_EXTERNAL ? p.actions[i] : (i64)-1, zero_obs);
    emit_board_mask_fast<S>(e, zero_obs, p.out, i - (threadIdx.x & 31), stage[threadIdx.x >> 5]);
}

// ------------------------------------------------------------------------- self-play
// The per-env machine behind PPOEnvs.step + subproc_worker (envs.py:188-227,
// ppo_run_self_play.py:244-368): which protagonist tuple each env owes the learner and when
// it is complete.  mode 0 = start of a round, mode 1 = one policy.act worth of plies.
#define OTH_STREAM_PROTAGONIST 2u
struct SelfPlayParams {
    oth_state st;
    oth_selfplay sp;
    i64 B;
    int flags;
    int mode;
    const i64 *actions;
    const float *value, *logp, *hidden;
    u64 seed;
    i64 base;
    int init_rand;
    oth_outputs out;
    int *need_count;
};

template <int S>
OTH_DEV void sp_store_saved(const oth_selfplay &sp, i64 i, i64 B, const EnvRegs<S> &e, u32 meta) {
    constexpr int W = Geo<S>::W;
    BB_FOR(k) {
        sp.sav_black[k * B + i] = e.black.w[k];
        sp.sav_white[k * B + i] = e.white.w[k];
        sp.sav_legal[k * B + i] = e.legal.w[k];
    }
    sp.sav_meta[i] = meta;
}

template <int S>
__global__ void __launch_bounds__(OTH_BLOCK) k_selfplay(SelfPlayParams p) {
    constexpr int W = Geo<S>::W;
    __shared__ WarpStage<S> stage[OTH_WARPS];
    const i64 i = (i64)blockIdx.x * OTH_BLOCK + threadIdx.x;
    const bool active = i < p.B;
    EnvRegs<S> e = empty_env<S>();
    bool need = false;
    if (active) {
        load_env<S>(e, p.st, i, p.B);
        u32 ph = p.sp.phase[i];
        bool emit = false, zero_tuple = false;
        int tr_type = 0, tr_done = 0;
        float tr_reward = 0.f, o0 = 0.f, o1 = 0.f, o2 = 0.f;
        i64 tr_action = 0;
        if (p.mode == 0) {
            ph &= ~(u32)OTH_SP_EMITTED;
            if (ph & OTH_SP_FINISHED) {                            // 'over' message (:273-274)
                emit = true; zero_tuple = true; tr_type = 1; tr_done = 1;
                BB_FOR(k) p.sp.sav_choices[k * p.B + i] = e.legal.w[k];   // env.possible_moves, stale list included
            }
        } else if (!(ph & OTH_SP_EMITTED)) {
            const bool wm = e.meta & OTH_META_TURN_WHITE;
            const int mover = wm ? 1 : -1, P = p.sp.protagonist[i];
            const i64 a_pol = p.actions[i];
            if (mover == P) {                                      // :285-290 / :325-330
                if (!(ph & OTH_SP_PREV_P) || (p.flags & OTH_F_SP_FRESH_STATE))
                    sp_store_saved<S>(p.sp, i, p.B, e, wm ? OTH_META_TURN_WHITE : 0u);
                BB_FOR(k) p.sp.sav_choices[k * p.B + i] = e.legal.w[k];
                o0 = p.value ? p.value[i] : 0.f; o1 = p.logp ? p.logp[i] : 0.f; o2 = p.hidden ? p.hidden[i] : 0.f;
                p.sp.pend_action[i] = a_pol;
                p.sp.pend_outputs[3 * i] = o0; p.sp.pend_outputs[3 * i + 1] = o1; p.sp.pend_outputs[3 * i + 2] = o2;
                ph |= OTH_SP_PENDING;
            }
            i64 action = a_pol;
            if (meta_rand_cnt(e.meta) < meta_rand_max(e.meta)) {   // opening override (othello.py:70-73)
                action = random_pick<S>(e.legal, philox_draw(p.seed, p.base + i, OTH_STREAM_ENV, e.ctr));
                e.ctr++;
                e.meta += 1u << OTH_META_RANDCNT_SHIFT;
            }
            const StepResult r = env_step<S>(e, action, p.flags);
            const bool next_is_p = ((e.meta & OTH_META_TURN_WHITE) ? 1 : -1) == P;
            if (mover == P) {
                if (r.done || next_is_p) {                         // :292-295,313-317 / :332-335,350-354
                    emit = true; tr_reward = (float)r.reward; tr_action = a_pol;
                }
                ph |= OTH_SP_PREV_P;
            } else {
                if (r.done || next_is_p) {
                    if (ph & OTH_SP_PENDING) {                     // :305-309 / :344-347
                        emit = true; tr_reward = (float)(-r.reward);
                        tr_action = p.sp.pend_action[i];
                        o0 = p.sp.pend_outputs[3 * i]; o1 = p.sp.pend_outputs[3 * i + 1]; o2 = p.sp.pend_outputs[3 * i + 2];
                    } else if (r.done) {                           // :318-321
                        emit = true; zero_tuple = true;
                        BB_FOR(k) p.sp.sav_choices[k * p.B + i] = 0ull;
                    }
                }
                ph &= ~(u32)OTH_SP_PREV_P;
            }
            tr_done = r.done;
            if (r.done) ph |= OTH_SP_FINISHED;
            store_env<S>(e, p.st, i, p.B);
        }
        if (emit) {
            if (zero_tuple) { p.sp.sav_meta[i] = OTH_META_TERMINATED; tr_action = 0; tr_reward = 0.f; o0 = o1 = o2 = 0.f; tr_done = 1; }
            p.sp.tr_action[i] = tr_action;
            p.sp.tr_reward[i] = tr_reward;
            p.sp.tr_done[i] = (uint8_t)tr_done;
            p.sp.tr_type[i] = (uint8_t)tr_type;
            p.sp.tr_outputs[3 * i] = o0; p.sp.tr_outputs[3 * i + 1] = o1; p.sp.tr_outputs[3 * i + 2] = o2;
            ph = (ph | OTH_SP_EMITTED) & ~(u32)OTH_SP_PENDING;
        }
        p.sp.phase[i] = (uint8_t)ph;
        need = !(ph & OTH_SP_EMITTED);
        put_status(p.out, i, (e.meta & OTH_META_TERMINATED) ? 1 : 0, (e.meta & OTH_META_TURN_WHITE) ? 1 : -1, 0, false);
    }
    const u32 nb = __ballot_sync(0xFFFFFFFFu, need);
    if ((threadIdx.x & 31) == 0 && nb != 0u && p.need_count != nullptr) atomicAdd(p.need_count, __popc(nb));
    const bool fin = active && (e.meta & OTH_META_TERMINATED);
    emit_dense<S>(e, active, fin, fin, p.flags, p.out, i - (threadIdx.x & 31), p.B, stage[threadIdx.x >> 5]);
}

template <int S>
__global__ void __launch_bounds__(OTH_BLOCK) k_selfplay_new_game(oth_state s, oth_selfplay sp, i64 B, u64 seed, i64 base,
                                                                 int init_rand, u32 game_index, const int8_t *protagonist_in) {
    const i64 i = (i64)blockIdx.x * OTH_BLOCK + threadIdx.x;
    if (i >= B) return;
    EnvRegs<S> e;
    e.ctr = s.ctr[i];
    env_reset<S>(e, seed, base + i, init_rand);
    store_env<S>(e, s, i, B);
    sp.phase[i] = 0;
    sp.protagonist[i] = protagonist_in != nullptr
        ? (int8_t)(protagonist_in[i] > 0 ? 1 : -1)
        : (int8_t)((philox_draw(seed, base + i, OTH_STREAM_PROTAGONIST, game_index) & 1u) ? 1 : -1);   // :265-266
}

// -------------------------------------------------------------------- teacher / student
// The per-env machine behind PPOTeacherStudentEnvs.step + othello_teacher_vs_student (envs.py:63-140,
// ppo_run_teacher_vs_student.py:425-560), see include/othello_b200.h.  mode 0 = start of a call,
// mode 1 = one policy[colour].act worth of plies.
struct TeacherStudentParams {
    oth_state st;
    oth_teacher_student ts;
    i64 B;
    int flags;
    int mode;
    int teacher;                 // -1 | +1
    float teacher_done_reward;   // sum_k(win_avg[k] - last_win_avg[k])   (:460-461)
    int colour;                  // mode 1: the colour being served
    const i64 *actions;
    const float *value, *logp, *hidden;
    u64 seed;
    i64 base;
    int init_rand;
    oth_outputs out;
    int *need_count;             // [2]: envs waiting for a black / a white action
};

// copies colour c's saved decision into the tr_* row of env i (:457-470 send_transition)
template <int S>
OTH_DEV void ts_emit(const TeacherStudentParams &p, i64 i, int c, int raw_reward, bool done) {
    constexpr int W = Geo<S>::W;
    const oth_teacher_student &t = p.ts;
    const i64 B = p.B;
    BB_FOR(k) {
        const i64 s = ((i64)c * W + k) * B + i;
        t.tr_black[k * B + i] = t.sav_black[s];
        t.tr_white[k * B + i] = t.sav_white[s];
        t.tr_legal[k * B + i] = t.sav_legal[s];
        t.tr_choices[k * B + i] = t.sav_choices[s];
    }
    t.tr_meta[i] = t.sav_meta[(i64)c * B + i];
    t.tr_action[i] = t.sav_action[(i64)c * B + i];
    const int colour = c == 0 ? -1 : 1;
    const bool is_teacher = colour == p.teacher;
    t.tr_reward[i] = is_teacher ? (done ? p.teacher_done_reward : 0.f) : (float)raw_reward;
    t.tr_done[i] = done ? 1 : 0;
    t.tr_type[i] = (uint8_t)(c == 0 ? OTH_TS_TYPE_BLACK : OTH_TS_TYPE_WHITE);
    t.tr_who[i] = is_teacher ? 0 : 1;
#pragma unroll
    for (int j = 0; j < 3; j++) t.tr_outputs[3 * i + j] = t.sav_outputs[3 * ((i64)c * B + i) + j];
}

template <int S>
__global__ void __launch_bounds__(OTH_BLOCK) k_teacher_student(TeacherStudentParams p) {
    constexpr int W = Geo<S>::W;
    __shared__ WarpStage<S> stage[OTH_WARPS];
    const i64 i = (i64)blockIdx.x * OTH_BLOCK + threadIdx.x;
    const bool active = i < p.B;
    const oth_teacher_student &t = p.ts;
    EnvRegs<S> e = empty_env<S>();
    bool need_b = false, need_w = false;
    if (active) {
        load_env<S>(e, p.st, i, p.B);
        u32 ph = t.phase[i];
        if (p.mode == 0) {
            ph &= ~(u32)OTH_TS_EMITTED;
            if (ph & OTH_TS_FINISHED) {
                if (ph & OTH_TS_QUEUED) {                          // the last mover's own tuple (:505-509 / :546-551)
                    ts_emit<S>(p, i, (ph & OTH_TS_QUEUED_WHITE) ? 1 : 0, t.queued_reward[i], true);
                    ph &= ~(u32)(OTH_TS_QUEUED | OTH_TS_QUEUED_WHITE);
                } else {                                           // 'over' (:470-471); envs.py:124-130 counts it as white
                    BB_FOR(k) { t.tr_choices[k * p.B + i] = 0ull; }
                    t.tr_meta[i] = OTH_META_TERMINATED;
                    t.tr_action[i] = 0; t.tr_reward[i] = 0.f; t.tr_done[i] = 1;
                    t.tr_type[i] = OTH_TS_TYPE_OVER;
                    t.tr_who[i] = p.teacher == 1 ? 0 : 1;
                    t.tr_outputs[3 * i] = 0.f; t.tr_outputs[3 * i + 1] = 0.f; t.tr_outputs[3 * i + 2] = 0.f;
                }
                ph |= OTH_TS_EMITTED;
            }
        } else if (!(ph & OTH_TS_EMITTED)) {
            const bool wm = e.meta & OTH_META_TURN_WHITE;
            const int mover = wm ? 1 : -1;
            if (mover == p.colour) {
                const int c = wm ? 1 : 0, o = 1 - c;
                const u32 prev_c = wm ? OTH_TS_PREV_WHITE : OTH_TS_PREV_BLACK;
                const u32 pend_c = wm ? OTH_TS_PEND_WHITE : OTH_TS_PEND_BLACK;
                const u32 pend_o = wm ? OTH_TS_PEND_BLACK : OTH_TS_PEND_WHITE;
                const i64 a_pol = p.actions[i];
                // c's decision: state at the start of c's run, the rest refreshed every ply
                if (!(ph & prev_c) || (p.flags & OTH_F_SP_FRESH_STATE)) {
                    BB_FOR(k) {
                        const i64 s = ((i64)c * W + k) * p.B + i;
                        t.sav_black[s] = e.black.w[k]; t.sav_white[s] = e.white.w[k]; t.sav_legal[s] = e.legal.w[k];
                    }
                    t.sav_meta[(i64)c * p.B + i] = wm ? OTH_META_TURN_WHITE : 0u;
                }
                BB_FOR(k) { t.sav_choices[((i64)c * W + k) * p.B + i] = e.legal.w[k]; }
                t.sav_action[(i64)c * p.B + i] = a_pol;
                const i64 so = 3 * ((i64)c * p.B + i);
                t.sav_outputs[so] = p.value ? p.value[i] : 0.f;
                t.sav_outputs[so + 1] = p.logp ? p.logp[i] : 0.f;
                t.sav_outputs[so + 2] = p.hidden ? p.hidden[i] : 0.f;
                ph |= pend_c;
                i64 action = a_pol;
                if (meta_rand_cnt(e.meta) < meta_rand_max(e.meta)) {   // opening override (othello.py:70-73)
                    action = random_pick<S>(e.legal, philox_draw(p.seed, p.base + i, OTH_STREAM_ENV, e.ctr));
                    e.ctr++;
                    e.meta += 1u << OTH_META_RANDCNT_SHIFT;
                }
                const StepResult r = env_step<S>(e, action, p.flags);
                const bool still_c = ((e.meta & OTH_META_TURN_WHITE) != 0) == wm;
                if (!r.done && still_c) {                          // the other side passes (:486-488 / :529-531)
                    ts_emit<S>(p, i, c, r.reward, false);
                    ph |= OTH_TS_EMITTED;
                } else if (ph & pend_o) {                          // :498-503 / :541-545
                    ts_emit<S>(p, i, o, -r.reward, r.done != 0);
                    ph |= OTH_TS_EMITTED;
                    if (r.done) {
                        t.queued_reward[i] = r.reward;
                        ph |= OTH_TS_QUEUED | (wm ? OTH_TS_QUEUED_WHITE : 0u);
                    }
                } else if (r.done) {                               // :505-509 while `init` is still True
                    ts_emit<S>(p, i, c, r.reward, true);
                    ph |= OTH_TS_EMITTED;
                }
                ph = (ph & ~(u32)(OTH_TS_PREV_BLACK | OTH_TS_PREV_WHITE)) | prev_c;
                if (r.done) ph |= OTH_TS_FINISHED;
                store_env<S>(e, p.st, i, p.B);
            }
        }
        t.phase[i] = (uint8_t)ph;
        const bool need = !(ph & OTH_TS_EMITTED);
        need_w = need && (e.meta & OTH_META_TURN_WHITE);
        need_b = need && !(e.meta & OTH_META_TURN_WHITE);
        put_status(p.out, i, (e.meta & OTH_META_TERMINATED) ? 1 : 0, (e.meta & OTH_META_TURN_WHITE) ? 1 : -1, 0, false);
    }
    const u32 nb = __ballot_sync(0xFFFFFFFFu, need_b), nw = __ballot_sync(0xFFFFFFFFu, need_w);
    if ((threadIdx.x & 31) == 0 && p.need_count != nullptr) {
        if (nb != 0u) atomicAdd(&p.need_count[0], __popc(nb));
        if (nw != 0u) atomicAdd(&p.need_count[1], __popc(nw));
    }
    const bool fin = active && (e.meta & OTH_META_TERMINATED);
    emit_dense<S>(e, active, fin, fin, p.flags, p.out, i - (threadIdx.x & 31), p.B, stage[threadIdx.x >> 5]);
}

template <int S>
__global__ void __launch_bounds__(OTH_BLOCK) k_ts_new_game(oth_state s, oth_teacher_student ts, i64 B, u64 seed, i64 base,
                                                           int init_rand) {
    const i64 i = (i64)blockIdx.x * OTH_BLOCK + threadIdx.x;
    if (i >= B) return;
    EnvRegs<S> e;
    e.ctr = s.ctr[i];
    env_reset<S>(e, seed, base + i, init_rand);
    store_env<S>(e, s, i, B);
    ts.phase[i] = 0;
    ts.queued_reward[i] = 0;
}

// --------------------------------------------------------------------------- policy
template <int S>
__global__ void __launch_bounds__(OTH_BLOCK) k_policy(oth_state s, i64 B, int kind, int depth, u64 seed, u64 counter,
                                                      i64 base, i64 *actions_out) {
    const i64 i = (i64)blockIdx.x * OTH_BLOCK + threadIdx.x;
    if (i >= B) return;
    EnvRegs<S> e;
    load_env<S>(e, s, i, B);
    int a = -1;
    if (!(e.meta & OTH_META_TERMINATED) && any(e.legal)) {
        const bool wm = e.meta & OTH_META_TURN_WHITE;
        u32 rnd = 0;
        if (kind == OTH_POLICY_RANDOM)
            rnd = philox_draw(seed, base + i, OTH_STREAM_POLICY + (u32)(counter >> 32), (u32)counter);
        a = policy_pick<S, false>(kind, depth, bb_select(wm, e.white, e.black), bb_select(wm, e.black, e.white), e.legal, rnd);
    }
    actions_out[i] = a;
}

// batched maximin (depth >= 2) on single-word boards: the warp stays convergent and searches its 32 positions
// lane-cooperatively (oth_coop_search.cuh) instead of 32 divergent per-thread DFS walks
template <int S>
__global__ void __launch_bounds__(OTH_BLOCK) k_policy_coop(oth_state s, i64 B, int depth, i64 *actions_out) {
    static_assert(Geo<S>::W == 1, "cooperative search: single-word boards");
    __shared__ CoopScratch coop[OTH_WARPS];
    const i64 i = (i64)blockIdx.x * OTH_BLOCK + threadIdx.x;
    EnvRegs<S> e = empty_env<S>();
    bool need = false;
    if (i < B) {
        load_env<S>(e, s, i, B);
        need = !(e.meta & OTH_META_TERMINATED) && any(e.legal);
    }
    const bool wm = e.meta & OTH_META_TURN_WHITE;
    const u64 own = wm ? e.white.w[0] : e.black.w[0], opp = wm ? e.black.w[0] : e.white.w[0];
    CoopScratch &sc = coop[threadIdx.x >> 5];
    int a = -1;
    if (depth == 2) a = coop_maximin_pick<S, 2>(need, own, opp, e.legal.w[0], sc, false);
    else if (depth == 3) a = coop_maximin_pick<S, 3>(need, own, opp, e.legal.w[0], sc, false);
    else a = coop_maximin_pick<S, 4>(need, own, opp, e.legal.w[0], sc, false);
    if (i < B) actions_out[i] = need ? a : -1;
}

// generic-depth maximin / alpha-beta (oth_search.cuh): per thread, or AB_GW lanes per position on single-word boards
template <int S>
__global__ void __launch_bounds__(OTH_BLOCK) k_search(oth_state s, i64 B, int depth, int prune, int cooperative,
                                                      i64 *actions_out, i64 *nodes_out) {
    const i64 i = (i64)blockIdx.x * OTH_BLOCK + threadIdx.x;
    EnvRegs<S> e = empty_env<S>();
    bool need = false;
    if (i < B) {
        load_env<S>(e, s, i, B);
        need = !(e.meta & OTH_META_TERMINATED) && any(e.legal);
    }
    const bool wm = e.meta & OTH_META_TURN_WHITE;
    const BB<Geo<S>::W> own = bb_select(wm, e.white, e.black), opp = bb_select(wm, e.black, e.white);
    unsigned long long nodes = 0;
    int a = -1;
    if constexpr (Geo<S>::W == 1) {
        if (cooperative) {                                       // warp-uniform: every lane takes part
            a = coop_search_pick<S>(need, own.w[0], opp.w[0], e.legal.w[0], depth, prune != 0, nodes);
            if (i < B) { actions_out[i] = need ? a : -1; if (nodes_out) nodes_out[i] = need ? (i64)nodes : 0; }
            return;
        }
    }
    if (need) a = search_pick<S>(own, opp, e.legal, depth, prune != 0, nodes);
    if (i < B) { actions_out[i] = a; if (nodes_out) nodes_out[i] = (i64)nodes; }
}

// -------------------------------------------------------------------------- rollout
struct RolloutParams {
    i64 B;
    int flags;
    oth_policy_spec black, white;
    int init_rand;
    int games_per_env;
    u64 seed;
    i64 base;
    unsigned long long *tally;
    int8_t *game_result;
    int16_t *game_plies;
    int16_t *game_counts;
    int16_t *trace;
    int trace_stride;
    PhiloxKeys keys;     // philox_round_keys(seed), filled by the launcher
};

// Whole games in registers, one thread per game, every decision per thread (random / greedy pairings, and
// every multi-word board size).
template <int S, bool DEEP>
__global__ void __launch_bounds__(OTH_BLOCK) k_rollout(RolloutParams p) {
    const i64 i = (i64)blockIdx.x * OTH_BLOCK + threadIdx.x;
    u32 t_black = 0, t_draw = 0, t_white = 0, t_plies = 0;
    if (i < p.B) {
        EnvRegs<S> e;
        e.ctr = 0;
        PhiloxCache rng;
        philox_cache_reset(rng);
        for (int g = 0; g < p.games_per_env; g++) {
            const i64 game = i * p.games_per_env + g;
            env_reset_cached<S>(e, rng, p.keys, p.base + i, p.init_rand);
            int ply = 0;
            StepResult r;
            r.done = 0; r.winner = 0;
            while (!r.done) {
                const bool wm = e.meta & OTH_META_TURN_WHITE;
                const oth_policy_spec spec = wm ? p.white : p.black;
                const bool opening = meta_rand_cnt(e.meta) < meta_rand_max(e.meta);
                int a;
                if (opening || spec.kind == OTH_POLICY_RANDOM) {
                    a = random_pick<S>(e.legal, philox_draw_cached(rng, p.keys, p.base + i, OTH_STREAM_ENV, e.ctr));
                    e.ctr++;
                    if (opening) e.meta += 1u << OTH_META_RANDCNT_SHIFT;
                } else {
                    a = policy_pick<S, DEEP>(spec.kind, spec.depth, bb_select(wm, e.white, e.black),
                                             bb_select(wm, e.black, e.white), e.legal, 0u);
                }
                if (p.trace != nullptr && ply < p.trace_stride) p.trace[game * p.trace_stride + ply] = (int16_t)a;
                r = env_step<S>(e, a, p.flags);
                ply++;
            }
            t_plies += ply;
            t_black += r.winner < 0;
            t_draw += r.winner == 0;
            t_white += r.winner > 0;
            if (p.game_result) p.game_result[game] = (int8_t)r.winner;
            if (p.game_plies) p.game_plies[game] = (int16_t)ply;
            if (p.game_counts) {
                p.game_counts[2 * game] = (int16_t)popc(e.black);
                p.game_counts[2 * game + 1] = (int16_t)popc(e.white);
            }
        }
    }
    // warp tree-reduce, one atomic per counter per warp into the (NCCL send) buffer
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        t_black += __shfl_down_sync(0xFFFFFFFFu, t_black, o);
        t_draw += __shfl_down_sync(0xFFFFFFFFu, t_draw, o);
        t_white += __shfl_down_sync(0xFFFFFFFFu, t_white, o);
        t_plies += __shfl_down_sync(0xFFFFFFFFu, t_plies, o);
    }
    if ((threadIdx.x & 31) == 0 && p.tally != nullptr) {
        if (t_black) atomicAdd(&p.tally[0], (unsigned long long)t_black);
        if (t_draw) atomicAdd(&p.tally[1], (unsigned long long)t_draw);
        if (t_white) atomicAdd(&p.tally[2], (unsigned long long)t_white);
        if (t_plies) atomicAdd(&p.tally[3], (unsigned long long)t_plies);
    }
}

// Whole games in registers, one thread per game.  The warp stays convergent per ply so that maximin
// decisions (depth >= 2, single-word boards) can be searched lane-cooperatively (oth_coop_search.cuh);
// random / opening / greedy moves are picked per thread.
template <int S>
__global__ void __launch_bounds__(OTH_BLOCK) k_rollout_coop(RolloutParams p) {
    constexpr int W = Geo<S>::W;
    const i64 i = (i64)blockIdx.x * OTH_BLOCK + threadIdx.x;
    const bool in_range = i < p.B;
    u32 t_black = 0, t_draw = 0, t_white = 0, t_plies = 0;
    EnvRegs<S> e = empty_env<S>();
    PhiloxCache rng;
    philox_cache_reset(rng);
    for (int g = 0; g < p.games_per_env; g++) {
        const i64 game = i * p.games_per_env + g;
        if (in_range) env_reset_cached<S>(e, rng, p.keys, p.base + i, p.init_rand);
        int ply = 0;
        StepResult r;
        r.done = in_range ? 0 : 1; r.winner = 0; r.reward = 0; r.action = -1;
        while (__any_sync(0xFFFFFFFFu, !r.done)) {
            const bool live = !r.done;
            const bool wm = e.meta & OTH_META_TURN_WHITE;
            const oth_policy_spec spec = wm ? p.white : p.black;
            const BB<W> own = bb_select(wm, e.white, e.black), opp = bb_select(wm, e.black, e.white);
            int a = -1;
            int coop_depth = 0;                                     // > 0: this lane's move comes from the shared search
            if (live) {
                const bool opening = meta_rand_cnt(e.meta) < meta_rand_max(e.meta);
                if (opening || spec.kind == OTH_POLICY_RANDOM) {
                    a = random_pick<S>(e.legal, philox_draw_cached(rng, p.keys, p.base + i, OTH_STREAM_ENV, e.ctr));
                    e.ctr++;
                    if (opening) e.meta += 1u << OTH_META_RANDCNT_SHIFT;
                } else if (W == 1 && spec.kind == OTH_POLICY_MAXIMIN && spec.depth >= 2 && spec.depth <= 4 && any(e.legal)) {
                    coop_depth = spec.depth;
                } else {
                    a = policy_pick<S, false>(spec.kind, spec.depth, own, opp, e.legal, 0u);
                }
            }
            if constexpr (W == 1) {
                __shared__ CoopScratch coop[OTH_WARPS];
                CoopScratch &sc = coop[threadIdx.x >> 5];
                const bool small_lists = p.flags & OTH_F_SEARCH_SMALL_LISTS;
                if (__any_sync(0xFFFFFFFFu, coop_depth == 2)) {
                    const int c = coop_maximin_pick<S, 2>(coop_depth == 2, own.w[0], opp.w[0], e.legal.w[0], sc, small_lists);
                    if (coop_depth == 2) a = c;
                }
                if (__any_sync(0xFFFFFFFFu, coop_depth == 3)) {
                    const int c = coop_maximin_pick<S, 3>(coop_depth == 3, own.w[0], opp.w[0], e.legal.w[0], sc, small_lists);
                    if (coop_depth == 3) a = c;
                }
                if (__any_sync(0xFFFFFFFFu, coop_depth == 4)) {
                    const int c = coop_maximin_pick<S, 4>(coop_depth == 4, own.w[0], opp.w[0], e.legal.w[0], sc, small_lists);
                    if (coop_depth == 4) a = c;
                }
            }
            if (live) {
                if (p.trace != nullptr && ply < p.trace_stride) p.trace[game * p.trace_stride + ply] = (int16_t)a;
                r = env_step<S>(e, a, p.flags);
                ply++;
            }
        }
        if (in_range) {
            t_plies += ply;
            t_black += r.winner < 0;
            t_draw += r.winner == 0;
            t_white += r.winner > 0;
            if (p.game_result) p.game_result[game] = (int8_t)r.winner;
            if (p.game_plies) p.game_plies[game] = (int16_t)ply;
            if (p.game_counts) {
                p.game_counts[2 * game] = (int16_t)popc(e.black);
                p.game_counts[2 * game + 1] = (int16_t)popc(e.white);
            }
        }
    }
    // warp tree-reduce, one atomic per counter per warp into the (NCCL send) buffer
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        t_black += __shfl_down_sync(0xFFFFFFFFu, t_black, o);
        t_draw += __shfl_down_sync(0xFFFFFFFFu, t_draw, o);
        t_white += __shfl_down_sync(0xFFFFFFFFu, t_white, o);
        t_plies += __shfl_down_sync(0xFFFFFFFFu, t_plies, o);
    }
    if ((threadIdx.x & 31) == 0 && p.tally != nullptr) {
        if (t_black) atomicAdd(&p.tally[0], (unsigned long long)t_black);
        if (t_draw) atomicAdd(&p.tally[1], (unsigned long long)t_draw);
        if (t_white) atomicAdd(&p.tally[2], (unsigned long long)t_white);
        if (t_plies) atomicAdd(&p.tally[3], (unsigned long long)t_plies);
    }
}

// ------------------------------------------------------------------------ set_board
template <int S>
__global__ void __launch_bounds__(OTH_BLOCK) k_set_board(oth_state s, i64 B, const int8_t *cells, const int8_t *perspective,
                                                         const int8_t *turn, int recompute_legal, int clear_terminal) {
    constexpr int N = Geo<S>::N, W = Geo<S>::W;
    const i64 i = (i64)blockIdx.x * OTH_BLOCK + threadIdx.x;
    if (i >= B) return;
    EnvRegs<S> e;
    load_env<S>(e, s, i, B);
    if (cells != nullptr) {
        const int persp = perspective ? perspective[i] : 1;
        e.black = bb_zero<W>();
        e.white = bb_zero<W>();
        for (int a = 0; a < N; a++) {
            const int v = cells[i * N + a] * persp;     // othello.py:386-389
            const u64 bit = 1ull << (a & 63);
            BB_FOR(k) {
                if ((a >> 6) == k) {
                    if (v > 0) e.white.w[k] |= bit;
                    if (v < 0) e.black.w[k] |= bit;
                }
            }
        }
    }
    if (turn != nullptr) e.meta = (e.meta & ~OTH_META_TURN_WHITE) | (turn[i] > 0 ? OTH_META_TURN_WHITE : 0u);
    if (clear_terminal) e.meta &= ~(OTH_META_TERMINATED | (3u << OTH_META_WINNER_SHIFT));
    if (recompute_legal) {                               // set_player_turn (othello.py:464-466)
        const bool wm = e.meta & OTH_META_TURN_WHITE;
        e.legal = legal<S>(bb_select(wm, e.white, e.black), bb_select(wm, e.black, e.white));
    }
    store_env<S>(e, s, i, B);
}

// --------------------------------------------------------------------- update_board
// OthelloBaseEnv.update_board (othello.py:391-410) on its own: the side to move places a
// disc on `action` and turns every bracketed run; no validity check, turn / cached list /
// terminal flags untouched.
template <int S>
__global__ void __launch_bounds__(OTH_BLOCK) k_update_board(oth_state s, i64 B, const i64 *actions) {
    constexpr int N = Geo<S>::N, W = Geo<S>::W;
    const i64 i = (i64)blockIdx.x * OTH_BLOCK + threadIdx.x;
    if (i >= B) return;
    const i64 action = actions[i];
    if ((u64)action >= (u64)N) return;
    EnvRegs<S> e;
    load_env<S>(e, s, i, B);
    const bool wm = e.meta & OTH_META_TURN_WHITE;
    BB<W> own = bb_select(wm, e.white, e.black), opp = bb_select(wm, e.black, e.white);
    const BB<W> mv = bb_bit<W>((int)action);
    opp = andnot(opp, mv);                 // the reference overwrites the cell whatever it held
    const BB<W> f = flips_at<S>((int)action, mv, own, opp);
    own = own | f | mv;
    opp = andnot(opp, f);
    e.white = bb_select(wm, own, opp);
    e.black = bb_select(wm, opp, own);
    store_env<S>(e, s, i, B);
}

// --------------------------------------------------------------------- host launchers
struct OthSizeOps {
    int S, W;
    cudaError_t (*reset)(oth_state, i64, u64, i64, int, int, const uint8_t *, cudaStream_t);
    cudaError_t (*observe)(oth_state, i64, int, oth_outputs, cudaStream_t);
    cudaError_t (*step)(const StepParams &, cudaStream_t);
    cudaError_t (*policy)(oth_state, i64, int, int, u64, u64, i64, i64 *, cudaStream_t);
    cudaError_t (*rollout)(const RolloutParams &, cudaStream_t);
    cudaError_t (*set_board)(oth_state, i64, const int8_t *, const int8_t *, const int8_t *, int, int, cudaStream_t);
    cudaError_t (*update_board)(oth_state, i64, const i64 *, cudaStream_t);
    cudaError_t (*selfplay)(const SelfPlayParams &, cudaStream_t);
    cudaError_t (*selfplay_new_game)(oth_state, oth_selfplay, i64, u64, i64, int, u32, const int8_t *, cudaStream_t);
    cudaError_t (*teacher_student)(const TeacherStudentParams &, cudaStream_t);
    cudaError_t (*ts_new_game)(oth_state, oth_teacher_student, i64, u64, i64, int, cudaStream_t);
    cudaError_t (*search)(oth_state, i64, int, int, int, i64 *, i64 *, cudaStream_t);
};

static inline unsigned oth_grid(i64 B) { return (unsigned)((B + OTH_BLOCK - 1) / OTH_BLOCK); }

template <int S>
struct SizeLaunch {
    static cudaError_t reset(oth_state s, i64 B, u64 seed, i64 base, int init_rand, int reset_rng,
                             const uint8_t *only_where, cudaStream_t q) {
        k_reset<S><<<oth_grid(B), OTH_BLOCK, 0, q>>>(s, B, seed, base, init_rand, reset_rng, only_where);
        return cudaGetLastError();
    }
    static cudaError_t observe(oth_state s, i64 B, int flags, oth_outputs out, cudaStream_t q) {
        k_observe<S><<<oth_grid(B), OTH_BLOCK, 0, q>>>(s, B, flags, out);
        return cudaGetLastError();
    }
    template <int SRC>
    static void step_src(const StepParams &p, cudaStream_t q) {
        const oth_outputs &o = p.out;
        bool fast = false;
        if constexpr ((S * S) % 4 == 0) {
            fast = p.B % OTH_BLOCK == 0 && o.packed != nullptr && o.obs != nullptr && o.legal_mask != nullptr &&
                   o.obs_layout == OTH_OBS_BOARD && o.obs_dtype == OTH_DTYPE_I8 && !o.reward && !o.done &&
                   !o.player_turn && !o.winner && !o.actions_out;
            if (fast) {
                if (small_batch(p.B)) k_step_fast<S, SRC, OTH_BLOCK_SMALL><<<(unsigned)(p.B / OTH_BLOCK_SMALL), OTH_BLOCK_SMALL, 0, q>>>(p);
                else k_step_fast<S, SRC><<<(unsigned)(p.B / OTH_BLOCK), OTH_BLOCK, 0, q>>>(p);
            }
        }
        if (!fast) {
            if (small_batch(p.B)) k_step<S, SRC, OTH_BLOCK_SMALL><<<(unsigned)((p.B + OTH_BLOCK_SMALL - 1) / OTH_BLOCK_SMALL), OTH_BLOCK_SMALL, 0, q>>>(p);
            else k_step<S, SRC><<<oth_grid(p.B), OTH_BLOCK, 0, q>>>(p);
        }
    }
    // A small batch is cut into 64-env CTAs when that lowers the load of the fullest SM: 65,536 envs are 512 CTAs of
    // 128 on 148 SMs -- 3.46 per SM, so the SMs that get 4 (512 envs) set the time -- but 1,024 CTAs of 64 put at most
    // 7 (448 envs) on one.  Equal loads keep 128 (measured: 32,768 and 131,072 envs are 2 % faster with 128), and so
    // do large batches (64-thread CTAs measured 2 % slower at 2^20 envs).  profiles/r04i_small_batches.md
    static bool small_batch(i64 B) {
        static int sms = 0;
        if (sms == 0) {
            int dev = 0, n = 0;
            if (cudaGetDevice(&dev) == cudaSuccess && cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && n > 0) sms = n;
            else sms = 148;
        }
        if (OTH_SMALL_BATCH_CTAS_PER_SM <= 0 || B >= (i64)sms * OTH_SMALL_BATCH_CTAS_PER_SM * OTH_BLOCK) return false;
        const i64 big = (B + OTH_BLOCK - 1) / OTH_BLOCK, small = (B + OTH_BLOCK_SMALL - 1) / OTH_BLOCK_SMALL;
        return ((small + sms - 1) / sms) * OTH_BLOCK_SMALL < ((big + sms - 1) / sms) * OTH_BLOCK;
    }
    static cudaError_t step(const StepParams &p0, cudaStream_t q) {
        StepParams p = p0;
        p.keys = philox_round_keys(p.seed);
        if (p.action_source == OTH_ACT_RANDOM) step_src<OTH_ACT_RANDOM>(p, q);
        else if (p.action_source == OTH_ACT_GREEDY) step_src<OTH_ACT_GREEDY>(p, q);
        else step_src<OTH_ACT_EXTERNAL>(p, q);
        return cudaGetLastError();
    }
    static cudaError_t policy(oth_state s, i64 B, int kind, int depth, u64 seed, u64 counter, i64 base, i64 *out,
                              cudaStream_t q) {
        if constexpr (Geo<S>::W == 1) {
            // measured on 2^18 mid-game 8x8 positions, cooperative vs per-thread DFS: depth 2 0.51 vs 0.73 ms,
            // depth 3 4.8 vs 12.6 ms, depth 4 60 vs 202 ms (profiles/r01p_policy.md)
            if (kind == OTH_POLICY_MAXIMIN && depth >= 2 && depth <= 4) {
                k_policy_coop<S><<<oth_grid(B), OTH_BLOCK, 0, q>>>(s, B, depth, out);
                return cudaGetLastError();
            }
        }
        if ((kind == OTH_POLICY_MAXIMIN && depth > 4) || kind == OTH_POLICY_ALPHABETA) {   // same choice, pruned search
            k_search<S><<<oth_grid(B), OTH_BLOCK, 0, q>>>(s, B, depth, 1, Geo<S>::W == 1 ? 1 : 0, out, nullptr);
            return cudaGetLastError();
        }
        k_policy<S><<<oth_grid(B), OTH_BLOCK, 0, q>>>(s, B, kind, depth, seed, counter, base, out);
        return cudaGetLastError();
    }
    static cudaError_t rollout(const RolloutParams &p0, cudaStream_t q) {
        RolloutParams p = p0;
        p.keys = philox_round_keys(p.seed);
        const bool deep = (p.black.kind >= OTH_POLICY_MAXIMIN && p.black.depth > 4) || p.black.kind == OTH_POLICY_ALPHABETA ||
                          (p.white.kind >= OTH_POLICY_MAXIMIN && p.white.depth > 4) || p.white.kind == OTH_POLICY_ALPHABETA;
        const bool searches = !deep && ((p.black.kind == OTH_POLICY_MAXIMIN && p.black.depth >= 2) ||
                                        (p.white.kind == OTH_POLICY_MAXIMIN && p.white.depth >= 2));
        if (deep) k_rollout<S, true><<<oth_grid(p.B), OTH_BLOCK, 0, q>>>(p);
        else if (Geo<S>::W == 1 && searches && !(p.flags & OTH_F_SEARCH_PER_THREAD)) k_rollout_coop<S><<<oth_grid(p.B), OTH_BLOCK, 0, q>>>(p);
        else k_rollout<S, false><<<oth_grid(p.B), OTH_BLOCK, 0, q>>>(p);
        return cudaGetLastError();
    }
    static cudaError_t set_board(oth_state s, i64 B, const int8_t *cells, const int8_t *persp, const int8_t *turn,
                                 int recompute, int clear_terminal, cudaStream_t q) {
        k_set_board<S><<<oth_grid(B), OTH_BLOCK, 0, q>>>(s, B, cells, persp, turn, recompute, clear_terminal);
        return cudaGetLastError();
    }
    static cudaError_t update_board(oth_state s, i64 B, const i64 *actions, cudaStream_t q) {
        k_update_board<S><<<oth_grid(B), OTH_BLOCK, 0, q>>>(s, B, actions);
        return cudaGetLastError();
    }
    static cudaError_t selfplay(const SelfPlayParams &p, cudaStream_t q) {
        k_selfplay<S><<<oth_grid(p.B), OTH_BLOCK, 0, q>>>(p);
        return cudaGetLastError();
    }
    static cudaError_t selfplay_new_game(oth_state s, oth_selfplay sp, i64 B, u64 seed, i64 base, int init_rand,
                                         u32 game_index, const int8_t *protagonist_in, cudaStream_t q) {
        k_selfplay_new_game<S><<<oth_grid(B), OTH_BLOCK, 0, q>>>(s, sp, B, seed, base, init_rand, game_index, protagonist_in);
        return cudaGetLastError();
    }
    static cudaError_t teacher_student(const TeacherStudentParams &p, cudaStream_t q) {
        k_teacher_student<S><<<oth_grid(p.B), OTH_BLOCK, 0, q>>>(p);
        return cudaGetLastError();
    }
    static cudaError_t ts_new_game(oth_state s, oth_teacher_student ts, i64 B, u64 seed, i64 base, int init_rand,
                                   cudaStream_t q) {
        k_ts_new_game<S><<<oth_grid(B), OTH_BLOCK, 0, q>>>(s, ts, B, seed, base, init_rand);
        return cudaGetLastError();
    }
    static cudaError_t search(oth_state s, i64 B, int depth, int prune, int cooperative, i64 *actions, i64 *nodes,
                              cudaStream_t q) {
        k_search<S><<<oth_grid(B), OTH_BLOCK, 0, q>>>(s, B, depth, prune, Geo<S>::W == 1 ? cooperative : 0, actions, nodes);
        return cudaGetLastError();
    }
    static const OthSizeOps *ops() {
        static const OthSizeOps o = {S, Geo<S>::W, &reset, &observe, &step, &policy, &rollout, &set_board, &update_board,
                                     &selfplay, &selfplay_new_game, &teacher_student, &ts_new_game, &search};
        return &o;
    }
};
