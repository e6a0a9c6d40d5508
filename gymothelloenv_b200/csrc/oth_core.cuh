// oth_core.cuh -- bitboard rules engine for S x S Othello, S compile-time (sm_100a).
//
// One colour = S*S bits in W = ceil(S*S/64) 64-bit words, bit a <-> action a = r*S + c
// (reference geometry: othello.py:342,392-393).  Everything here lives in registers of
// ONE thread; the kernels in oth_kernels.cuh map one thread to one env.
//
// Reference behaviour implemented (paths relative to the reference root):
//   legal()      othello.py:273-343  get_num_killed_enemy / get_possible_actions
//   flips()      othello.py:391-410  update_board
//   env_step()   othello.py:412-501  step / set_player_turn / determine_winner / count_disks
//   env_reset()  othello.py:256-271 + :62-64 (opening budget)
//   greedy_pick  simple_policies.py:69-92
//   maximin_pick simple_policies.py:111-163
// The algorithm is NOT the reference's cell-by-cell ray walk: east/west moves and all flips come
// from carry propagation (an add ripples through a run of opponent discs), the other move axes
// from Kogge-Stone occluded fills (shift + mask); popcounts give the scores.
#pragma once
#include <stdint.h>
#include "../../include/othello_b200.h"

typedef unsigned long long u64;
typedef unsigned int u32;
typedef long long i64;

#define OTH_DEV __device__ __forceinline__

// ----------------------------------------------------------------------------------
// geometry: compile-time masks
// ----------------------------------------------------------------------------------
constexpr u64 geo_word_full(int S, int wi) {
    u64 m = 0;
    for (int b = 0; b < 64; b++) {
        int a = wi * 64 + b;
        if (a < S * S) m |= 1ull << b;
    }
    return m;
}
// cells whose column is neither 0 nor S-1: the only cells a horizontal / diagonal run of
// bracketed discs can occupy, and masking the propagator with it makes shifts wrap-free
constexpr u64 geo_word_inner(int S, int wi) {
    u64 m = 0;
    for (int b = 0; b < 64; b++) {
        int a = wi * 64 + b;
        if (a < S * S) {
            int c = a % S;
            if (c != 0 && c != S - 1) m |= 1ull << b;
        }
    }
    return m;
}
constexpr u64 geo_word_of_cells(int S, int wi, int a0, int a1, int a2, int a3) {
    u64 m = 0;
    int cells[4] = {a0, a1, a2, a3};
    for (int k = 0; k < 4; k++)
        if (cells[k] >= 0 && cells[k] / 64 == wi) m |= 1ull << (cells[k] % 64);
    return m;
}
template <int S, int WI>
struct GeoWord {
    static constexpr int c = S / 2;
    static constexpr u64 full = geo_word_full(S, WI);
    static constexpr u64 inner = geo_word_inner(S, WI);
    // othello.py:256-263: WHITE (c-1,c-1),(c,c); BLACK (c,c-1),(c-1,c)
    static constexpr u64 start_white = geo_word_of_cells(S, WI, (c - 1) * S + (c - 1), c * S + c, -1, -1);
    static constexpr u64 start_black = geo_word_of_cells(S, WI, c * S + (c - 1), (c - 1) * S + c, -1, -1);
    // BLACK's first legal set: (c-2,c-1),(c-1,c-2),(c,c+1),(c+1,c)
    static constexpr u64 start_legal =
        geo_word_of_cells(S, WI, (c - 2) * S + (c - 1), (c - 1) * S + (c - 2), c * S + (c + 1), (c + 1) * S + c);
};

template <int S>
struct Geo {
    static constexpr int N = S * S;
    static constexpr int W = (N + 63) / 64;
    static_assert(S >= OTH_MIN_BOARD_SIZE && S <= 16, "board size out of range");
};

template <int W>
struct BB {
    u64 w[W];
};

#define BB_FOR(i) _Pragma("unroll") for (int i = 0; i < W; i++)

template <int W> OTH_DEV BB<W> operator&(const BB<W> &a, const BB<W> &b) { BB<W> r; BB_FOR(i) r.w[i] = a.w[i] & b.w[i]; return r; }
template <int W> OTH_DEV BB<W> operator|(const BB<W> &a, const BB<W> &b) { BB<W> r; BB_FOR(i) r.w[i] = a.w[i] | b.w[i]; return r; }
template <int W> OTH_DEV BB<W> operator^(const BB<W> &a, const BB<W> &b) { BB<W> r; BB_FOR(i) r.w[i] = a.w[i] ^ b.w[i]; return r; }
template <int W> OTH_DEV BB<W> andnot(const BB<W> &a, const BB<W> &b) { BB<W> r; BB_FOR(i) r.w[i] = a.w[i] & ~b.w[i]; return r; }
template <int W> OTH_DEV bool any(const BB<W> &a) { u64 x = 0; BB_FOR(i) x |= a.w[i]; return x != 0; }
template <int W> OTH_DEV int popc(const BB<W> &a) { int n = 0; BB_FOR(i) n += __popcll(a.w[i]); return n; }
template <int W> OTH_DEV BB<W> bb_zero() { BB<W> r; BB_FOR(i) r.w[i] = 0; return r; }
template <int W> OTH_DEV BB<W> bb_select(bool c, const BB<W> &a, const BB<W> &b) { BB<W> r; BB_FOR(i) r.w[i] = c ? a.w[i] : b.w[i]; return r; }
template <int W> OTH_DEV BB<W> bb_bit(int a) {   // a in [0, 64*W)
    BB<W> r;
    BB_FOR(i) r.w[i] = (a >> 6) == i ? (1ull << (a & 63)) : 0ull;
    return r;
}
template <int W> OTH_DEV bool bb_test(const BB<W> &x, int a) {
    u64 v = 0;
    BB_FOR(i) v = (a >> 6) == i ? x.w[i] : v;
    return (v >> (a & 63)) & 1ull;
}

#define GEO_BB(NAME, FIELD)                                                        \
    template <int S> OTH_DEV BB<Geo<S>::W> NAME() {                                \
        BB<Geo<S>::W> r;                                                           \
        r.w[0] = GeoWord<S, 0>::FIELD;                                             \
        if constexpr (Geo<S>::W > 1) r.w[1] = GeoWord<S, 1>::FIELD;                \
        if constexpr (Geo<S>::W > 2) r.w[2] = GeoWord<S, 2>::FIELD;                \
        if constexpr (Geo<S>::W > 3) r.w[3] = GeoWord<S, 3>::FIELD;                \
        return r;                                                                  \
    }
GEO_BB(geo_full, full)
GEO_BB(geo_inner, inner)
GEO_BB(geo_start_white, start_white)
GEO_BB(geo_start_black, start_black)
GEO_BB(geo_start_legal, start_legal)

// whole-board shifts by a compile-time distance (towards higher / lower bit indices)
template <int K, int W>
OTH_DEV BB<W> shl(const BB<W> &a) {
    constexpr int q = K / 64, b = K % 64;
    BB<W> r;
    BB_FOR(i) {
        u64 hi = (i - q >= 0) ? a.w[(i - q >= 0) ? i - q : 0] : 0ull;
        u64 lo = (i - q - 1 >= 0) ? a.w[(i - q - 1 >= 0) ? i - q - 1 : 0] : 0ull;
        r.w[i] = (b == 0) ? hi : ((hi << b) | (lo >> ((64 - b) & 63)));
    }
    return r;
}
template <int K, int W>
OTH_DEV BB<W> shr(const BB<W> &a) {
    constexpr int q = K / 64, b = K % 64;
    BB<W> r;
    BB_FOR(i) {
        u64 lo = (i + q < W) ? a.w[(i + q < W) ? i + q : 0] : 0ull;
        u64 hi = (i + q + 1 < W) ? a.w[(i + q + 1 < W) ? i + q + 1 : 0] : 0ull;
        r.w[i] = (b == 0) ? lo : ((lo >> b) | (hi << ((64 - b) & 63)));
    }
    return r;
}
template <int K, bool LEFT, int W>
OTH_DEV BB<W> sh(const BB<W> &a) {
    if constexpr (LEFT) return shl<K, W>(a);
    else return shr<K, W>(a);
}

// ---- single-word boards (S <= 8): legal moves -----------------------------------------
// East: adding the cells just east of own discs to the inner-column opponent mask makes
// the carry ripple through each bracket-able run and land on the cell behind it (edge
// columns are 0 in the mask and a row is S contiguous bits, so a carry never leaves its
// row).  West = the same on bit-reversed boards (BREV runs off the ALU pipe; the reversed
// board sits in the top S*S bits, rows still contiguous).  The three other axes use Kogge-Stone
// occluded fills (steps 1, 2, 2 cover the longest run, S-2 <= 6), upward on the normal boards and
// -- for the three downward directions -- upward on the reversed boards.
// One direction only, towards higher bit indices: a Kogge-Stone occluded fill, reach 2, 4, 6 cells.
// A 64-bit LEFT shift is one IMAD.SHL (low half, FMA pipe) + one SHF (high half, ALU pipe); a RIGHT shift is two
// SHF on the ALU pipe, the binding unit of k_step.  legal1 therefore runs the downward directions as upward fills
// on the bit-reversed boards (BREV is on the XU pipe) and reverses the union once.
// REACH = the longest run a fill must cover (S - 2): rounds reach 1, 2, 4, 6 cells; smaller boards stop earlier.
template <int D, int REACH>
OTH_DEV u64 fill_up1(u64 P, u64 m) {
    u64 f = m & (P << D);
    f |= m & (f << D);
    if constexpr (REACH > 2) {
        const u64 pl = m & (m << D);
        f |= pl & (f << (2 * D));
        if constexpr (REACH > 4) f |= pl & (f << (2 * D));
    }
    return f << D;
}
template <int S>
OTH_DEV u64 legal1(u64 P, u64 O) {
    static_assert(Geo<S>::W == 1, "single-word boards only");
    const u64 Oi = O & GeoWord<S, 0>::inner;
    const u64 Pr = __brevll(P), Or = __brevll(O), Oir = __brevll(Oi);
    u64 up = (Oi + ((P << 1) & Oi)) & ~Oi;                 // east: carry through the opponent run
    up |= fill_up1<S, S - 2>(P, O);
    up |= fill_up1<S - 1, S - 2>(P, Oi);
    up |= fill_up1<S + 1, S - 2>(P, Oi);
    u64 dn = (Oir + ((Pr << 1) & Oir)) & ~Oir;             // west = east on the reversed board, and so on
    dn |= fill_up1<S, S - 2>(Pr, Or);
    dn |= fill_up1<S - 1, S - 2>(Pr, Oir);
    dn |= fill_up1<S + 1, S - 2>(Pr, Oir);
    const u64 moves = up | __brevll(dn);
    if constexpr (S == 8) return moves & ~(P | O);
    else return moves & ~(P | O) & GeoWord<S, 0>::full;
}

// ---- multi-word boards (S = 9..16, W = 2..4): whole-board arithmetic --------------------
// The single-word tricks carry over once add / subtract / bit-reverse work on the whole W-word
// board: a row is still S contiguous bits (it may straddle a word boundary, which is why the
// carry must run through all words), and the bit-reversed board still has contiguous rows.
#ifdef OTH_HOST_EMUL   // tests/host_emul: the same carry chains in portable C++ (never part of the product build)
template <int W> OTH_DEV BB<W> bb_add(const BB<W> &a, const BB<W> &b) {
    BB<W> r; unsigned c = 0;
    BB_FOR(i) { const unsigned __int128 t = (unsigned __int128)a.w[i] + b.w[i] + c; r.w[i] = (u64)t; c = (unsigned)(t >> 64); }
    return r;
}
template <int W> OTH_DEV BB<W> bb_sub(const BB<W> &a, const BB<W> &b) {
    BB<W> r; unsigned c = 0;
    BB_FOR(i) { const unsigned __int128 t = (unsigned __int128)a.w[i] - b.w[i] - c; r.w[i] = (u64)t; c = (unsigned)(t >> 64) & 1u; }
    return r;
}
#else
template <int W> OTH_DEV BB<W> bb_add(const BB<W> &a, const BB<W> &b) {
    BB<W> r;
    if constexpr (W == 1) {
        r.w[0] = a.w[0] + b.w[0];
    } else if constexpr (W == 2) {
        asm("add.cc.u64 %0, %2, %4;\n\taddc.u64 %1, %3, %5;"
            : "=l"(r.w[0]), "=l"(r.w[1]) : "l"(a.w[0]), "l"(a.w[1]), "l"(b.w[0]), "l"(b.w[1]));
    } else if constexpr (W == 3) {
        asm("add.cc.u64 %0, %3, %6;\n\taddc.cc.u64 %1, %4, %7;\n\taddc.u64 %2, %5, %8;"
            : "=l"(r.w[0]), "=l"(r.w[1]), "=l"(r.w[2])
            : "l"(a.w[0]), "l"(a.w[1]), "l"(a.w[2]), "l"(b.w[0]), "l"(b.w[1]), "l"(b.w[2]));
    } else {
        static_assert(W == 4, "bb_add: W <= 4");
        asm("add.cc.u64 %0, %4, %8;\n\taddc.cc.u64 %1, %5, %9;\n\taddc.cc.u64 %2, %6, %10;\n\taddc.u64 %3, %7, %11;"
            : "=l"(r.w[0]), "=l"(r.w[1]), "=l"(r.w[2]), "=l"(r.w[3])
            : "l"(a.w[0]), "l"(a.w[1]), "l"(a.w[2]), "l"(a.w[3]), "l"(b.w[0]), "l"(b.w[1]), "l"(b.w[2]), "l"(b.w[3]));
    }
    return r;
}
template <int W> OTH_DEV BB<W> bb_sub(const BB<W> &a, const BB<W> &b) {
    BB<W> r;
    if constexpr (W == 1) {
        r.w[0] = a.w[0] - b.w[0];
    } else if constexpr (W == 2) {
        asm("sub.cc.u64 %0, %2, %4;\n\tsubc.u64 %1, %3, %5;"
            : "=l"(r.w[0]), "=l"(r.w[1]) : "l"(a.w[0]), "l"(a.w[1]), "l"(b.w[0]), "l"(b.w[1]));
    } else if constexpr (W == 3) {
        asm("sub.cc.u64 %0, %3, %6;\n\tsubc.cc.u64 %1, %4, %7;\n\tsubc.u64 %2, %5, %8;"
            : "=l"(r.w[0]), "=l"(r.w[1]), "=l"(r.w[2])
            : "l"(a.w[0]), "l"(a.w[1]), "l"(a.w[2]), "l"(b.w[0]), "l"(b.w[1]), "l"(b.w[2]));
    } else {
        static_assert(W == 4, "bb_sub: W <= 4");
        asm("sub.cc.u64 %0, %4, %8;\n\tsubc.cc.u64 %1, %5, %9;\n\tsubc.cc.u64 %2, %6, %10;\n\tsubc.u64 %3, %7, %11;"
            : "=l"(r.w[0]), "=l"(r.w[1]), "=l"(r.w[2]), "=l"(r.w[3])
            : "l"(a.w[0]), "l"(a.w[1]), "l"(a.w[2]), "l"(a.w[3]), "l"(b.w[0]), "l"(b.w[1]), "l"(b.w[2]), "l"(b.w[3]));
    }
    return r;
}
#endif
template <int W> OTH_DEV BB<W> bb_one(u64 v) { BB<W> r; BB_FOR(i) r.w[i] = i == 0 ? v : 0ull; return r; }
// bit j -> bit 64*W - 1 - j
template <int W> OTH_DEV BB<W> bb_brev(const BB<W> &a) { BB<W> r; BB_FOR(i) r.w[i] = __brevll(a.w[W - 1 - i]); return r; }

// East / west moves of a multi-word board by carry propagation (see legal1); Oi = opponent discs in
// the inner columns, in the same (normal or reversed) bit order as P.
template <int W> OTH_DEV BB<W> carry_moves(const BB<W> &P, const BB<W> &Oi) {
    return andnot(bb_add<W>(Oi, shl<1, W>(P) & Oi), Oi);
}
// The three other axes: paired Kogge-Stone fills sharing the doubled propagators between the two
// opposite directions; reach after the rounds: 2, 4, 8, 12, 16 cells (the longest run is S-2 <= 14).
template <int S, int D>
OTH_DEV BB<Geo<S>::W> pair_moves(const BB<Geo<S>::W> &P, const BB<Geo<S>::W> &m) {
    constexpr int W = Geo<S>::W;
    BB<W> fl = m & shl<D, W>(P), fr = m & shr<D, W>(P);
    fl = fl | (m & shl<D, W>(fl));
    fr = fr | (m & shr<D, W>(fr));
    const BB<W> pl = m & shl<D, W>(m), pr = shr<D, W>(pl);
    fl = fl | (pl & shl<2 * D, W>(fl));
    fr = fr | (pr & shr<2 * D, W>(fr));
    if constexpr (S - 2 > 4) {
        const BB<W> pl2 = pl & shl<2 * D, W>(pl), pr2 = pr & shr<2 * D, W>(pr);
        fl = fl | (pl2 & shl<4 * D, W>(fl));
        fr = fr | (pr2 & shr<4 * D, W>(fr));
        if constexpr (S - 2 > 8) {
            fl = fl | (pl2 & shl<4 * D, W>(fl));
            fr = fr | (pr2 & shr<4 * D, W>(fr));
        }
        if constexpr (S - 2 > 12) {
            fl = fl | (pl2 & shl<4 * D, W>(fl));
            fr = fr | (pr2 & shr<4 * D, W>(fr));
        }
    }
    return shl<D, W>(fl) | shr<D, W>(fr);
}

// othello.py:313-343: empty cells from which >=1 direction holds a run of opponent discs
// closed by an own disc.
template <int S>
OTH_DEV BB<Geo<S>::W> legal(const BB<Geo<S>::W> &own, const BB<Geo<S>::W> &opp) {
    constexpr int W = Geo<S>::W;
    if constexpr (W == 1) {
        BB<W> r;
        r.w[0] = legal1<S>(own.w[0], opp.w[0]);
        return r;
    } else {
        const BB<W> Oi = opp & geo_inner<S>();
        BB<W> m = carry_moves<W>(own, Oi);
        m = m | bb_brev<W>(carry_moves<W>(bb_brev<W>(own), bb_brev<W>(Oi)));
        m = m | pair_moves<S, S>(own, opp);
        m = m | pair_moves<S, S - 1>(own, Oi);
        m = m | pair_moves<S, S + 1>(own, Oi);
        return andnot(m & geo_full<S>(), own | opp);
    }
}

// ---- single-word boards (S <= 8): flips by carry propagation ---------------------------
// For a ray of cells above the move (towards higher bit indices) with mask m:
//   x = O | ~m ; t = x + 1 : the +1 carry ripples through every non-ray bit and through the
//   contiguous opponent discs at the start of the ray and stops on the first ray cell that
//   is not an opponent disc; t & ~x is that cell.  If it holds an own disc the run below it
//   is bracketed (ray_stop1 returns that closing disc, else 0).
// The four rays towards lower indices are the same computation on bit-reversed boards.
// Rays are mv * pattern (mv is a power of two, so the 64-bit IMADs are shifts that run on
// the FMA pipe instead of the saturated ALU pipe); ray cells past the last board bit hold
// neither colour, so they can only end a run unbracketed.  Masking the opponent with the
// inner columns for the horizontal / diagonal rays stops a run at the board edge (no wrap).
constexpr u64 geo_ray_pattern(int S, int D) {
    u64 m = 0;
    for (int k = 1; k < S; k++)
        if (k * D < 64) m |= 1ull << (k * D);
    return m;
}
template <int S>
struct GeoRays {
    static constexpr u64 e = geo_ray_pattern(S, 1), s = geo_ray_pattern(S, S);
    static constexpr u64 se = geo_ray_pattern(S, S + 1), sw = geo_ray_pattern(S, S - 1);
};
// the closing own disc of one ray (0 if the run is not bracketed): see above
OTH_DEV u64 ray_stop1(u64 m, u64 P, u64 O) {
    const u64 x = O | ~m;
    const u64 t = x + 1ull;
    return t & ~x & P;
}
// The discs a ray turns, rebuilt from its closing disc WITHOUT a compare / select: on the bit-reversed board the
// closing disc lies BELOW the move and the turned discs are the ray cells from it up to the move, i.e.
// (brev(stop) * pattern) & (brev(mv) - 1) -- the product is 0 when the ray is open, the 64-bit multiply runs on the
// FMA pipe and the reversal on the XU pipe, so a ray costs 6 ALU-pipe instructions (O|~m, +1, &) plus its share of
// one OR / AND per half instead of 14.  The four rays of one orientation share the AND with the cells below the
// move; their result comes out in the OTHER orientation.
template <int S>
OTH_DEV u64 flips1_half(u64 mv, u64 mv_other, u64 P, u64 O, u64 Oi) {
    u64 f = __brevll(ray_stop1(mv * GeoRays<S>::e, P, Oi)) * GeoRays<S>::e;
    f |= __brevll(ray_stop1(mv * GeoRays<S>::s, P, O)) * GeoRays<S>::s;
    f |= __brevll(ray_stop1(mv * GeoRays<S>::se, P, Oi)) * GeoRays<S>::se;
    f |= __brevll(ray_stop1(mv * GeoRays<S>::sw, P, Oi)) * GeoRays<S>::sw;
    return f & (mv_other - 1ull);          // mv_other = brev(mv); mv == 0 gives stops 0, hence f == 0
}
template <int S>
OTH_DEV u64 flips1(u64 mv, u64 P, u64 O) {
    static_assert(Geo<S>::W == 1, "single-word boards only");
    const u64 Oi = O & GeoWord<S, 0>::inner;
    const u64 mvr = __brevll(mv);
    const u64 dn_rev = flips1_half<S>(mv, mvr, P, O, Oi);                                 // rays towards higher indices
    const u64 up = flips1_half<S>(mvr, mv, __brevll(P), __brevll(O), __brevll(Oi));      // rays towards lower indices
    return up | __brevll(dn_rev);
}

// ---- single-word boards: the flip count of EVERY cell at once (bit-sliced) --------------
// GreedyPolicy and the last ply of MaxiMinPolicy.search only need the LARGEST number of discs any one move turns
// (simple_policies.py:69-92; :127-137 at depth 1: value = own count + flips + 1, resp. opponent count - flips).
// Instead of one flips1() per move, each of the 8 directions is walked once for all cells: t_k = the k-th disc of an
// opponent run that starts next to an own disc; a move on the cell behind t_k turns exactly k discs in that direction.
// The run length k (1..S-2) of a cell is kept as three bit planes; the eight directions are added with bit-sliced
// adders -- opposite directions first: a line through a cell holds at most S-1 other cells and every bracketed run
// needs its closing disc, so the two sums together stay <= S-2 and still fit three planes; four such sums
// (<= 4(S-2) <= 24) fit five -- and the maximum over a set of cells falls out of the planes from the top bit down.
// About 400 instructions whatever the number of moves (one flips1 + popcount per move costs ~125), and no loop whose
// trip count differs between the lanes of a warp.  Planes are only meaningful on EMPTY cells.
struct FlipPlanes { u64 c[5]; };
template <int D, bool UP> OTH_DEV u64 sh1(u64 x) { return UP ? (x << D) : (x >> D); }
OTH_DEV u64 maj3(u64 a, u64 b, u64 c) { return (a & b) | (c & (a | b)); }
// one direction: bit planes of the run length seen from the cell behind the run
template <int S, int D, bool UP>
OTH_DEV void run_planes1(u64 P, u64 m, u64 &b0, u64 &b1, u64 &b2) {
    constexpr int R = S - 2;
    u64 t[6];
    t[0] = sh1<D, UP>(P) & m;
#pragma unroll
    for (int k = 1; k < 6; k++) t[k] = k < R ? (sh1<D, UP>(t[k - 1]) & m) : 0ull;
    b0 = sh1<D, UP>(t[0] | t[2] | t[4]);
    b1 = sh1<D, UP>(t[1] | t[2] | t[5]);
    b2 = sh1<D, UP>(t[3] | t[4] | t[5]);
}
// both directions of one axis, added (sum <= S-2: no fourth plane)
template <int S, int D>
OTH_DEV void axis_planes1(u64 P, u64 m, u64 *s) {
    u64 a0, a1, a2, b0, b1, b2;
    run_planes1<S, D, true>(P, m, a0, a1, a2);
    run_planes1<S, D, false>(P, m, b0, b1, b2);
    const u64 c0 = a0 & b0;
    s[0] = a0 ^ b0;
    s[1] = a1 ^ b1 ^ c0;
    s[2] = a2 ^ b2 ^ maj3(a1, b1, c0);
}
template <int S>
OTH_DEV FlipPlanes flip_planes1(u64 P, u64 O) {
    static_assert(Geo<S>::W == 1, "single-word boards only");
    const u64 Oi = O & GeoWord<S, 0>::inner;
    u64 h[3], v[3], d[3], e[3];
    axis_planes1<S, 1>(P, Oi, h);
    axis_planes1<S, S>(P, O, v);
    axis_planes1<S, S + 1>(P, Oi, d);
    axis_planes1<S, S - 1>(P, Oi, e);
    u64 x[4], y[4], c;                                    // x = h + v, y = d + e: four planes each
    x[0] = h[0] ^ v[0]; c = h[0] & v[0];
    x[1] = h[1] ^ v[1] ^ c; c = maj3(h[1], v[1], c);
    x[2] = h[2] ^ v[2] ^ c; x[3] = maj3(h[2], v[2], c);
    y[0] = d[0] ^ e[0]; c = d[0] & e[0];
    y[1] = d[1] ^ e[1] ^ c; c = maj3(d[1], e[1], c);
    y[2] = d[2] ^ e[2] ^ c; y[3] = maj3(d[2], e[2], c);
    FlipPlanes f;                                         // x + y: five planes
    f.c[0] = x[0] ^ y[0]; c = x[0] & y[0];
    f.c[1] = x[1] ^ y[1] ^ c; c = maj3(x[1], y[1], c);
    f.c[2] = x[2] ^ y[2] ^ c; c = maj3(x[2], y[2], c);
    f.c[3] = x[3] ^ y[3] ^ c; f.c[4] = maj3(x[3], y[3], c);
    return f;
}
// the largest count among the cells of `cand`; cand shrinks to the cells that reach it
OTH_DEV int planes_max(const FlipPlanes &f, u64 &cand) {
    int v = 0;
#pragma unroll
    for (int p = 4; p >= 0; p--) {
        const u64 t = cand & f.c[p];
        if (t != 0ull) { cand = t; v |= 1 << p; }
    }
    return v;
}
// legal moves of P (as legal1) and the most discs any of them turns (0 when there is no move)
template <int S>
OTH_DEV int max_flips1(u64 P, u64 O, u64 &moves) {
    const FlipPlanes f = flip_planes1<S>(P, O);
    moves = (f.c[0] | f.c[1] | f.c[2] | f.c[3] | f.c[4]) & ~(P | O) & GeoWord<S, 0>::full;
    u64 cand = moves;
    return planes_max(f, cand);
}

// ---- multi-word boards: flips by carry propagation -------------------------------------
// whole-board shift towards higher bit indices by a VARIABLE distance 0 <= a < 64*W
template <int W> OTH_DEV BB<W> bb_shl_var(const BB<W> &x, int a) {
    const int q = a >> 6, b = a & 63;
    BB<W> s;                                             // x << b
    BB_FOR(i) {
        const u64 lo = i > 0 ? x.w[i > 0 ? i - 1 : 0] : 0ull;
        s.w[i] = (x.w[i] << b) | ((lo >> 1) >> (63 - b));
    }
    BB<W> r;                                             // then q whole words
    BB_FOR(i) {
        u64 v = 0ull;
        BB_FOR(j) { if (j <= i) v = (q == i - j) ? s.w[j] : v; }
        r.w[i] = v;
    }
    return r;
}
constexpr u64 geo_word_ray_pattern(int S, int D, int wi) {   // cells D, 2D, ..., (S-1)D
    u64 m = 0;
    for (int k = 1; k < S; k++) {
        const int c = k * D;
        if (c / 64 == wi) m |= 1ull << (c % 64);
    }
    return m;
}
template <int S, int D, int WI>
struct GeoRayWord {
    static constexpr u64 v = geo_word_ray_pattern(S, D, WI);
};
template <int S, int D> OTH_DEV BB<Geo<S>::W> geo_ray() {
    BB<Geo<S>::W> r;
    r.w[0] = GeoRayWord<S, D, 0>::v;
    if constexpr (Geo<S>::W > 1) r.w[1] = GeoRayWord<S, D, 1>::v;
    if constexpr (Geo<S>::W > 2) r.w[2] = GeoRayWord<S, D, 2>::v;
    if constexpr (Geo<S>::W > 3) r.w[3] = GeoRayWord<S, D, 3>::v;
    return r;
}
// one ray with mask m (cells above the move, nearest first): the multi-word form of ray_stop1 -- the flipped run is
// the ray cells the carry cleared, m & O & ~t, when the closing cell holds an own disc
template <int W> OTH_DEV BB<W> ray_flips(const BB<W> &m, const BB<W> &P, const BB<W> &O) {
    BB<W> x;
    BB_FOR(i) x.w[i] = O.w[i] | ~m.w[i];
    const BB<W> t = bb_add<W>(x, bb_one<W>(1ull));
    BB<W> stop;
    BB_FOR(i) stop.w[i] = t.w[i] & ~x.w[i] & P.w[i];       // 0 or the single closing own disc
    const bool closed = any(stop);
    BB<W> r;
    BB_FOR(i) r.w[i] = closed ? (m.w[i] & O.w[i] & ~t.w[i]) : 0ull;   // ray cells the carry rippled through
    return r;
}
// the four rays towards higher bit indices from cell a (mv = its bit): east needs no ray mask -- the
// carry out of mv<<1 ripples through the inner-column opponent run and lands on the cell behind it
template <int S>
OTH_DEV BB<Geo<S>::W> flips_up(int a, const BB<Geo<S>::W> &mv, const BB<Geo<S>::W> &P, const BB<Geo<S>::W> &O,
                               const BB<Geo<S>::W> &Oi) {
    constexpr int W = Geo<S>::W;
    const BB<W> first = shl<1, W>(mv);
    const BB<W> sum = bb_add<W>(Oi, first);
    const BB<W> land = andnot(sum, Oi) & P;
    BB<W> f = bb_select(any(land), andnot(Oi, sum), bb_zero<W>());   // the discs the carry rippled through
    // (ray masks as (every D-th bit << a mod D) & above(a) avoid this variable whole-board shift but measured 1-3 %
    // slower: profiles/r03e_experiments.md)
    f = f | ray_flips<W>(bb_shl_var<W>(geo_ray<S, S>(), a), P, O);
    f = f | ray_flips<W>(bb_shl_var<W>(geo_ray<S, S + 1>(), a), P, Oi);
    f = f | ray_flips<W>(bb_shl_var<W>(geo_ray<S, S - 1>(), a), P, Oi);
    return f;
}

// othello.py:391-410: discs turned by placing an own disc on cell a (mv = its single bit, or 0 for
// "no move": the result is then 0, which is how invalid moves stay branch-free).
template <int S>
OTH_DEV BB<Geo<S>::W> flips_at(int a, const BB<Geo<S>::W> &mv, const BB<Geo<S>::W> &own, const BB<Geo<S>::W> &opp) {
    constexpr int W = Geo<S>::W;
    if constexpr (W == 1) {
        BB<W> f;
        f.w[0] = flips1<S>(mv.w[0], own.w[0], opp.w[0]);
        return f;
    } else {
        const BB<W> Oi = opp & geo_inner<S>();
        const BB<W> up = flips_up<S>(a, mv, own, opp, Oi);
        const BB<W> dn = flips_up<S>(64 * W - 1 - a, bb_brev<W>(mv), bb_brev<W>(own), bb_brev<W>(opp), bb_brev<W>(Oi));
        const BB<W> f = up | bb_brev<W>(dn);
        return bb_select(any(mv), f, bb_zero<W>());
    }
}
// position of the n-th (0-based) set bit, n < popc(x)
OTH_DEV int nth_bit64(u64 x, int n) {
    int pos = 0;
    u32 v = (u32)x;
    int c = __popc(v);
    if (n >= c) { n -= c; pos = 32; v = (u32)(x >> 32); }
    c = __popc(v & 0xFFFFu); if (n >= c) { n -= c; pos += 16; v >>= 16; }
    c = __popc(v & 0xFFu);   if (n >= c) { n -= c; pos += 8;  v >>= 8; }
    c = __popc(v & 0xFu);    if (n >= c) { n -= c; pos += 4;  v >>= 4; }
    c = __popc(v & 0x3u);    if (n >= c) { n -= c; pos += 2;  v >>= 2; }
    c = v & 1u;              if (n >= c) { pos += 1; }
    return pos;
}
template <int W>
OTH_DEV int nth_bit(const BB<W> &x, int n) {
    int base = 0;
    u64 word = x.w[0];
    bool found = false;
    BB_FOR(i) {
        if (i + 1 < W) {
            int c = __popcll(x.w[i]);
            if (!found) {
                if (n < c) found = true;
                else { n -= c; base = (i + 1) * 64; word = x.w[(i + 1 < W) ? i + 1 : 0]; }
            }
        }
    }
    return base + nth_bit64(word, n);
}
// lowest set bit index, x != 0
template <int W>
OTH_DEV int lowest_bit(const BB<W> &x) {
    int r = 0;
    bool found = false;
    BB_FOR(i) {
        if (!found && x.w[i]) { r = i * 64 + __ffsll((long long)x.w[i]) - 1; found = true; }
    }
    return r;
}
// flips of the single cell `mv` when its index is not at hand (mv == 0 gives 0)
template <int S>
OTH_DEV BB<Geo<S>::W> flips(const BB<Geo<S>::W> &mv, const BB<Geo<S>::W> &own, const BB<Geo<S>::W> &opp) {
    return flips_at<S>(lowest_bit(mv), mv, own, opp);
}

template <int W>
OTH_DEV void clear_bit(BB<W> &x, int a) {
    BB_FOR(i) x.w[i] &= ~((a >> 6) == i ? (1ull << (a & 63)) : 0ull);
}

// ----------------------------------------------------------------------------------
// Philox4x32-10 (Salmon et al., SC'11), counter-based: the stream of env `gid` is
// block(ctr>>2)[ctr&3] with key = seed, counter = {block index, stream id, gid}.
// ----------------------------------------------------------------------------------
struct Philox4 { u32 v[4]; };
OTH_DEV Philox4 philox4x32_10(u32 c0, u32 c1, u32 c2, u32 c3, u32 k0, u32 k1) {
    const u32 M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
    for (int r = 0; r < 10; r++) {
        u32 hi0 = __umulhi(M0, c0), lo0 = M0 * c0;
        u32 hi1 = __umulhi(M1, c2), lo1 = M1 * c2;
        u32 n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
        c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
        k0 += W0; k1 += W1;
    }
    Philox4 o;
    o.v[0] = c0; o.v[1] = c1; o.v[2] = c2; o.v[3] = c3;
    return o;
}
// The ten round keys depend on the seed only: the hot kernel takes them precomputed by the
// host as launch parameters, so every round's key is a constant-bank operand of the XOR
// instead of two per-thread IADDs (20 ALU-pipe instructions per draw).
struct PhiloxKeys { u32 k0[10], k1[10]; };
__host__ __device__ inline PhiloxKeys philox_round_keys(u64 seed) {
    PhiloxKeys K;
    u32 a = (u32)seed, b = (u32)(seed >> 32);
    for (int r = 0; r < 10; r++) { K.k0[r] = a; K.k1[r] = b; a += 0x9E3779B9u; b += 0xBB67AE85u; }
    return K;
}
OTH_DEV Philox4 philox4x32_10(u32 c0, u32 c1, u32 c2, u32 c3, const PhiloxKeys &K) {
    const u32 M0 = 0xD2511F53u, M1 = 0xCD9E8D57u;
#pragma unroll
    for (int r = 0; r < 10; r++) {
        u32 hi0 = __umulhi(M0, c0), lo0 = M0 * c0;
        u32 hi1 = __umulhi(M1, c2), lo1 = M1 * c2;
        u32 n0 = hi1 ^ c1 ^ K.k0[r], n2 = hi0 ^ c3 ^ K.k1[r];
        c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
    }
    Philox4 o;
    o.v[0] = c0; o.v[1] = c1; o.v[2] = c2; o.v[3] = c3;
    return o;
}
#define OTH_STREAM_ENV 0u      // opening budget, opening moves, fused random policy
#define OTH_STREAM_POLICY 1u   // oth_policy(random)
OTH_DEV u32 philox_draw(u64 seed, i64 gid, u32 stream, u32 ctr) {
    Philox4 b = philox4x32_10(ctr >> 2, stream, (u32)(u64)gid, (u32)((u64)gid >> 32), (u32)seed, (u32)(seed >> 32));
    u32 k = ctr & 3u;
    return k == 0 ? b.v[0] : (k == 1 ? b.v[1] : (k == 2 ? b.v[2] : b.v[3]));
}
OTH_DEV u32 philox_draw(const PhiloxKeys &K, i64 gid, u32 stream, u32 ctr) {
    Philox4 b = philox4x32_10(ctr >> 2, stream, (u32)(u64)gid, (u32)((u64)gid >> 32), K);
    u32 k = ctr & 3u;
    return k == 0 ? b.v[0] : (k == 1 ? b.v[1] : (k == 2 ? b.v[2] : b.v[3]));
}
// Whole games in one kernel draw consecutive words of one stream: keep the 4-word block in registers and run the ten
// rounds only when the draw index crosses a block boundary (every 4th draw).  Same stream, 1/4 of the Philox work.
struct PhiloxCache {
    Philox4 b;
    u32 block;          // index of the cached block, 0xFFFFFFFF = none
};
OTH_DEV void philox_cache_reset(PhiloxCache &c) { c.block = 0xFFFFFFFFu; c.b.v[0] = c.b.v[1] = c.b.v[2] = c.b.v[3] = 0u; }
OTH_DEV u32 philox_draw_cached(PhiloxCache &c, const PhiloxKeys &K, i64 gid, u32 stream, u32 ctr) {
    if ((ctr >> 2) != c.block) {
        c.b = philox4x32_10(ctr >> 2, stream, (u32)(u64)gid, (u32)((u64)gid >> 32), K);
        c.block = ctr >> 2;
    }
    const u32 k = ctr & 3u;
    return k == 0 ? c.b.v[0] : (k == 1 ? c.b.v[1] : (k == 2 ? c.b.v[2] : c.b.v[3]));
}
// uniform integer in [0, n) from 32 random bits (Lemire multiply-shift; bias <= n/2^32)
OTH_DEV int bounded(u32 r, int n) { return (int)(((u64)r * (u64)(u32)n) >> 32); }

// ----------------------------------------------------------------------------------
// env in registers
// ----------------------------------------------------------------------------------
template <int S>
struct EnvRegs {
    BB<Geo<S>::W> black, white, legal;
    u32 meta, ctr;
};

OTH_DEV int meta_rand_cnt(u32 m) { return (m >> OTH_META_RANDCNT_SHIFT) & 0xFF; }
OTH_DEV int meta_rand_max(u32 m) { return (m >> OTH_META_RANDMAX_SHIFT) & 0xFF; }
OTH_DEV int meta_winner(u32 m) { int c = (m >> OTH_META_WINNER_SHIFT) & 3; return c == 1 ? 1 : (c == 2 ? -1 : 0); }
OTH_DEV u32 winner_code(int w) { return w > 0 ? 1u : (w < 0 ? 2u : 0u); }

// OthelloBaseEnv.reset (othello.py:265-271) + opening budget (othello.py:62-64)
template <int S, class KEY>
OTH_DEV void env_reset(EnvRegs<S> &e, const KEY &seed, i64 gid, int init_rand_steps) {
    e.black = geo_start_black<S>();
    e.white = geo_start_white<S>();
    e.legal = geo_start_legal<S>();
    int half = init_rand_steps / 2;
    int budget = 0;
    if (half > 0) {   // rnd.randint(0, init//2 + 1) * 2
        budget = 2 * bounded(philox_draw(seed, gid, OTH_STREAM_ENV, e.ctr), half + 1);
        e.ctr++;
    }
    e.meta = (u32)budget << OTH_META_RANDMAX_SHIFT;   // turn = BLACK, not terminated, winner 0, cnt 0
}

// the same with the rollout kernels' cached Philox block
template <int S>
OTH_DEV void env_reset_cached(EnvRegs<S> &e, PhiloxCache &c, const PhiloxKeys &K, i64 gid, int init_rand_steps) {
    e.black = geo_start_black<S>();
    e.white = geo_start_white<S>();
    e.legal = geo_start_legal<S>();
    int half = init_rand_steps / 2;
    int budget = 0;
    if (half > 0) {
        budget = 2 * bounded(philox_draw_cached(c, K, gid, OTH_STREAM_ENV, e.ctr), half + 1);
        e.ctr++;
    }
    e.meta = (u32)budget << OTH_META_RANDMAX_SHIFT;
}

struct StepResult {
    int reward;   // othello.py:444-461
    int done;     // self.terminated
    int winner;   // -1/0/+1 when done
    int action;   // the action that was applied (or attempted)
};

// OthelloBaseEnv.step (othello.py:412-462) for a NON-terminated env.
template <int S>
OTH_DEV StepResult env_step(EnvRegs<S> &e, i64 action, int flags) {
    constexpr int N = Geo<S>::N, W = Geo<S>::W;
    const bool white_moves = e.meta & OTH_META_TURN_WHITE;
    const int mover = white_moves ? 1 : -1;
    BB<W> own = bb_select(white_moves, e.white, e.black);
    BB<W> opp = bb_select(white_moves, e.black, e.white);

    const bool in_range = (u64)action < (u64)N;
    const int a = in_range ? (int)action : 0;
    const bool valid = in_range && bb_test(e.legal, a);        // against the CACHED list (:417)
    BB<W> mv = bb_select(valid, bb_bit<W>(a), bb_zero<W>());
    BB<W> f = flips_at<S>(a, mv, own, opp);                    // update_board (:391-410)
    own = own | f | mv;
    opp = andnot(opp, f);

    const bool sudden = !valid && (flags & OTH_F_SUDDEN_DEATH); // :427
    const BB<W> vacant = andnot(geo_full<S>(), own | opp);
    const bool full = !any(vacant);                             // :425-426
    bool terminated = sudden || full;
    bool next_white = white_moves;
    BB<W> next_legal = e.legal;                                 // stale on sudden/full terminal (App. A.6)
    if (!terminated) {
        BB<W> lo = legal<S>(opp, own);                          // set_player_turn(-turn) (:436)
        if (any(lo)) {
            next_legal = lo;
            next_white = !white_moves;
        } else {
            BB<W> lm = legal<S>(own, opp);                      // pass (:437-438)
            next_legal = lm;
            terminated = !any(lm);                              // double pass (:439-442)
        }
    }
    const int own_cnt = popc(own), opp_cnt = popc(opp);
    int winner = 0, reward = 0;
    if (terminated) {
        const int white_cnt = white_moves ? own_cnt : opp_cnt, black_cnt = white_moves ? opp_cnt : own_cnt;
        winner = sudden ? -mover : (white_cnt > black_cnt ? 1 : (black_cnt > white_cnt ? -1 : 0)); // :473-501
        if (flags & OTH_F_DISK_REWARD) {
            reward = sudden ? -N : (opp_cnt == 0 ? N : own_cnt - opp_cnt);   // :446-459
        } else {
            reward = winner * mover;                                         // :461
        }
    }
    e.white = bb_select(white_moves, own, opp);
    e.black = bb_select(white_moves, opp, own);
    e.legal = next_legal;
    e.meta = (e.meta & 0xFFFFFF00u) | (next_white ? OTH_META_TURN_WHITE : 0u) |
             (terminated ? OTH_META_TERMINATED : 0u) | (winner_code(winner) << OTH_META_WINNER_SHIFT);
    StepResult r;
    r.reward = reward; r.done = terminated; r.winner = winner; r.action = (int)action;
    return r;
}

// ---- multi-word boards (S = 9..16): the flip count of every cell at once, on whole-board arithmetic ------------
// The scheme of flip_planes1 with BB<W> operands: run lengths reach S-2 <= 14 (four planes per direction, and per
// axis: the two directions of a line still sum to <= S-2), the four axes add up to <= 4(S-2) <= 56 (six planes).
// Not inlined: one copy per board size (about 1 000 instructions per 64-bit word), called by greedy and by the last
// ply of the searches.
#ifndef OTH_MW_PLANES
#define OTH_MW_PLANES 1      // 0: multi-word boards score one move at a time (A/B builds)
#endif
template <int W> OTH_DEV BB<W> bb_maj(const BB<W> &a, const BB<W> &b, const BB<W> &c) { return (a & b) | (c & (a | b)); }
template <int W> struct FlipPlanesW { BB<W> c[6]; };
template <int S, int D, bool UP>
OTH_DEV void run_planes(const BB<Geo<S>::W> &P, const BB<Geo<S>::W> &m, BB<Geo<S>::W> *b) {
    constexpr int W = Geo<S>::W, R = S - 2;
    BB<W> t = sh<D, UP, W>(P) & m;                        // t = the k-th disc of a run, k = 1
    b[0] = t; b[1] = bb_zero<W>(); b[2] = b[1]; b[3] = b[1];
#pragma unroll
    for (int k = 2; k <= R; k++) {
        t = sh<D, UP, W>(t) & m;
#pragma unroll
        for (int j = 0; j < 4; j++)
            if ((k >> j) & 1) b[j] = b[j] | t;
    }
#pragma unroll
    for (int j = 0; j < 4; j++) b[j] = sh<D, UP, W>(b[j]);
}
// o = a + b over NA planes each; the carry out becomes plane NA when NO > NA
template <int W, int NA, int NO>
OTH_DEV void planes_add(const BB<W> *a, const BB<W> *b, BB<W> *o) {
    BB<W> c = a[0] & b[0];
    o[0] = a[0] ^ b[0];
#pragma unroll
    for (int i = 1; i < NA; i++) {
        o[i] = a[i] ^ b[i] ^ c;
        if (i + 1 < NA || NO > NA) c = bb_maj<W>(a[i], b[i], c);
    }
    if constexpr (NO > NA) o[NA] = c;
}
template <int S, int D>
OTH_DEV void axis_planes(const BB<Geo<S>::W> &P, const BB<Geo<S>::W> &m, BB<Geo<S>::W> *s) {
    constexpr int W = Geo<S>::W;
    BB<W> a[4], b[4];
    run_planes<S, D, true>(P, m, a);
    run_planes<S, D, false>(P, m, b);
    planes_add<W, 4, 4>(a, b, s);
}
template <int S>
__device__ __noinline__ FlipPlanesW<Geo<S>::W> flip_planes(const BB<Geo<S>::W> &P, const BB<Geo<S>::W> &O) {
    constexpr int W = Geo<S>::W;
    const BB<W> Oi = O & geo_inner<S>();
    BB<W> h[4], v[4], x[5], y[5];
    axis_planes<S, 1>(P, Oi, h);
    axis_planes<S, S>(P, O, v);
    planes_add<W, 4, 5>(h, v, x);
    axis_planes<S, S + 1>(P, Oi, h);
    axis_planes<S, S - 1>(P, Oi, v);
    planes_add<W, 4, 5>(h, v, y);
    FlipPlanesW<W> f;
    planes_add<W, 5, 6>(x, y, f.c);
    return f;
}
// the largest count among the cells of `cand`; cand shrinks to the cells that reach it
template <int W>
OTH_DEV int planes_max(const FlipPlanesW<W> &f, BB<W> &cand) {
    int v = 0;
#pragma unroll
    for (int p = 5; p >= 0; p--) {
        const BB<W> t = cand & f.c[p];
        if (any(t)) { cand = t; v |= 1 << p; }
    }
    return v;
}
// legal moves of P (as legal<S>) and the most discs any of them turns (0 when there is no move), any board size
template <int S>
OTH_DEV int max_flips(const BB<Geo<S>::W> &P, const BB<Geo<S>::W> &O, BB<Geo<S>::W> &moves) {
    constexpr int W = Geo<S>::W;
    if constexpr (W == 1) {
        return max_flips1<S>(P.w[0], O.w[0], moves.w[0]);
    } else {
        const FlipPlanesW<W> f = flip_planes<S>(P, O);
        moves = andnot((f.c[0] | f.c[1] | f.c[2] | f.c[3] | f.c[4] | f.c[5]) & geo_full<S>(), P | O);
        BB<W> cand = moves;
        return planes_max<W>(f, cand);
    }
}

// RandomPolicy.get_action / the opening override: uniform over the cached list
// (simple_policies.py:37-41, othello.py:70-72)
template <int S>
OTH_DEV int random_pick(const BB<Geo<S>::W> &moves, u32 rnd) {
    int n = popc(moves);
    return n > 0 ? nth_bit(moves, bounded(rnd, n)) : -1;
}

// GreedyPolicy.get_action (simple_policies.py:69-92): own disc count after each move =
// own + flips + 1, np.argmax => most flips, lowest index on ties.
template <int S>
OTH_DEV int greedy_pick(const BB<Geo<S>::W> &own, const BB<Geo<S>::W> &opp, BB<Geo<S>::W> moves) {
    constexpr int W = Geo<S>::W;
    if constexpr (W == 1) {
        // all moves at once (flip_planes1).  A cached list that names OCCUPIED cells can only come from
        // set_board_state without a refresh of the list (othello.py:380-389 keeps possible_moves): the loop below
        // reproduces update_board on such cells.
        if ((moves.w[0] & (own.w[0] | opp.w[0])) == 0ull) {
            if (moves.w[0] == 0ull) return -1;
            const FlipPlanes f = flip_planes1<S>(own.w[0], opp.w[0]);
            u64 cand = moves.w[0];
            planes_max(f, cand);
            return __ffsll((long long)cand) - 1;           // np.argmax: the first (lowest) of the best moves
        }
    }
    if constexpr (W > 1 && OTH_MW_PLANES) {
        if (!any(moves & (own | opp))) {
            if (!any(moves)) return -1;
            const FlipPlanesW<W> f = flip_planes<S>(own, opp);
            BB<W> cand = moves;
            planes_max<W>(f, cand);
            return lowest_bit(cand);
        }
    }
    int best = -1, best_cnt = -1;
    while (any(moves)) {
        int a = lowest_bit(moves);
        clear_bit(moves, a);
        int cnt = popc(flips_at<S>(a, bb_bit<W>(a), own, opp));
        if (cnt > best_cnt) { best_cnt = cnt; best = a; }
    }
    return best;
}

// MaxiMinPolicy.search below the root (simple_policies.py:111-155).  `pl` is to move.
// Value = disc count of the searching side `me`; leaf when the depth budget is spent or
// the side to move has no move (covers terminal boards and the forced-pass rule :139-144).
template <int S, int D>
struct MaxiMin {
    static __device__ int value(const BB<Geo<S>::W> &pl, const BB<Geo<S>::W> &op, bool me_is_pl) {
        constexpr int W = Geo<S>::W;
        if constexpr (D == 0) {
            return popc(me_is_pl ? pl : op);
        } else if constexpr (D == 1 && (W == 1 || OTH_MW_PLANES)) {   // children are leaves: every move scored at once
            BB<W> moves;
            const int k = max_flips<S>(pl, op, moves);
            if (!any(moves)) return popc(me_is_pl ? pl : op);
            return me_is_pl ? popc(pl) + k + 1 : popc(op) - k;
        } else {
            BB<W> moves = legal<S>(pl, op);
            if (!any(moves)) return popc(me_is_pl ? pl : op);
            int best = me_is_pl ? -1 : 0x7FFFFFFF;
            const int pl_cnt = popc(pl), op_cnt = popc(op);
            while (any(moves)) {
                int a = lowest_bit(moves);
                clear_bit(moves, a);
                BB<W> mv = bb_bit<W>(a);
                BB<W> f = flips_at<S>(a, mv, pl, op);
                int v;
                if constexpr (D == 1) {      // children are leaves: count without expanding them
                    int k = popc(f);
                    v = me_is_pl ? pl_cnt + k + 1 : op_cnt - k;
                } else {
                    v = MaxiMin<S, D - 1>::value(andnot(op, f), pl | f | mv, !me_is_pl);
                }
                best = me_is_pl ? max(best, v) : min(best, v);
            }
            return best;
        }
    }
    // root: iterate the CACHED list (env.possible_moves), first argmax (:151-155,157-163)
    static __device__ int pick(const BB<Geo<S>::W> &own, const BB<Geo<S>::W> &opp, BB<Geo<S>::W> moves) {
        constexpr int W = Geo<S>::W;
        static_assert(D >= 1, "depth");
        int best = -1, best_v = -1;
        const int own_cnt = popc(own);
        while (any(moves)) {
            int a = lowest_bit(moves);
            clear_bit(moves, a);
            BB<W> mv = bb_bit<W>(a);
            BB<W> f = flips_at<S>(a, mv, own, opp);
            int v;
            if constexpr (D == 1) v = own_cnt + popc(f) + 1;
            else v = MaxiMin<S, D - 1>::value(andnot(opp, f), own | f | mv, false);
            if (v > best_v) { best_v = v; best = a; }
        }
        return best;
    }
};

// (policy_pick lives in oth_search.cuh: depths above 4 run the pruned generic-depth search)
