// oth_coop_search.cuh -- lane-cooperative maximin search for single-word boards (S <= 8).
//
// MaxiMinPolicy.search (simple_policies.py:111-163) as a per-thread DFS wastes most of a warp: every
// nested `for move in possible_moves` loop runs to the LONGEST trip count among the 32 games of the warp
// (measured on B200: 10 of 32 lanes active per instruction for maximin-3, 14 for maximin-2).  Here the lanes of a
// warp share the work of the games that need a decision, and the LAST ply of every line is never enumerated: the
// value of a position whose children are leaves is own count + 1 + (most discs any move turns), resp. opponent
// count - (most discs ...), and max_flips1 (oth_core.cuh) scores all its moves at once.  A work item is therefore
// "apply one move, score the position behind it one ply deep" (~550 instructions, the same for every item):
//   depth 2   items = the root moves of ALL games of the warp that need a decision, numbered through the warp
//             (prefix sum of the move counts) and dealt round-robin to the 32 lanes; a game's best (value, lowest
//             move) is folded with one shared-memory atomicMax per item on a packed key
//   depth 3   the same pooling on two levels: the root moves of a batch of games are dealt round-robin (one lane = one
//             (game, root move): apply it, compute the reply set -> level-1 node list in shared memory), then the
//             (node, reply) pairs of the whole batch are split into equal consecutive chunks over the lanes and each
//             pair's value is folded into its node with atomicMin; every lane takes the first argmax of ITS game
//   depth 4   the root's moves in sequence, each searched to depth 3 by the whole warp (roles swapped), one game at
//             a time
// Values only travel upwards through min / max, so the result does not depend on the order of evaluation; the one
// place where ties matter is the root (np.argmax first occurrence, :151-155), reduced with (value, lowest index).
// More root moves in one position than the lists hold (never seen on 8x8): that decision falls back to the per-thread
// DFS; OTH_F_SEARCH_SMALL_LISTS shrinks the limits so that the tests run batching and fallback constantly.
#pragma once
#include "oth_core.cuh"

constexpr int COOP_CAP1 = 32;       // depth 4: root moves of the ONE game being searched
constexpr int COOP_CAPN = 256;      // depth 3: level-1 nodes of ALL games of a batch (32 games x ~10 root moves fit twice over)
constexpr int COOP_INF = 0x3FFFFFFF;

// per warp: the level-1 node list (depth >= 3); depth 2 uses val / first per GAME LANE
struct CoopScratch {
    u64 pl[COOP_CAPN], op[COOP_CAPN], mv[COOP_CAPN];   // side to move, other side, its legal moves
    int val[COOP_CAPN];
    short first[COOP_CAPN + 2];                        // exclusive prefix sums of the move counts
    unsigned short item[COOP_CAPN];                    // depth 3: the batch's root moves, (game lane << 6) | cell, in item order
};

// exclusive prefix sum of x over the warp, and the total
OTH_DEV int coop_excl_scan(int x, int lane, int &total) {
    int incl = x;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xFFFFFFFFu, incl, o);
        if (lane >= o) incl += t;
    }
    total = __shfl_sync(0xFFFFFFFFu, incl, 31);
    return incl - x;
}
// largest p in [0, n) with first[p] <= q   (first[0] = 0, non-decreasing)
OTH_DEV int coop_find_node(const short *first, int n, int q) {
    int lo = 0, hi = n - 1;
    while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (first[mid] <= q) lo = mid; else hi = mid - 1;
    }
    return lo;
}
OTH_DEV u64 coop_drop_lowest(u64 m, int k) {
    for (; k > 0; k--) m &= m - 1ull;
    return m;
}

// One ply below a position in which `cpl` is to move and `cop` is the other side, for the searching side `me`:
// cpl_is_me ? own count + 1 + most flips : own count - most flips; a side without a move makes the position a leaf
// (simple_policies.py:117-126,139-144), which is the same formula with "most flips" = 0 and no disc placed.
template <int S>
OTH_DEV int coop_one_ply(u64 cpl, u64 cop, bool cpl_is_me) {
    u64 L;
    const int k = max_flips1<S>(cpl, cop, L);
    if (cpl_is_me) return __popcll(cpl) + (L != 0ull ? k + 1 : 0);
    return __popcll(cop) - k;
}

// Every lane lists ITS game's moves at its place in the warp-wide item numbering (`first` = the exclusive prefix sum
// of the move counts): the lane that later takes item q reads (game, cell) with one load instead of searching the
// prefix sums and extracting the k-th set bit.
OTH_DEV void coop_list_moves(unsigned short *item, int lane, int first, u64 moves) {
    int q = first;
    for (u64 m = moves; m != 0ull; m &= m - 1ull) item[q++] = (unsigned short)((lane << 6) | (__ffsll((long long)m) - 1));
}

// ---- depth 2: every needing game of the warp at once ------------------------------------------------------------
template <int S>
OTH_DEV int coop_pick_depth2(bool need, u64 own, u64 opp, u64 moves, CoopScratch &sc) {
    const int lane = threadIdx.x & 31;
    const int n = need ? __popcll(moves) : 0;
    int total;
    const int first = coop_excl_scan(n, lane, total);
    sc.val[lane] = -1;
    // the item list lies over the node arrays, which depth 2 does not use: pl + op = 2 048 entries, and 32 games have
    // at most 32 x 60 moves between them
    static_assert(sizeof(sc.pl) + sizeof(sc.op) >= 2 * 32 * 64, "item list capacity");
    unsigned short *item = reinterpret_cast<unsigned short *>(sc.pl);
    if (n > 0) coop_list_moves(item, lane, first, moves);
    __syncwarp();
    for (int q0 = 0; q0 < total; q0 += 32) {                         // warp-uniform
        const int q = q0 + lane;
        const bool on = q < total;
        const int it = on ? item[q] : 0;
        const int g = it >> 6, a = it & 63;                          // the game (lane) and cell of item q
        const u64 P = __shfl_sync(0xFFFFFFFFu, own, g), O = __shfl_sync(0xFFFFFFFFu, opp, g);
        if (on) {
            const u64 mv = 1ull << a;
            const u64 f = flips1<S>(mv, P, O);
            const int v = coop_one_ply<S>(O & ~f, P | f | mv, false);
            atomicMax(&sc.val[g], v * 64 + (63 - a));                // most discs, then the LOWEST move (np.argmax)
        }
    }
    __syncwarp();
    const int key = sc.val[lane];
    __syncwarp();                                                    // the scratch is reused by the next call
    return (need && key >= 0) ? 63 - (key & 63) : -1;
}

// ---- depth 3: one game at a time, the whole warp ----------------------------------------------------------------
// Searches 3 plies below a node whose side to move (P) has the non-empty move set M.  top_is_me: that side is the
// searching side.  Returns the node's value; *best_index = index (in ascending move order) of the first move
// achieving it.  Warp-uniform arguments and results.  ok = false: more moves than the list holds, nothing computed.
template <int S>
OTH_DEV int coop_search3(u64 P, u64 O, u64 M, bool top_is_me, CoopScratch &sc, int lane, int *best_index, bool &ok, int cap1) {
    const int n1 = __popcll(M);
    ok = n1 <= cap1;
    if (!ok) return 0;
    const bool child_is_me = !top_is_me;
    // level 1: one lane per move of the top node
    {
        int local = 0;
        if (lane < n1) {
            const u64 m = coop_drop_lowest(M, lane);
            const u64 mv = m & (0ull - m);
            const u64 f = flips1<S>(mv, P, O);
            const u64 cpl = O & ~f, cop = P | f | mv;                // the other side moves next
            const u64 L = legal1<S>(cpl, cop);
            sc.pl[lane] = cpl; sc.op[lane] = cop; sc.mv[lane] = L;
            // a side with no move makes the node a leaf (simple_policies.py:117-126,139-144)
            sc.val[lane] = (L == 0ull) ? __popcll(child_is_me ? cpl : cop) : (child_is_me ? -1 : COOP_INF);
            local = __popcll(L);
        }
        int total;
        const int run = coop_excl_scan(local, lane, total);
        if (lane < n1) sc.first[lane] = (short)run;
        if (lane == 0) sc.first[n1] = (short)total;
        __syncwarp();
    }
    // level 2: (node, move) pairs in equal consecutive chunks; each pair = apply the move, score one more ply
    const int n2 = sc.first[n1];
    {
        const int chunk = (n2 + 31) >> 5;
        const int q0 = min(lane * chunk, n2), q1 = min(q0 + chunk, n2);
        if (q0 < q1) {
            int p = coop_find_node(sc.first, n1, q0);
            u64 pl = sc.pl[p], op = sc.op[p];
            u64 m = coop_drop_lowest(sc.mv[p], q0 - sc.first[p]);
            int best = child_is_me ? -1 : COOP_INF;
            for (int q = q0; q < q1; q++) {
                while (m == 0ull) {
                    if (child_is_me) atomicMax(&sc.val[p], best); else atomicMin(&sc.val[p], best);
                    p++; pl = sc.pl[p]; op = sc.op[p]; m = sc.mv[p];
                    best = child_is_me ? -1 : COOP_INF;
                }
                const u64 mv = m & (0ull - m);
                m &= m - 1ull;
                const u64 f = flips1<S>(mv, pl, op);
                const int v = coop_one_ply<S>(op & ~f, pl | f | mv, top_is_me);   // behind it the top's side moves again
                best = child_is_me ? max(best, v) : min(best, v);
            }
            if (child_is_me) atomicMax(&sc.val[p], best); else atomicMin(&sc.val[p], best);
        }
        __syncwarp();
    }
    // top: max (min) over the moves, first occurrence on ties
    int bv = top_is_me ? -1 : COOP_INF, bi = 0x7FFF;
    if (lane < n1) { bv = sc.val[lane]; bi = lane; }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const int ov = __shfl_xor_sync(0xFFFFFFFFu, bv, o), oi = __shfl_xor_sync(0xFFFFFFFFu, bi, o);
        const bool better = top_is_me ? (ov > bv) : (ov < bv);
        if (better || (ov == bv && oi < bi)) { bv = ov; bi = oi; }      // idle lanes hold -1 / INF: never better
    }
    __syncwarp();                                                    // the list is reused by the next game
    if (best_index) *best_index = bi;
    return bv;
}

// ---- depth 3, every needing game of the warp at once ----------------------------------------------------------------
// Level 1: the root moves of all games of a batch are numbered through the warp and dealt round-robin (one lane = one
// (game, root move): apply it, compute the reply set); level 2: the (node, reply) pairs of the whole batch in equal
// consecutive chunks, each pair scored one ply deep and folded into its node with atomicMin; then every lane takes the
// first maximum over ITS game's nodes.  A batch = as many games, in lane order, as fit `cap` level-1 nodes; a game
// with more root moves than that is searched by its own lane alone (per-thread DFS).
template <int S>
OTH_DEV int coop_pick_depth3(bool need, u64 own, u64 opp, u64 moves, CoopScratch &sc, int cap) {
    const int lane = threadIdx.x & 31;
    const int n1 = need ? __popcll(moves) : 0;
    int result = -1;
    bool pending = n1 > 0;
    if (pending && n1 > cap) {                                       // never on 8x8 with the real capacity
        BB<1> o, p, mm;
        o.w[0] = own; p.w[0] = opp; mm.w[0] = moves;
        result = MaxiMin<S, 3>::pick(o, p, mm);
        pending = false;
    }
    while (__any_sync(0xFFFFFFFFu, pending)) {                       // warp-uniform
        // the batch: pending games in lane order while their root moves fit
        int total;
        const int start = coop_excl_scan(pending ? n1 : 0, lane, total);
        const bool mine = pending && start + n1 <= cap;              // a prefix of the pending games (counts are > 0)
        const int T1 = __reduce_max_sync(0xFFFFFFFFu, mine ? start + n1 : 0);
        if (mine) coop_list_moves(sc.item, lane, start, moves);
        __syncwarp();
        // level 1
        int run = 0;                                                 // reply pairs before this round's nodes
        for (int q0 = 0; q0 < T1; q0 += 32) {
            const int q = q0 + lane;
            const bool on = q < T1;
            const int it = on ? sc.item[q] : 0;
            const int g = it >> 6;                                   // the game (lane) of item q
            const u64 P = __shfl_sync(0xFFFFFFFFu, own, g), O = __shfl_sync(0xFFFFFFFFu, opp, g);
            int cnt = 0;
            if (on) {
                const u64 mv = 1ull << (it & 63);
                const u64 f = flips1<S>(mv, P, O);
                const u64 cpl = O & ~f, cop = P | f | mv;            // the opponent moves next
                const u64 L = legal1<S>(cpl, cop);
                sc.pl[q] = cpl; sc.op[q] = cop; sc.mv[q] = L;
                sc.val[q] = (L == 0ull) ? __popcll(cop) : COOP_INF;  // no reply: a leaf (simple_policies.py:139-144)
                cnt = __popcll(L);
            }
            int round_total;
            const int ex = coop_excl_scan(cnt, lane, round_total);
            if (on) sc.first[q] = (short)(run + ex);
            run += round_total;
        }
        if (lane == 0) sc.first[T1] = (short)run;
        __syncwarp();
        // level 2
        {
            const int T2 = run;
            const int chunk = (T2 + 31) >> 5;
            const int q0 = min(lane * chunk, T2), q1 = min(q0 + chunk, T2);
            if (q0 < q1) {
                int p = coop_find_node(sc.first, T1, q0);
                u64 pl = sc.pl[p], op = sc.op[p];
                u64 m = coop_drop_lowest(sc.mv[p], q0 - sc.first[p]);
                int best = COOP_INF;
                for (int q = q0; q < q1; q++) {
                    while (m == 0ull) {
                        atomicMin(&sc.val[p], best);
                        p++; pl = sc.pl[p]; op = sc.op[p]; m = sc.mv[p];
                        best = COOP_INF;
                    }
                    const u64 mv = m & (0ull - m);
                    m &= m - 1ull;
                    const u64 f = flips1<S>(mv, pl, op);
                    best = min(best, coop_one_ply<S>(op & ~f, pl | f | mv, true));   // the searching side moves again
                }
                atomicMin(&sc.val[p], best);
            }
            __syncwarp();
        }
        // every game of the batch: first maximum over its root moves (np.argmax, :151-155)
        if (mine) {
            int bv = -1, bi = 0;
            for (int j = 0; j < n1; j++) {
                const int v = sc.val[start + j];
                if (v > bv) { bv = v; bi = j; }
            }
            result = nth_bit64(moves, bi);
            pending = false;
        }
        __syncwarp();                                                // the lists are reused by the next batch
    }
    return result;
}

// MaxiMinPolicy.get_action (depth 3 or 4) for one game by the whole warp (warp-uniform arguments): the chosen action.
template <int S, int D>
OTH_DEV int coop_pick_one(u64 own, u64 opp, u64 moves, CoopScratch &sc, int lane, int cap1) {
    static_assert(D == 3 || D == 4, "coop_pick_one depth");
    bool ok = true;
    int a = -1;
    if constexpr (D == 3) {
        int bi = 0;
        coop_search3<S>(own, opp, moves, true, sc, lane, &bi, ok, cap1);
        if (ok) a = nth_bit64(moves, bi);
    } else {                                                         // D == 4: root moves in sequence
        int best_v = -1;
        u64 m = moves;
        while (m != 0ull && ok) {
            const u64 mv = m & (0ull - m);
            m &= m - 1ull;
            const u64 f = flips1<S>(mv, own, opp);
            const u64 cpl = opp & ~f, cop = own | f | mv;
            const u64 L = legal1<S>(cpl, cop);
            int v;
            if (L == 0ull) v = __popcll(cop);
            else v = coop_search3<S>(cpl, cop, L, false, sc, lane, nullptr, ok, cap1);
            if (ok && v > best_v) { best_v = v; a = __ffsll((long long)mv) - 1; }
        }
    }
    if (!ok) {                                                       // capacity overflow: per-thread DFS, redundantly
        BB<1> o, p, mm;
        o.w[0] = own; p.w[0] = opp; mm.w[0] = moves;
        a = MaxiMin<S, D>::pick(o, p, mm);
    }
    return a;
}

// Warp-convergent entry: every lane passes its own game; lanes with need = true get their action back.
// depth in 2..4 must be the same for all needing lanes of one call.
template <int S, int D>
__device__ __noinline__ int coop_maximin_pick(bool need, u64 own, u64 opp, u64 moves, CoopScratch &sc, bool small_lists) {
    static_assert(Geo<S>::W == 1, "cooperative search: single-word boards");
    if constexpr (D == 2) {
        return coop_pick_depth2<S>(need, own, opp, moves, sc);
    } else if constexpr (D == 3) {
        return coop_pick_depth3<S>(need, own, opp, moves, sc, small_lists ? 8 : COOP_CAPN);   // OTH_F_SEARCH_SMALL_LISTS
    } else {
        const int lane = threadIdx.x & 31;
        u32 todo = __ballot_sync(0xFFFFFFFFu, need);
        const int cap1 = small_lists ? 3 : COOP_CAP1;                // OTH_F_SEARCH_SMALL_LISTS
        int result = -1;
        while (todo != 0u) {                                         // warp-uniform
            const int src = __ffs(todo) - 1;
            todo &= todo - 1u;
            const u64 P = __shfl_sync(0xFFFFFFFFu, own, src), O = __shfl_sync(0xFFFFFFFFu, opp, src);
            const u64 M = __shfl_sync(0xFFFFFFFFu, moves, src);
            const int a = coop_pick_one<S, D>(P, O, M, sc, lane, cap1);
            if (lane == src) result = a;
        }
        __syncwarp();
        return result;
    }
}
