// oth_coop_search.cuh -- lane-cooperative maximin search for single-word boards (S <= 8).
//
// MaxiMinPolicy.search (simple_policies.py:111-163) as a per-thread DFS wastes most of a warp: every
// nested `for move in possible_moves` loop runs to the LONGEST trip count among the 32 games of the warp
// (measured on B200: 10 of 32 lanes active per instruction for maximin-3, 14 for maximin-2).  Here a group
// of GW lanes searches ONE game at a time, level by level:
//   level 1   one item per root move: apply it, compute the reply set            -> node list 1
//   level 2   (depth >= 3) one item per (root move, reply) pair: same             -> node list 2
//   leaves    one item per move of every node of the deepest list: flips + popcount, folded into the
//             node's value with an order-independent min / max
// Items of a level are split into equal CONSECUTIVE chunks over the lanes (one binary search per lane to
// find its first node, then a walk), so every lane does the same number of rules evaluations.
// Values only travel upwards through min / max, so the result does not depend on the order of evaluation;
// the one place where ties matter is the root (np.argmax first occurrence, :151-155), reduced with
// (value, lowest index).  Depth 4 = the root's moves in sequence, each searched to depth 3 cooperatively.
// More (root move, reply) pairs than list 2 holds: the root moves are processed in batches that fit.  More than
// COOP_CAP1 root moves (never seen on 8x8): that decision falls back to the per-thread DFS;
// OTH_F_SEARCH_SMALL_LISTS shrinks both limits so that the tests run the batching and the fallback constantly.
#pragma once
#include "oth_core.cuh"

constexpr int COOP_CAP1 = 32;
constexpr int COOP_CAP2 = 256;
constexpr int COOP_INF = 0x3FFFFFFF;

struct CoopLevel1 {
    u64 pl[COOP_CAP1], op[COOP_CAP1], mv[COOP_CAP1];   // side to move, other side, its legal moves
    int val[COOP_CAP1];
    short first[COOP_CAP1 + 1];                        // exclusive prefix sums of popc(mv)
};
struct CoopLevel2 {
    u64 pl[COOP_CAP2], op[COOP_CAP2], mv[COOP_CAP2];
    int val[COOP_CAP2];
    short first[COOP_CAP2 + 1];
};
// per warp: either one 32-lane group with both lists (depth >= 3) or 32 / COOP_GW2 narrow groups with list 1 only
#ifndef COOP_GW2
#define COOP_GW2 8          // lanes per game for depth 2 (4: rollouts +0.6 %, batched policy -4 %; profiles/r01p_policy.md)
#endif
union CoopScratch {
    struct { CoopLevel1 l1; CoopLevel2 l2; } wide;
    CoopLevel1 narrow[32 / COOP_GW2];
};

template <int GW> OTH_DEV u32 coop_group_mask(int lane) {
    if constexpr (GW == 32) return 0xFFFFFFFFu;
    else return ((1u << GW) - 1u) << (lane & ~(GW - 1));
}
// exclusive prefix sum of x over the GW lanes of a group, and the group total
template <int GW> OTH_DEV int coop_excl_scan(int x, int gl, u32 gmask, int &total) {
    int incl = x;
#pragma unroll
    for (int o = 1; o < GW; o <<= 1) {
        const int t = __shfl_up_sync(gmask, incl, o, GW);
        if (gl >= o) incl += t;
    }
    total = __shfl_sync(gmask, incl, GW - 1, GW);
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

// Expands the moves of `n_src` parent nodes (arrays spl/sop/smv, item prefix sfirst, n_items items) into the child
// list (dpl/dop/dmv/dval/dfirst): child q = the q-th (parent, move) pair in (parent, ascending move) order.
// child_is_me: the side to move at the children is the searching side.  Returns the total number of moves of all
// children (the next level's item count).
// The parents may be a BATCH of a longer list: sfirst keeps the list's absolute item numbering and item_base is the
// absolute number of the batch's first item (= sfirst[0]); children are stored from index 0.
template <int S, int GW>
OTH_DEV int coop_expand(const u64 *spl, const u64 *sop, const u64 *smv, const short *sfirst, int n_src, int n_items,
                        int item_base, u64 *dpl, u64 *dop, u64 *dmv, int *dval, short *dfirst, bool child_is_me, int gl,
                        u32 gmask) {
    const int chunk = (n_items + GW - 1) / GW;
    const int q0 = min(gl * chunk, n_items), q1 = min(q0 + chunk, n_items);
    int local = 0;
    if (q0 < q1) {
        int p = coop_find_node(sfirst, n_src, q0 + item_base);
        u64 pl = spl[p], op = sop[p];
        u64 m = coop_drop_lowest(smv[p], q0 + item_base - sfirst[p]);
        for (int q = q0; q < q1; q++) {
            while (m == 0ull) { p++; pl = spl[p]; op = sop[p]; m = smv[p]; }
            const u64 mv = m & (0ull - m);
            m &= m - 1ull;
            const u64 f = flips1<S>(mv, pl, op);
            const u64 cpl = op & ~f, cop = pl | f | mv;          // the other side moves next
            const u64 L = legal1<S>(cpl, cop);
            dpl[q] = cpl; dop[q] = cop; dmv[q] = L;
            // a side with no move makes the node a leaf (simple_policies.py:117-126,139-144)
            dval[q] = (L == 0ull) ? __popcll(child_is_me ? cpl : cop) : (child_is_me ? -1 : COOP_INF);
            local += __popcll(L);
        }
    }
    int total;
    int run = coop_excl_scan<GW>(local, gl, gmask, total);
    for (int q = q0; q < q1; q++) { dfirst[q] = (short)run; run += __popcll(dmv[q]); }
    if (gl == 0) dfirst[n_items] = (short)total;
    __syncwarp(gmask);
    return total;
}

// Scores every move of every node of a list by popcount (the depth-1 rule of MaxiMin::value) and folds the scores
// into val[node]: max when the side to move there is the searching side, else min.
template <int S, int GW>
OTH_DEV void coop_leaves(const u64 *npl, const u64 *nop, const u64 *nmv, int *nval, const short *nfirst, int n_nodes,
                         int n_items, bool mover_is_me, int gl, u32 gmask) {
    const int chunk = (n_items + GW - 1) / GW;
    const int q0 = min(gl * chunk, n_items), q1 = min(q0 + chunk, n_items);
    if (q0 < q1) {
        int p = coop_find_node(nfirst, n_nodes, q0);
        u64 pl = npl[p], op = nop[p];
        u64 m = coop_drop_lowest(nmv[p], q0 - nfirst[p]);
        int base = mover_is_me ? __popcll(pl) + 1 : __popcll(op);
        int best = mover_is_me ? -1 : COOP_INF;
        for (int q = q0; q < q1; q++) {
            while (m == 0ull) {
                if (mover_is_me) atomicMax(&nval[p], best); else atomicMin(&nval[p], best);
                p++; pl = npl[p]; op = nop[p]; m = nmv[p];
                base = mover_is_me ? __popcll(pl) + 1 : __popcll(op);
                best = mover_is_me ? -1 : COOP_INF;
            }
            const u64 mv = m & (0ull - m);
            m &= m - 1ull;
            const int k = __popcll(flips1<S>(mv, pl, op));
            best = mover_is_me ? max(best, base + k) : min(best, base - k);
        }
        if (mover_is_me) atomicMax(&nval[p], best); else atomicMin(&nval[p], best);
    }
    __syncwarp(gmask);
}

// Searches DEPTH (2 or 3) plies below a node whose side to move (pl) has the non-empty move set M.
// top_is_me: that side is the searching side.  Returns the node's value; *best_index = index (in ascending
// move order) of the first move achieving it.  Group-uniform arguments and results.
// ok = false: capacity overflow, nothing computed.
template <int S, int DEPTH, int GW>
OTH_DEV int coop_search(u64 P, u64 O, u64 M, bool top_is_me, CoopLevel1 &l1, CoopLevel2 *l2, int gl, u32 gmask,
                        int *best_index, bool &ok, int cap1, int cap2) {
    static_assert(DEPTH == 2 || DEPTH == 3, "coop_search depth");
    const int n1 = __popcll(M);
    ok = n1 <= cap1;
    if (!ok) return 0;
    // level 1: the parent "list" is the single top node, held in registers, so it is expanded inline
    {
        const int chunk = (n1 + GW - 1) / GW;
        const int q0 = min(gl * chunk, n1), q1 = min(q0 + chunk, n1);
        int local = 0;
        u64 m = coop_drop_lowest(M, q0);
        const bool child_is_me = !top_is_me;
        for (int q = q0; q < q1; q++) {
            const u64 mv = m & (0ull - m);
            m &= m - 1ull;
            const u64 f = flips1<S>(mv, P, O);
            const u64 cpl = O & ~f, cop = P | f | mv;
            const u64 L = legal1<S>(cpl, cop);
            l1.pl[q] = cpl; l1.op[q] = cop; l1.mv[q] = L;
            l1.val[q] = (L == 0ull) ? __popcll(child_is_me ? cpl : cop) : (child_is_me ? -1 : COOP_INF);
            local += __popcll(L);
        }
        int total;
        int run = coop_excl_scan<GW>(local, gl, gmask, total);
        for (int q = q0; q < q1; q++) { l1.first[q] = (short)run; run += __popcll(l1.mv[q]); }
        if (gl == 0) l1.first[n1] = (short)total;
        __syncwarp(gmask);
    }
    const int n2 = l1.first[n1];
    if constexpr (DEPTH == 2) {
        coop_leaves<S, GW>(l1.pl, l1.op, l1.mv, l1.val, l1.first, n1, n2, !top_is_me, gl, gmask);
    } else {
        // list 2 holds cap2 pairs: level-1 nodes are taken in batches [j0, j1) of consecutive root moves whose
        // replies fit (one node has at most 64 replies, so with the real capacity a batch is never empty)
        int j0 = 0;
        while (j0 < n1) {                                            // group-uniform
            const int base = l1.first[j0];
            int j1 = j0 + 1;
            while (j1 < n1 && l1.first[j1 + 1] - base <= cap2) j1++;
            const int items = l1.first[j1] - base;
            ok = items <= cap2;                                      // only the shrunken test capacity can fail here
            if (!ok) return 0;
            if (items > 0) {
                const int n3 = coop_expand<S, GW>(l1.pl + j0, l1.op + j0, l1.mv + j0, l1.first + j0, j1 - j0, items, base,
                                                  l2->pl, l2->op, l2->mv, l2->val, l2->first, top_is_me, gl, gmask);
                coop_leaves<S, GW>(l2->pl, l2->op, l2->mv, l2->val, l2->first, items, n3, top_is_me, gl, gmask);
                // fold list 2 into list 1: the side to move at a level-1 node is the other side
                for (int j = j0 + gl; j < j1; j += GW) {
                    const int a = l1.first[j] - base, b = l1.first[j + 1] - base;
                    if (a < b) {
                        int v = top_is_me ? COOP_INF : -1;
                        for (int p = a; p < b; p++) v = top_is_me ? min(v, l2->val[p]) : max(v, l2->val[p]);
                        l1.val[j] = v;
                    }
                }
                __syncwarp(gmask);
            }
            j0 = j1;
        }
    }
    // top: max (min) over the root moves, first occurrence on ties
    int bv = top_is_me ? -1 : COOP_INF, bi = 0x7FFF;
    for (int j = gl; j < n1; j += GW) {
        const int v = l1.val[j];
        if (top_is_me ? (v > bv) : (v < bv)) { bv = v; bi = j; }
    }
#pragma unroll
    for (int o = GW / 2; o > 0; o >>= 1) {
        const int ov = __shfl_xor_sync(gmask, bv, o, GW), oi = __shfl_xor_sync(gmask, bi, o, GW);
        const bool better = top_is_me ? (ov > bv) : (ov < bv);
        if (better || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
    }
    __syncwarp(gmask);                                               // lists are reused by the next game
    if (best_index) *best_index = bi;
    return bv;
}

// MaxiMinPolicy.get_action for one game by a group of GW lanes (group-uniform arguments): the chosen action.
template <int S, int D, int GW>
OTH_DEV int coop_pick_one(u64 own, u64 opp, u64 moves, CoopLevel1 &l1, CoopLevel2 *l2, int gl, u32 gmask, int cap1,
                          int cap2) {
    bool ok = true;
    int a = -1;
    if constexpr (D == 2 || D == 3) {
        int bi = 0;
        coop_search<S, D, GW>(own, opp, moves, true, l1, l2, gl, gmask, &bi, ok, cap1, cap2);
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
            else v = coop_search<S, 3, GW>(cpl, cop, L, false, l1, l2, gl, gmask, nullptr, ok, cap1, cap2);
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
    constexpr int GW = (D == 2) ? COOP_GW2 : 32;
    const int lane = threadIdx.x & 31, gl = lane & (GW - 1), gbase = lane & ~(GW - 1);
    const u32 gmask = coop_group_mask<GW>(lane);
    const u32 needing = __ballot_sync(0xFFFFFFFFu, need);
    u32 todo = (GW == 32) ? needing : ((needing >> gbase) & ((1u << GW) - 1u));
    CoopLevel1 &l1 = (GW == 32) ? sc.wide.l1 : sc.narrow[(lane / GW) % (32 / COOP_GW2)];
    CoopLevel2 *l2 = (GW == 32) ? &sc.wide.l2 : nullptr;
    const int cap1 = small_lists ? 3 : COOP_CAP1, cap2 = small_lists ? 12 : COOP_CAP2;   // OTH_F_SEARCH_SMALL_LISTS
    int result = -1;
    while (todo != 0u) {                                             // group-uniform
        const int src = gbase + __ffs(todo) - 1;
        todo &= todo - 1u;
        const u64 P = __shfl_sync(gmask, own, src), O = __shfl_sync(gmask, opp, src), M = __shfl_sync(gmask, moves, src);
        const int a = coop_pick_one<S, D, GW>(P, O, M, l1, l2, gl, gmask, cap1, cap2);
        if (lane == src) result = a;
    }
    __syncwarp();
    return result;
}
