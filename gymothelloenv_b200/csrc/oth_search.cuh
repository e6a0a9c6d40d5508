// oth_search.cuh -- maximin of ANY depth up to OTH_MAX_SEARCH_DEPTH, with optional alpha-beta pruning.
//
// MaxiMinPolicy.search (simple_policies.py:111-163) visits every node; README.md:55-56 notes it "can be improved by
// pruning (not included)".  This is that pruning, built so that the ROOT CHOICE is bit-identical to the unpruned
// search: values are the searching side's disc count; the root takes the first argmax (np.argmax, :151-155), so
// root move j is searched with the window (best value so far, +inf) -- it returns its exact value whenever that
// beats the best so far and some value <= it otherwise (never chosen, as a tie would not be either); below the root
// the usual cut-offs apply (max node: value >= beta, min node: value <= alpha).  Ties below the root only affect
// which equal value travels up, never the value.
// One non-inlined function per remaining ply (static call chain: ptxas knows the stack size; no device recursion),
// all sharing non-inlined legal / flips helpers so a board size's code does not grow with the depth.
#pragma once
#include "oth_core.cuh"

constexpr int AB_INF = 0x3FFFFFFF;

template <int S>
__device__ __noinline__ BB<Geo<S>::W> legal_ni(const BB<Geo<S>::W> &pl, const BB<Geo<S>::W> &op) { return legal<S>(pl, op); }
template <int S>
__device__ __noinline__ BB<Geo<S>::W> flips_ni(int a, const BB<Geo<S>::W> &pl, const BB<Geo<S>::W> &op) {
    return flips_at<S>(a, bb_bit<Geo<S>::W>(a), pl, op);
}

// R = plies still to search below this node (>= 1); `pl` is to move; me_is_pl: the side to move is the searching
// side (a max node).  nodes += positions scored (every child of every expanded node: the work unit of the search).
template <int S, int R>
struct AlphaBeta {
    static __device__ __noinline__ int value(const BB<Geo<S>::W> &pl, const BB<Geo<S>::W> &op, bool me_is_pl, int alpha,
                                             int beta, bool prune, unsigned long long &nodes) {
        constexpr int W = Geo<S>::W;
        if constexpr (R == 1 && (W == 1 || OTH_MW_PLANES)) {
            // children are leaves: all of them are scored at once (max_flips) -- the exact value of the node, which
            // a cut-off could only have replaced by a bound on the same side of the window
            BB<W> mv;
            const int k = max_flips<S>(pl, op, mv);
            if (!any(mv)) return popc(me_is_pl ? pl : op);
            nodes += (unsigned long long)popc(mv);
            return me_is_pl ? popc(pl) + k + 1 : popc(op) - k;
        }
        BB<W> moves = legal_ni<S>(pl, op);
        if (!any(moves)) return popc(me_is_pl ? pl : op);        // no move: a leaf (simple_policies.py:117-126,139-144)
        const int pl_cnt = popc(pl), op_cnt = popc(op);
        int best = me_is_pl ? -1 : AB_INF;
        while (any(moves)) {
            const int a = lowest_bit(moves);
            clear_bit(moves, a);
            const BB<W> f = flips_ni<S>(a, pl, op);
            int v;
            nodes++;
            if constexpr (R == 1) {                              // children are leaves: count without expanding them
                const int k = popc(f);
                v = me_is_pl ? pl_cnt + k + 1 : op_cnt - k;
            } else {
                v = AlphaBeta<S, R - 1>::value(andnot(op, f), pl | f | bb_bit<W>(a), !me_is_pl, alpha, beta, prune, nodes);
            }
            if (me_is_pl) {
                best = max(best, v);
                if (prune) { if (best >= beta) break; alpha = max(alpha, best); }
            } else {
                best = min(best, v);
                if (prune) { if (best <= alpha) break; beta = min(beta, best); }
            }
        }
        return best;
    }
};

// value of the position after a root move: the OTHER side is to move, `rest` = depth - 1 plies remain
template <int S>
__device__ int search_below_root(const BB<Geo<S>::W> &pl, const BB<Geo<S>::W> &op, int rest, int alpha, bool prune,
                                 unsigned long long &nodes) {
    switch (rest) {
    case 1: return AlphaBeta<S, 1>::value(pl, op, false, alpha, AB_INF, prune, nodes);
    case 2: return AlphaBeta<S, 2>::value(pl, op, false, alpha, AB_INF, prune, nodes);
    case 3: return AlphaBeta<S, 3>::value(pl, op, false, alpha, AB_INF, prune, nodes);
    case 4: return AlphaBeta<S, 4>::value(pl, op, false, alpha, AB_INF, prune, nodes);
    case 5: return AlphaBeta<S, 5>::value(pl, op, false, alpha, AB_INF, prune, nodes);
    case 6: return AlphaBeta<S, 6>::value(pl, op, false, alpha, AB_INF, prune, nodes);
    default: return AlphaBeta<S, 7>::value(pl, op, false, alpha, AB_INF, prune, nodes);
    }
}

// MaxiMinPolicy.get_action at any depth 1..8: iterates the CACHED move list in ascending order, first argmax.
// One root move `a` (the work item of the lane-cooperative variant): its value for the window (alpha, +inf).
template <int S>
__device__ int search_root_move(const BB<Geo<S>::W> &own, const BB<Geo<S>::W> &opp, int a, int depth, int alpha, bool prune,
                                unsigned long long &nodes) {
    constexpr int W = Geo<S>::W;
    const BB<W> f = flips_ni<S>(a, own, opp);
    nodes++;
    if (depth <= 1) return popc(own) + popc(f) + 1;
    return search_below_root<S>(andnot(opp, f), own | f | bb_bit<W>(a), depth - 1, alpha, prune, nodes);
}
template <int S>
__device__ int search_pick(const BB<Geo<S>::W> &own, const BB<Geo<S>::W> &opp, BB<Geo<S>::W> moves, int depth, bool prune,
                           unsigned long long &nodes) {
    int best = -1, best_v = -1;
    while (any(moves)) {
        const int a = lowest_bit(moves);
        clear_bit(moves, a);
        const int v = search_root_move<S>(own, opp, a, depth, best_v, prune, nodes);
        if (v > best_v) { best_v = v; best = a; }
    }
    return best;
}

// policy.get_action for one position.  Maximin depth 2..4: the template-unrolled DFS of oth_core.cuh; deeper (and
// OTH_POLICY_ALPHABETA at any depth): the pruned search -- the SAME choice, never a silently shallower one.
// DEEP = false leaves the generic-depth search out of the caller (the tournament kernels for depths <= 4 must not
// pay its call chain in registers and stack); the launchers send deeper specs to DEEP = true instantiations only.
template <int S, bool DEEP>
__device__ int policy_pick(int kind, int depth, const BB<Geo<S>::W> &own, const BB<Geo<S>::W> &opp,
                           const BB<Geo<S>::W> &moves, u32 rnd) {
    if (kind == OTH_POLICY_RANDOM) return random_pick<S>(moves, rnd);
    if (kind == OTH_POLICY_GREEDY || depth <= 1) return greedy_pick<S>(own, opp, moves);   // README.md:54
    if (kind == OTH_POLICY_MAXIMIN || !DEEP) {
        if (depth == 2) return MaxiMin<S, 2>::pick(own, opp, moves);
        if (depth == 3) return MaxiMin<S, 3>::pick(own, opp, moves);
        if (depth == 4) return MaxiMin<S, 4>::pick(own, opp, moves);
    }
    if constexpr (DEEP) {
        unsigned long long nodes = 0;
        return search_pick<S>(own, opp, moves, depth, true, nodes);
    }
    return -1;      // a depth this instantiation does not search: never a silently shallower move (an invalid action)
}

#ifndef OTH_HOST_EMUL
// ---- lane-cooperative alpha-beta (single-word boards): a group of AB_GW lanes searches ONE position -------------
// Round A ("eldest brother"): the replies to the first root move are split over the lanes, each searched with a full
// window; their minimum is the first move's exact value v0.  Round B: the remaining root moves go one per lane with
// the window (best so far, +inf), AB_GW at a time; a lane's result is exact whenever it beats the bound, so the
// group's (max value, lowest move) reduction reproduces the sequential first argmax.
constexpr int AB_GW = 8;
template <int S>
__device__ int search_reply(u64 cpl, u64 cop, int b, int depth, bool prune, unsigned long long &nodes) {
    // position after root move: cpl (the opponent) is to move and plays b; `depth - 2` plies remain below
    BB<1> pl, op;
    pl.w[0] = cpl; op.w[0] = cop;
    const BB<1> g = flips_ni<S>(b, pl, op);
    nodes++;
    if (depth <= 2) return __popcll(cop) - popc(g);
    BB<1> npl, nop;                                              // the searching side moves again
    npl.w[0] = cop & ~g.w[0]; nop.w[0] = cpl | g.w[0] | (1ull << b);
    switch (depth - 2) {
    case 1: return AlphaBeta<S, 1>::value(npl, nop, true, -1, AB_INF, prune, nodes);
    case 2: return AlphaBeta<S, 2>::value(npl, nop, true, -1, AB_INF, prune, nodes);
    case 3: return AlphaBeta<S, 3>::value(npl, nop, true, -1, AB_INF, prune, nodes);
    case 4: return AlphaBeta<S, 4>::value(npl, nop, true, -1, AB_INF, prune, nodes);
    case 5: return AlphaBeta<S, 5>::value(npl, nop, true, -1, AB_INF, prune, nodes);
    default: return AlphaBeta<S, 6>::value(npl, nop, true, -1, AB_INF, prune, nodes);
    }
}
OTH_DEV u64 ab_nth_lowest(u64 m, int k) {                        // the k-th lowest set bit as a mask, 0 if none
    for (; k > 0 && m; k--) m &= m - 1ull;
    return m & (0ull - m);
}
// every lane passes its own position; lanes with need = true get their action (and node count) back
template <int S>
__device__ __noinline__ int coop_search_pick(bool need, u64 own, u64 opp, u64 moves, int depth, bool prune,
                                             unsigned long long &nodes_out) {
    static_assert(Geo<S>::W == 1, "cooperative search: single-word boards");
    const int lane = threadIdx.x & 31, gl = lane & (AB_GW - 1), gbase = lane & ~(AB_GW - 1);
    const u32 gmask = ((1u << AB_GW) - 1u) << gbase;
    u32 todo = (__ballot_sync(0xFFFFFFFFu, need) >> gbase) & ((1u << AB_GW) - 1u);
    int result = -1;
    while (todo != 0u) {                                         // group-uniform
        const int src = gbase + __ffs(todo) - 1;
        todo &= todo - 1u;
        const u64 P = __shfl_sync(gmask, own, src), O = __shfl_sync(gmask, opp, src), M = __shfl_sync(gmask, moves, src);
        unsigned long long nodes = 0;
        BB<1> bp, bo;
        bp.w[0] = P; bo.w[0] = O;
        // ---- round A: the first root move
        const int a0 = __ffsll((long long)M) - 1;
        const u64 f0 = flips_ni<S>(a0, bp, bo).w[0];
        if (gl == 0) nodes++;
        int best = a0, best_v;
        if (depth <= 1) {
            best_v = __popcll(P) + __popcll(f0) + 1;
        } else {
            const u64 cpl = O & ~f0, cop = P | f0 | (1ull << a0);
            BB<1> c1, c2;
            c1.w[0] = cpl; c2.w[0] = cop;
            const u64 L = legal_ni<S>(c1, c2).w[0];
            if (L == 0ull) {
                best_v = __popcll(cop);
            } else {
                int local = AB_INF;
                u64 rest = L;
                while (rest != 0ull) {                           // group-uniform trip count
                    const u64 mine = ab_nth_lowest(rest, gl);
                    if (mine) local = min(local, search_reply<S>(cpl, cop, __ffsll((long long)mine) - 1, depth, prune, nodes));
                    for (int k = 0; k < AB_GW && rest; k++) rest &= rest - 1ull;
                }
#pragma unroll
                for (int o = AB_GW / 2; o > 0; o >>= 1) local = min(local, __shfl_xor_sync(gmask, local, o, AB_GW));
                best_v = local;
            }
        }
        // ---- round B: the other root moves, AB_GW at a time, window (best so far, +inf)
        u64 rest = M & (M - 1ull);
        while (rest != 0ull) {
            const u64 mine = ab_nth_lowest(rest, gl);
            int v = -1, a = 0x7FFF;
            if (mine) {
                a = __ffsll((long long)mine) - 1;
                v = search_root_move<S>(bp, bo, a, depth, best_v, prune, nodes);
            }
#pragma unroll
            for (int o = AB_GW / 2; o > 0; o >>= 1) {
                const int ov = __shfl_xor_sync(gmask, v, o, AB_GW), oa = __shfl_xor_sync(gmask, a, o, AB_GW);
                if (ov > v || (ov == v && oa < a)) { v = ov; a = oa; }
            }
            if (v > best_v) { best_v = v; best = a; }
            for (int k = 0; k < AB_GW && rest; k++) rest &= rest - 1ull;
        }
        // node total of the group
#pragma unroll
        for (int o = AB_GW / 2; o > 0; o >>= 1) nodes += __shfl_xor_sync(gmask, nodes, o, AB_GW);
        if (lane == src) { result = best; nodes_out = nodes; }
    }
    __syncwarp();
    return result;
}
#endif
