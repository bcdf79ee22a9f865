// go_bb.cuh -- warp-per-position bitboard Go engine (rules, feature planes, ladder reader).
//
// Replaces, for the device, the group/liberty bookkeeping of the reference's
// AlphaGo/go.py (GameState, go.py:9-581) and the feature functions of
// AlphaGo/preprocessing/preprocessing.py:14-205.  It is NOT a translation: the
// reference keeps Python sets per point and deep-copies the state for every
// look-ahead; here a position is two 19-row bitboards spread over the lanes of
// one warp (lane r owns row x = r, bit y of its word is point (x, y)), groups
// and liberties are fixed points of shift/shuffle dilations, and look-ahead
// copies are register moves.
//
// All control flow in this file is WARP-UNIFORM: every branch depends only on
// values that are identical in all 32 lanes (ballots, reductions, broadcast
// scalars).  That is what lets the same source also be compiled as plain C++
// with a 32-row array standing in for the warp (RAG_HOSTSIM; used only by
// tests/hostsim to check this file's logic against the oracle on machines
// without a GPU -- it is never linked into the product library).
//
// Point index inside this file: q = x*32 + y  ("q-index"; x = q >> 5, y = q & 31).
// The reference's flat index p = x*S + y (util.py:11-13) is used only at the I/O edge.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__) && !defined(RAG_HOSTSIM)
#define RAG_DEVICE_BUILD 1
#define RAG_FN __device__ __forceinline__
#define RAG_FN_NOINLINE __device__ __noinline__
#else
#define RAG_DEVICE_BUILD 0
#define RAG_FN static inline
#define RAG_FN_NOINLINE static
#endif

namespace rag {

constexpr int kMaxBoard = 19;
constexpr int kLadderDepth = 80;          // remaining_attempts default, go.py:329,396
constexpr unsigned kFullMask = 0xffffffffu;

// Canonical plane numbering (DEFAULT_FEATURES order, preprocessing.py:260-263, then `legal`).
enum Plane {
    kPlaneBoard = 0, kPlaneOnes = 3, kPlaneTurnsSince = 4, kPlaneLiberties = 12,
    kPlaneCaptureSize = 20, kPlaneSelfAtari = 28, kPlaneLibertiesAfter = 36,
    kPlaneLadderCapture = 44, kPlaneLadderEscape = 45, kPlaneSensibleness = 46,
    kPlaneZeros = 47, kPlaneLegal = 48,
    kPlaneColor = 49,      // not a reference feature: all ones iff BLACK is to move (value net's 49th plane)
    kNumPlanes = 50
};

// ------------------------------------------------------------------ BB: a board-sized bit set
#if RAG_DEVICE_BUILD

struct BB { uint32_t w; };                       // this lane's row

struct Geo {                                     // per-warp constants
    int S;                                       // board size (<= 19)
    uint32_t lane;                               // threadIdx.x & 31
    uint32_t rowmask;                            // (1<<S)-1 on lanes < S, else 0
};

RAG_FN Geo make_geo(int S) {
    Geo g; g.S = S; g.lane = threadIdx.x & 31u;
    g.rowmask = (g.lane < (uint32_t)S) ? ((1u << S) - 1u) : 0u;
    return g;
}

RAG_FN BB bb_zero() { return BB{0u}; }
RAG_FN BB bb_board(const Geo& g) { return BB{g.rowmask}; }
RAG_FN BB operator&(BB a, BB b) { return BB{a.w & b.w}; }
RAG_FN BB operator|(BB a, BB b) { return BB{a.w | b.w}; }
RAG_FN BB operator^(BB a, BB b) { return BB{a.w ^ b.w}; }
RAG_FN BB andnot(BB a, BB b) { return BB{a.w & ~b.w}; }             // a & ~b
RAG_FN bool bb_any(BB a) { return __any_sync(kFullMask, a.w != 0u) != 0; }
RAG_FN bool bb_equal(BB a, BB b) { return __all_sync(kFullMask, a.w == b.w) != 0; }
RAG_FN int bb_count(BB a) { return (int)__reduce_add_sync(kFullMask, (unsigned)__popc(a.w)); }
// several "is this set non-empty?" questions answered by ONE warp reduction: bit i of the result is bb_any(x_i)
RAG_FN unsigned bb_any4(BB a, BB b, BB c, BB d) {
    return __reduce_or_sync(kFullMask, (a.w ? 1u : 0u) | (b.w ? 2u : 0u) | (c.w ? 4u : 0u) | (d.w ? 8u : 0u));
}
RAG_FN BB bb_single(int q, const Geo& g) { return BB{(g.lane == (uint32_t)(q >> 5)) ? (1u << (q & 31)) : 0u}; }
RAG_FN bool bb_test(BB a, int q) { return ((__shfl_sync(kFullMask, a.w, q >> 5) >> (q & 31)) & 1u) != 0u; }
// lowest point in x-major order, or -1
RAG_FN int bb_first(BB a) {
    unsigned rows = __ballot_sync(kFullMask, a.w != 0u);
    if (rows == 0u) return -1;
    int x = __ffs(rows) - 1;
    uint32_t w = __shfl_sync(kFullMask, a.w, x);
    return (x << 5) | (__ffs(w) - 1);
}
// r-th (0-based) point of the set in x-major order; r must be < bb_count(a)
RAG_FN int bb_select(BB a, int r) {
    unsigned c = (unsigned)__popc(a.w), incl = c;
    const unsigned lane = threadIdx.x & 31u;
    #pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        unsigned t = __shfl_up_sync(kFullMask, incl, o);
        if (lane >= (unsigned)o) incl += t;
    }
    unsigned excl = incl - c;
    bool mine = (unsigned)r >= excl && (unsigned)r < incl;
    int x = __ffs(__ballot_sync(kFullMask, mine)) - 1;
    uint32_t w = __shfl_sync(kFullMask, a.w, x);
    unsigned k = (unsigned)r - __shfl_sync(kFullMask, excl, x);
    return (x << 5) | (int)__fns(w, 0, (int)k + 1);
}
// 4-neighbourhood of a set (without the set itself); go.py:103-116
RAG_FN BB bb_nbrs(BB a, const Geo& g) {
    uint32_t up = __shfl_up_sync(kFullMask, a.w, 1);        // row x-1
    uint32_t dn = __shfl_down_sync(kFullMask, a.w, 1);      // row x+1
    if (g.lane == 0u) up = 0u;
    if (g.lane == 31u) dn = 0u;
    return BB{(up | dn | (a.w << 1) | (a.w >> 1)) & g.rowmask};
}
// value at (x,y) taken from (x-1,y) / (x+1,y); rows outside the board read as `fill`
RAG_FN BB bb_from_north(BB a, const Geo& g, bool fill) {
    uint32_t v = __shfl_up_sync(kFullMask, a.w, 1);
    if (g.lane == 0u) v = fill ? kFullMask : 0u;
    return BB{v & g.rowmask};
}
RAG_FN BB bb_from_south(BB a, const Geo& g, bool fill) {
    uint32_t v = __shfl_down_sync(kFullMask, a.w, 1);
    if (g.lane == (uint32_t)(g.S - 1)) v = fill ? kFullMask : 0u;
    return BB{v & g.rowmask};
}
RAG_FN BB bb_from_west(BB a, const Geo& g, bool fill) { return BB{((a.w << 1) | (fill ? 1u : 0u)) & g.rowmask}; }
RAG_FN BB bb_from_east(BB a, const Geo& g, bool fill) {
    return BB{((a.w >> 1) | (fill ? (1u << (g.S - 1)) : 0u)) & g.rowmask};
}
// planes: 32 words per plane, word r = row r
RAG_FN void plane_or(uint32_t* planes, int k, BB a, const Geo& g) { planes[k * 32 + g.lane] |= a.w; }
RAG_FN BB bb_load(const uint32_t* rows32, const Geo& g) { return BB{rows32[g.lane] & g.rowmask}; }
RAG_FN void bb_store(uint32_t* rows32, BB a, const Geo& g) { rows32[g.lane] = a.w; }

#else  // ------------------------------------------------------------- host stand-in for one warp

struct BB { uint32_t w[32]; };
struct Geo { int S; uint32_t rowmask; };
// host-only operation counters (tools/ladder_study.py): how many warp collectives would the device issue?
struct OpCounts { long long nbrs, any, count, first, test, flood_calls, flood_iters, from_dir; };
static OpCounts g_ops = {0, 0, 0, 0, 0, 0, 0, 0};

RAG_FN Geo make_geo(int S) { Geo g; g.S = S; g.rowmask = (1u << S) - 1u; return g; }
RAG_FN uint32_t rm(const Geo& g, int r) { return r < g.S ? g.rowmask : 0u; }
RAG_FN BB bb_zero() { BB r; for (int i = 0; i < 32; i++) r.w[i] = 0; return r; }
RAG_FN BB bb_board(const Geo& g) { BB r; for (int i = 0; i < 32; i++) r.w[i] = rm(g, i); return r; }
RAG_FN BB operator&(BB a, BB b) { for (int i = 0; i < 32; i++) a.w[i] &= b.w[i]; return a; }
RAG_FN BB operator|(BB a, BB b) { for (int i = 0; i < 32; i++) a.w[i] |= b.w[i]; return a; }
RAG_FN BB operator^(BB a, BB b) { for (int i = 0; i < 32; i++) a.w[i] ^= b.w[i]; return a; }
RAG_FN BB andnot(BB a, BB b) { for (int i = 0; i < 32; i++) a.w[i] &= ~b.w[i]; return a; }
RAG_FN bool bb_any(BB a) { g_ops.any++; for (int i = 0; i < 32; i++) if (a.w[i]) return true; return false; }
RAG_FN bool bb_equal(BB a, BB b) { g_ops.any++; for (int i = 0; i < 32; i++) if (a.w[i] != b.w[i]) return false; return true; }
RAG_FN int bb_count(BB a) { g_ops.count++; int c = 0; for (int i = 0; i < 32; i++) c += __builtin_popcount(a.w[i]); return c; }
RAG_FN unsigned bb_any4(BB a, BB b, BB c, BB d) {
    return (bb_any(a) ? 1u : 0u) | (bb_any(b) ? 2u : 0u) | (bb_any(c) ? 4u : 0u) | (bb_any(d) ? 8u : 0u);
}
RAG_FN BB bb_single(int q, const Geo&) { BB r = bb_zero(); r.w[q >> 5] = 1u << (q & 31); return r; }
RAG_FN bool bb_test(BB a, int q) { g_ops.test++; return (a.w[q >> 5] >> (q & 31)) & 1u; }
RAG_FN int bb_first(BB a) {
    g_ops.first++;
    for (int i = 0; i < 32; i++) if (a.w[i]) return (i << 5) | __builtin_ctz(a.w[i]);
    return -1;
}
RAG_FN int bb_select(BB a, int r) {
    for (int i = 0; i < 32; i++) {
        int c = __builtin_popcount(a.w[i]);
        if (r < c) { uint32_t w = a.w[i]; while (r--) w &= w - 1; return (i << 5) | __builtin_ctz(w); }
        r -= c;
    }
    return -1;
}
RAG_FN BB bb_nbrs(BB a, const Geo& g) {
    BB r;
    g_ops.nbrs++;
    for (int i = 0; i < 32; i++) {
        uint32_t up = i > 0 ? a.w[i - 1] : 0u, dn = i < 31 ? a.w[i + 1] : 0u;
        r.w[i] = (up | dn | (a.w[i] << 1) | (a.w[i] >> 1)) & rm(g, i);
    }
    return r;
}
RAG_FN BB bb_from_north(BB a, const Geo& g, bool fill) {
    g_ops.from_dir++;
    BB r; for (int i = 0; i < 32; i++) r.w[i] = (i > 0 ? a.w[i - 1] : (fill ? kFullMask : 0u)) & rm(g, i); return r;
}
RAG_FN BB bb_from_south(BB a, const Geo& g, bool fill) {
    BB r;
    for (int i = 0; i < 32; i++) r.w[i] = ((i == g.S - 1) ? (fill ? kFullMask : 0u) : (i < 31 ? a.w[i + 1] : 0u)) & rm(g, i);
    return r;
}
RAG_FN BB bb_from_west(BB a, const Geo& g, bool fill) {
    BB r; for (int i = 0; i < 32; i++) r.w[i] = ((a.w[i] << 1) | (fill ? 1u : 0u)) & rm(g, i); return r;
}
RAG_FN BB bb_from_east(BB a, const Geo& g, bool fill) {
    BB r; for (int i = 0; i < 32; i++) r.w[i] = ((a.w[i] >> 1) | (fill ? (1u << (g.S - 1)) : 0u)) & rm(g, i); return r;
}
RAG_FN void plane_or(uint32_t* planes, int k, BB a, const Geo&) { for (int i = 0; i < 32; i++) planes[k * 32 + i] |= a.w[i]; }
RAG_FN BB bb_load(const uint32_t* rows32, const Geo& g) { BB r; for (int i = 0; i < 32; i++) r.w[i] = rows32[i] & rm(g, i); return r; }
RAG_FN void bb_store(uint32_t* rows32, BB a, const Geo&) { for (int i = 0; i < 32; i++) rows32[i] = a.w[i]; }

#endif

// ------------------------------------------------------------------ groups and liberties

// Connected component(s) of `mask` containing `seed` (seed must lie inside mask).
// Device-side twin of group_sets[x][y] (go.py:49,125-157): a fixed point of dilation.
RAG_FN BB bb_flood(BB seed, BB mask, const Geo& g) {
    BB cur = seed;
#if !RAG_DEVICE_BUILD
    g_ops.flood_calls++;
#endif
    for (;;) {
#if !RAG_DEVICE_BUILD
        g_ops.flood_iters++;
#endif
        BB nx = (cur | bb_nbrs(cur, g)) & mask;
        nx = (nx | bb_nbrs(nx, g)) & mask;
        if (bb_equal(nx, cur)) return cur;
        cur = nx;
    }
}

RAG_FN int q_neighbor(int q, int dir, int S) {
    // reference order (x-1,y),(x+1,y),(x,y-1),(x,y+1) (go.py:108); -1 when off board
    int x = q >> 5, y = q & 31;
    if (dir == 0) return x > 0 ? q - 32 : -1;
    if (dir == 1) return x + 1 < S ? q + 32 : -1;
    if (dir == 2) return y > 0 ? q - 1 : -1;
    return y + 1 < S ? q + 1 : -1;
}

RAG_FN int q_diagonal(int q, int i, int S) {
    // reference order (x-1,y-1),(x+1,y+1),(x+1,y-1),(x-1,y+1) (go.py:122-123)
    int x = (q >> 5) + ((i == 1 || i == 2) ? 1 : -1);
    int y = (q & 31) + ((i == 1 || i == 3) ? 1 : -1);
    if (x < 0 || y < 0 || x >= S || y >= S) return -1;
    return (x << 5) | y;
}

// ------------------------------------------------------------------ do_move  (go.py:521-581)

// Mover `M` plays q against `O`.  Returns false (nothing changed) when the move is not
// legal by go.py:267-285 (occupied, ko, suicide; superko is not enforced on this path).
// Captures are resolved in the reference's neighbour order so that the ko point, including
// the "stale ko" of SURVEY H3, comes out the same (go.py:542-561).
RAG_FN bool bb_play_ex(const Geo& g, BB& M, BB& O, int& ko, int q, int* captured, BB& ptn_out) {
    if (q == ko) return false;
    const BB pt = bb_single(q, g);
    const BB M2 = M | pt;
    BB O2 = O;
    int newko = -1, ncap = 0;
    const BB ptn = bb_nbrs(pt, g);
    ptn_out = ptn;
    const BB empty2 = andnot(bb_board(g), M2 | O);
    // Only an enemy neighbour with no empty point next to that very stone can be captured;
    // most moves have none, and then no group needs to be traced at all.
    const BB maybe = andnot(ptn & O, bb_nbrs(empty2, g));
    // occupied? / capture candidates? / an empty neighbour? / a friendly neighbour?  -- one reduction
    const unsigned f = bb_any4((M | O) & pt, maybe, ptn & empty2, ptn & M);
    if (f & 1u) return false;
    if (f & 2u) {
        for (int d = 0; d < 4; d++) {
            int nq = q_neighbor(q, d, g.S);
            if (nq < 0 || !bb_test(maybe & O2, nq)) continue;
            BB grp = bb_flood(bb_single(nq, g), O2, g);
            BB empty = andnot(bb_board(g), M2 | O2);
            if (bb_any(bb_nbrs(grp, g) & empty)) continue;
            O2 = andnot(O2, grp);                                   // _remove_group, go.py:163-184
            int sz = bb_count(grp);
            ncap += sz;
            if (sz == 1) {                                          // go.py:553-561
                // recapture would take exactly the played stone: it is alone and its only
                // liberty is the point just emptied
                BB e2 = andnot(bb_board(g), M2 | O2);
                if (!(f & 8u) && bb_count(ptn & e2) == 1) newko = nq;
            }
        }
    }
    // suicide (go.py:219-240): a capture or an empty neighbour leaves the stone a liberty;
    // otherwise the group it joined must have one somewhere
    if (ncap == 0 && !(f & 4u)) {
        BB me = bb_flood(pt, M2, g);
        if (!bb_any(bb_nbrs(me, g) & empty2)) return false;
    }
    M = M2; O = O2; ko = newko;
    if (captured) *captured = ncap;
    return true;
}

RAG_FN bool bb_play(const Geo& g, BB& M, BB& O, int& ko, int q, int* captured) {
    BB ptn;
    return bb_play_ex(g, M, O, ko, q, captured, ptn);
}

// ------------------------------------------------------------------ eyes  (go.py:287-327)

struct EyeStack { uint16_t pos[kMaxBoard * kMaxBoard]; uint8_t idx[kMaxBoard * kMaxBoard]; uint8_t bad[kMaxBoard * kMaxBoard]; };

RAG_FN bool q_interior(int q, int S) { int x = q >> 5, y = q & 31; return x > 0 && y > 0 && x + 1 < S && y + 1 < S; }

// is_eye(q, owner) with owner's stones in `own`; `eyeish` = empty points whose 4-neighbours are
// all `own` (go.py:287-297), precomputed bit-parallel.  Exact DFS incl. the ancestor-stack rule (H6).
RAG_FN_NOINLINE bool bb_is_eye(const Geo& g, BB opp, BB empty, BB eyeish, int q, EyeStack& st) {
    if (!bb_test(eyeish, q)) return false;
    int depth = 0;
    st.pos[0] = (uint16_t)q; st.idx[0] = 0; st.bad[0] = 0;
    for (;;) {
        int p = st.pos[depth];
        int allow = q_interior(p, g.S) ? 1 : 0;
        bool ret_valid = false, ret = false;
        if (st.idx[depth] == 4) { ret_valid = true; ret = true; }
        else {
            int d = q_diagonal(p, st.idx[depth]++, g.S);
            if (d < 0) continue;
            bool badinc = false;
            if (bb_test(opp, d)) badinc = true;
            else if (bb_test(empty, d)) {
                bool on_stack = false;
                for (int k = 0; k < depth; k++) on_stack = on_stack || (st.pos[k] == d);
                if (!on_stack) {
                    if (!bb_test(eyeish, d)) badinc = true;
                    else { depth++; st.pos[depth] = (uint16_t)d; st.idx[depth] = 0; st.bad[depth] = 0; continue; }
                }
            }
            if (badinc && ++st.bad[depth] > allow) { ret_valid = true; ret = false; }
        }
        while (ret_valid) {                        // return `ret` from frame `depth`
            if (depth == 0) return ret;
            depth--;
            ret_valid = false;
            if (!ret) {
                int pa = st.pos[depth];
                if (++st.bad[depth] > (q_interior(pa, g.S) ? 1 : 0)) { ret_valid = true; ret = false; }
            }
        }
    }
}

// ------------------------------------------------------------------ per-position analysis

struct Analysis {
    BB empty, legal, sensible;
    BB own_atari, opp_atari;       // stones of groups with exactly one liberty
    BB opp_lib2;                   // opponent stones of groups with exactly two liberties
};

// One flood per group: stones of groups with exactly 1 / exactly 2 liberties, and (when
// `planes` is given) the one-hot min(liberty_counts, 8) planes of preprocessing.py:40-56.
RAG_FN void bb_scan_groups(const Geo& g, BB own, BB opp, BB empty, uint32_t* planes, BB& lib1, BB& lib2) {
    lib1 = bb_zero(); lib2 = bb_zero();
    BB rem = own | opp;
    for (;;) {
        int s = bb_first(rem);
        if (s < 0) break;
        BB grp = bb_flood(bb_single(s, g), bb_test(own, s) ? own : opp, g);
        int L = bb_count(bb_nbrs(grp, g) & empty);
        if (planes) {
            int k = (L < 8 ? L : 8) - 1;
            if (k >= 0) plane_or(planes, kPlaneLiberties + k, grp, g);
        }
        if (L == 1) lib1 = lib1 | grp;
        if (L == 2) lib2 = lib2 | grp;
        rem = andnot(rem, grp);
    }
}

// Legality of every point at once (go.py:219-285): an empty point is playable iff it has an
// empty neighbour, or joins an own group that keeps another liberty, or takes the last
// liberty of an enemy group; and it is not the ko point.
RAG_FN BB bb_legal(const Geo& g, BB own, BB empty, BB own_atari, BB opp_atari, int ko_q) {
    BB legal = empty & (bb_nbrs(empty, g) | bb_nbrs(andnot(own, own_atari), g) | bb_nbrs(opp_atari, g));
    if (ko_q >= 0) legal = andnot(legal, bb_single(ko_q, g));
    return legal;
}

// Own true eyes among `among` (go.py:299-327).
RAG_FN BB bb_own_eyes(const Geo& g, BB own, BB opp, BB empty, BB among, EyeStack& est) {
    BB eyeish = empty & bb_from_north(own, g, true) & bb_from_south(own, g, true) &
                bb_from_west(own, g, true) & bb_from_east(own, g, true);
    BB eyes = bb_zero();
    BB rem = eyeish & among;
    for (;;) {
        int q = bb_first(rem);
        if (q < 0) break;
        BB pt = bb_single(q, g);
        if (bb_is_eye(g, opp, empty, eyeish, q, est)) eyes = eyes | pt;
        rem = andnot(rem, pt);
    }
    return eyes;
}

// Legal and sensible moves only (what a playout step needs; go.py:465-480).
RAG_FN_NOINLINE void bb_moves(const Geo& g, BB own, BB opp, int ko_q, EyeStack& est, BB& legal, BB& sensible) {
    BB empty = andnot(bb_board(g), own | opp), lib1, lib2;
    bb_scan_groups(g, own, opp, empty, nullptr, lib1, lib2);
    legal = bb_legal(g, own, empty, own & lib1, opp & lib1, ko_q);
    sensible = andnot(legal, bb_own_eyes(g, own, opp, empty, legal, est));
}

// Fills the canonical planes that need no look-ahead (everything but 44/45) and returns the
// masks the ladder reader starts from.  `planes` = kNumPlanes x 32 words, zeroed by the caller.
// own/opp are relative to the side to move (preprocessing.py:19-20); age0..2 = bits of
// min(stone_ages, 7) (preprocessing.py:33-36); ko_q = q-index of the ko point or -1.
// `forbid`: points that are illegal for a reason this function cannot see (positional superko).
RAG_FN_NOINLINE Analysis bb_features(const Geo& g, BB own, BB opp, BB age0, BB age1, BB age2, int ko_q,
                                     uint32_t* planes, EyeStack& est, BB forbid) {
    Analysis A;
    const BB board = bb_board(g);
    const BB stones = own | opp;
    const BB empty = andnot(board, stones);
    A.empty = empty;

    // board / ones (preprocessing.py:14-22, 214-217); zeros stays clear
    plane_or(planes, kPlaneBoard + 0, own, g);
    plane_or(planes, kPlaneBoard + 1, opp, g);
    plane_or(planes, kPlaneBoard + 2, empty, g);
    plane_or(planes, kPlaneOnes, board, g);

    // turns_since (preprocessing.py:25-37)
    for (int k = 0; k < 8; k++) {
        BB m = stones;
        m = (k & 1) ? (m & age0) : andnot(m, age0);
        m = (k & 2) ? (m & age1) : andnot(m, age1);
        m = (k & 4) ? (m & age2) : andnot(m, age2);
        plane_or(planes, kPlaneTurnsSince + k, m, g);
    }

    // liberties (preprocessing.py:40-56)
    BB lib1, lib2;
    bb_scan_groups(g, own, opp, empty, planes, lib1, lib2);
    A.own_atari = own & lib1;
    A.opp_atari = opp & lib1;
    A.opp_lib2 = opp & lib2;

    BB legal = andnot(bb_legal(g, own, empty, A.own_atari, A.opp_atari, ko_q), forbid);
    A.legal = legal;
    plane_or(planes, kPlaneLegal, legal, g);

    // sensibleness (preprocessing.py:190-196): legal and not an own true eye
    A.sensible = andnot(legal, bb_own_eyes(g, own, opp, empty, legal, est));
    plane_or(planes, kPlaneSensibleness, A.sensible, g);

    // candidate-move planes (preprocessing.py:59-169).
    // Quiet moves -- no own neighbour, no capturable neighbour -- in one bit-parallel pass:
    // captures nothing, group = the stone alone, liberties = its empty neighbours.
    BB touching = bb_nbrs(own, g) | bb_nbrs(A.opp_atari, g);
    {
        BB quiet = andnot(legal, touching);
        BB a = bb_from_north(empty, g, false), b = bb_from_south(empty, g, false);
        BB c = bb_from_west(empty, g, false), d = bb_from_east(empty, g, false);
        BB s1 = a ^ b ^ c, c1 = (a & b) | (c & (a ^ b));          // a+b+c = s1 + 2*c1
        BB s0 = s1 ^ d, c2 = s1 & d;                               // +d
        BB two = c1 ^ c2, four = c1 & c2;                          // count = s0 + 2*two + 4*four
        BB n1 = andnot(andnot(s0, two), four), n2 = andnot(andnot(two, s0), four);
        BB n3 = andnot(s0 & two, four), n4 = four;
        plane_or(planes, kPlaneCaptureSize + 0, quiet, g);
        plane_or(planes, kPlaneLibertiesAfter + 0, quiet & n1, g);
        plane_or(planes, kPlaneLibertiesAfter + 1, quiet & n2, g);
        plane_or(planes, kPlaneLibertiesAfter + 2, quiet & n3, g);
        plane_or(planes, kPlaneLibertiesAfter + 3, quiet & n4, g);
        plane_or(planes, kPlaneSelfAtari + 0, quiet & n1, g);
    }
    // The rest one by one: merged group and captured stones by flood fill.
    {
        BB rem = legal & touching;
        for (;;) {
            int q = bb_first(rem);
            if (q < 0) break;
            BB pt = bb_single(q, g);
            rem = andnot(rem, pt);
            BB ptn = bb_nbrs(pt, g);
            BB merged = bb_any(ptn & own) ? bb_flood(pt, own | pt, g) : pt;
            BB seeds = ptn & A.opp_atari;
            BB cap = bb_any(seeds) ? bb_flood(seeds, opp, g) : bb_zero();
            int ncap = bb_count(cap);
            BB libs = andnot(bb_nbrs(merged, g) & (empty | cap), pt);
            int la = bb_count(libs);
            plane_or(planes, kPlaneCaptureSize + (ncap < 7 ? ncap : 7), pt, g);
            plane_or(planes, kPlaneLibertiesAfter + ((la - 1) < 7 ? ((la - 1) & 7) : 7), pt, g);
            if (la == 1) {
                int sz = bb_count(merged);
                plane_or(planes, kPlaneSelfAtari + ((sz - 1) < 7 ? (sz - 1) : 7), pt, g);
            }
        }
    }
    return A;
}

// ------------------------------------------------------------------ full do_move incl. stone ages

// Colours are absolute here (black/white); `player` = +1 black to move, -1 white.
struct FullState { BB black, white, age0, age1, age2; int ko_q; int player; };

// GameState.do_move(action) for the side to move (go.py:521-581); q < 0 = pass.
// Ages are the 3-bit saturating twin of stone_ages: +1 for every stone on any legal move
// incl. pass (go.py:533), 0 for the new stone (go.py:539), cleared on capture (go.py:175).
// Returns false and leaves the state alone on an illegal move (IllegalMove, go.py:573-575).
RAG_FN bool bb_do_move(const Geo& g, FullState& s, int q) {
    BB M = s.player > 0 ? s.black : s.white, O = s.player > 0 ? s.white : s.black;
    int ko = s.ko_q;
    BB stones_before = M | O;
    if (q >= 0) { if (!bb_play(g, M, O, ko, q, nullptr)) return false; }
    else ko = -1;
    BB sat = s.age0 & s.age1 & s.age2;
    BB n0 = andnot(stones_before, s.age0) | sat;
    BB c0 = s.age0;
    BB n1 = (s.age1 ^ c0) | sat;
    BB n2 = (s.age2 ^ (s.age1 & c0)) | sat;
    BB keep = M | O;                                   // survivors + the new stone
    if (q >= 0) keep = andnot(keep, bb_single(q, g));  // new stone: age 0
    s.age0 = n0 & stones_before & keep; s.age1 = n1 & stones_before & keep; s.age2 = n2 & stones_before & keep;
    if (s.player > 0) { s.black = M; s.white = O; } else { s.white = M; s.black = O; }
    s.ko_q = ko; s.player = -s.player;
    return true;
}

// ------------------------------------------------------------------ ladder reader (go.py:329-463)

struct LadderStack {
    BB hunter[kLadderDepth];
    BB prey[kLadderDepth];
    BB group[kLadderDepth];      // the prey group itself, carried along instead of re-traced
    BB pending[kLadderDepth];
    int16_t ko[kLadderDepth];
};

enum { kLadderValue0 = 0, kLadderValue1 = 1, kLadderPush = 2 };

#if !RAG_DEVICE_BUILD
// host-only instrumentation for tests/hostsim experiments
struct LadderStats { long long nodes; int max_depth; long long per_depth[96]; int cur_depth; };
static LadderStats g_ladder_stats = {0, 0, {0}, 0};
#define RAG_LADDER_COUNT_NODE() (g_ladder_stats.nodes++, g_ladder_stats.per_depth[g_ladder_stats.cur_depth]++)
static void (*g_ladder_node_hook)(const uint32_t* h32, const uint32_t* p32, int ko, int is_capture, int mv) = nullptr;
#define RAG_LADDER_DEPTH(d) do { if ((d) > g_ladder_stats.max_depth) g_ladder_stats.max_depth = (d); } while (0)
#else
#define RAG_LADDER_COUNT_NODE() ((void)0)
#define RAG_LADDER_DEPTH(d) ((void)0)
#endif

// One call of is_ladder_capture (is_capture) or is_ladder_escape on a SINGLE prey group.
// Plays `mv`, then either decides at once or leaves the set of replies to try in `pending`.
//   capture node (go.py:345-391): hunter plays; replies = prey's liberties + the last liberty of
//       every hunter group in atari that touches the prey; value = no reply escapes.
//   escape node  (go.py:413-460): prey plays; >=3 liberties -> 1, 1 liberty -> 0, else
//       replies = the two liberties; value = no reply is a ladder capture.
// `remaining` is the reference's remaining_attempts; legality is tested before it (H11).
RAG_FN int ladder_expand_played(const Geo& g, bool is_capture, BB& H, BB& P, BB& G, int mv, int remaining,
                                BB& pending, BB ptn);

RAG_FN int ladder_expand(const Geo& g, bool is_capture, BB& H, BB& P, BB& G, int& ko, int mv,
                         int remaining, BB& pending) {
    RAG_LADDER_COUNT_NODE();
#if !RAG_DEVICE_BUILD
    if (g_ladder_node_hook) g_ladder_node_hook(H.w, P.w, ko, is_capture ? 1 : 0, mv);
#endif
    BB ptn;
    bool ok = is_capture ? bb_play_ex(g, H, P, ko, mv, nullptr, ptn) : bb_play_ex(g, P, H, ko, mv, nullptr, ptn);
    if (!ok) return kLadderValue0;
    return ladder_expand_played(g, is_capture, H, P, G, mv, remaining, pending, ptn);
}

// The part of ladder_expand after the (legal) move `mv` has been played; `ptn` = its neighbours.
RAG_FN int ladder_expand_played(const Geo& g, bool is_capture, BB& H, BB& P, BB& G, int mv, int remaining,
                                BB& pending, BB ptn) {
    if (remaining <= 0) return is_capture ? kLadderValue1 : kLadderValue0;
    const BB empty = andnot(bb_board(g), H | P);
    if (!is_capture) {
        // The prey side moved.  Its group G grows only if the new stone touches it (a reply can
        // also be a capture elsewhere); other groups of its colour the stone touches join too.
        // A hunter move never changes G: the group keeps a liberty, so it is not what is captured.
        const BB extra = andnot(ptn & P, G);
        const unsigned f = bb_any4(ptn & G, extra, bb_zero(), bb_zero());
        if (f & 1u) {
            G = G | bb_single(mv, g);
            if (f & 2u) G = G | bb_flood(extra, P, g);
        }
    }
    const BB gn = bb_nbrs(G, g);
    const BB libs = gn & empty;
    if (is_capture) {
        pending = libs;
        // hunter groups in atari that touch the prey (go.py:381-384).  A stone with two empty
        // neighbours of its own cannot be in such a group, so only the others are traced.
        const BB a = bb_from_north(empty, g, false), b = bb_from_south(empty, g, false);
        const BB c = bb_from_west(empty, g, false), d = bb_from_east(empty, g, false);
        const BB ge2 = (a & b) | (c & d) | ((a | b) & (c | d));
        BB rem = andnot(gn & H, ge2);
        for (;;) {
            int h = bb_first(rem);
            if (h < 0) break;
            BB hg = bb_flood(bb_single(h, g), H, g);
            BB hl = bb_nbrs(hg, g) & empty;
            if (bb_count(hl) == 1) pending = pending | hl;
            rem = andnot(rem, hg);
        }
        return kLadderPush;
    }
    int L = bb_count(libs);
    if (L >= 3) return kLadderValue1;
    if (L == 1) return kLadderValue0;
    pending = libs;
    return kLadderPush;
}

constexpr int kLadderAborted = 2;

// Depth-first search with an explicit stack; a node's value is "no child has value 1".
// `remaining0` is the reference's remaining_attempts at the first move (80 at the top level);
// `budget` bounds the number of moves tried (< 0: unbounded) -- when it runs out the search
// stops with kLadderAborted and `nodes` tells how far it got.  Returns 0, 1 or kLadderAborted.
// `frame_is_capture` is the kind of the node whose replies are in `pending` (frame 0); its
// replies are tried with remaining_attempts = remaining_children.
struct NeverCancel {
#if RAG_DEVICE_BUILD
    __device__ __forceinline__
#endif
    bool operator()() const { return false; }
};

// `cancel()` is asked every 16 moves; when it says yes the search stops with kLadderAborted (used
// by the all-warps evaluator to drop subtrees whose result no longer matters).
template <class Cancel>
#if RAG_DEVICE_BUILD
__device__ __noinline__
#else
static
#endif
int ladder_dfs(const Geo& g, BB H, BB P, BB G, int ko, BB pending, bool frame_is_capture,
               int remaining_children, int budget, int nodes, LadderStack& st, Cancel cancel) {
    int d = 0;
    st.hunter[0] = H; st.prey[0] = P; st.group[0] = G; st.pending[0] = pending; st.ko[0] = (int16_t)ko;
    for (;;) {
        int v;
        int m = bb_first(st.pending[d]);
        if (m < 0) v = 1;                                    // every reply failed
        else {
            if (budget >= 0 && nodes >= budget) return kLadderAborted;
            if ((nodes & 15) == 15 && cancel()) return kLadderAborted;
            nodes++;
            st.pending[d] = andnot(st.pending[d], bb_single(m, g));
            bool node_is_capture = frame_is_capture ^ ((d & 1) != 0);
            H = st.hunter[d]; P = st.prey[d]; G = st.group[d]; ko = st.ko[d];
#if !RAG_DEVICE_BUILD
            g_ladder_stats.cur_depth = d + 1;
#endif
            int res = ladder_expand(g, !node_is_capture, H, P, G, ko, m, remaining_children - d, pending);
            if (res == kLadderPush) {
                d++;
                RAG_LADDER_DEPTH(d);
                st.hunter[d] = H; st.prey[d] = P; st.group[d] = G; st.pending[d] = pending; st.ko[d] = (int16_t)ko;
                continue;
            }
            if (res == kLadderValue0) continue;              // this reply does not refute; try the next
            v = 0;
        }
        for (;;) {                                           // frame d returns v
            if (d == 0) return v;
            d--;
            if (v == 0) break;                               // parent keeps looking
            v = 0;                                           // a child of value 1 refutes the parent
        }
    }
}

// is_ladder_capture / is_ladder_escape of `first_move` on the prey group G (go.py:329-463).
RAG_FN int ladder_search(const Geo& g, BB H, BB P, BB G, int ko, int first_move,
                         bool root_is_capture, int remaining0, int budget, LadderStack& st) {
    BB pending = bb_zero();
    int res = ladder_expand(g, root_is_capture, H, P, G, ko, first_move, remaining0, pending);
    if (res != kLadderPush) return res;
    return ladder_dfs(g, H, P, G, ko, pending, root_is_capture, remaining0 - 1, budget, 1, st, NeverCancel());
}

// The prey groups is_ladder_capture(action) / is_ladder_escape(action) look at when called with
// prey=None (go.py:355-361, 422-428): adjacent opponent groups with exactly two liberties
// (capture) / own groups with exactly one (escape), each once.  Returns how many.
RAG_FN int ladder_root_preys(const Geo& g, BB own, BB opp, int action_q, bool is_escape, BB groups_out[4]) {
    const BB side = is_escape ? own : opp;
    const BB empty = andnot(bb_board(g), own | opp);
    const int want = is_escape ? 1 : 2;
    BB done = bb_zero();
    int n = 0;
    for (int d = 0; d < 4; d++) {
        int nq = q_neighbor(action_q, d, g.S);
        if (nq < 0 || !bb_test(andnot(side, done), nq)) continue;
        BB grp = bb_flood(bb_single(nq, g), side, g);
        done = done | grp;
        if (bb_count(bb_nbrs(grp, g) & empty) == want) groups_out[n++] = grp;
    }
    return n;
}

// Top level of is_ladder_capture(action) / is_ladder_escape(action): true if the search succeeds
// for ANY of those groups.  Returns 0, 1, or kLadderAborted when a search ran out of `budget`.
RAG_FN_NOINLINE int ladder_root(const Geo& g, BB own, BB opp, int ko, int action_q, bool is_escape,
                                int budget, LadderStack& st) {
    BB preys[4];
    const int np = ladder_root_preys(g, own, opp, action_q, is_escape, preys);
    for (int i = 0; i < np; i++) {
        // capture: side to move hunts `opp`; escape: side to move is the prey
        int r = is_escape ? ladder_search(g, opp, own, preys[i], ko, action_q, false, kLadderDepth, budget, st)
                          : ladder_search(g, own, opp, preys[i], ko, action_q, true, kLadderDepth, budget, st);
        if (r != kLadderValue0) return r;
    }
    return kLadderValue0;
}

// ------------------------------------------------------------------ positional superko  (go.py:242-265)
//
// Only for states created with enforce_superko=True (the GTP front end).  is_legal then also asks
// is_positional_superko: if the point was played before by the side to move (by the parity of its
// index in `history`) or is a handicap point, play it on a copy and look the resulting Zobrist hash
// up in previous_hashes.  copy() carries hash, hash set and history along (go.py:193-199), so the
// ladder reader's look-ahead enforces it too, against the hashes of its own path as well.

typedef unsigned long long u64;

struct SuperkoRoot {                // constant during one search
    const u64* zob;                 // hash_lookup by q-index: [0][1024] BLACK, [1][1024] WHITE  (go.py:58-62)
    const u64* prev;                // previous_hashes of the position
    int nprev;
    BB handi;                       // handicaps
    bool has_handicap;
    bool enforce;                   // enforce_superko of this position (a batch may mix both kinds)
};

struct SuperkoCtx {                 // of one state (the position, or a node of a ladder search)
    u64 hash;                       // current_hash
    BB even, odd;                   // points of history[0::2] / history[1::2]
    int hist_len;                   // len(history)
};

// XOR of table[q] over the points of `stones`
RAG_FN u64 bb_zobrist(const Geo& g, BB stones, const u64* table) {
#if RAG_DEVICE_BUILD
    u64 h = 0;
    for (uint32_t w = stones.w; w; w &= w - 1) h ^= table[(g.lane << 5) | (uint32_t)(__ffs(w) - 1)];
    const uint32_t lo = __reduce_xor_sync(kFullMask, (uint32_t)h), hi = __reduce_xor_sync(kFullMask, (uint32_t)(h >> 32));
    return ((u64)hi << 32) | lo;
#else
    u64 h = 0;
    for (int x = 0; x < 32; x++)
        for (uint32_t w = stones.w[x]; w; w &= w - 1) h ^= table[(x << 5) | __builtin_ctz(w)];
    (void)g;
    return h;
#endif
}

// h in previous_hashes of a state whose own moves so far produced path[0 .. npath) ?
RAG_FN bool sk_seen(const Geo& g, const SuperkoRoot& r, u64 h, const u64* path, int npath) {
#if RAG_DEVICE_BUILD
    bool hit = false;
    for (int i = (int)g.lane; i < r.nprev; i += 32) hit |= r.prev[i] == h;
    for (int i = (int)g.lane; i < npath; i += 32) hit |= path[i] == h;
    return __any_sync(kFullMask, hit) != 0;
#else
    (void)g;
    for (int i = 0; i < r.nprev; i++) if (r.prev[i] == h) return true;
    for (int i = 0; i < npath; i++) if (path[i] == h) return true;
    return false;
#endif
}

// the points is_positional_superko looks at for the side to move (go.py:248-256)
RAG_FN BB sk_mine(const SuperkoRoot& r, const SuperkoCtx& c, bool mover_is_black) {
    const bool first_slice = r.has_handicap ? !mover_is_black : mover_is_black;
    return r.handi | (first_slice ? c.even : c.odd);
}

// Moves of the side to move (stones M, colour given) that is_legal rejects for superko alone.
RAG_FN_NOINLINE BB sk_forbidden_moves(const Geo& g, const SuperkoRoot& r, const SuperkoCtx& c, BB M, BB O, int ko_q,
                                      bool mover_is_black) {
    BB out = bb_zero();
    if (!r.enforce) return out;
    BB rem = andnot(sk_mine(r, c, mover_is_black), M | O);
    for (;;) {
        const int q = bb_first(rem);
        if (q < 0) break;
        const BB pt = bb_single(q, g);
        rem = andnot(rem, pt);
        BB M2 = M, O2 = O;
        int k2 = ko_q, ncap = 0;
        if (!bb_play(g, M2, O2, k2, q, &ncap)) continue;              // illegal anyway
        u64 h = c.hash ^ r.zob[(mover_is_black ? 0 : 1024) + q];
        if (ncap) h ^= bb_zobrist(g, andnot(O, O2), r.zob + (mover_is_black ? 1024 : 0));
        if (sk_seen(g, r, h, nullptr, 0)) out = out | pt;
    }
    return out;
}

struct LadderStackSk {
    LadderStack base;
    u64 hash[kLadderDepth];
    BB even[kLadderDepth], odd[kLadderDepth];
    u64 path[kLadderDepth + 2];     // hashes after each move of the current line
};

// ladder_expand for a state that enforces superko: `c` is the state's bookkeeping before the move and
// becomes the child's; `h_out` is the hash after the move.
RAG_FN int ladder_expand_sk(const Geo& g, bool is_capture, BB& H, BB& P, BB& G, int& ko, int mv, int remaining,
                            BB& pending, const SuperkoRoot& r, SuperkoCtx& c, bool hunter_is_black,
                            const u64* path, int npath) {
    RAG_LADDER_COUNT_NODE();
    const BB Hb = H, Pb = P;
    BB ptn;
    int ncap = 0;
    const bool ok = is_capture ? bb_play_ex(g, H, P, ko, mv, &ncap, ptn) : bb_play_ex(g, P, H, ko, mv, &ncap, ptn);
    if (!ok) return kLadderValue0;
    const bool mover_is_black = is_capture ? hunter_is_black : !hunter_is_black;
    u64 h = c.hash ^ r.zob[(mover_is_black ? 0 : 1024) + mv];
    if (ncap) h ^= bb_zobrist(g, is_capture ? andnot(Pb, P) : andnot(Hb, H), r.zob + (mover_is_black ? 1024 : 0));
    if (r.enforce && bb_test(sk_mine(r, c, mover_is_black), mv) && sk_seen(g, r, h, path, npath)) return kLadderValue0;   // go.py:282-283
    c.hash = h;
    if (c.hist_len & 1) c.odd = c.odd | bb_single(mv, g); else c.even = c.even | bb_single(mv, g);
    c.hist_len++;
    return ladder_expand_played(g, is_capture, H, P, G, mv, remaining, pending, ptn);
}

// ladder_search + ladder_dfs with the superko bookkeeping carried per frame.
RAG_FN_NOINLINE int ladder_search_sk(const Geo& g, BB H, BB P, BB G, int ko, int first_move, bool root_is_capture,
                                     int remaining0, const SuperkoRoot& r, SuperkoCtx c, bool hunter_is_black,
                                     LadderStackSk& sk) {
    LadderStack& st = sk.base;
    BB pending = bb_zero();
    int res = ladder_expand_sk(g, root_is_capture, H, P, G, ko, first_move, remaining0, pending, r, c, hunter_is_black, sk.path, 0);
    if (res != kLadderPush) return res;
    const int remaining_children = remaining0 - 1;
    int d = 0;
    st.hunter[0] = H; st.prey[0] = P; st.group[0] = G; st.pending[0] = pending; st.ko[0] = (int16_t)ko;
    sk.hash[0] = c.hash; sk.even[0] = c.even; sk.odd[0] = c.odd; sk.path[0] = c.hash;
    const int len0 = c.hist_len;                         // len(history) of frame 0's state
    for (;;) {
        int v;
        const int m = bb_first(st.pending[d]);
        if (m < 0) v = 1;
        else {
            st.pending[d] = andnot(st.pending[d], bb_single(m, g));
            const bool node_is_capture = root_is_capture ^ ((d & 1) != 0);
            H = st.hunter[d]; P = st.prey[d]; G = st.group[d]; ko = st.ko[d];
            c.hash = sk.hash[d]; c.even = sk.even[d]; c.odd = sk.odd[d]; c.hist_len = len0 + d;
            res = ladder_expand_sk(g, !node_is_capture, H, P, G, ko, m, remaining_children - d, pending, r, c,
                                   hunter_is_black, sk.path, d + 1);
            if (res == kLadderPush) {
                d++;
                st.hunter[d] = H; st.prey[d] = P; st.group[d] = G; st.pending[d] = pending; st.ko[d] = (int16_t)ko;
                sk.hash[d] = c.hash; sk.even[d] = c.even; sk.odd[d] = c.odd; sk.path[d] = c.hash;
                continue;
            }
            if (res == kLadderValue0) continue;
            v = 0;
        }
        for (;;) {
            if (d == 0) return v;
            d--;
            if (v == 0) break;
            v = 0;
        }
    }
}

}  // namespace rag
