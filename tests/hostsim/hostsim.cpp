// tests/hostsim/hostsim.cpp -- TEST INFRASTRUCTURE ONLY (never linked into the product).
//
// Compiles rocalphago_b200/csrc/go_bb.cuh as plain C++ with a 32-row array standing in for
// the 32 lanes of a warp (RAG_HOSTSIM), so the warp-uniform device logic -- floods, do_move,
// eye DFS, candidate-move planes, ladder reader -- can be checked against the oracle and the
// golden fixtures on a machine without a GPU.  It mirrors what features_kernel +
// ladder_kernel do per position: canonical planes, then one ladder_root per candidate move.
#define RAG_HOSTSIM 1
#include "../../rocalphago_b200/csrc/go_bb.cuh"
#include <string.h>
#include <stdlib.h>

using namespace rag;
static int g_skip_ladders = 0;
extern "C" void hostsim_skip_ladders(int v) { g_skip_ladders = v; }

extern "C" int hostsim_features(int n, int S, const int8_t* board, const int16_t* age, const int32_t* ko,
                                const int32_t* player, uint8_t* out, int64_t* stats, int64_t* item_nodes, int64_t item_cap)
{
    const int np = S * S;
    Geo g = make_geo(S);
    EyeStack* est = (EyeStack*)malloc(sizeof(EyeStack));
    LadderStack* lst = (LadderStack*)malloc(sizeof(LadderStack));
    for (int b = 0; b < n; b++) {
        BB own = bb_zero(), opp = bb_zero(), a0 = bb_zero(), a1 = bb_zero(), a2 = bb_zero();
        for (int x = 0; x < S; x++)
            for (int y = 0; y < S; y++) {
                int v = board[(size_t)b * np + x * S + y];
                if (v == 0) continue;
                if (v == player[b]) own.w[x] |= 1u << y; else opp.w[x] |= 1u << y;
                int ag = age[(size_t)b * np + x * S + y];
                if (ag > 7) ag = 7;
                if (ag < 0) ag = 0;
                if (ag & 1) a0.w[x] |= 1u << y;
                if (ag & 2) a1.w[x] |= 1u << y;
                if (ag & 4) a2.w[x] |= 1u << y;
            }
        int ko_q = ko[b] < 0 ? -1 : ((ko[b] / S) << 5) | (ko[b] % S);
        uint32_t planes[kNumPlanes * 32];
        memset(planes, 0, sizeof planes);
        Analysis A = bb_features(g, own, opp, a0, a1, a2, ko_q, planes, *est, bb_zero());
        for (int x = 0; x < S; x++)                            // stones without an age are in no turns_since plane
            for (int y = 0; y < S; y++)
                if (board[(size_t)b * np + x * S + y] != 0 && age[(size_t)b * np + x * S + y] < 0)
                    for (int k = 0; k < 8; k++) planes[(kPlaneTurnsSince + k) * 32 + x] &= ~(1u << y);
        BB cand_cap = A.legal & bb_nbrs(A.opp_lib2, g);
        BB cand_esc = A.legal & bb_nbrs(A.own_atari, g);
        if (stats) { stats[0] += bb_count(cand_cap); stats[1] += bb_count(cand_esc); }
        for (int kind = 0; kind < 2 && !g_skip_ladders; kind++) {
            BB rem = kind ? cand_esc : cand_cap;
            for (;;) {
                int q = bb_first(rem);
                if (q < 0) break;
                rem = andnot(rem, bb_single(q, g));
                long long n0 = g_ladder_stats.nodes; g_ladder_stats.max_depth = 0;
                bool r = ladder_root(g, own, opp, ko_q, q, kind == 1, -1, *lst) == kLadderValue1;
                if (stats) {
                    long long dn = g_ladder_stats.nodes - n0;
                    stats[2] += dn; stats[3] += 1;
                    if (dn > stats[4]) { stats[4] = dn; stats[5] = b; stats[6] = q; stats[7] = kind; stats[8] = g_ladder_stats.max_depth; }
                    if (item_nodes && stats[3] <= item_cap) item_nodes[stats[3] - 1] = dn;
                }
                if (r) planes[(kind ? kPlaneLadderEscape : kPlaneLadderCapture) * 32 + (q >> 5)] |= 1u << (q & 31);
            }
        }
        uint8_t* o = out + (size_t)b * 49 * np;
        for (int k = 0; k < 49; k++)
            for (int x = 0; x < S; x++)
                for (int y = 0; y < S; y++) o[k * np + x * S + y] = (planes[k * 32 + x] >> y) & 1u;
    }
    free(est); free(lst);
    return 0;
}

// The same with positional superko enforced (go.py:242-265): per position the Zobrist hash, the
// earlier hashes and the history masks, as og_features_superko takes them.  zob: [2][S*S], [0] BLACK.
extern "C" int hostsim_features_superko(int n, int S, const int8_t* board, const int16_t* age, const int32_t* ko,
                                        const int32_t* player, uint8_t* out, const uint64_t* zob_p,
                                        const int32_t* has_handicap, const int32_t* hist_len, const uint64_t* hash,
                                        const uint64_t* prev, const int32_t* prev_off, const uint8_t* masks)
{
    const int np = S * S;
    Geo g = make_geo(S);
    EyeStack* est = (EyeStack*)malloc(sizeof(EyeStack));
    LadderStackSk* lst = (LadderStackSk*)malloc(sizeof(LadderStackSk));
    u64* zob = (u64*)calloc(2048, sizeof(u64));
    for (int c = 0; c < 2; c++)
        for (int x = 0; x < S; x++)
            for (int y = 0; y < S; y++) zob[c * 1024 + (x << 5) + y] = zob_p[(size_t)c * np + x * S + y];
    for (int b = 0; b < n; b++) {
        BB own = bb_zero(), opp = bb_zero(), a0 = bb_zero(), a1 = bb_zero(), a2 = bb_zero();
        SuperkoRoot r; SuperkoCtx c;
        r.handi = bb_zero(); c.even = bb_zero(); c.odd = bb_zero();
        for (int x = 0; x < S; x++)
            for (int y = 0; y < S; y++) {
                const size_t p = (size_t)b * np + x * S + y;
                if (masks[((size_t)b * 3 + 0) * np + x * S + y]) c.even.w[x] |= 1u << y;
                if (masks[((size_t)b * 3 + 1) * np + x * S + y]) c.odd.w[x] |= 1u << y;
                if (masks[((size_t)b * 3 + 2) * np + x * S + y]) r.handi.w[x] |= 1u << y;
                int v = board[p];
                if (v == 0) continue;
                if (v == player[b]) own.w[x] |= 1u << y; else opp.w[x] |= 1u << y;
                int ag = age[p];
                if (ag > 7) ag = 7;
                if (ag < 0) ag = 0;
                if (ag & 1) a0.w[x] |= 1u << y;
                if (ag & 2) a1.w[x] |= 1u << y;
                if (ag & 4) a2.w[x] |= 1u << y;
            }
        r.zob = zob; r.prev = (const u64*)prev + prev_off[b]; r.nprev = prev_off[b + 1] - prev_off[b];
        r.has_handicap = has_handicap[b] != 0;
        r.enforce = true;
        c.hash = hash[b]; c.hist_len = hist_len[b];
        const bool black = player[b] > 0;
        int ko_q = ko[b] < 0 ? -1 : ((ko[b] / S) << 5) | (ko[b] % S);
        uint32_t planes[kNumPlanes * 32];
        memset(planes, 0, sizeof planes);
        const BB forbid = sk_forbidden_moves(g, r, c, own, opp, ko_q, black);
        Analysis A = bb_features(g, own, opp, a0, a1, a2, ko_q, planes, *est, forbid);
        BB cand[2] = { A.legal & bb_nbrs(A.opp_lib2, g), A.legal & bb_nbrs(A.own_atari, g) };
        for (int kind = 0; kind < 2; kind++) {
            BB rem = cand[kind];
            for (;;) {
                int q = bb_first(rem);
                if (q < 0) break;
                rem = andnot(rem, bb_single(q, g));
                BB preys[4];
                const int npreys = ladder_root_preys(g, own, opp, q, kind == 1, preys);
                int res = kLadderValue0;
                for (int i = 0; i < npreys && res == kLadderValue0; i++)
                    res = kind == 1 ? ladder_search_sk(g, opp, own, preys[i], ko_q, q, false, kLadderDepth, r, c, !black, *lst)
                                    : ladder_search_sk(g, own, opp, preys[i], ko_q, q, true, kLadderDepth, r, c, black, *lst);
                if (res == kLadderValue1) planes[(kind ? kPlaneLadderEscape : kPlaneLadderCapture) * 32 + (q >> 5)] |= 1u << (q & 31);
            }
        }
        uint8_t* o = out + (size_t)b * 49 * np;
        for (int k = 0; k < 49; k++)
            for (int x = 0; x < S; x++)
                for (int y = 0; y < S; y++) o[k * np + x * S + y] = (planes[k * 32 + x] >> y) & 1u;
    }
    free(est); free(lst); free(zob);
    return 0;
}

// do_move on one state per call (for the device do_move / playout logic)
extern "C" int hostsim_play(int S, int8_t* board, int32_t* ko, int32_t* player, int move)
{
    Geo g = make_geo(S);
    BB M = bb_zero(), O = bb_zero();
    for (int x = 0; x < S; x++)
        for (int y = 0; y < S; y++) {
            int v = board[x * S + y];
            if (v == *player) M.w[x] |= 1u << y; else if (v != 0) O.w[x] |= 1u << y;
        }
    int ko_q = *ko < 0 ? -1 : ((*ko / S) << 5) | (*ko % S);
    if (move >= 0) {
        int q = ((move / S) << 5) | (move % S);
        if (!bb_play(g, M, O, ko_q, q, nullptr)) return -1;
    } else ko_q = -1;
    for (int x = 0; x < S; x++)
        for (int y = 0; y < S; y++)
            board[x * S + y] = (int8_t)(((M.w[x] >> y) & 1u) ? *player : (((O.w[x] >> y) & 1u) ? -*player : 0));
    *ko = ko_q < 0 ? -1 : (ko_q >> 5) * S + (ko_q & 31);
    *player = -*player;
    return 0;
}

// ---- experiment: how many DISTINCT (position, side, move) nodes does one ladder search visit?
#include <unordered_set>
#include <string>
static std::unordered_set<std::string>* g_seen = nullptr;
static void seen_hook(const uint32_t* h, const uint32_t* p, int ko, int is_capture, int mv) {
    std::string k((const char*)h, 19 * 4);
    k.append((const char*)p, 19 * 4);
    k.append((const char*)&ko, 4); k.append((const char*)&is_capture, 4); k.append((const char*)&mv, 4);
    g_seen->insert(k);
}
extern "C" long long hostsim_ladder_distinct(int S, const int8_t* board, int ko, int player, int q, int kind, long long* nodes, long long* per_depth)
{
    for (int i = 0; i < 96; i++) g_ladder_stats.per_depth[i] = 0;
    g_ladder_stats.cur_depth = 0;
    Geo g = make_geo(S);
    BB own = bb_zero(), opp = bb_zero();
    for (int x = 0; x < S; x++) for (int y = 0; y < S; y++) {
        int v = board[x * S + y];
        if (v == player) own.w[x] |= 1u << y; else if (v) opp.w[x] |= 1u << y;
    }
    int ko_q = ko < 0 ? -1 : ((ko / S) << 5) | (ko % S);
    std::unordered_set<std::string> seen; g_seen = &seen; g_ladder_node_hook = seen_hook;
    LadderStack* lst = (LadderStack*)malloc(sizeof(LadderStack));
    long long n0 = g_ladder_stats.nodes;
    ladder_root(g, own, opp, ko_q, q, kind == 1, -1, *lst);
    *nodes = g_ladder_stats.nodes - n0;
    if (per_depth) for (int i = 0; i < 96; i++) per_depth[i] = g_ladder_stats.per_depth[i];
    g_ladder_node_hook = nullptr; free(lst);
    return (long long)seen.size();
}

extern "C" void hostsim_opcounts(long long* out, int reset)
{
    out[0] = g_ops.nbrs; out[1] = g_ops.any; out[2] = g_ops.count; out[3] = g_ops.first; out[4] = g_ops.test;
    out[5] = g_ops.flood_calls; out[6] = g_ops.flood_iters; out[7] = g_ops.from_dir; out[8] = g_ladder_stats.nodes;
    if (reset) { g_ops = OpCounts{0, 0, 0, 0, 0, 0, 0, 0}; }
}
