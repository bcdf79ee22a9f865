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

extern "C" int hostsim_features(int n, int S, const int8_t* board, const int16_t* age, const int32_t* ko,
                                const int32_t* player, uint8_t* out, int64_t* stats)
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
        Analysis A = bb_features(g, own, opp, a0, a1, a2, ko_q, planes, *est);
        BB cand_cap = A.legal & bb_nbrs(A.opp_lib2, g);
        BB cand_esc = A.legal & bb_nbrs(A.own_atari, g);
        if (stats) { stats[0] += bb_count(cand_cap); stats[1] += bb_count(cand_esc); }
        for (int kind = 0; kind < 2; kind++) {
            BB rem = kind ? cand_esc : cand_cap;
            for (;;) {
                int q = bb_first(rem);
                if (q < 0) break;
                rem = andnot(rem, bb_single(q, g));
                bool r = ladder_root(g, own, opp, ko_q, q, kind == 1, *lst);
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
