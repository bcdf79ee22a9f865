// features.cu -- batched Preprocess.state_to_tensor on packed game states (sm_100a).
//
// Replaces AlphaGo/preprocessing/preprocessing.py:14-294 and the parts of AlphaGo/go.py
// it leans on (is_legal :267-285, is_eye :299-327, is_ladder_capture/escape :329-463,
// do_move :521-581) for a batch held in HBM.  Three kernels per batch:
//
//   analyze_kernel  one warp per position (lane r = board row r): every plane that needs no
//                   look-ahead, as 49 x 32 row words per position in the workspace, plus one
//                   work item per ladder candidate (position, move, capture|escape) appended
//                   to a queue in the workspace.
//   ladder_kernel   persistent warps pull items off the queue and run the bounded-depth
//                   ladder search (explicit stack in local memory), OR-ing hits into planes
//                   44/45 of the workspace.  A queue instead of "one warp per position"
//                   because ladder cost varies by 100x between positions.
//   emit_*_kernel   bit-planes -> the layout the caller wants (uint8/float NCHW for the
//                   Preprocess API, zero-padded 16-bit pixel rows for the first convolution).
//                   Pure streaming; this is the HBM-bound kernel of the feature path.
//
// Also here: rag_do_move and rag_random_playouts (device do_move / playout generator).
#include "go_bb.cuh"
#include "rag_common.h"
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <string.h>
#include <mutex>
#include <thread>
#include <vector>

namespace rag {

char* last_error_buffer() {
    static thread_local char buf[512] = {0};
    return buf;
}

constexpr int kWarpsPerBlock = 4;
constexpr int kPlaneWords = kNumPlanes * 32;           // 1600 words = 6400 B per position
constexpr int kCounterBytes = 256;
constexpr int kMaxPositionsPerLaunch = 1 << 21;        // queue item = pos<<11 | kind<<10 | q

// A reply handed to another warp by a long ladder search (see ladder_kernel).
struct StealTask {
    uint32_t H[32], P[32], G[32];    // hunter stones, prey-side stones, prey group: the position the reply is played in
    int ko, move, is_capture, remaining;
    int root;                        // queue index of the (position, move) item this reply belongs to: its slot counts the item's moves
    int pad[27];
};
static_assert(sizeof(StealTask) == 512, "StealTask layout");
enum { kTaskFree = 0, kTaskPosted = 1, kTaskTaken = 2, kTaskDone0 = 3, kTaskDone1 = 4, kTaskCancelled = 5, kTaskBudget = 6 };
enum { kCtlQueued = 0, kCtlTaken = 1, kCtlPosted = 2, kCtlBusy = 4, kCtlAllocCursor = 5, kCtlTimeout = 6, kCtlCutoffs = 7, kCtlMaxNodes = 8 };

struct Workspace {
    unsigned* counters;      // kCtl* slots
    uint32_t* planes;        // [n][50][32]
    uint32_t* queue;         // [n * 2 * S*S]  ladder work items
    size_t queue_cap;
    StealTask* tasks;        // [task_cap]
    int* task_status;        // [task_cap]  kTaskFree .. kTaskCancelled, kept apart so that idle warps can scan it cheaply
    unsigned task_cap;
};

static unsigned steal_task_cap(int n) {
    int c = n / 2;
    return (unsigned)(c < 64 ? 64 : (c > 2048 ? 2048 : c));
}

static size_t workspace_fixed_bytes(int n) {          // counters + planes + queue, rounded for the task pool's alignment
    size_t b = (size_t)kCounterBytes + (size_t)n * kPlaneWords * 4 + (size_t)n * 2 * kMaxBoard * kMaxBoard * 4;
    return (b + 511) / 512 * 512;
}

static size_t workspace_bytes(int n) {
    return workspace_fixed_bytes(n) + (size_t)steal_task_cap(n) * (sizeof(StealTask) + 4);
}

static Workspace carve(void* base, int n) {
    Workspace w;
    char* p = (char*)base;
    w.counters = (unsigned*)p;
    w.planes = (uint32_t*)(p + kCounterBytes);
    w.queue = (uint32_t*)(p + kCounterBytes + (size_t)n * kPlaneWords * 4);
    w.queue_cap = (size_t)n * 2 * kMaxBoard * kMaxBoard;
    w.task_cap = steal_task_cap(n);
    w.tasks = (StealTask*)(p + workspace_fixed_bytes(n));
    w.task_status = (int*)(p + workspace_fixed_bytes(n) + (size_t)w.task_cap * sizeof(StealTask));
    return w;
}

// ---- packed state access --------------------------------------------------------------------

constexpr int kAgelessWord = 168;          // meta plane words 8..26: rows of the stones that have no age
struct LoadedState { BB black, white, age0, age1, age2; int ko_p, player; };

__device__ __forceinline__ LoadedState load_state(const uint32_t* __restrict__ st, const Geo& g) {
    LoadedState s;
    s.black = bb_load(st + 0, g);  s.white = bb_load(st + 32, g);
    s.age0 = bb_load(st + 64, g);  s.age1 = bb_load(st + 96, g);  s.age2 = bb_load(st + 128, g);
    s.ko_p = (int)__ldg(st + 160); s.player = (int)__ldg(st + 161);
    return s;
}

__device__ __forceinline__ int p_to_q(int p, int S) { return p < 0 ? -1 : (((p / S) << 5) | (p % S)); }
__device__ __forceinline__ int q_to_p(int q, int S) { return q < 0 ? -1 : ((q >> 5) * S + (q & 31)); }

// ---- kernel 1: planes without look-ahead + ladder work items ---------------------------------

// Two warps (16 KB of shared memory) per CTA, so that a CTA still fits on an SM whose shared memory
// is otherwise taken by a convolution CTA of another stream (conv.cu: ~208 of 228 KB): the analysis
// of the next batch then runs beside the convolutions instead of in the gaps between their layers.
constexpr int kAnalyzeWarps = 2;

// What is_positional_superko reads, per position (rag_superko in the header), as device pointers.
struct SuperkoArgs {
    const u64* zob; const u64* hash; const u64* prev; const int32_t* prev_off;
    const uint32_t* masks; const int32_t* flags;
};

__device__ __forceinline__ void load_superko(const SuperkoArgs& a, int pos, const Geo& g, SuperkoRoot& r, SuperkoCtx& c) {
    const uint32_t* m = a.masks + (size_t)pos * 96;
    const int flags = a.flags[pos];
    c.even = bb_load(m, g); c.odd = bb_load(m + 32, g); r.handi = bb_load(m + 64, g);
    c.hash = a.hash[pos]; c.hist_len = flags >> 8;
    r.zob = a.zob; r.prev = a.prev + a.prev_off[pos]; r.nprev = a.prev_off[pos + 1] - a.prev_off[pos];
    r.has_handicap = (flags & 2) != 0;
    r.enforce = (flags & 1) != 0;
}

template <bool kSuperko>
__global__ void __launch_bounds__(kAnalyzeWarps * 32)
analyze_kernel(const uint32_t* __restrict__ states, int n, int S, int with_ladders, Workspace ws, SuperkoArgs ska)
{
    __shared__ uint32_t sm_planes[kAnalyzeWarps][kPlaneWords];
    __shared__ EyeStack sm_eye[kAnalyzeWarps];
    const int warp = threadIdx.x >> 5;
    const int pos = blockIdx.x * kAnalyzeWarps + warp;
    if (pos >= n) return;
    const Geo g = make_geo(S);
    uint32_t* planes = sm_planes[warp];
    for (int k = 0; k < kNumPlanes; k++) planes[k * 32 + g.lane] = 0u;
    __syncwarp();

    const LoadedState s = load_state(states + (size_t)pos * RAG_STATE_WORDS, g);
    const BB own = s.player > 0 ? s.black : s.white;
    const BB opp = s.player > 0 ? s.white : s.black;
    BB forbid = bb_zero();
    if (kSuperko) {
        SuperkoRoot r; SuperkoCtx c;
        load_superko(ska, pos, g, r, c);
        forbid = sk_forbidden_moves(g, r, c, own, opp, p_to_q(s.ko_p, S), s.player > 0);
    }
    const Analysis A = bb_features(g, own, opp, s.age0, s.age1, s.age2, p_to_q(s.ko_p, S), planes, sm_eye[warp], forbid);
    // stones without an age (stone_ages -1 under a stone: what GameState.copy() leaves behind, go.py:186-217)
    // are in no turns_since plane (preprocessing.py:33-36)
    const uint32_t ageless = g.lane < 24u ? __ldg(states + (size_t)pos * RAG_STATE_WORDS + kAgelessWord + g.lane) : 0u;
    if (__any_sync(kFullMask, ageless != 0u))
        for (int k = 0; k < 8; k++) planes[(kPlaneTurnsSince + k) * 32 + g.lane] &= ~ageless;
    if (s.player > 0) plane_or(planes, kPlaneColor, bb_board(g), g);
    __syncwarp();

    uint32_t* out = ws.planes + (size_t)pos * kPlaneWords;
    for (int k = 0; k < kNumPlanes; k++) out[k * 32 + g.lane] = planes[k * 32 + g.lane];

    if (with_ladders) {
        // is_ladder_capture needs an adjacent enemy group with exactly 2 liberties, is_ladder_escape
        // an adjacent own group with exactly 1 (go.py:358-361, 425-428); only those moves are searched
        BB cand[2] = { A.legal & bb_nbrs(A.opp_lib2, g), A.legal & bb_nbrs(A.own_atari, g) };
        const int total = bb_count(cand[0]) + bb_count(cand[1]);
        if (total > 0) {
            unsigned base = 0;
            if (g.lane == 0) base = atomicAdd(&ws.counters[kCtlQueued], (unsigned)total);
            base = __shfl_sync(kFullMask, base, 0);
            for (int kind = 0; kind < 2; kind++) {
                BB rem = cand[kind];
                for (;;) {
                    int q = bb_first(rem);
                    if (q < 0) break;
                    rem = andnot(rem, bb_single(q, g));
                    if (g.lane == 0 && base < ws.queue_cap)
                        ws.queue[base] = ((uint32_t)pos << 11) | ((uint32_t)kind << 10) | (uint32_t)q;
                    base++;
                }
            }
        }
    }
}

// ---- kernel 2: ladder reader -----------------------------------------------------------------
//
// One warp per work item, depth-first (go_bb.cuh).  Most searches try a handful of moves; about one
// in a thousand -- long chases along the first line of an open board -- tries thousands, and on a
// single warp that one search used to set the time of the whole batch.  Such a search is a spine
// with side chases hanging off it, so once a search has tried kStealAfter moves it starts to HAND
// OVER replies: when it opens a node with several replies it keeps the first and posts the others
// as tasks (position + move) that any idle warp may take.  The node's value is still "no reply has
// value 1": a reply of value 1 settles it at once (the handed-over replies are cancelled); otherwise
// the search takes back what nobody took and waits for the rest.  Warps that run out of work items
// serve the task ring until nobody is busy.  Nothing here changes a result -- only who computes it.

constexpr int kStealAfterDefault = 96;   // moves tried before a search starts handing over replies
constexpr int kMaxPosted = 3;            // replies handed over per node
// The reference's recursion (go.py:329-463) is bounded in depth (80) but not in size: on a rare position
// (one in ~4 million random-playout positions met so far) the tree of replies is astronomically large and
// the reference itself does not return in any practical time.  Here the search for one (position, move) item
// gives up -- "not a ladder" -- once it has tried this many moves in total, over all the warps that work on
// it (they count in steps of 16 on the item's queue slot, which is free once the item has been read).  The
// largest search met in 8.4 million random-playout positions (tools/ladder_longest.py) tried 46,544 moves.
// rag_features_ladder_cutoffs counts the items cut off.
constexpr int kLadderNodeBudget = 1 << 18;
constexpr int kLadderBudget = 3;         // return value of a search that ran out of budget

__device__ __forceinline__ int ld_vol(const int* p) { return *(const volatile int*)p; }
__device__ __forceinline__ unsigned ld_vol(const unsigned* p) { return *(const volatile unsigned*)p; }

struct StealFrames {                     // which tasks each stack frame has out (uniform across lanes)
    int16_t slot[kLadderDepth][kMaxPosted];
    uint8_t count[kLadderDepth];
    uint8_t offered[kLadderDepth];       // a frame hands replies over at most once
};

// lane 0 decides, the answer is broadcast
#define LANE0_INT(expr) __shfl_sync(kFullMask, (g.lane == 0) ? (int)(expr) : 0, 0)

__device__ __forceinline__ int steal_post(const Workspace& ws, const Geo& g, BB H, BB P, BB G, int ko, int move,
                                          bool is_capture, int remaining, int root) {
    int slot = -1;
    if (g.lane == 0) {
        for (int probe = 0; probe < 8 && slot < 0; probe++) {
            unsigned c = atomicAdd(&ws.counters[kCtlAllocCursor], 1u) % ws.task_cap;
            if (atomicCAS(&ws.task_status[c], kTaskFree, kTaskTaken) == kTaskFree) slot = (int)c;   // reserved while it is filled
        }
    }
    slot = __shfl_sync(kFullMask, slot, 0);
    if (slot < 0) return -1;
    StealTask* t = ws.tasks + slot;
    t->H[g.lane] = H.w; t->P[g.lane] = P.w; t->G[g.lane] = G.w;
    if (g.lane == 0) { t->ko = ko; t->move = move; t->is_capture = is_capture ? 1 : 0; t->remaining = remaining; t->root = root; }
    __threadfence();
    __syncwarp();
    if (g.lane == 0) {
        atomicAdd(&ws.counters[kCtlPosted], 1u);
        atomicExch(&ws.task_status[slot], kTaskPosted);
    }
    return slot;
}

// A task leaves the "posted" state exactly once (taken, taken back, or cancelled); whoever moves
// it keeps the count of posted tasks right.
__device__ __forceinline__ bool steal_claim(const Workspace& ws, int slot, int new_status) {      // lane 0 only
    if (atomicCAS(&ws.task_status[slot], kTaskPosted, new_status) != kTaskPosted) return false;
    atomicSub(&ws.counters[kCtlPosted], 1u);
    return true;
}

// Cancels a task this warp posted.  Whoever holds it last frees the slot.
__device__ __forceinline__ void steal_cancel(const Workspace& ws, int slot) {      // lane 0 only
    int* st = ws.task_status + slot;
    for (;;) {
        const int v = ld_vol(st);
        if (v == kTaskPosted) { if (steal_claim(ws, slot, kTaskFree)) return; }
        else if (v == kTaskTaken) { if (atomicCAS(st, kTaskTaken, kTaskCancelled) == kTaskTaken) return; }
        else { atomicExch(st, kTaskFree); return; }                  // done: nobody else will touch it
    }
}

struct TaskCancelled {                    // "has the task I am working on been cancelled?"
    const int* status;                    // null for a search that works on a queue item
    __device__ __forceinline__ bool operator()() const {
        int c = 0;
        if ((threadIdx.x & 31u) == 0u) c = status && ld_vol(status) == kTaskCancelled;
        return __shfl_sync(kFullMask, c, 0) != 0;
    }
};

// ladder_dfs (go_bb.cuh) with hand-over of replies.  `nodes` = moves tried so far by this search.
__device__ __noinline__ int ladder_dfs_steal(const Workspace& ws, const Geo& g, BB H, BB P, BB G, int ko, BB pending,
                                             bool frame_is_capture, int remaining_children, int nodes, int steal_after,
                                             LadderStack& st, StealFrames& sf, const int* own_status, long long t_deadline, int root)
{
    TaskCancelled cancelled{own_status};
    int d = 0;
    st.hunter[0] = H; st.prey[0] = P; st.group[0] = G; st.pending[0] = pending; st.ko[0] = (int16_t)ko;
    sf.count[0] = 0; sf.offered[0] = 0;
    for (;;) {
        int v = -1;
        // Once the search is long, every frame it stands on -- also the old, shallow ones it comes
        // back to, whose remaining replies are the biggest pieces of work -- offers its extra replies.
        if (nodes >= steal_after && !sf.offered[d]) {
            sf.offered[d] = 1;
            // keep the first reply, post up to kMaxPosted others
            BB rest = andnot(st.pending[d], bb_single(bb_first(st.pending[d]) >= 0 ? bb_first(st.pending[d]) : 0, g));
            const bool node_is_capture = frame_is_capture ^ ((d & 1) != 0);
            while (sf.count[d] < kMaxPosted) {
                const int m = bb_first(rest);
                if (m < 0) break;
                const int slot = steal_post(ws, g, st.hunter[d], st.prey[d], st.group[d], st.ko[d], m, !node_is_capture,
                                            remaining_children - d, root);
                if (slot < 0) break;
                rest = andnot(rest, bb_single(m, g));
                st.pending[d] = andnot(st.pending[d], bb_single(m, g));
                sf.slot[d][sf.count[d]++] = (int16_t)slot;
            }
        }
        const int m = bb_first(st.pending[d]);
        if (m >= 0) {
            int spent = 0;
            if ((nodes & 15) == 15) spent = LANE0_INT(atomicAdd(&ws.queue[root], 16u) + 16u);     // the item's total so far
            if ((nodes & 15) == 15 && cancelled()) { v = kLadderAborted; }
            else if (spent >= kLadderNodeBudget) { v = kLadderBudget; }
            else {
                nodes++;
                st.pending[d] = andnot(st.pending[d], bb_single(m, g));
                const bool node_is_capture = frame_is_capture ^ ((d & 1) != 0);
                H = st.hunter[d]; P = st.prey[d]; G = st.group[d]; ko = st.ko[d];
                const int res = ladder_expand(g, !node_is_capture, H, P, G, ko, m, remaining_children - d, pending);
                if (res == kLadderPush) {
                    d++;
                    st.hunter[d] = H; st.prey[d] = P; st.group[d] = G; st.pending[d] = pending; st.ko[d] = (int16_t)ko;
                    sf.count[d] = 0; sf.offered[d] = 0;
                    continue;
                }
                if (res == kLadderValue0) continue;                  // this reply does not refute; try the next
                v = 0;
            }
        } else if (sf.count[d] > 0) {
            // own replies are exhausted (all 0): collect one handed-over reply
            const int slot = sf.slot[d][sf.count[d] - 1];
            int* tstat = ws.task_status + slot;
            const int mv = LANE0_INT(ld_vol(&ws.tasks[slot].move));   // read before the slot can be reused
            const int took_back = LANE0_INT(steal_claim(ws, slot, kTaskFree));
            if (took_back) {                                          // nobody wanted it: do it here after all
                st.pending[d] = st.pending[d] | bb_single(mv, g);
                sf.count[d]--;
                continue;
            }
            int stt;
            for (;;) {                                                // somebody is on it: wait for the value
                stt = LANE0_INT(ld_vol(tstat));
                if (stt == kTaskDone0 || stt == kTaskDone1 || stt == kTaskBudget) break;
                if (cancelled()) { stt = -1; break; }
                __nanosleep(200);
                // never hang the device: past the deadline the wait is given up, the search unwinds as
                // aborted, and the workspace is marked (rag_features_status reports it)
                if (LANE0_INT(clock64() > t_deadline || ld_vol(&ws.counters[kCtlTimeout]) != 0u)) {
                    if (g.lane == 0) atomicExch(&ws.counters[kCtlTimeout], 1u);
                    stt = -1;
                    break;
                }
            }
            if (stt < 0) v = kLadderAborted;
            else {
                if (g.lane == 0) atomicExch(tstat, kTaskFree);
                sf.count[d]--;
                if (stt == kTaskDone0) continue;
                v = stt == kTaskBudget ? kLadderBudget : 0;           // 0: that reply has value 1
            }
        } else v = 1;                                                 // every reply failed

        // frame d returns v (0, 1, or kLadderAborted: unwind everything)
        for (;;) {
            if (g.lane == 0) for (int i = 0; i < sf.count[d]; i++) steal_cancel(ws, sf.slot[d][i]);
            sf.count[d] = 0;
            if (d == 0) return v;
            d--;
            if (v == kLadderAborted || v == kLadderBudget) continue;
            if (v == 0) break;                                        // parent keeps looking
            v = 0;                                                    // a child of value 1 refutes the parent
        }
    }
}

// is_ladder_capture / is_ladder_escape of one move on one prey group, with hand-over.
__device__ __forceinline__ int ladder_search_steal(const Workspace& ws, const Geo& g, BB H, BB P, BB G, int ko, int move,
                                                   bool is_capture, int remaining, int nodes, int steal_after, LadderStack& st,
                                                   StealFrames& sf, const int* own_status, long long t_deadline, int root) {
    BB pending = bb_zero();
    const int res = ladder_expand(g, is_capture, H, P, G, ko, move, remaining, pending);
    if (res != kLadderPush) return res;
    return ladder_dfs_steal(ws, g, H, P, G, ko, pending, is_capture, remaining - 1, nodes + 1, steal_after, st, sf, own_status, t_deadline, root);
}

// kBesideConv only makes two copies of the kernel, so that each can have its own shared-memory / L1
// preference (match_conv_carveout): 0 = launched by the one-call entry points, nothing else on the
// GPU, keeps the default split (large L1 for the search stacks); 1 = the pipeline's separate ladder
// stage, which runs underneath the convolutions of another stream.
template <int kBesideConv>
__global__ void __launch_bounds__(kWarpsPerBlock * 32)
ladder_kernel(const uint32_t* __restrict__ states, int S, int steal_after, int helper_ctas, long long watchdog, Workspace ws)
{
    const Geo g = make_geo(S);
    LadderStack st;
    StealFrames sf;
    const long long t_deadline = clock64() + watchdog;
    const unsigned total = min((unsigned)ws.queue_cap, ws.counters[kCtlQueued]);

    // ---- work items
    for (;;) {
        unsigned idx = 0;
        if (g.lane == 0) idx = atomicAdd(&ws.counters[kCtlTaken], 1u);
        idx = __shfl_sync(kFullMask, idx, 0);
        if (idx >= total) break;
        if (g.lane == 0) atomicAdd(&ws.counters[kCtlBusy], 1u);
        const uint32_t item = LANE0_INT(ld_vol(&ws.queue[idx]));
        if (g.lane == 0) ws.queue[idx] = 0u;                       // from here on the slot counts the moves this item's search tries
        const int pos = (int)(item >> 11), kind = (int)((item >> 10) & 1u), q = (int)(item & 1023u);
        const LoadedState s = load_state(states + (size_t)pos * RAG_STATE_WORDS, g);
        const BB own = s.player > 0 ? s.black : s.white;
        const BB opp = s.player > 0 ? s.white : s.black;
        const int ko = p_to_q(s.ko_p, S);
        const bool is_escape = kind == 1;
#ifdef RAG_LADDER_TIMING
        const long long t_item = clock64();
#endif
        // top level (go.py:355-361, 422-428): any qualifying neighbour group
        BB preys[4];
        const int np = ladder_root_preys(g, own, opp, q, is_escape, preys);
        int r = kLadderValue0;
        for (int i = 0; i < np && r == kLadderValue0; i++)
            r = is_escape ? ladder_search_steal(ws, g, opp, own, preys[i], ko, q, false, kLadderDepth, 0, steal_after, st, sf, nullptr, t_deadline, (int)idx)
                          : ladder_search_steal(ws, g, own, opp, preys[i], ko, q, true, kLadderDepth, 0, steal_after, st, sf, nullptr, t_deadline, (int)idx);
#ifdef RAG_LADDER_TIMING
        if (g.lane == 0) ws.queue[idx] = (uint32_t)((clock64() - t_item) >> 4);     // tools/ladder_profile.py
#endif
        if (g.lane == 0) {
            if (r == kLadderValue1)
                atomicOr(ws.planes + (size_t)pos * kPlaneWords + (kPlaneLadderCapture + kind) * 32 + (q >> 5), 1u << (q & 31));
            if (r == kLadderBudget) atomicAdd(&ws.counters[kCtlCutoffs], 1u);
            const unsigned spent = ld_vol(&ws.queue[idx]);
            if (spent > ld_vol(&ws.counters[kCtlMaxNodes])) atomicMax(&ws.counters[kCtlMaxNodes], spent);
            __threadfence();
            atomicSub(&ws.counters[kCtlBusy], 1u);
        }
    }

    // ---- out of items: serve handed-over replies until nobody is busy.  One warp of each of the
    // first `helper_ctas` CTAs stays for that (there are never more than a few dozen tasks in
    // flight); every other CTA exits and gives its registers back, so that the kernels of the next
    // batch can start underneath a long search.  Idle warps sleep between looks, longer each time,
    // so that they take neither issue slots nor L2 bandwidth from the searches.
    if (steal_after >= (1 << 29) || (threadIdx.x >> 5) != 0 || (int)blockIdx.x >= helper_ctas) return;
    unsigned scan = blockIdx.x * 37u;                                              // every warp starts its scans elsewhere
    unsigned nap = 500;
    for (;;) {
        int slot = -1;
        for (;;) {
            const unsigned posted = LANE0_INT(ld_vol(&ws.counters[kCtlPosted]));
            if (posted != 0u) {
                // one sweep over the status array, 32 entries at a time
                for (unsigned k = 0; k < ws.task_cap && slot < 0; k += 32) {
                    const unsigned i = (scan + k + g.lane) % ws.task_cap;
                    const unsigned hit = __ballot_sync(kFullMask, ld_vol(&ws.task_status[i]) == kTaskPosted);
                    if (hit) {
                        const int cand = (int)((scan + k + (unsigned)(__ffs(hit) - 1)) % ws.task_cap);
                        int ok = 0;
                        if (g.lane == 0) {
                            atomicAdd(&ws.counters[kCtlBusy], 1u);                 // before the claim: never "idle" while holding a task
                            ok = steal_claim(ws, cand, kTaskTaken) ? 1 : 0;
                            if (!ok) atomicSub(&ws.counters[kCtlBusy], 1u);
                        }
                        if (__shfl_sync(kFullMask, ok, 0)) slot = cand;
                    }
                }
                scan += 97u;
                if (slot >= 0) break;
            }
            const int done = LANE0_INT(ld_vol(&ws.counters[kCtlBusy]) == 0u && ld_vol(&ws.counters[kCtlPosted]) == 0u &&
                                       ld_vol(&ws.counters[kCtlTaken]) >= total);
            if (done) break;
            __nanosleep(nap);
            if (nap < 8000) nap += nap / 2;
            if (LANE0_INT(clock64() > t_deadline || ld_vol(&ws.counters[kCtlTimeout]) != 0u)) {   // never hang the device
                if (g.lane == 0) atomicExch(&ws.counters[kCtlTimeout], 1u);
                break;
            }
        }
        if (slot < 0) break;
        nap = 500;
        __threadfence();
        StealTask* t = ws.tasks + slot;
        int* tstat = ws.task_status + slot;
        const BB H{__ldcg(&t->H[g.lane])}, P{__ldcg(&t->P[g.lane])}, G{__ldcg(&t->G[g.lane])};
        const int ko = LANE0_INT(ld_vol(&t->ko)), move = LANE0_INT(ld_vol(&t->move));
        const int is_capture = LANE0_INT(ld_vol(&t->is_capture)), remaining = LANE0_INT(ld_vol(&t->remaining));
        const int root = LANE0_INT(ld_vol(&t->root));
        const int r = ladder_search_steal(ws, g, H, P, G, ko, move, is_capture != 0, remaining, 0, steal_after, st, sf, tstat, t_deadline, root);
        if (g.lane == 0) {
            __threadfence();
            if (r == kLadderAborted) atomicExch(tstat, kTaskFree);
            else if (atomicCAS(tstat, kTaskTaken, r == kLadderBudget ? kTaskBudget : r == kLadderValue1 ? kTaskDone1 : kTaskDone0) != kTaskTaken)
                atomicExch(tstat, kTaskFree);                         // cancelled while it ran: nobody waits for it
            atomicSub(&ws.counters[kCtlBusy], 1u);
        }
        __syncwarp();
    }
}

// The ladder reader for positions that enforce positional superko: the same queue, one warp per item,
// ladder_search_sk (go_bb.cuh) carrying hash, history masks and the hashes of its own line.  No reply
// hand-over here: this path serves a GTP front end, a few positions at a time.
__global__ void __launch_bounds__(kWarpsPerBlock * 32)
ladder_superko_kernel(const uint32_t* __restrict__ states, int S, Workspace ws, SuperkoArgs ska)
{
    const Geo g = make_geo(S);
    LadderStackSk st;
    const unsigned total = min((unsigned)ws.queue_cap, ws.counters[kCtlQueued]);
    for (;;) {
        unsigned idx = 0;
        if (g.lane == 0) idx = atomicAdd(&ws.counters[kCtlTaken], 1u);
        idx = __shfl_sync(kFullMask, idx, 0);
        if (idx >= total) break;
        const uint32_t item = ws.queue[idx];
        const int pos = (int)(item >> 11), kind = (int)((item >> 10) & 1u), q = (int)(item & 1023u);
        const LoadedState s = load_state(states + (size_t)pos * RAG_STATE_WORDS, g);
        const BB own = s.player > 0 ? s.black : s.white;
        const BB opp = s.player > 0 ? s.white : s.black;
        const int ko = p_to_q(s.ko_p, S);
        const bool is_escape = kind == 1, black = s.player > 0;
        SuperkoRoot r; SuperkoCtx c;
        load_superko(ska, pos, g, r, c);
        BB preys[4];
        const int np = ladder_root_preys(g, own, opp, q, is_escape, preys);
        int res = kLadderValue0;
        for (int i = 0; i < np && res == kLadderValue0; i++)
            res = is_escape ? ladder_search_sk(g, opp, own, preys[i], ko, q, false, kLadderDepth, r, c, !black, st)
                            : ladder_search_sk(g, own, opp, preys[i], ko, q, true, kLadderDepth, r, c, black, st);
        if (g.lane == 0 && res == kLadderValue1)
            atomicOr(ws.planes + (size_t)pos * kPlaneWords + (kPlaneLadderCapture + kind) * 32 + (q >> 5), 1u << (q & 31));
    }
}

// ---- kernel 3: emit ---------------------------------------------------------------------------

struct PlaneMap { int F; uint8_t src[RAG_MAX_OUT_PLANES]; };

// NCHW: out[pos][f][x*S+y]; each thread writes V consecutive elements (16 bytes).
template <typename T, int V>
__global__ void __launch_bounds__(256)
emit_nchw_kernel(const uint32_t* __restrict__ planes, long long total, int S, PlaneMap map, T* __restrict__ out)
{
    const int S2 = S * S, per_pos = map.F * S2;
    for (long long e0 = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * V; e0 < total;
         e0 += (long long)gridDim.x * blockDim.x * V) {
        long long pos = e0 / per_pos;
        int rem = (int)(e0 - pos * per_pos);
        int f = rem / S2, p = rem - f * S2;
        int x = p / S, y = p - x * S;
        const uint32_t* pl = planes + pos * kPlaneWords;
        uint32_t row = __ldg(pl + map.src[f] * 32 + x);
        __align__(16) T vals[V];
        #pragma unroll
        for (int j = 0; j < V; j++) {
            vals[j] = (T)((row >> y) & 1u);
            if (++y == S) {
                y = 0;
                if (++x == S) {
                    x = 0;
                    if (++f == map.F) { f = 0; pos++; pl += kPlaneWords; }
                }
                if (e0 + j + 1 < total) row = __ldg(pl + map.src[f] * 32 + x);
            }
        }
        if (e0 + V <= total) {
            *reinterpret_cast<uint4*>(out + e0) = *reinterpret_cast<const uint4*>(vals);
        } else {
            for (int j = 0; e0 + j < total; j++) out[e0 + j] = vals[j];
        }
    }
}

// uint8 NCHW for boards of 15 lines and more, the layout of Preprocess.state_to_tensor -- the HBM-bound
// kernel of the feature path (17,328 B out per position at F = 48, S = 19).  The output of the whole batch
// is one flat byte stream; a thread makes 16 consecutive bytes of it in registers and stores them with
// one instruction, so a warp writes 512 contiguous bytes at a time.  16 bytes never span more than two
// board rows (S >= 15): bits y.. of the row word the chunk starts in, then the low bits of the next row
// word (next row, next plane, or the next position's first plane); four bits become four bytes with one
// multiply (nibble * 0x00204081 & 0x01010101).  A CTA walks kEmitChunks consecutive chunks per thread, so
// the position of its first byte costs one 64-bit division and everything after that is 32-bit.
constexpr int kEmitChunks = 8;

template <int kS>
__global__ void __launch_bounds__(256)
emit_u8_flat_kernel(const uint32_t* __restrict__ planes, long long total, int n, int S_rt, PlaneMap map, uint8_t* __restrict__ out)
{
    const int S = kS ? kS : S_rt;
    const int S2 = S * S, per_pos = map.F * S2;
    const long long cta_base = (long long)blockIdx.x * (256 * kEmitChunks * 16);
    long long pos64 = cta_base / per_pos;
    int pos = (int)pos64;
    int off = (int)(cta_base - pos64 * per_pos) + (int)threadIdx.x * 16;          // byte offset inside position `pos`
    long long e0 = cta_base + (long long)threadIdx.x * 16;
    #pragma unroll 2
    for (int k = 0; k < kEmitChunks; k++, off += 256 * 16, e0 += 256 * 16) {
        if (e0 >= total) return;
        while (off >= per_pos) { off -= per_pos; pos++; }
        const int f = off / S2, r = off - f * S2;
        const int x = r / S, y = r - x * S;
        const uint32_t* pl = planes + (size_t)pos * kPlaneWords;
        const uint32_t row_mask = (1u << S) - 1u;
        const uint32_t w0 = __ldg(pl + map.src[f] * 32 + x) & row_mask;
        uint32_t w1 = 0;
        if (S - y < 16) {                                   // the chunk runs into the next row word
            int x1 = x + 1, f1 = f;
            const uint32_t* pl1 = pl;
            if (x1 == S) { x1 = 0; if (++f1 == map.F) { f1 = 0; pl1 += kPlaneWords; } }
            if (pl1 == pl || pos + 1 < n) w1 = __ldg(pl1 + map.src[f1] * 32 + x1);
        }
        const uint32_t bits = ((w0 >> y) | (w1 << (S - y))) & 0xFFFFu;
        uint4 v;
        v.x = ((bits & 15u) * 0x00204081u) & 0x01010101u;
        v.y = (((bits >> 4) & 15u) * 0x00204081u) & 0x01010101u;
        v.z = (((bits >> 8) & 15u) * 0x00204081u) & 0x01010101u;
        v.w = ((bits >> 12) * 0x00204081u) & 0x01010101u;
        if (e0 + 16 <= total) {
            *reinterpret_cast<uint4*>(out + e0) = v;       // (streaming / write-through hints measured no faster)
        } else {
            const uint32_t w[4] = {v.x, v.y, v.z, v.w};
            for (int j = 0; e0 + j < total; j++) out[e0 + j] = (uint8_t)(w[j >> 2] >> (8 * (j & 3)));
        }
    }
}

// uint8 NCHW for small boards (and outputs that are not 16-byte aligned): one CTA per position.  The F x S*S bytes
// of a position are assembled in shared memory (each thread expands (plane, row) words into S
// bytes), then leave as one contiguous, 16-byte-vectorised stream -- the HBM-bound kernel of the
// feature path (17,328 B out and 3.6 KB of row words in per position at F = 48).
__global__ void __launch_bounds__(256)
emit_u8_kernel(const uint32_t* __restrict__ planes, int n, int S, PlaneMap map, uint8_t* __restrict__ out)
{
    extern __shared__ __align__(16) uint8_t sm_bytes[];
    const int S2 = S * S, per_pos = map.F * S2;
    for (int pos = blockIdx.x; pos < n; pos += gridDim.x) {
        const uint32_t* pl = planes + (size_t)pos * kPlaneWords;
        for (int t = threadIdx.x; t < map.F * S; t += blockDim.x) {
            const int f = t / S, x = t - f * S;
            uint32_t w = __ldg(pl + map.src[f] * 32 + x);
            uint8_t* dst = sm_bytes + f * S2 + x * S;
            for (int y = 0; y < S; y++) dst[y] = (uint8_t)((w >> y) & 1u);
        }
        __syncthreads();
        uint8_t* o = out + (size_t)pos * per_pos;
        // the position's block starts 16-byte aligned iff per_pos * pos is; handle the general case
        const int head = (int)((16 - ((size_t)o & 15)) & 15);
        const int body = per_pos > head ? (per_pos - head) / 16 : 0;
        if ((int)threadIdx.x < head && (int)threadIdx.x < per_pos) o[threadIdx.x] = sm_bytes[threadIdx.x];
        for (int c = threadIdx.x; c < body; c += blockDim.x) {
            const uint8_t* sp = sm_bytes + head + c * 16;
            uint4 v;
            if (head == 0) v = *reinterpret_cast<const uint4*>(sp);
            else {
                uint32_t w4[4];
                #pragma unroll
                for (int k = 0; k < 4; k++)
                    w4[k] = (uint32_t)sp[4 * k] | ((uint32_t)sp[4 * k + 1] << 8) | ((uint32_t)sp[4 * k + 2] << 16) | ((uint32_t)sp[4 * k + 3] << 24);
                v = make_uint4(w4[0], w4[1], w4[2], w4[3]);
            }
            *reinterpret_cast<uint4*>(o + head + c * 16) = v;
        }
        for (int i = head + body * 16 + threadIdx.x; i < per_pos; i += blockDim.x) o[i] = sm_bytes[i];
        __syncthreads();
    }
}

// Conv input ("pixel rows", see conv.cu): dense 16-bit rows of C channels, row pos*S*S + x*S + y,
// planes as 1.0/0.0, channels >= F zero.  A warp owns a board row x: lane l keeps the row words of
// channels 2l and 2l+1 (and 2l+64, 2l+65 when C = 128) and writes one 4-byte pair per point, so every
// store instruction of the warp is one full 128-byte line of a pixel row.
__global__ void __launch_bounds__(256)
emit_conv_kernel(const uint32_t* __restrict__ planes, int n, int S, int C, PlaneMap map,
                 uint16_t one, uint16_t* __restrict__ out)
{
    const int S2 = S * S;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, warps = blockDim.x >> 5;
    const uint32_t one_lo = one, one_hi = (uint32_t)one << 16;
    for (int pos = blockIdx.x; pos < n; pos += gridDim.x) {
        const uint32_t* pl = planes + (size_t)pos * kPlaneWords;
        uint32_t* obase = reinterpret_cast<uint32_t*>(out + (size_t)pos * S2 * C);
        for (int x = warp; x < S; x += warps) {
            for (int half = 0; half < C / 64; half++) {
                const int c0 = half * 64 + 2 * lane;
                uint32_t w0 = 0, w1 = 0;
                if (c0 < map.F) w0 = __ldg(pl + map.src[c0] * 32 + x);
                if (c0 + 1 < map.F) w1 = __ldg(pl + map.src[c0 + 1] * 32 + x);
                uint32_t* orow = obase + (size_t)x * S * (C / 2) + half * 32 + lane;
                for (int y = 0; y < S; y++) {
                    const uint32_t v = (((w0 >> y) & 1u) ? one_lo : 0u) | (((w1 >> y) & 1u) ? one_hi : 0u);
                    orow[(size_t)y * (C / 2)] = v;
                }
            }
        }
    }
}

// ---- do_move / playouts -----------------------------------------------------------------------

// `ageless`: this lane's row of the stones that have no age (every word of the meta plane is written by
// exactly one lane: 160..167 by lanes 0..7, 168..191 by lanes 0..23)
__device__ __forceinline__ void store_state(uint32_t* st, const FullState& s, int S, const Geo& g, uint32_t ageless = 0u) {
    bb_store(st + 0, s.black, g);  bb_store(st + 32, s.white, g);
    bb_store(st + 64, s.age0, g);  bb_store(st + 96, s.age1, g);  bb_store(st + 128, s.age2, g);
    uint32_t m = 0;
    if (g.lane == 0) m = (uint32_t)q_to_p(s.ko_q, S);
    if (g.lane == 1) m = (uint32_t)s.player;
    if (g.lane == 2) m = (uint32_t)S;
    if (g.lane < 8u) st[160 + g.lane] = m;
    if (g.lane < 24u) st[kAgelessWord + g.lane] = ageless;
}

__global__ void __launch_bounds__(kWarpsPerBlock * 32)
do_move_kernel(uint32_t* states, int n, int S, const int32_t* __restrict__ moves, int32_t* status)
{
    const int pos = blockIdx.x * kWarpsPerBlock + (threadIdx.x >> 5);
    if (pos >= n) return;
    const Geo g = make_geo(S);
    uint32_t* st = states + (size_t)pos * RAG_STATE_WORDS;
    const LoadedState l = load_state(st, g);
    FullState s{l.black, l.white, l.age0, l.age1, l.age2, p_to_q(l.ko_p, S), l.player};
    const int mv = moves[pos];
    bool ok = mv < S * S;
    const uint32_t ageless = g.lane < 24u ? st[kAgelessWord + g.lane] : 0u;
    if (ok) ok = bb_do_move(g, s, mv < 0 ? -1 : p_to_q(mv, S));
    // stones without an age stay so until they are captured (go.py:533 ages only stone_ages >= 0)
    if (ok) store_state(st, s, S, g, ageless & (s.black.w | s.white.w));
    if (g.lane == 0 && status) status[pos] = ok ? 0 : -1;
}

__device__ __forceinline__ uint64_t splitmix64(uint64_t& x) {
    uint64_t z = (x += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

// One warp owns `quota` consecutive output slots and plays games until they are full.
__global__ void __launch_bounds__(kWarpsPerBlock * 32)
playout_kernel(uint32_t* out, int n_positions, int S, uint64_t seed, int max_ply, int stride, int quota,
               int32_t* move_log, int log_stride)
{
    __shared__ EyeStack sm_eye[kWarpsPerBlock];
    const int warp = threadIdx.x >> 5;
    const int w = blockIdx.x * kWarpsPerBlock + warp;
    const int slot0 = w * quota;
    if (slot0 >= n_positions) return;
    const int slot_end = min(n_positions, slot0 + quota);
    const Geo g = make_geo(S);
    int slot = slot0, logged = 0;
    for (int game = 0; slot < slot_end; game++) {
        uint64_t rng = seed ^ (0xD1B54A32D192ED03ull * (uint64_t)(w + 1)) ^ (0x8CB92BA72F3D8DD7ull * (uint64_t)(game + 1));
        FullState s{bb_zero(), bb_zero(), bb_zero(), bb_zero(), bb_zero(), -1, +1};
        if (move_log && g.lane == 0 && logged < log_stride) move_log[(size_t)w * log_stride + logged] = -2;
        logged++;
        int passes = 0;
        for (int t = 0; t <= max_ply && slot < slot_end; t++) {
            if (t % stride == (w + game) % stride) {
                store_state(out + (size_t)slot * RAG_STATE_WORDS, s, S, g);
                slot++;
            }
            BB own = s.player > 0 ? s.black : s.white, opp = s.player > 0 ? s.white : s.black;
            BB legal, sensible;
            bb_moves(g, own, opp, s.ko_q, sm_eye[warp], legal, sensible);
            const int cnt = bb_count(sensible);
            int q = -1;
            if (cnt > 0) q = bb_select(sensible, (int)(splitmix64(rng) % (uint64_t)cnt));
            bb_do_move(g, s, q);
            if (move_log && g.lane == 0 && logged < log_stride)
                move_log[(size_t)w * log_stride + logged] = q_to_p(q, S);
            logged++;
            passes = q < 0 ? passes + 1 : 0;
            if (passes >= 2) break;
        }
    }
}

// Rollout to the end and winner (mcts.py:171-189 with uniformly random sensible moves; go.py:482-506).
// One warp per position: plays until the side to move has no sensible move or `limit` moves were made,
// then scores by area -- stones plus empty points whose neighbours all have one colour, komi to WHITE
// (no passes are made here, so the pass counters of get_winner do not change).  result = +1 if the
// player to move at the START wins, -1 if the opponent does, 0 on a tie.
__global__ void __launch_bounds__(kWarpsPerBlock * 32)
rollout_kernel(const uint32_t* __restrict__ states, int n, int S, uint64_t seed, int limit, float komi, int32_t* result)
{
    __shared__ EyeStack sm_eye[kWarpsPerBlock];
    const int warp = threadIdx.x >> 5;
    const int pos = blockIdx.x * kWarpsPerBlock + warp;
    if (pos >= n) return;
    const Geo g = make_geo(S);
    const LoadedState l = load_state(states + (size_t)pos * RAG_STATE_WORDS, g);
    FullState s{l.black, l.white, l.age0, l.age1, l.age2, p_to_q(l.ko_p, S), l.player};
    const int player0 = s.player;
    uint64_t rng = seed ^ (0xD1B54A32D192ED03ull * (uint64_t)(pos + 1));
    for (int t = 0; t < limit; t++) {
        const BB own = s.player > 0 ? s.black : s.white, opp = s.player > 0 ? s.white : s.black;
        BB legal, sensible;
        bb_moves(g, own, opp, s.ko_q, sm_eye[warp], legal, sensible);
        const int cnt = bb_count(sensible);
        if (cnt == 0) break;
        bb_do_move(g, s, bb_select(sensible, (int)(splitmix64(rng) % (uint64_t)cnt)));
    }
    const BB empty = andnot(bb_board(g), s.black | s.white);
    const BB eb = empty & bb_from_north(s.black, g, true) & bb_from_south(s.black, g, true) &
                  bb_from_west(s.black, g, true) & bb_from_east(s.black, g, true);
    const BB ew = andnot(empty & bb_from_north(s.white, g, true) & bb_from_south(s.white, g, true) &
                         bb_from_west(s.white, g, true) & bb_from_east(s.white, g, true), eb);
    const float black = (float)(bb_count(s.black) + bb_count(eb));
    const float white = (float)(bb_count(s.white) + bb_count(ew)) + komi;
    const int winner = black > white ? +1 : (white > black ? -1 : 0);
    if (g.lane == 0) result[pos] = winner == 0 ? 0 : (winner == player0 ? +1 : -1);
}

// ---- scratch that nobody allocates (rag_features_static) -----------------------------------------
//
// Same layout as the caller-provided workspace (counters, bit-planes, ladder queue), sized for
// kStaticMaxPositions positions, but living in the library's own device image: every GPU that loads
// the library carries these ~2.4 MB until the process exits, and no allocation call is ever needed.
constexpr int kStaticMaxPositions = 256;
constexpr size_t kStaticScratchBytes =                              // = workspace_bytes(kStaticMaxPositions)
    ((size_t)kCounterBytes + (size_t)kStaticMaxPositions * (kPlaneWords * 4 + 2 * kMaxBoard * kMaxBoard * 4) + 511) / 512 * 512
    + (size_t)(kStaticMaxPositions / 2) * (sizeof(StealTask) + 4);
static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchBytes];

// The kernels that run next to the convolutions in a pipeline (analysis, the ladder reader of
// rag_features_ladders, conv-layout emit) ask for the same shared-memory / L1 split as the convolution kernel uses (all of it shared
// memory).  An SM has one split at a time and changes it only when it is empty: with different
// preferences, a convolution CTA could not start on an SM where a ladder warp was still searching,
// and -- the tiles of a layer being dealt out statically -- every layer then waited for a second
// round of CTAs.  (Measured at batch 256: forward 0.70 ms alone, 1.07 ms beside the ladder kernels, 0.73 ms with
// matching splits.)
static int match_conv_carveout() {
    static std::mutex mu;
    static bool done[64] = {false};
    int dev = -1;
    RAG_CUDA_CHECK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return fail(RAG_EUNSUPPORTED, "device ordinal %d not supported", dev);
    std::lock_guard<std::mutex> lock(mu);
    if (done[dev]) return RAG_OK;
    // all shared memory, like the convolutions.  (Measured: asking for the 196 KB split instead -- enough for the
    // CTA-pair convolution, 60 KB of L1 for the search stacks -- made the pipelined step 4 % slower, not faster.)
    static const int pct = getenv("RAG_FEATURE_CARVEOUT") ? atoi(getenv("RAG_FEATURE_CARVEOUT")) : (int)cudaSharedmemCarveoutMaxShared;
    RAG_CUDA_CHECK(cudaFuncSetAttribute(analyze_kernel<false>, cudaFuncAttributePreferredSharedMemoryCarveout, pct));
    RAG_CUDA_CHECK(cudaFuncSetAttribute(ladder_kernel<1>, cudaFuncAttributePreferredSharedMemoryCarveout, pct));
    RAG_CUDA_CHECK(cudaFuncSetAttribute(emit_conv_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, pct));
    done[dev] = true;
    return RAG_OK;
}

// In host code `g_static_scratch` is only a handle for the runtime, not an address a kernel could
// use; the device address has to be asked for, per device, with that device current.
static int static_scratch_address(void** out) {
    static std::mutex mu;
    static void* per_device[64] = {nullptr};
    int dev = -1;
    RAG_CUDA_CHECK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return fail(RAG_EUNSUPPORTED, "device ordinal %d not supported", dev);
    std::lock_guard<std::mutex> lock(mu);
    if (!per_device[dev]) {
        void* p = nullptr;
        RAG_CUDA_CHECK(cudaGetSymbolAddress(&p, g_static_scratch));
        if (!p) return fail(RAG_ECUDA, "cudaGetSymbolAddress returned a null address on device %d", dev);
        per_device[dev] = p;
    }
    *out = per_device[dev];
    return RAG_OK;
}

// batch_eval_state's host loop (policy.py:46-52) on the device: for every position the moves of
// get_legal_moves() in the reference's order -- legal non-eye points x-major, then legal own eyes
// x-major (go.py:465-480) -- with the network's probabilities renormalised over them
// (_select_moves_and_normalize, policy.py:15-25).  One warp per position, lane r = board row r.
__global__ void __launch_bounds__(128)
move_dist_kernel(const float* __restrict__ probs, const uint32_t* __restrict__ planes, int n, int S, int include_eyes,
                 int16_t* __restrict__ moves, float* __restrict__ out, int32_t* __restrict__ counts)
{
    const int pos = blockIdx.x * 4 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (pos >= n) return;
    const int P = S * S;
    const uint32_t* pl = planes + (size_t)pos * kPlaneWords;
    const uint32_t row_mask = (1u << S) - 1u;
    uint32_t first = 0, second = 0;
    if (lane < S) {
        first = __ldg(pl + kPlaneSensibleness * 32 + lane) & row_mask;
        if (include_eyes) second = __ldg(pl + kPlaneLegal * 32 + lane) & row_mask & ~first;
    }
    const float* pr = probs + (size_t)pos * P;
    int16_t* mv = moves + (size_t)pos * P;
    float* po = out + (size_t)pos * P;
    int base = 0;
    float sum = 0.f;
    #pragma unroll
    for (int pass = 0; pass < 2; pass++) {
        uint32_t w = pass == 0 ? first : second;
        const int mine = __popc(w);
        int incl = mine;
        for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
        int idx = base + incl - mine;
        while (w) {
            const int y = __ffs(w) - 1;
            w &= w - 1;
            const int p = lane * S + y;
            const float v = pr[p];
            mv[idx] = (int16_t)p;
            po[idx] = v;
            sum += v;
            idx++;
        }
        base += __shfl_sync(0xffffffffu, incl, 31);
    }
    for (int o = 16; o; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    __syncwarp();
    for (int i = lane; i < base; i += 32) po[i] = po[i] / sum;
    if (lane == 0) counts[pos] = base;
}

// ---- move selection for batched players (ai.py:33-45, 84-133) ----------------------------------
//
// One warp per position: the network's probabilities restricted to the sensible moves (plane 46 of
// the analysis: legal and not an own eye -- get_legal_moves(include_eyes=False)), renormalised,
// raised to beta = 1/temperature (apply_temperature, ai.py:57-66), then either the most probable
// move (greedy; first in x-major order on ties, as max() over the move list picks) or a draw.
// -1 = pass, when no sensible move exists.
__global__ void __launch_bounds__(128)
select_moves_kernel(const float* __restrict__ probs, const uint32_t* __restrict__ planes, int n, int S, float beta,
                    int greedy, unsigned long long seed, int32_t* __restrict__ moves)
{
    const int pos = blockIdx.x * 4 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (pos >= n) return;
    const int P = S * S;
    const uint32_t* sens = planes + (size_t)pos * kPlaneWords + kPlaneSensibleness * 32;
    const float* pr = probs + (size_t)pos * P;
    // log-space weights of my points (12 per lane cover 19x19)
    float lw[12];
    float mx = -INFINITY;
    #pragma unroll
    for (int j = 0; j < 12; j++) {
        const int i = lane + 32 * j;
        lw[j] = -INFINITY;
        if (i < P) {
            const int x = i / S, y = i - x * S;
            if ((__ldg(sens + x) >> y) & 1u) lw[j] = beta * logf(fmaxf(pr[i], 1e-37f));
        }
        mx = fmaxf(mx, lw[j]);
    }
    for (int o = 16; o; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if (mx == -INFINITY) { if (lane == 0) moves[pos] = -1; return; }          // nothing sensible: pass
    if (greedy) {
        int best = 1 << 30;
        #pragma unroll
        for (int j = 0; j < 12; j++) if (lw[j] == mx) best = min(best, lane + 32 * j);
        for (int o = 16; o; o >>= 1) best = min(best, __shfl_xor_sync(0xffffffffu, best, o));
        if (lane == 0) moves[pos] = best;
        return;
    }
    float w[12], mine = 0.f;
    #pragma unroll
    for (int j = 0; j < 12; j++) { w[j] = lw[j] == -INFINITY ? 0.f : expf(lw[j] - mx); mine += w[j]; }
    float incl = mine;                                                         // inclusive scan over lanes
    for (int o = 1; o < 32; o <<= 1) { float t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
    const float total = __shfl_sync(0xffffffffu, incl, 31);
    // one uniform number per (seed, position): splitmix64
    unsigned long long z = seed + 0x9E3779B97F4A7C15ull * (unsigned long long)(pos + 1);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull; z = (z ^ (z >> 27)) * 0x94D049BB133111EBull; z ^= z >> 31;
    const float u = (float)((z >> 40) + 0.5) * (1.0f / 16777216.0f) * total;   // in (0, total)
    // the points are ordered lane-major here (lane 0 owns 0, 32, 64 ...), which is a fixed
    // permutation of the move list: the draw is exact for the distribution either way
    const float before = incl - mine;
    const bool in_me = u >= before && u < incl;
    const unsigned who = __ballot_sync(0xffffffffu, in_me);
    int chosen = -1;
    const int src_lane = who ? __ffs(who) - 1 : -1;
    if (lane == src_lane) {
        float acc = before;
        #pragma unroll
        for (int j = 0; j < 12; j++) { if (w[j] > 0.f) { chosen = lane + 32 * j; acc += w[j]; if (u < acc) break; } }
    }
    if (src_lane < 0) {                                                        // rounding at the very end: last sensible point
        int last = -1;
        #pragma unroll
        for (int j = 0; j < 12; j++) if (w[j] > 0.f) last = max(last, lane + 32 * j);
        for (int o = 16; o; o >>= 1) last = max(last, __shfl_xor_sync(0xffffffffu, last, o));
        chosen = last;
        if (lane == 0) moves[pos] = chosen;
        return;
    }
    chosen = __shfl_sync(0xffffffffu, chosen, src_lane);
    if (lane == 0) moves[pos] = chosen;
}

static int g_steal_after = kStealAfterDefault;      // rag_set_ladder_handover
static long long g_ladder_watchdog = 20000000000ll; // rag_set_ladder_watchdog: SM clock cycles a ladder kernel may wait for handed-over work
static int g_helper_ctas = 32;                      // rag_set_ladder_helpers

static int check_size(int size) {
    if (size < 2 || size > kMaxBoard) return fail(RAG_EUNSUPPORTED, "board size %d not in [2, 19]", size);
    return RAG_OK;
}

static int build_map(const int32_t* h_plane_src, int F, PlaneMap* map) {
    if (!h_plane_src || F <= 0 || F > RAG_MAX_OUT_PLANES)
        return fail(RAG_EINVAL, "plane list must have 1..%d entries (got %d)", RAG_MAX_OUT_PLANES, F);
    map->F = F;
    for (int i = 0; i < F; i++) {
        if (h_plane_src[i] < 0 || h_plane_src[i] >= kNumPlanes)
            return fail(RAG_EINVAL, "plane_src[%d] = %d is not a canonical plane (0..%d)", i, h_plane_src[i], kNumPlanes - 1);
        map->src[i] = (uint8_t)h_plane_src[i];
    }
    return RAG_OK;
}

}  // namespace rag

using namespace rag;

extern "C" {

const char* rag_last_error(void) { return last_error_buffer(); }
int rag_abi_version(void) { return 6; }

// packs positions [b0, b1); returns 0 or the index + 1 of the first bad position (reason in *why)
static int pack_range(int b0, int b1, int size, const int8_t* h_board, const int16_t* h_age, const int32_t* h_ko,
                      const int32_t* h_player, uint32_t* h_packed, int* why)
{
    const int np = size * size;
    for (int b = b0; b < b1; b++) {
        uint32_t* st = h_packed + (size_t)b * RAG_STATE_WORDS;
        memset(st, 0, RAG_STATE_BYTES);
        const int8_t* bd = h_board + (size_t)b * np;
        const int16_t* ag = h_age + (size_t)b * np;
        for (int x = 0; x < size; x++) {
            uint32_t black = 0, white = 0, a0 = 0, a1 = 0, a2 = 0, ageless = 0;
            for (int y = 0; y < size; y++) {
                const int v = bd[x * size + y];
                if (v == 0) continue;
                if (v != 1 && v != -1) { *why = 1; return b + 1; }
                const uint32_t bit = 1u << y;
                if (v > 0) black |= bit; else white |= bit;
                int a = ag[x * size + y];
                if (a < 0) ageless |= bit;                       // a stone GameState.copy() left without an age
                a = a < 0 ? 0 : (a > 7 ? 7 : a);
                if (a & 1) a0 |= bit;
                if (a & 2) a1 |= bit;
                if (a & 4) a2 |= bit;
            }
            st[x] = black; st[32 + x] = white; st[64 + x] = a0; st[96 + x] = a1; st[128 + x] = a2; st[kAgelessWord + x] = ageless;
        }
        if (h_ko[b] >= np) { *why = 2; return b + 1; }
        if (h_player[b] != 1 && h_player[b] != -1) { *why = 3; return b + 1; }
        st[160] = (uint32_t)(h_ko[b] < 0 ? -1 : h_ko[b]);
        st[161] = (uint32_t)h_player[b];
        st[162] = (uint32_t)size;
    }
    return 0;
}

int rag_pack_states(int n, int size, const int8_t* h_board, const int16_t* h_age, const int32_t* h_ko,
                    const int32_t* h_player, uint32_t* h_packed)
{
    if (n < 0 || !h_board || !h_age || !h_ko || !h_player || !h_packed) return fail(RAG_EINVAL, "null argument");
    if (int rc = check_size(size)) return rc;
    // large batches (a 4,096-state batch_eval_state call) are split over a few host threads
    int nt = n >= 1024 ? (int)std::thread::hardware_concurrency() : 1;
    nt = nt < 1 ? 1 : (nt > 8 ? 8 : nt);
    std::vector<int> bad(nt, 0), why(nt, 0);
    if (nt == 1) bad[0] = pack_range(0, n, size, h_board, h_age, h_ko, h_player, h_packed, &why[0]);
    else {
        std::vector<std::thread> th;
        const int per = (n + nt - 1) / nt;
        for (int t = 0; t < nt; t++) {
            const int b0 = t * per, b1 = b0 + per < n ? b0 + per : n;
            th.emplace_back([=, &bad, &why]() { if (b0 < b1) bad[t] = pack_range(b0, b1, size, h_board, h_age, h_ko, h_player, h_packed, &why[t]); });
        }
        for (auto& t : th) t.join();
    }
    for (int t = 0; t < nt; t++)
        if (bad[t]) {
            const int b = bad[t] - 1;
            if (why[t] == 1) return fail(RAG_EINVAL, "board[%d] holds a value other than -1, 0, +1", b);
            if (why[t] == 2) return fail(RAG_EINVAL, "ko[%d] = %d out of range", b, h_ko[b]);
            return fail(RAG_EINVAL, "player[%d] = %d", b, h_player[b]);
        }
    return RAG_OK;
}

int rag_unpack_states(int n, int size, const uint32_t* h_packed, int8_t* h_board, int16_t* h_age,
                      int32_t* h_ko, int32_t* h_player)
{
    if (n < 0 || !h_packed || !h_board || !h_age || !h_ko || !h_player) return fail(RAG_EINVAL, "null argument");
    if (int rc = check_size(size)) return rc;
    const int np = size * size;
    for (int b = 0; b < n; b++) {
        const uint32_t* st = h_packed + (size_t)b * RAG_STATE_WORDS;
        for (int x = 0; x < size; x++)
            for (int y = 0; y < size; y++) {
                int bl = (st[x] >> y) & 1u, wh = (st[32 + x] >> y) & 1u;
                h_board[(size_t)b * np + x * size + y] = (int8_t)(bl ? 1 : (wh ? -1 : 0));
                int a = ((st[64 + x] >> y) & 1u) | (((st[96 + x] >> y) & 1u) << 1) | (((st[128 + x] >> y) & 1u) << 2);
                const int ageless = (st[kAgelessWord + x] >> y) & 1u;
                h_age[(size_t)b * np + x * size + y] = (int16_t)(((bl || wh) && !ageless) ? a : -1);
            }
        h_ko[b] = (int32_t)st[160];
        h_player[b] = (int32_t)st[161];
    }
    return RAG_OK;
}

size_t rag_features_workspace_bytes(int n) { return n < 0 ? 0 : workspace_bytes(n); }

int rag_features_analyze(const uint32_t* d_states, int n, int size, int with_ladders, void* d_workspace,
                         size_t workspace_bytes_given, void* stream)
{
    if (n < 0 || (n > 0 && (!d_states || !d_workspace))) return fail(RAG_EINVAL, "null argument");
    if (int rc = check_size(size)) return rc;
    if (n > kMaxPositionsPerLaunch) return fail(RAG_EUNSUPPORTED, "n = %d exceeds %d positions per call", n, kMaxPositionsPerLaunch);
    if (n == 0) return RAG_OK;
    if (workspace_bytes_given < workspace_bytes(n))
        return fail(RAG_ENOMEM, "workspace has %zu bytes, %zu needed for n = %d", workspace_bytes_given, workspace_bytes(n), n);
    cudaStream_t s = (cudaStream_t)stream;
    if (int rc = match_conv_carveout()) return rc;
    Workspace ws = carve(d_workspace, n);
    RAG_CUDA_CHECK(cudaMemsetAsync(ws.counters, 0, kCounterBytes, s));
    RAG_CUDA_CHECK(cudaMemsetAsync(ws.tasks, 0, (size_t)ws.task_cap * (sizeof(StealTask) + 4), s));     // all tasks free, ring empty
    analyze_kernel<false><<<(n + kAnalyzeWarps - 1) / kAnalyzeWarps, kAnalyzeWarps * 32, 0, s>>>(d_states, n, size, with_ladders, ws, SuperkoArgs());
    RAG_CUDA_CHECK(cudaGetLastError());
    if (with_ladders == 1) {
        int dev = 0, sms = 148;
        RAG_CUDA_CHECK(cudaGetDevice(&dev));
        RAG_CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
        // Persistent warps, each looping until the queue is drained.  This one-call path has the
        // SMs to itself: 8 CTAs (32 warps) per SM, measured 3 % faster than 4.  (The pipeline's
        // separate stage, rag_features_ladders, stays at 4 so that a convolution CTA of another
        // stream still finds registers beside it.)
        const int cap = 8;
        long long want = ((long long)n * 16 + kWarpsPerBlock - 1) / kWarpsPerBlock;
        int blocks = (int)(want < (long long)sms * cap ? (want < 1 ? 1 : want) : (long long)sms * cap);
        ladder_kernel<0><<<blocks, kWarpsPerBlock * 32, 0, s>>>(d_states, size, g_steal_after, g_helper_ctas, g_ladder_watchdog, ws);
        RAG_CUDA_CHECK(cudaGetLastError());
    }
    return RAG_OK;
}

int rag_features_ladders(const uint32_t* d_states, int n, int size, void* d_workspace, size_t workspace_bytes_given,
                         void* stream)
{
    if (n < 0 || (n > 0 && (!d_states || !d_workspace))) return fail(RAG_EINVAL, "null argument");
    if (int rc = check_size(size)) return rc;
    if (n > kMaxPositionsPerLaunch) return fail(RAG_EUNSUPPORTED, "n = %d exceeds %d positions per call", n, kMaxPositionsPerLaunch);
    if (n == 0) return RAG_OK;
    if (workspace_bytes_given < workspace_bytes(n))
        return fail(RAG_ENOMEM, "workspace has %zu bytes, %zu needed for n = %d", workspace_bytes_given, workspace_bytes(n), n);
    if (int rc = match_conv_carveout()) return rc;
    Workspace ws = carve(d_workspace, n);
    int dev = 0, sms = 148;
    RAG_CUDA_CHECK(cudaGetDevice(&dev));
    RAG_CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    long long want = ((long long)n * 16 + kWarpsPerBlock - 1) / kWarpsPerBlock;
    static const int per_sm = getenv("RAG_LADDER_CTAS_PER_SM") ? atoi(getenv("RAG_LADDER_CTAS_PER_SM")) : 4;     // experiment knob
    int blocks = (int)(want < (long long)sms * per_sm ? (want < 1 ? 1 : want) : (long long)sms * per_sm);
    ladder_kernel<1><<<blocks, kWarpsPerBlock * 32, 0, (cudaStream_t)stream>>>(d_states, size, g_steal_after, g_helper_ctas, g_ladder_watchdog, ws);
    RAG_CUDA_CHECK(cudaGetLastError());
    return RAG_OK;
}

int rag_features_analyze_superko(const uint32_t* d_states, int n, int size, int with_ladders, const rag_superko* sk,
                                 void* d_workspace, size_t workspace_bytes_given, void* stream)
{
    if (!sk) return fail(RAG_EINVAL, "null superko description");
    if (n < 0 || (n > 0 && (!d_states || !d_workspace))) return fail(RAG_EINVAL, "null argument");
    if (n > 0 && (!sk->d_zobrist || !sk->d_hash || !sk->d_prev || !sk->d_prev_off || !sk->d_masks || !sk->d_flags))
        return fail(RAG_EINVAL, "null pointer in the superko description");
    if (int rc = check_size(size)) return rc;
    if (with_ladders != 0 && with_ladders != 1) return fail(RAG_EINVAL, "with_ladders must be 0 or 1 here");
    if (n > kMaxPositionsPerLaunch) return fail(RAG_EUNSUPPORTED, "n = %d exceeds %d positions per call", n, kMaxPositionsPerLaunch);
    if (n == 0) return RAG_OK;
    if (workspace_bytes_given < workspace_bytes(n))
        return fail(RAG_ENOMEM, "workspace has %zu bytes, %zu needed for n = %d", workspace_bytes_given, workspace_bytes(n), n);
    cudaStream_t s = (cudaStream_t)stream;
    Workspace ws = carve(d_workspace, n);
    SuperkoArgs a;
    a.zob = (const u64*)sk->d_zobrist; a.hash = (const u64*)sk->d_hash; a.prev = (const u64*)sk->d_prev;
    a.prev_off = sk->d_prev_off; a.masks = sk->d_masks; a.flags = sk->d_flags;
    RAG_CUDA_CHECK(cudaMemsetAsync(ws.counters, 0, kCounterBytes, s));
    analyze_kernel<true><<<(n + kAnalyzeWarps - 1) / kAnalyzeWarps, kAnalyzeWarps * 32, 0, s>>>(d_states, n, size, with_ladders, ws, a);
    RAG_CUDA_CHECK(cudaGetLastError());
    if (with_ladders) {
        int dev = 0, sms = 148;
        RAG_CUDA_CHECK(cudaGetDevice(&dev));
        RAG_CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
        long long want = ((long long)n * 16 + kWarpsPerBlock - 1) / kWarpsPerBlock;
        int blocks = (int)(want < (long long)sms * 4 ? (want < 1 ? 1 : want) : (long long)sms * 4);
        ladder_superko_kernel<<<blocks, kWarpsPerBlock * 32, 0, s>>>(d_states, size, ws, a);
        RAG_CUDA_CHECK(cudaGetLastError());
    }
    return RAG_OK;
}

int rag_features_emit(const void* d_workspace, int n, int size, const int32_t* h_plane_src, int F, int layout,
                      int conv_dtype, int conv_channels, void* d_out, void* stream)
{
    if (n < 0 || (n > 0 && (!d_workspace || !d_out))) return fail(RAG_EINVAL, "null argument");
    if (int rc = check_size(size)) return rc;
    PlaneMap map;
    if (int rc = build_map(h_plane_src, F, &map)) return rc;
    if (n == 0) return RAG_OK;
    cudaStream_t s = (cudaStream_t)stream;
    const uint32_t* planes = (const uint32_t*)((const char*)d_workspace + kCounterBytes);
    const long long total = (long long)n * F * size * size;
    if (layout == RAG_LAYOUT_NCHW_U8) {
        const size_t smem = (size_t)F * size * size + 16;
        if (size >= 15 && ((size_t)d_out & 15) == 0) {
            const long long per_cta = 256ll * kEmitChunks * 16;
            const long long blocks = (total + per_cta - 1) / per_cta;
            if (blocks > 0x7fffffffll) return fail(RAG_EUNSUPPORTED, "batch too large for one emit launch");
            if (size == 19) emit_u8_flat_kernel<19><<<(int)blocks, 256, 0, s>>>(planes, total, n, size, map, (uint8_t*)d_out);
            else emit_u8_flat_kernel<0><<<(int)blocks, 256, 0, s>>>(planes, total, n, size, map, (uint8_t*)d_out);
        } else if (smem <= 48 * 1024) {
            int blocks = n < 148 * 8 ? n : 148 * 8;
            emit_u8_kernel<<<blocks, 256, smem, s>>>(planes, n, size, map, (uint8_t*)d_out);
        } else {
            long long threads = (total + 15) / 16;
            int blocks = (int)((threads + 255) / 256);
            emit_nchw_kernel<uint8_t, 16><<<blocks, 256, 0, s>>>(planes, total, size, map, (uint8_t*)d_out);
        }
    } else if (layout == RAG_LAYOUT_NCHW_F32) {
        long long threads = (total + 3) / 4;
        int blocks = (int)((threads + 255) / 256);
        emit_nchw_kernel<float, 4><<<blocks, 256, 0, s>>>(planes, total, size, map, (float*)d_out);
    } else if (layout == RAG_LAYOUT_CONV_IN) {
        if (conv_channels != 64 && conv_channels != 128) return fail(RAG_EINVAL, "conv_channels must be 64 or 128");
        if (F > conv_channels) return fail(RAG_EINVAL, "%d planes do not fit %d channels", F, conv_channels);
        uint16_t one = conv_dtype == RAG_DTYPE_FP16 ? (uint16_t)0x3C00 : (uint16_t)0x3F80;
        int blocks = n < 148 * 8 ? n : 148 * 8;
        emit_conv_kernel<<<blocks, 256, 0, s>>>(planes, n, size, conv_channels, map, one, (uint16_t*)d_out);
    } else {
        return fail(RAG_EINVAL, "unknown layout %d", layout);
    }
    RAG_CUDA_CHECK(cudaGetLastError());
    return RAG_OK;
}

int rag_features(const uint32_t* d_states, int n, int size, const int32_t* h_plane_src, int F, int layout,
                 int conv_dtype, int conv_channels, void* d_out, void* d_workspace,
                 size_t workspace_bytes_given, void* stream)
{
    int with_ladders = 0;
    for (int i = 0; h_plane_src && i < F; i++)
        if (h_plane_src[i] == kPlaneLadderCapture || h_plane_src[i] == kPlaneLadderEscape) with_ladders = 1;
    if (int rc = rag_features_analyze(d_states, n, size, with_ladders, d_workspace, workspace_bytes_given, stream)) return rc;
    return rag_features_emit(d_workspace, n, size, h_plane_src, F, layout, conv_dtype, conv_channels, d_out, stream);
}

int rag_set_ladder_helpers(int ctas) { int old = g_helper_ctas; g_helper_ctas = ctas < 0 ? 1 << 30 : ctas; return old; }
long long rag_set_ladder_watchdog(long long cycles) { long long old = g_ladder_watchdog; if (cycles > 0) g_ladder_watchdog = cycles; return old; }

int rag_features_status(const void* d_workspace, void* stream)
{
    if (!d_workspace) return fail(RAG_EINVAL, "null argument");
    unsigned flag = 0;
    RAG_CUDA_CHECK(cudaMemcpyAsync(&flag, (const char*)d_workspace + kCtlTimeout * sizeof(unsigned), sizeof flag, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    RAG_CUDA_CHECK(cudaStreamSynchronize((cudaStream_t)stream));
    if (flag) return fail(RAG_ECUDA, "the ladder reader gave up waiting for handed-over work (watchdog of %lld cycles): the ladder planes of this "
                                     "workspace are incomplete; run the batch again", g_ladder_watchdog);
    return RAG_OK;
}

int rag_features_ladder_longest(const void* d_workspace, void* stream)
{
    if (!d_workspace) return fail(RAG_EINVAL, "null argument");
    unsigned count = 0;
    RAG_CUDA_CHECK(cudaMemcpyAsync(&count, (const char*)d_workspace + kCtlMaxNodes * sizeof(unsigned), sizeof count, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    RAG_CUDA_CHECK(cudaStreamSynchronize((cudaStream_t)stream));
    return (int)(count > 0x7fffffffu ? 0x7fffffffu : count);
}

int rag_features_ladder_cutoffs(const void* d_workspace, void* stream)
{
    if (!d_workspace) return fail(RAG_EINVAL, "null argument");
    unsigned count = 0;
    RAG_CUDA_CHECK(cudaMemcpyAsync(&count, (const char*)d_workspace + kCtlCutoffs * sizeof(unsigned), sizeof count, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    RAG_CUDA_CHECK(cudaStreamSynchronize((cudaStream_t)stream));
    return (int)(count > 0x7fffffffu ? 0x7fffffffu : count);
}

int rag_set_ladder_handover(int moves) { int old = g_steal_after; g_steal_after = moves < 0 ? 1 << 30 : moves; return old; }

int rag_features_static_max_positions(void) { return kStaticMaxPositions; }

int rag_features_static(const uint32_t* d_states, int n, int size, const int32_t* h_plane_src, int F, int layout,
                        int conv_dtype, int conv_channels, void* d_out, void* stream)
{
    if (n > kStaticMaxPositions)
        return fail(RAG_ENOMEM, "rag_features_static handles at most %d positions per call (got %d); "
                                "split the batch or use rag_features with your own workspace", kStaticMaxPositions, n);
    static_assert(kStaticScratchBytes >= (size_t)kCounterBytes, "scratch too small");
    if (workspace_bytes(kStaticMaxPositions) != kStaticScratchBytes)
        return fail(RAG_EINVAL, "internal: static scratch size out of date");
    void* scratch = nullptr;
    if (int rc = static_scratch_address(&scratch)) return rc;
    // The array still holds whatever the previous call left.  What has to start from zero is
    // cleared on the stream by the analysis step itself (the queue counters); the bit-planes and
    // queue slots of positions < n are fully rewritten before anything reads them.
    return rag_features(d_states, n, size, h_plane_src, F, layout, conv_dtype, conv_channels, d_out,
                        scratch, kStaticScratchBytes, stream);
}

long long rag_conv_rows(int n, int size) { return conv_rows(n, size); }

int rag_move_distributions(const float* d_probs, const void* d_workspace, int n, int size, int include_eyes,
                           int16_t* d_moves, float* d_out, int32_t* d_counts, void* stream)
{
    if (n < 0 || (n > 0 && (!d_probs || !d_workspace || !d_moves || !d_out || !d_counts))) return fail(RAG_EINVAL, "null argument");
    if (int rc = check_size(size)) return rc;
    if (n == 0) return RAG_OK;
    const uint32_t* planes = (const uint32_t*)((const char*)d_workspace + kCounterBytes);
    move_dist_kernel<<<(n + 3) / 4, 128, 0, (cudaStream_t)stream>>>(d_probs, planes, n, size, include_eyes, d_moves, d_out, d_counts);
    RAG_CUDA_CHECK(cudaGetLastError());
    return RAG_OK;
}

int rag_select_moves(const float* d_probs, const void* d_workspace, int n, int size, float temperature, int greedy,
                     uint64_t seed, int32_t* d_moves, void* stream)
{
    if (n < 0 || (n > 0 && (!d_probs || !d_workspace || !d_moves))) return fail(RAG_EINVAL, "null argument");
    if (int rc = check_size(size)) return rc;
    if (!(temperature > 0.f)) return fail(RAG_EINVAL, "temperature must be positive (got %g)", (double)temperature);
    if (n == 0) return RAG_OK;
    const uint32_t* planes = (const uint32_t*)((const char*)d_workspace + kCounterBytes);
    select_moves_kernel<<<(n + 3) / 4, 128, 0, (cudaStream_t)stream>>>(d_probs, planes, n, size, 1.0f / temperature, greedy,
                                                                    (unsigned long long)seed, d_moves);
    RAG_CUDA_CHECK(cudaGetLastError());
    return RAG_OK;
}

int rag_do_move(uint32_t* d_states, int n, int size, const int32_t* d_moves, int32_t* d_status, void* stream)
{
    if (n < 0 || (n > 0 && (!d_states || !d_moves))) return fail(RAG_EINVAL, "null argument");
    if (int rc = check_size(size)) return rc;
    if (n == 0) return RAG_OK;
    do_move_kernel<<<(n + kWarpsPerBlock - 1) / kWarpsPerBlock, kWarpsPerBlock * 32, 0, (cudaStream_t)stream>>>(
        d_states, n, size, d_moves, d_status);
    RAG_CUDA_CHECK(cudaGetLastError());
    return RAG_OK;
}

int rag_random_playouts(uint32_t* d_states_out, int n_positions, int size, uint64_t seed, int max_ply, int stride,
                        int32_t* d_move_log, int log_stride, void* stream)
{
    if (n_positions < 0 || (n_positions > 0 && !d_states_out)) return fail(RAG_EINVAL, "null argument");
    if (int rc = check_size(size)) return rc;
    if (max_ply < 0 || stride < 1) return fail(RAG_EINVAL, "max_ply %d / stride %d", max_ply, stride);
    if (n_positions == 0) return RAG_OK;
    const int quota = max_ply / stride + 1;
    const int warps = (n_positions + quota - 1) / quota;
    playout_kernel<<<(warps + kWarpsPerBlock - 1) / kWarpsPerBlock, kWarpsPerBlock * 32, 0, (cudaStream_t)stream>>>(
        d_states_out, n_positions, size, seed, max_ply, stride, quota, d_move_log, d_move_log ? log_stride : 0);
    RAG_CUDA_CHECK(cudaGetLastError());
    return RAG_OK;
}

int rag_rollout_winners(const uint32_t* d_states, int n, int size, int limit, uint64_t seed, float komi,
                        int32_t* d_result, void* stream)
{
    if (n < 0 || (n > 0 && (!d_states || !d_result))) return fail(RAG_EINVAL, "null argument");
    if (int rc = check_size(size)) return rc;
    if (limit < 0) return fail(RAG_EINVAL, "limit must not be negative");
    if (n == 0) return RAG_OK;
    rollout_kernel<<<(n + kWarpsPerBlock - 1) / kWarpsPerBlock, kWarpsPerBlock * 32, 0, (cudaStream_t)stream>>>(
        d_states, n, size, seed, limit, komi, d_result);
    RAG_CUDA_CHECK(cudaGetLastError());
    return RAG_OK;
}

int rag_playout_quota(int max_ply, int stride) { return stride < 1 ? 0 : max_ply / stride + 1; }

}  // extern "C"
