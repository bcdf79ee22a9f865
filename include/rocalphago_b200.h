/*
 * rocalphago_b200.h -- C ABI of librocalphago_b200.so (sm_100a).
 *
 * The drop-in boundary for RocAlphaGo's batched position-evaluation path:
 *     AlphaGo/preprocessing/preprocessing.py   Preprocess.state_to_tensor     (:287-294)
 *     AlphaGo/models/nn_util.py                NeuralNetBase.forward          (:34-58)
 *     AlphaGo/models/policy.py                 CNNPolicy.batch_eval_state     (:27-52)
 *     AlphaGo/models/value.py                  value network forward          (:12-28)
 * The reference is pure Python and has no FFI of its own; these are the entry
 * points a ctypes/cffi stub in those four files binds (INTEGRATION.md shows the
 * stub).  rocalphago_b200/_cabi.py is that stub for this repository's own
 * Python host layer.
 *
 * Conventions
 *   - every function returns 0 on success and a negative RAG_E* code on failure;
 *     rag_last_error() gives the message of the calling thread's last failure
 *   - pointers named d_* are DEVICE pointers owned by the caller; h_* are host
 *     pointers; nothing here allocates or frees caller memory
 *   - `stream` is a cudaStream_t passed as void* (0 = the legacy default stream);
 *     calls enqueue work and return without synchronising unless stated
 *   - board size S <= 19; flat point index p = x*S + y (AlphaGo/util.py:11-13)
 */
#ifndef ROCALPHAGO_B200_H
#define ROCALPHAGO_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RAG_OK            0
#define RAG_EINVAL       (-1)   /* bad argument */
#define RAG_ECUDA        (-2)   /* a CUDA runtime/driver call failed */
#define RAG_ENOMEM       (-3)   /* workspace too small / allocation failed */
#define RAG_EUNSUPPORTED (-4)   /* shape or device not supported */

/* message for the last error returned on this thread ("" if none) */
const char* rag_last_error(void);
/* ABI version, bumped on any signature change */
int rag_abi_version(void);

/* ------------------------------------------------------------------ packed game states
 *
 * A position on the device is RAG_STATE_WORDS uint32: six planes of 32 row words
 * (row x in word x, point (x, y) in bit y): black, white, three bit-planes of
 * min(stone_ages, 7), and a meta plane (word 0 = ko point p or -1, word 1 =
 * current_player +1/-1, word 2 = board size, words 8..26 = rows of the stones that have NO age:
 * stone_ages is -1 under every stone of a GameState.copy() (go.py:186-217 does not copy it), such
 * stones are in no turns_since plane and are never aged (go.py:533) -- what MCTS feeds the networks).  This is the device twin of the
 * fields of GameState the 48 planes depend on (go.py:18-22,55).
 */
#define RAG_STATE_WORDS 192
#define RAG_STATE_BYTES (RAG_STATE_WORDS * 4)

/* Host-side packing of GameState fields (go.py:18-22,55) into the device format.
 * board int8 [n][S*S] in {+1 black, -1 white, 0}; age int16 [n][S*S] (stone_ages,
 * -1 on empties, and -1 under a stone that has no age); ko int32 [n] (x*S+y or -1); player int32 [n] (+1/-1).
 * h_packed: n * RAG_STATE_WORDS uint32 (pinned memory recommended). */
int rag_pack_states(int n, int size, const int8_t* h_board, const int16_t* h_age,
                    const int32_t* h_ko, const int32_t* h_player, uint32_t* h_packed);
/* The inverse (ages come back saturated at 7). */
int rag_unpack_states(int n, int size, const uint32_t* h_packed, int8_t* h_board, int16_t* h_age,
                      int32_t* h_ko, int32_t* h_player);

/* ------------------------------------------------------------------ feature planes
 *
 * Preprocess.state_to_tensor (preprocessing.py:287-294) for a batch.  Planes are
 * named by their index in the canonical list
 *   0-2 board, 3 ones, 4-11 turns_since, 12-19 liberties, 20-27 capture_size,
 *   28-35 self_atari_size, 36-43 liberties_after, 44 ladder_capture,
 *   45 ladder_escape, 46 sensibleness, 47 zeros, 48 legal
 * (preprocessing.py:209-263) plus 49 color (all ones iff BLACK is to move: the value
 * network's 49th input plane, which the reference leaves undefined -- value.py:16); h_plane_src lists, for each of the F output planes,
 * which canonical plane it is, so any feature subset / order / repetition works.
 */
#define RAG_NUM_PLANES 50
#define RAG_MAX_OUT_PLANES 128

#define RAG_LAYOUT_NCHW_U8   0   /* uint8  [n][F][S*S]            (game_converter.py:76's storage dtype) */
#define RAG_LAYOUT_NCHW_F32  1   /* float  [n][F][S*S]            (what NeuralNetBase.forward is fed)   */
#define RAG_LAYOUT_CONV_IN   2   /* 16-bit [rag_conv_rows(n,S)][conv_channels] dense pixel rows (NHWC): row
                                    b*S*S + x*S + y holds point (x, y) of board b; the first conv
                                    layer's TMA source; element type chosen by `conv_dtype`            */

#define RAG_DTYPE_BF16 0
#define RAG_DTYPE_FP16 1
/* fp16 tensor-core operands, every product formed from split operands (x = hi + lo): hi*hi + lo*hi +
 * hi*lo with fp32 accumulation -- fp32-like results at a third of the speed.  For callers that need the
 * reference's fp32 arg-max on nearly flat outputs (random-init networks); not for residual networks. */
#define RAG_DTYPE_FP16X3 2
/* fp16 operands, with the split above applied only to the 32 (64) input channels of every layer after
 * the first that are expected to carry the most energy (filters are reordered at load time so that
 * these come first): one (two) extra 64-deep K step(s) per filter tap, i.e. 4/3 (5/3) of the fp16
 * work for a 192-filter layer, for 0.6x (0.45x) of fp16's rounding noise in the logits.  The mode
 * that keeps the reference's fp32 arg-max on >= 99.5 % of random-init positions at near-fp16 speed. */
#define RAG_DTYPE_FP16C32 3
#define RAG_DTYPE_FP16C64 4

/* bytes of scratch rag_features / rag_features_analyze need for n positions */
size_t rag_features_workspace_bytes(int n);

/* Step 1: analyse n packed states into canonical bit-planes held in the workspace.
 * with_ladders = 0 skips the ladder reader (planes 44/45 stay zero); 1 runs it; 2 only queues
 * the ladder candidates, for a later rag_features_ladders on the same workspace (lets a caller
 * put the short analysis and the long, latency-bound ladder search on different streams).
 * The workspace must stay untouched until the emits that read it have run. */
int rag_features_analyze(const uint32_t* d_states, int n, int size, int with_ladders,
                         void* d_workspace, size_t workspace_bytes, void* stream);
/* Tuning/test knob (process-wide): after how many tried moves a ladder search starts handing
 * replies over to idle warps (default 96; 0 = from the first move; < 0 = never).  Results do not
 * depend on it.  Returns the previous value. */
int rag_set_ladder_handover(int moves);
/* Tuning knob (process-wide): how many CTAs of the ladder kernel keep one warp waiting for handed-over
 * replies once the queue is empty (default 32; < 0 = all of them).  Results do not depend on it.  Returns the
 * previous value. */
int rag_set_ladder_helpers(int ctas);

/* The ladder kernel never spins without bound: a warp that has waited longer than the watchdog (SM clock
 * cycles, default 2e10 ~ 10 s; rag_set_ladder_watchdog returns the previous value, <= 0 only reads it)
 * for work it handed to another warp gives the search up and marks the workspace.
 * rag_features_status waits for `stream` and returns RAG_ECUDA with a message if a ladder kernel that
 * wrote this workspace did so (its ladder planes are then incomplete), RAG_OK otherwise.  The Python
 * layer calls it wherever it synchronises anyway.  No reference counterpart (go.py:329-463 recurses on
 * one thread); the liveness argument is in DESIGN.md. */
long long rag_set_ladder_watchdog(long long cycles);
int rag_features_status(const void* d_workspace, void* stream);
/* The reference's ladder recursion (go.py:329-463) is bounded in depth (80) but not in size; on a rare
 * position (one in ~4 million random-playout positions so far) it does not return in any practical time.
 * The device search gives up -- the move is then "not a ladder" -- once one line of it has tried 65,536
 * moves (the longest search of the parity sets tries a few thousand).  Returns, after waiting for
 * `stream`, how many of the (position, move) searches of the ladder kernel that last wrote this workspace
 * were cut off that way (>= 0), or a negative RAG_E* code. */
int rag_features_ladder_cutoffs(const void* d_workspace, void* stream);
/* Measurement aid: the largest number of moves any one line of a ladder search had tried in the ladder kernel
 * that last wrote this workspace (what the budget above is compared with). */
int rag_features_ladder_longest(const void* d_workspace, void* stream);

/* Positional superko (go.py:242-265, 282-283) for states created with enforce_superko=True: what
 * is_positional_superko reads, per position, in device memory.  q = x*32 + y.
 *   d_zobrist  uint64 [2][1024]: hash_lookup by q, [0] BLACK, [1] WHITE           (go.py:58-62)
 *   d_hash     uint64 [n]: current_hash                                            (go.py:63)
 *   d_prev     uint64 [..]: previous_hashes of all positions, position b's in
 *              [d_prev_off[b], d_prev_off[b+1]);  d_prev_off int32 [n+1]          (go.py:64)
 *   d_masks    uint32 [n][3][32] row words (as in a packed state): the points of history[0::2],
 *              of history[1::2], and of handicaps                                 (go.py:248-256)
 *   d_flags    int32 [n]: bit 0 enforce_superko, bit 1 len(handicaps) > 0, bits 8.. len(history)
 * rag_features_analyze_superko is rag_features_analyze(with_ladders = 0 or 1) with these rules added:
 * superko-forbidden moves are illegal in every plane and move list, and the ladder look-ahead
 * enforces superko on its own line as GameState.copy() makes the reference do (go.py:193-199). */
typedef struct rag_superko {
    const uint64_t* d_zobrist;
    const uint64_t* d_hash;
    const uint64_t* d_prev;
    const int32_t*  d_prev_off;
    const uint32_t* d_masks;
    const int32_t*  d_flags;
} rag_superko;
int rag_features_analyze_superko(const uint32_t* d_states, int n, int size, int with_ladders, const rag_superko* sk,
                                 void* d_workspace, size_t workspace_bytes, void* stream);

/* Step 1b: the ladder reader over the candidates queued by rag_features_analyze(with_ladders = 2). */
int rag_features_ladders(const uint32_t* d_states, int n, int size, void* d_workspace,
                         size_t workspace_bytes, void* stream);
/* Step 2: write the listed planes from the workspace in the requested layout.
 * conv_dtype / conv_channels are read only for RAG_LAYOUT_CONV_IN (conv_channels = 64 or 128 >= F;
 * channels past F are written as zero). */
int rag_features_emit(const void* d_workspace, int n, int size, const int32_t* h_plane_src, int F,
                      int layout, int conv_dtype, int conv_channels, void* d_out, void* stream);
/* Both steps (the ladder reader runs iff plane 44 or 45 is listed). */
int rag_features(const uint32_t* d_states, int n, int size, const int32_t* h_plane_src, int F,
                 int layout, int conv_dtype, int conv_channels, void* d_out,
                 void* d_workspace, size_t workspace_bytes, void* stream);

/* rag_features for callers that must not allocate.  (No counterpart in the reference: its
 * Python allocates a fresh numpy array per call.)
 *
 * Some hosts cannot give this library a workspace: an engine that hands out all of its device
 * memory at start-up and forbids cudaMalloc afterwards, a plug-in running inside another
 * framework's memory pool, or a latency-critical loop (a GTP engine answering under a clock)
 * that cannot afford the device-wide synchronisation an allocation may imply.  For them the
 * scratch of the analysis and ladder kernels lives in the library itself: a fixed array in device
 * memory, about 2.4 MB on every GPU that loads the library, kept until the process exits.
 *
 *   - at most rag_features_static_max_positions() (= 256) positions per call; more is refused
 *     with RAG_ENOMEM and a message -- split the batch or use rag_features
 *   - the caller still passes the states and the output buffer; nothing is allocated or freed,
 *     here or in any routine this calls
 *   - the first call on each device looks the array's device address up (one runtime query,
 *     with that device current); make that first call during start-up, before any stream
 *     capture is begun
 *   - all calls share the one array per GPU: two calls whose work overlaps on the GPU (different
 *     streams, different threads) race on it -- keep them on one stream or serialise them
 *     yourself.  Contents left by an earlier call never leak into a later result.
 * Arguments and results are exactly those of rag_features. */
int rag_features_static(const uint32_t* d_states, int n, int size, const int32_t* h_plane_src, int F,
                        int layout, int conv_dtype, int conv_channels, void* d_out, void* stream);
int rag_features_static_max_positions(void);

/* Rows a RAG_LAYOUT_CONV_IN buffer for n boards must have: n*S*S rounded up to a multiple of 256
 * (rows past n*S*S are never written nor used). */
long long rag_conv_rows(int n, int size);

/* ------------------------------------------------------------------ batched players
 *
 * GreedyPolicyPlayer.get_move (ai.py:18-45) / ProbabilisticPolicyPlayer.get_moves (ai.py:107-133)
 * for a batch, on the device: d_probs float [n][S*S] from a policy forward, d_workspace the
 * analysis of the same n states (its sensibleness plane = get_legal_moves(include_eyes=False)).
 * The probabilities are restricted to the sensible moves, renormalised, raised to 1/temperature,
 * and either the most probable move is taken (greedy != 0) or one is drawn (seeded: splitmix64 of
 * seed and position index; the reference draws from numpy's global generator, so only the
 * distribution can match).  d_moves[i] = x*S+y, or -1 (pass) when no sensible move exists. */
int rag_select_moves(const float* d_probs, const void* d_workspace, int n, int size, float temperature,
                     int greedy, uint64_t seed, int32_t* d_moves, void* stream);

/* batch_eval_state's per-state host loop (policy.py:46-52) for a batch, on the device: for position i the
 * moves of get_legal_moves() in the reference's order -- legal non-eye points x-major, then (include_eyes
 * != 0) legal own eyes x-major (go.py:465-480) -- as flat points x*S+y in d_moves[i*S*S ..], with the
 * probabilities of d_probs renormalised over exactly those moves (_select_moves_and_normalize,
 * policy.py:15-25) in d_out[i*S*S ..]; d_counts[i] = how many.  Entries past the count are not written.
 * d_workspace: the analysis of the same n states. */
int rag_move_distributions(const float* d_probs, const void* d_workspace, int n, int size, int include_eyes,
                           int16_t* d_moves, float* d_out, int32_t* d_counts, void* stream);

/* ------------------------------------------------------------------ search tree
 *
 * The tree of AlphaGo/mcts.py (TreeNode :11-92, MCTS._playout :137-169, update_with_move :208-217) as host-side
 * native code, advanced in waves: `count` playouts move one ply per rag_tree_wave_step; between steps the caller
 * evaluates the positions standing on unexpanded nodes (rag_tree_wave_leaves -> policy batch -> rag_tree_wave_expand)
 * and plays the chosen moves (rag_do_move); rag_tree_wave_backup adds the leaf values (mcts.py:72-80; virtual visits of
 * the wave, `virtual_loss` per visit, are taken back).  With count = 1 the arithmetic is the reference's, bit for bit.
 * Moves are flat points x*S+y.  Host memory only; one thread at a time per tree. */
typedef struct rag_tree rag_tree;
int rag_tree_create(double c_puct, rag_tree** out);
int rag_tree_destroy(rag_tree* tree);
int rag_tree_reset(rag_tree* tree);
int rag_tree_wave_begin(rag_tree* tree, int count, double virtual_loss);
int rag_tree_wave_leaves(rag_tree* tree, int32_t* rows);                     /* returns k <= count */
int rag_tree_wave_expand(rag_tree* tree, int k, const int32_t* rows, const int64_t* counts, const int64_t* moves,
                         const float* priors);
int rag_tree_wave_step(rag_tree* tree, int32_t* moves_out);                  /* returns playouts alive; -2 = no move */
int rag_tree_wave_backup(rag_tree* tree, const double* values);
int rag_tree_advance(rag_tree* tree, int action);                            /* 1: subtree kept, 0: new tree */
int rag_tree_root(const rag_tree* tree);
long long rag_tree_size(const rag_tree* tree);
int rag_tree_node_children(const rag_tree* tree, int node);
int rag_tree_node_stats(const rag_tree* tree, int node, long long* visits, double* q, int32_t* parent, int32_t* index);
int rag_tree_node_arrays(const rag_tree* tree, int node, int32_t* actions, double* P, double* Q, double* u, int64_t* n,
                         int64_t* virt, int32_t* child);

/* ------------------------------------------------------------------ state updates
 *
 * GameState.do_move (go.py:521-581) on packed states, in place, one move per
 * state (p = x*S+y, or -1 = pass); d_status[i] = 0 played / -1 illegal (state
 * untouched; the reference raises IllegalMove). */
int rag_do_move(uint32_t* d_states, int n, int size, const int32_t* d_moves, int32_t* d_status,
                void* stream);

/* MCTS._evaluate_rollout (mcts.py:171-189) for a batch, with uniformly random sensible moves in place of
 * the greedy rollout policy, then GameState.get_winner (go.py:482-506, area scoring, komi to WHITE).
 * Each position is played on (the input is not modified) until the side to move has no sensible move
 * or `limit` moves were made.  d_result int32 [n]: +1 the player to move in d_states[i] wins, -1 the
 * opponent, 0 tie.  Draws are a pure function of (seed, i). */
int rag_rollout_winners(const uint32_t* d_states, int n, int size, int limit, uint64_t seed, float komi,
                        int32_t* d_result, void* stream);

/* Random playouts on the device (the synthetic workload of BASELINE.json /
 * SURVEY 8(d)).  Warp w owns output slots [w*quota, (w+1)*quota), quota =
 * rag_playout_quota(max_ply, stride) = max_ply/stride + 1, and plays games from
 * the empty board until they are full: every ply a uniformly random move from
 * get_legal_moves(include_eyes=False) (go.py:465-480; pass when there is none;
 * a game ends after two passes or max_ply), recording the position before ply t
 * whenever t % stride == (w + game) % stride.  Deterministic in `seed`.
 * d_move_log (optional, int32 [warps][log_stride]): every move warp w plays, in
 * order (p, -1 = pass), each game preceded by -2 -- enough to replay the games
 * elsewhere. */
int rag_random_playouts(uint32_t* d_states_out, int n_positions, int size, uint64_t seed, int max_ply,
                        int stride, int32_t* d_move_log, int log_stride, void* stream);
int rag_playout_quota(int max_ply, int stride);

/* ------------------------------------------------------------------ network forward
 *
 * CNNPolicy.create_network (policy.py:66-138) / value_trainer (value.py:12-28) as one object:
 * `layers` k x k 'same' convolutions with ReLU, a 1x1 convolution to one plane, then either
 * Flatten + Bias + softmax (policy) or Flatten + Dense(dense_units) + Dense(1, tanh) (value).
 * Convolutions run as tcgen05 implicit GEMMs on 16-bit operands (`dtype`) with fp32
 * accumulation; the 1x1 layer and the heads are fp32. */
#define RAG_MAX_LAYERS 32
#define RAG_HEAD_POLICY 0
#define RAG_HEAD_VALUE  1

typedef struct rag_net_desc {
    int board;                       /* params["board"]                     policy.py:82  */
    int in_planes;                   /* params["input_dim"]                 policy.py:98  */
    int layers;                      /* params["layers"] (k x k layers; the 1x1 is implied) */
    int filters[RAG_MAX_LAYERS];     /* filters_per_layer_K                 policy.py:112-114 */
    int widths[RAG_MAX_LAYERS];      /* filter_width_K (odd, <= 7)          policy.py:108-110 */
    int head;                        /* RAG_HEAD_POLICY / RAG_HEAD_VALUE    */
    int dense_units;                 /* value head: 256                     value.py:27   */
    int dtype;                       /* RAG_DTYPE_BF16 / RAG_DTYPE_FP16 operands, or a split mode: FP16X3, FP16C32, FP16C64 */
    int kernel_flip;                 /* 1: Theano semantics (true convolution, the reference's
                                        pinned backend); 0: cross-correlation (TensorFlow) */
    int resnet;                      /* 1: ResnetPolicy (policy.py:141-271): convolution 1 is linear, every
                                        later one is BatchNormalization -> ReLU -> convolution, units are
                                        closed by adding the unit's input, a last ReLU precedes the 1x1 */
    int unit_end[RAG_MAX_LAYERS];    /* resnet: 1 where convolution K closes a unit (n_skip_K, policy.py:220) */
} rag_net_desc;

typedef struct rag_net rag_net;

int rag_net_create(const rag_net_desc* desc, rag_net** out);
int rag_net_destroy(rag_net* net);

/* Keras model.get_weights() order and layouts: per conv layer W (out, in, kh, kw) and b (out,);
 * the 1x1 conv's W (1, in, 1, 1) and b (1,); then policy: Bias W (S*S,)  (nn_util.py:128-133);
 * value: Dense W (S*S, units), b (units,), Dense W (units, 1), b (1,)  (value.py:27-28).
 * Residual networks: every convolution after the first is preceded by its BatchNormalization's
 * gamma, beta, running_mean, running_std, each (S,) (Keras-1 default axis -1).
 * Host float arrays; copies them to the device (synchronous). */
int rag_net_num_weight_arrays(const rag_net* net);
int rag_net_load_weights(rag_net* net, const float* const* h_arrays, int n_arrays);
/* Optional, before rag_net_load_weights, for RAG_DTYPE_FP16C32 / FP16C64: which canonical feature
 * plane every input plane is (the list rag_eval_batch is given; Preprocess's feature order,
 * preprocessing.py:266-285).  Only sharpens the ranking of channels by expected energy -- results
 * are valid without it; without it every input plane counts the same. */
int rag_net_set_input_planes(rag_net* net, const int32_t* h_plane_src, int F);

size_t rag_net_workspace_bytes(const rag_net* net, int n);
/* layout of the input the *_rows entry expects: rag_conv_rows(n, board) rows of `channels` */
long long rag_net_input_rows(const rag_net* net, int n);
int rag_net_input_channels(const rag_net* net);

/* Tuning/test knob (process-wide): 0 = one CTA per tile (default); 1 = convolutions run on CTA pairs
 * (tcgen05 cta_group::2: half the weight tile per CTA, one M=256 UMMA per pair; also RAG_CONV_PAIR=1 in the
 * environment).  Results are identical and the pair kernel is ~13 % faster, but beside the pipeline's persistent
 * feature kernels its cluster launch sporadically never finishes on B200 / driver 580 (csrc/conv.cu): use 1
 * only when nothing else shares the GPU with the forward.  Returns the previous setting. */
int rag_set_conv_pair(int enable);
/* No wait inside the convolution kernel is unbounded: a barrier that has not moved for ~4 s (a layer takes
 * 1 ms) makes the kernel record where it stood in host-mapped memory and trap, so that a stall surfaces as
 * a CUDA error instead of a hang.  After such an error this writes the record as text into buf and returns
 * 1; 0 if there is none.  Needs no working CUDA context. */
int rag_conv_stall_report(char* buf, int size);

/* Measurement aid: with profiling on, every forward records a CUDA event on its stream before
 * each convolution launch and after the last; rag_net_layer_times then gives the device time
 * of each of the `layers` convolution kernels of the most recent forward (waits for it).
 * Between two layers the interval also covers the margin memsets of the next buffer. */
int rag_net_set_profiling(rag_net* net, int enable);
int rag_net_layer_times(rag_net* net, float* h_ms, int n_layers);

/* Byte offset, inside the workspace of a finished forward for n boards, of the fp32 output of
 * the 1x1 convolution ([n][S*S], before Bias/softmax or the dense head): for inspection/tests. */
size_t rag_net_logits_offset(const rag_net* net, int n);

/* NeuralNetBase.forward (nn_util.py:57-58): d_nchw float [n][in_planes][S][S] ->
 * d_out float [n][S*S] probabilities (policy) or [n] values in [-1, 1] (value). */
int rag_net_forward(rag_net* net, const float* d_nchw, int n, float* d_out, void* d_workspace,
                    size_t workspace_bytes, void* stream);
/* The same from pixel rows (RAG_LAYOUT_CONV_IN with this net's dtype / channels). */
int rag_net_forward_rows(rag_net* net, const void* d_in_rows, int n, float* d_out, void* d_workspace,
                         size_t workspace_bytes, void* stream);

/* CNNPolicy.batch_eval_state's device part (policy.py:42-45) in one call: packed states ->
 * feature planes (h_plane_src, F = in_planes) -> forward.  No host round trip in between. */
size_t rag_eval_workspace_bytes(const rag_net* net, int n);
int rag_eval_batch(rag_net* net, const uint32_t* d_states, int n, const int32_t* h_plane_src, int F,
                   float* d_out, void* d_workspace, size_t workspace_bytes, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* ROCALPHAGO_B200_H */
