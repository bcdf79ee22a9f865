// conv.cu -- the CNNPolicy / value-network forward as tcgen05 implicit-GEMM convolutions (sm_100a).
//
// Replaces the Keras/Theano graph built by AlphaGo/models/policy.py:94-138 (CNNPolicy.create_network)
// and AlphaGo/models/value.py:14-28 and run by AlphaGo/models/nn_util.py:48-58 (K.function).
//
// Data layout ("pixel rows"): an activation tensor is a dense 2-D matrix [n*S*S][channels] of
// 16-bit values, row b*S*S + x*S + y for point (x, y) of board b -- NHWC with nothing in between.
// A k x k 'same' convolution is then
//       out[r, :] = act( bias + sum_taps  in[r shifted by (dx, dy), :] . W_t^T )
// -- k*k GEMMs that share one accumulator -- and the shifted operand is produced by the TMA
// engine in im2col mode: one instruction fetches, for 128 consecutive output points, the input
// point (x + dx, y + dy) of each, crossing row and board boundaries and writing zeros where the
// point falls off the board.  No im2col buffer, no padded copies, no wasted MMA rows.
//
// Kernel (one persistent CTA per SM, 6 warps):
//   warp 0      TMA producer: per (tap, 64-channel chunk) one im2col A tile (shifted activations)
//               and one B tile (that tap's weights) into a ring of 128B-swizzled smem stages
//   warp 1      tcgen05.mma issuer (one elected lane): M=128 x N=filters x K=16 UMMAs, fp32
//               accumulators in TMEM, double-buffered so the epilogue of tile i overlaps the
//               MMAs of tile i+1
//   warps 2-5   epilogue: tcgen05.ld the accumulator, + bias, ReLU, then either 16-bit pixel rows
//               for the next layer, or (last layer) the fused 1x1 convolution: a dot product
//               with the head's 192 weights giving one fp32 logit per point (policy.py:125-130).
// Heads: Bias + softmax (nn_util.py:118-133, policy.py:134-136) or Dense(256) + Dense(1, tanh)
// (value.py:27-28) as small fp32 kernels over the logits.
#include "rag_common.h"
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <algorithm>
#include <vector>

namespace rag {

constexpr int kTileM = 128;           // rows (pixels) per accumulator tile
constexpr int kChunkK = 64;           // channels per smem stage (= one 128-byte swizzled row)
constexpr int kMaxBoardPoints = 361;
constexpr int kTmemCols = 512;
constexpr int kConvThreads = 192;
constexpr int kATileBytes = kTileM * kChunkK * 2;     // 16 KiB
constexpr int kOutBoxBytes = 32 * kChunkK * 2;        // 4 KiB: what one epilogue warp hands to one TMA store

struct ConvParams {
    int tiles, width, kchunks, N, stages, dtype;
    int S, n;                 // board size, boards
    int mode;                 // 0: write 16-bit pixel rows; 1: fused 1x1 head -> fp32 logits;
                              // 2 / 3: the same for a residual network (see the epilogue)
    int relu;
    long long rows;           // n * S * S
    const float* bias;        // [N]
    const float* head_w;      // [N]   (mode 1)
    float head_b;
    void* out;
    // residual networks (policy.py:141-271): L = conv + bias (+ residual); `linear_out` keeps L when it
    // is a unit's output; what the next convolution reads is relu(L * bn_scale[y] + bn_shift[y])
    const uint16_t* residual;     // [rows][N] or null
    uint16_t* linear_out;         // [rows][N] or null
    const float* bn_scale;        // [S] (indexed by y, the reference's BatchNormalization axis) or null
    const float* bn_shift;
    // split operands (RAG_DTYPE_FP16X3 / FP16C32 / FP16C64): besides the fp16 value `hi` of every channel an
    // activation row carries lo = fp16(x - hi) of its first `lo_cols` channels; the K walk reads extra A
    // chunks against extra weight columns so that those channels contribute hi.W_hi + lo.W_hi + hi.W_lo
    // (layouts in rag_net_create / rag_net_load_weights)
    unsigned long long a_map;     // 4 bits per K chunk: which 64-channel chunk of the A row it reads
    unsigned long long k_map;     // 4 bits per K chunk: how many K=16 steps of it hold real channels (1..4)
    int out_pitch;                // mode 0: channels per output row
    int lo_cols;                  // mode 0: leading channels whose lo half is written at column N + c
    int dup_hi;                   // mode 0: also copy hi of channels 0..31 to columns N + 32 .. N + 63
    int debug;                    // timing experiments only (RAG_CONV_DEBUG): 1 skip weight loads, 2 skip activation loads, 8 skip stores
};

// ---- PTX wrappers ----------------------------------------------------------------------------

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
// Where a wait that never ended was standing: written to host-mapped memory just before the kernel traps
// (rag_conv_stall_report reads it after the context is gone).  A correct run never touches it.
struct StallRecord { unsigned magic, code, block, thread, tile, stage, parity, pad; };
__device__ StallRecord* g_stall_record = nullptr;
constexpr long long kStallCycles = 6000000000ll;      // ~4 s at 1.5 GHz: a 3x3 layer takes 1 ms

__device__ __noinline__ void stall_trap(uint32_t code, uint32_t a, uint32_t b, uint32_t c) {
    StallRecord* r = g_stall_record;
    if (r && atomicCAS(&r->magic, 0u, 0x57A11u) == 0u) {
        r->code = code; r->block = blockIdx.x; r->thread = threadIdx.x; r->tile = a; r->stage = b; r->parity = c;
        __threadfence_system();
    }
    __trap();
}

__device__ __forceinline__ bool mbar_try(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t"
        "}" : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    return ok != 0;
}
// `code` says which wait this is (1 producer/empty, 2 MMA/accumulator free, 3 MMA/stage full, 4 epilogue/accumulator full)
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity, uint32_t code = 0, uint32_t a = 0, uint32_t b = 0) {
    if (mbar_try(bar, parity)) return;
    const long long t0 = clock64();
    for (;;) {
        #pragma unroll 1
        for (int i = 0; i < 1024; i++)
            if (mbar_try(bar, parity)) return;
        if (clock64() - t0 > kStallCycles) stall_trap(code, a, b, parity);
    }
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, int c0, int c1, uint32_t bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1) : "memory");
}
// im2col-mode load: `pixels` consecutive base points starting at (w, h, n), each displaced by the
// filter tap (w_off, h_off); off-board points read as zero
__device__ __forceinline__ void tma_load_im2col(uint32_t dst, const CUtensorMap* map, int c, int w, int h, int n,
                                                uint16_t w_off, uint16_t h_off, uint32_t bar) {
    asm volatile(
        "cp.async.bulk.tensor.4d.shared::cluster.global.im2col.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2], {%7, %8};"
        ::"r"(dst), "l"(map), "r"(bar), "r"(c), "r"(w), "r"(h), "r"(n), "h"(w_off), "h"(h_off) : "memory");
}
__device__ __forceinline__ void tma_load_im2col_pair(uint32_t dst, const CUtensorMap* map, int c, int w, int h, int n,
                                                     uint16_t w_off, uint16_t h_off, uint32_t leader_bar) {
    asm volatile(
        "cp.async.bulk.tensor.4d.cta_group::2.shared::cluster.global.im2col.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2], {%7, %8};"
        ::"r"(dst), "l"(map), "r"(leader_bar), "r"(c), "r"(w), "r"(h), "r"(n), "h"(w_off), "h"(h_off) : "memory");
}
// bulk tensor store smem -> global (rows past the tensor's extent are clipped by the TMA unit)
__device__ __forceinline__ void tma_store_2d(const CUtensorMap* map, int c0, int c1, uint32_t src) {
    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
                 ::"l"(map), "r"(src), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read0() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait0() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
// one lane of a converged warp (elect.sync): unlike `lane == 0`, the compiler knows exactly one thread is
// inside the branch and issues the TMA / MMA instructions straight from uniform registers
__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
    return pred != 0;
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tc_mma_f16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
        "}" ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void tc_ld32(uint32_t taddr, uint32_t (&v)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
          "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
          "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
          "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tc_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// K-major operand tile, 128-byte rows, SWIZZLE_128B, 8-row groups 1024 B apart (cute::UMMA::SmemDescriptor)
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t smem_addr) {
    uint64_t lo = (uint64_t)((smem_addr >> 4) & 0x3FFFu) | ((uint64_t)1 << 16);          // start, LBO (unused) = 1
    uint64_t hi = (uint64_t)(1024 >> 4) | ((uint64_t)1 << 14) | ((uint64_t)2 << 29);     // SBO, version 1, SW128
    return lo | (hi << 32);
}

// kind::f16 instruction descriptor (cute::UMMA::InstrDescriptor): fp32 accumulate, K-major A and B
__device__ __forceinline__ uint32_t umma_idesc(int dtype, int N, int M) {
    uint32_t fmt = dtype == RAG_DTYPE_BF16 ? 1u : 0u;
    return (1u << 4) | (fmt << 7) | (fmt << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

// ---- the convolution kernel --------------------------------------------------------------------
//
// kPair = false: one CTA per tile of 128 rows (cta_group::1).
// kPair = true : two CTAs of a cluster share one 256-row tile (cta_group::2).  Each CTA loads its
//     own 128 rows of A and HALF of the weight tile (N/2 rows of B); the leader CTA issues one
//     M=256 UMMA that reads both halves, and each CTA's TMEM receives its own 128 rows.  Per SM
//     that halves the shared-memory traffic of the weights -- the resource this kernel runs out of
//     first (each 40 KB stage is written once by TMA and read once by four UMMAs).

__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void tma_load_2d_pair(uint32_t dst, const CUtensorMap* map, int c0, int c1, uint32_t leader_bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(dst), "l"(map), "r"(leader_bar), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tc_commit_pair(uint32_t bar) {      // arrives on the same barrier in both CTAs
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(bar), "h"((uint16_t)3) : "memory");
}
__device__ __forceinline__ void tc_mma_f16_pair(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t"
        "}" ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void mbar_arrive_leader(uint32_t local_bar) {   // arrive on CTA 0's copy of the barrier
    uint32_t remote;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(local_bar), "r"(0));
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(remote) : "memory");
}
constexpr uint32_t kPeerBitMask = 0xFEFFFFFFu;   // shared::cluster address of the same offset in CTA 0 of the pair

template <bool kPair>
__global__ void __launch_bounds__(kConvThreads, 1)
conv_umma_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                 const __grid_constant__ CUtensorMap tmO, const ConvParams p)
{
    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw = smem_u32(smem_raw);
    const uint32_t base = (raw + 1023u) & ~1023u;                 // SWIZZLE_128B tiles need 1024-byte alignment
    uint8_t* smem = smem_raw + (base - raw);
    const uint32_t b_rows = kPair ? (uint32_t)p.N / 2u : (uint32_t)p.N;       // rows of B this CTA holds
    const uint32_t stage_bytes = kATileBytes + b_rows * 128u;
    uint8_t* out_stage = smem + (size_t)p.stages * stage_bytes;   // 4 epilogue warps x one 32-row x 64-channel store box
    uint8_t* after = out_stage + 4 * kOutBoxBytes;
    uint64_t* bars = reinterpret_cast<uint64_t*>(after);          // full[stages], empty[stages], tfull[2], tempty[2]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * p.stages + 4);
    float* s_bias = reinterpret_cast<float*>(tmem_slot + 4);
    float* s_head = s_bias + p.N;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = kPair ? cluster_ctarank() : 0u;
    const int n_workers = kPair ? (int)gridDim.x / 2 : (int)gridDim.x;        // tiles are dealt round-robin to CTAs / pairs
    const int worker = kPair ? (int)blockIdx.x / 2 : (int)blockIdx.x;
    const int rows_per_tile = kPair ? 2 * kTileM : kTileM;
    const uint32_t bar0 = smem_u32(bars);
    auto full_bar = [&](int s) { return bar0 + 8u * (uint32_t)s; };
    auto empty_bar = [&](int s) { return bar0 + 8u * (uint32_t)(p.stages + s); };
    auto tfull_bar = [&](int a) { return bar0 + 8u * (uint32_t)(2 * p.stages + a); };
    auto tempty_bar = [&](int a) { return bar0 + 8u * (uint32_t)(2 * p.stages + 2 + a); };

    if (threadIdx.x == 0) {
        for (int s = 0; s < p.stages; s++) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
        for (int a = 0; a < 2; a++) { mbar_init(tfull_bar(a), 1); mbar_init(tempty_bar(a), kPair ? 8 : 4); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        if (kPair) {
            asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(kTmemCols));
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
        } else {
            asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(kTmemCols));
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
        }
    }
    for (int i = threadIdx.x; i < p.N; i += blockDim.x) {
        s_bias[i] = p.bias[i];
        s_head[i] = (p.mode == 1 || p.mode == 3) ? p.head_w[i] : 0.f;
    }
    tc_fence_before();
    __syncthreads();
    if (kPair) cluster_sync_all();                                 // the peer's barriers exist before anyone signals them
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const int kblocks = p.width * p.width * p.kchunks;

    if (warp == 0) {
        // ===== TMA producer: the whole warp walks the ring, one elected lane issues =====
        {
            // (ring position and phase are carried along, never divided out: this lane's and the MMA lane's
            // instruction count per stage is on the kernel's critical path)
            uint32_t it = 0, ph = 0;
            int s = 0;
            const int pad = p.width / 2, S2 = p.S * p.S;
            for (int tile = worker; tile < p.tiles; tile += n_workers) {
                // first output point of my 128 rows -> base point of the im2col walk (shifted by -pad)
                const long long m0 = (long long)tile * rows_per_tile + (long long)rank * kTileM;
                const int b0 = (int)(m0 / S2), r0 = (int)(m0 - (long long)b0 * S2);
                const int x0 = r0 / p.S, y0 = r0 - x0 * p.S;
                for (int i = 0; i < p.width; i++) {
                    for (int j = 0; j < p.width; j++) {
                        const int t = i * p.width + j;
                        for (int c = 0; c < p.kchunks; c++, it++) {
                            mbar_wait(empty_bar(s), ph ^ 1u, 1u, (uint32_t)tile, (uint32_t)s);
                            const uint32_t sa = base + (uint32_t)s * stage_bytes;
                            const int ca = (int)((p.a_map >> (4 * c)) & 15ull);
                            // tensor dims are (C, y, x, board): y is the fast spatial index
                            if (!elect_one()) {
                                // nothing to issue on this lane
                            } else if (kPair) {
                                // both CTAs' bytes are counted on the leader's barrier
                                if (rank == 0) mbar_expect_tx(full_bar(s), 2u * stage_bytes);
                                const uint32_t lb = full_bar(s) & kPeerBitMask;
                                tma_load_im2col_pair(sa, &tmA, ca * kChunkK, y0 - pad, x0 - pad, b0, (uint16_t)j, (uint16_t)i, lb);
                                tma_load_2d_pair(sa + kATileBytes, &tmB, c * kChunkK, t * p.N + (int)(rank * b_rows), lb);
                            } else if (p.debug & 3) {
                                // timing experiment: after the ring's first pass leave stale operands in place
                                const bool la = !(p.debug & 2) || it < (uint32_t)p.stages, lb = !(p.debug & 1) || it < (uint32_t)p.stages;
                                const uint32_t bytes = (la ? (uint32_t)kATileBytes : 0u) + (lb ? b_rows * 128u : 0u);
                                if (bytes) mbar_expect_tx(full_bar(s), bytes); else mbar_arrive(full_bar(s));
                                if (la) tma_load_im2col(sa, &tmA, ca * kChunkK, y0 - pad, x0 - pad, b0, (uint16_t)j, (uint16_t)i, full_bar(s));
                                if (lb) tma_load_2d(sa + kATileBytes, &tmB, c * kChunkK, t * p.N, full_bar(s));
                            } else {
                                mbar_expect_tx(full_bar(s), stage_bytes);
                                tma_load_im2col(sa, &tmA, ca * kChunkK, y0 - pad, x0 - pad, b0, (uint16_t)j, (uint16_t)i, full_bar(s));
                                tma_load_2d(sa + kATileBytes, &tmB, c * kChunkK, t * p.N, full_bar(s));
                            }
                            __syncwarp();
                            if (++s == p.stages) { s = 0; ph ^= 1u; }
                        }
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer (the leader CTA's, for a pair): the warp walks, one elected lane issues =====
        if (rank == 0) {
            const uint32_t idesc = umma_idesc(p.dtype, p.N, rows_per_tile);
            uint32_t lt = 0, ph = 0;
            int s = 0;
            const uint64_t desc0 = umma_desc_sw128(base), desc_stage = (uint64_t)(stage_bytes >> 4), desc_b = (uint64_t)(kATileBytes >> 4);
            for (int tile = worker; tile < p.tiles; tile += n_workers, lt++) {
                const uint32_t a = lt & 1u, aph = (lt >> 1) & 1u;
                mbar_wait(tempty_bar((int)a), aph ^ 1u, 2u, (uint32_t)tile, a);
                tc_fence_after();
                const uint32_t tmem_d = tmem_base + a * (uint32_t)p.N;
                uint32_t accumulate = 0u;                         // the tile's first MMA overwrites the accumulator
                unsigned long long km = p.k_map;
                int c = 0;
                for (int kb = 0; kb < kblocks; kb++) {
                    mbar_wait(full_bar(s), ph, 3u, (uint32_t)tile, (uint32_t)s);
                    tc_fence_after();
                    // the stage's descriptors: the start-address field advances by stage_bytes / 16 per stage
                    // (the ring stays far below the 14-bit field's 256 KiB reach)
                    const uint64_t da = desc0 + (uint64_t)s * desc_stage, db = da + desc_b;
                    const int ksteps = (int)(km & 15ull);         // K=16 steps of this chunk that hold real channels
                    if (elect_one()) {
                        #pragma unroll
                        for (int k = 0; k < kChunkK / 16; k++) {  // +32 bytes along K inside the swizzled row
                            if (k < ksteps) {
                                if (kPair) tc_mma_f16_pair(tmem_d, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, k == 0 ? accumulate : 1u);
                                else tc_mma_f16(tmem_d, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, k == 0 ? accumulate : 1u);
                            }
                        }
                        if (kPair) tc_commit_pair(empty_bar(s)); else tc_commit(empty_bar(s));   // frees the stage (in both CTAs)
                    }
                    __syncwarp();
                    accumulate = 1u;
                    if (++s == p.stages) { s = 0; ph ^= 1u; }
                    km >>= 4;
                    if (++c == p.kchunks) { c = 0; km = p.k_map; }
                }
                if (elect_one()) { if (kPair) tc_commit_pair(tfull_bar((int)a)); else tc_commit(tfull_bar((int)a)); }   // accumulator complete
                __syncwarp();
            }
        }
    } else {
        // ===== epilogue =====
        const int quarter = warp & 3;                              // TMEM lanes this warp may read
        uint8_t* sbuf = out_stage + quarter * kOutBoxBytes;
        const uint32_t sbuf_u32 = smem_u32(sbuf);
        bool store_pending = false;
        auto store_done = [&]() {                                  // the last box has been read out of sbuf
            if (store_pending) { if (lane == 0) bulk_wait_read0(); __syncwarp(); store_pending = false; }
        };
        auto store_box = [&](int chan, int row0) {                 // every lane has written its row of the box
            fence_proxy_async();
            __syncwarp();
            if (lane == 0 && row0 < p.rows && !(p.debug & 8)) { tma_store_2d(&tmO, chan, row0, sbuf_u32); bulk_commit(); }
            store_pending = true;
        };
        uint32_t lt = 0;
        for (int tile = worker; tile < p.tiles; tile += n_workers, lt++) {
            const uint32_t a = lt & 1u, aph = (lt >> 1) & 1u;
            // my row of the tile = one point of one board; rows past the last board are not written
            const long long row = (long long)tile * rows_per_tile + (long long)rank * kTileM + quarter * 32 + lane;
            const bool real = row < p.rows && !(p.debug & 8);
            mbar_wait(tfull_bar((int)a), aph, 4u, (uint32_t)tile, a);
            tc_fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + a * (uint32_t)p.N;
            if (p.mode == 0) {
                // 16-bit pixel rows for the next layer, 64 channels at a time: the warp assembles its 32 rows
                // of the box in shared memory (128-byte rows, 16-byte chunks XOR-swizzled by row so that the
                // 32 lanes' stores spread over all banks) and one lane hands the box to the TMA unit, which
                // writes full lines and clips rows past the last board.  Per-thread global stores of 16 B at
                // a 384-byte stride cost the SM's L1/shared-memory pipe 32 passes per instruction -- the
                // pipe the tensor core's operands come through.
                const int row0 = (int)((long long)tile * rows_per_tile + (long long)rank * kTileM + quarter * 32);
                uint8_t* my_row = sbuf + lane * 128;
                const int sw = lane & 7;
                for (int bx = 0; bx < p.N; bx += kChunkK) {
                    const bool with_lo = bx < p.lo_cols;          // x = hi + lo, both fp16: the next layer multiplies both
                    uint32_t lo[32], dup[16];
                    #pragma unroll
                    for (int half = 0; half < 2; half++) {
                        const int c0 = bx + 32 * half;
                        uint32_t v[32];
                        tc_ld32(taddr + (uint32_t)c0, v);
                        tc_wait_ld();
                        uint32_t packed[16];
                        #pragma unroll
                        for (int j = 0; j < 16; j++) {
                            float f0 = __uint_as_float(v[2 * j]) + s_bias[c0 + 2 * j];
                            float f1 = __uint_as_float(v[2 * j + 1]) + s_bias[c0 + 2 * j + 1];
                            if (p.relu) { f0 = fmaxf(f0, 0.f); f1 = fmaxf(f1, 0.f); }
                            if (p.dtype == RAG_DTYPE_BF16) {
                                __nv_bfloat162 h = __floats2bfloat162_rn(f0, f1);
                                packed[j] = *reinterpret_cast<uint32_t*>(&h);
                            } else {
                                __half2 h = __floats2half2_rn(f0, f1);
                                packed[j] = *reinterpret_cast<uint32_t*>(&h);
                                if (with_lo) {
                                    const float2 back = __half22float2(h);
                                    __half2 l = __floats2half2_rn(f0 - back.x, f1 - back.y);
                                    lo[16 * half + j] = *reinterpret_cast<uint32_t*>(&l);
                                }
                            }
                            if (half == 0) dup[j] = packed[j];
                        }
                        if (half == 0) store_done();
                        #pragma unroll
                        for (int j = 0; j < 4; j++)
                            *reinterpret_cast<uint4*>(my_row + (((4 * half + j) ^ sw) << 4)) =
                                make_uint4(packed[4 * j], packed[4 * j + 1], packed[4 * j + 2], packed[4 * j + 3]);
                    }
                    store_box(bx, row0);
                    if (with_lo) {
                        store_done();
                        if (p.dup_hi) {            // FP16C32: [lo 0..31 | hi 0..31]
                            #pragma unroll
                            for (int j = 0; j < 4; j++) {
                                *reinterpret_cast<uint4*>(my_row + ((j ^ sw) << 4)) = make_uint4(lo[4 * j], lo[4 * j + 1], lo[4 * j + 2], lo[4 * j + 3]);
                                *reinterpret_cast<uint4*>(my_row + (((4 + j) ^ sw) << 4)) = make_uint4(dup[4 * j], dup[4 * j + 1], dup[4 * j + 2], dup[4 * j + 3]);
                            }
                        } else {
                            #pragma unroll
                            for (int j = 0; j < 8; j++)
                                *reinterpret_cast<uint4*>(my_row + ((j ^ sw) << 4)) = make_uint4(lo[4 * j], lo[4 * j + 1], lo[4 * j + 2], lo[4 * j + 3]);
                        }
                        store_box(p.N + bx, row0);
                    }
                }
            } else if (p.mode == 1) {
                float acc = 0.f;
                for (int c0 = 0; c0 < p.N; c0 += 32) {
                    uint32_t v[32];
                    tc_ld32(taddr + (uint32_t)c0, v);
                    tc_wait_ld();
                    #pragma unroll
                    for (int j = 0; j < 32; j++) {
                        float f = __uint_as_float(v[j]) + s_bias[c0 + j];
                        if (p.relu) f = fmaxf(f, 0.f);
                        acc = fmaf(f, s_head[c0 + j], acc);
                    }
                }
                if (real) reinterpret_cast<float*>(p.out)[row] = acc + p.head_b;      // Flatten: index b*S*S + x*S + y
            }
            if (p.mode >= 2) {
                // residual network layer: same accumulator walk with the extra terms
                float sc = 1.f, sh = 0.f;
                if (p.bn_scale) { const int y = (int)(row % p.S); sc = p.bn_scale[y]; sh = p.bn_shift[y]; }
                const bool fp16 = p.dtype != RAG_DTYPE_BF16;
                const uint16_t* rrow = p.residual ? p.residual + row * p.N : nullptr;
                uint16_t* lrow = p.linear_out ? p.linear_out + row * p.N : nullptr;
                uint16_t* orow = reinterpret_cast<uint16_t*>(p.out) + row * p.N;
                float acc = 0.f;
                for (int c0 = 0; c0 < p.N; c0 += 32) {
                    uint32_t v[32];
                    tc_ld32(taddr + (uint32_t)c0, v);
                    tc_wait_ld();
                    uint32_t res[16];
                    #pragma unroll
                    for (int j = 0; j < 16; j++) res[j] = 0u;
                    if (rrow && real) {
                        const uint4* src = reinterpret_cast<const uint4*>(rrow + c0);
                        #pragma unroll
                        for (int j = 0; j < 4; j++) { uint4 t = src[j]; res[4 * j] = t.x; res[4 * j + 1] = t.y; res[4 * j + 2] = t.z; res[4 * j + 3] = t.w; }
                    }
                    uint32_t lin[16], act[16];
                    #pragma unroll
                    for (int j = 0; j < 16; j++) {
                        float r0, r1;
                        if (fp16) { float2 t = __half22float2(*reinterpret_cast<__half2*>(&res[j])); r0 = t.x; r1 = t.y; }
                        else { float2 t = __bfloat1622float2(*reinterpret_cast<__nv_bfloat162*>(&res[j])); r0 = t.x; r1 = t.y; }
                        const float l0 = __uint_as_float(v[2 * j]) + s_bias[c0 + 2 * j] + r0;
                        const float l1 = __uint_as_float(v[2 * j + 1]) + s_bias[c0 + 2 * j + 1] + r1;
                        const float a0 = fmaxf(fmaf(l0, sc, sh), 0.f), a1 = fmaxf(fmaf(l1, sc, sh), 0.f);
                        if (fp16) {
                            __half2 h = __floats2half2_rn(l0, l1); lin[j] = *reinterpret_cast<uint32_t*>(&h);
                            __half2 k = __floats2half2_rn(a0, a1); act[j] = *reinterpret_cast<uint32_t*>(&k);
                        } else {
                            __nv_bfloat162 h = __floats2bfloat162_rn(l0, l1); lin[j] = *reinterpret_cast<uint32_t*>(&h);
                            __nv_bfloat162 k = __floats2bfloat162_rn(a0, a1); act[j] = *reinterpret_cast<uint32_t*>(&k);
                        }
                        acc = fmaf(a0, s_head[c0 + 2 * j], acc);
                        acc = fmaf(a1, s_head[c0 + 2 * j + 1], acc);
                    }
                    if (real) {
                        if (lrow) {
                            uint4* dst = reinterpret_cast<uint4*>(lrow + c0);
                            #pragma unroll
                            for (int j = 0; j < 4; j++) dst[j] = make_uint4(lin[4 * j], lin[4 * j + 1], lin[4 * j + 2], lin[4 * j + 3]);
                        }
                        if (p.mode == 2) {
                            uint4* dst = reinterpret_cast<uint4*>(orow + c0);
                            #pragma unroll
                            for (int j = 0; j < 4; j++) dst[j] = make_uint4(act[4 * j], act[4 * j + 1], act[4 * j + 2], act[4 * j + 3]);
                        }
                    }
                }
                if (p.mode == 3 && real) reinterpret_cast<float*>(p.out)[row] = acc + p.head_b;
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) { if (kPair) mbar_arrive_leader(tempty_bar((int)a)); else mbar_arrive(tempty_bar((int)a)); }
        }
        if (lane == 0) bulk_wait0();                               // outstanding stores own sbuf until they are done
    }
    tc_fence_before();
    __syncthreads();
    if (kPair) cluster_sync_all();                                 // nobody frees TMEM or leaves while the peer still uses it
    if (warp == 1) {
        tc_fence_after();
        if (kPair) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kTmemCols));
        else asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kTmemCols));
    }
}

// ---- small kernels around it ---------------------------------------------------------------------

// (N, F, S, S) float -> pixel rows (what NeuralNetBase.forward is handed, nn_util.py:57-58)
__global__ void __launch_bounds__(256)
nchw_to_rows_kernel(const float* __restrict__ in, int n, int F, int S, int C, int dtype, int split, uint16_t* __restrict__ out)
{
    const int S2 = S * S, chunks = C / 8;
    const long long total = (long long)n * S2 * chunks;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
        const int ch = (int)(t % chunks);
        const long long pix = t / chunks;
        const int pos = (int)(pix / S2), r = (int)(pix - (long long)pos * S2);
        __align__(16) uint16_t v[8];
        #pragma unroll
        for (int j = 0; j < 8; j++) {
            // split mode: channels [0, C/2) hold fp16(x), channels [C/2, C) hold fp16(x - fp16(x))
            const bool lo_half = split && ch * 8 >= C / 2;
            const int f = ch * 8 + j - (lo_half ? C / 2 : 0);
            float val = f < F ? in[((long long)pos * F + f) * S2 + r] : 0.f;
            if (lo_half) val -= __half2float(__float2half_rn(val));
            if (dtype == RAG_DTYPE_BF16) { __nv_bfloat16 h = __float2bfloat16_rn(val); v[j] = *reinterpret_cast<uint16_t*>(&h); }
            else { __half h = __float2half_rn(val); v[j] = *reinterpret_cast<uint16_t*>(&h); }
        }
        *reinterpret_cast<uint4*>(out + pix * C + ch * 8) = *reinterpret_cast<const uint4*>(v);
    }
}

// Flatten -> Bias -> softmax (policy.py:132-136): one warp per board
__global__ void __launch_bounds__(128)
softmax_head_kernel(const float* __restrict__ logits, const float* __restrict__ pos_bias, int n, int P, float* __restrict__ probs)
{
    const int b = blockIdx.x * 4 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (b >= n) return;
    const float* in = logits + (long long)b * P;
    float v[12];
    float mx = -INFINITY;
    #pragma unroll
    for (int j = 0; j < 12; j++) {
        const int i = lane + 32 * j;
        v[j] = i < P ? in[i] + pos_bias[i] : -INFINITY;
        mx = fmaxf(mx, v[j]);
    }
    for (int o = 16; o; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    float sum = 0.f;
    #pragma unroll
    for (int j = 0; j < 12; j++) { v[j] = (lane + 32 * j) < P ? expf(v[j] - mx) : 0.f; sum += v[j]; }
    for (int o = 16; o; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    const float inv = 1.f / sum;
    #pragma unroll
    for (int j = 0; j < 12; j++) { const int i = lane + 32 * j; if (i < P) probs[(long long)b * P + i] = v[j] * inv; }
}

// Flatten -> Dense(D, linear) -> Dense(1, tanh) (value.py:26-28): one block of D threads per board
__global__ void __launch_bounds__(256)
value_head_kernel(const float* __restrict__ logits, const float* __restrict__ w1, const float* __restrict__ b1,
                  const float* __restrict__ w2, const float* __restrict__ b2, int n, int P, int D, float* __restrict__ out)
{
    __shared__ float s_in[kMaxBoardPoints];
    __shared__ float s_red[8];
    const int b = blockIdx.x, j = threadIdx.x;
    for (int i = j; i < P; i += blockDim.x) s_in[i] = logits[(long long)b * P + i];
    __syncthreads();
    float part = 0.f;
    for (int jj = j; jj < D; jj += blockDim.x) {
        float h = b1[jj];
        for (int i = 0; i < P; i++) h = fmaf(s_in[i], w1[(long long)i * D + jj], h);     // Dense W is (in, out)
        part = fmaf(h, w2[jj], part);
    }
    for (int o = 16; o; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
    if ((j & 31) == 0) s_red[j >> 5] = part;
    __syncthreads();
    if (j == 0) {
        float t = b2[0];
        for (int w = 0; w < (int)(blockDim.x >> 5); w++) t += s_red[w];
        out[b] = tanhf(t);
    }
}

// ---- host side --------------------------------------------------------------------------------------

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
typedef CUresult (*EncodeIm2colFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const int*, const int*, cuuint32_t, cuuint32_t, const cuuint32_t*,
                                   CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static void* driver_symbol(const char* name) {
    void* sym = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint(name, &sym, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) return nullptr;
    return sym;
}

// weights: 2-D row-major 16-bit matrix [rows][cols], box = [box_rows][64 columns], 128-byte swizzle
static int make_map(CUtensorMap* map, const void* ptr, long long rows, int cols, int box_rows, int dtype) {
    static EncodeTiledFn enc = (EncodeTiledFn)driver_symbol("cuTensorMapEncodeTiled");
    if (!enc) return fail(RAG_ECUDA, "cuTensorMapEncodeTiled is not available from this driver");
    cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)cols * 2};
    cuuint32_t box[2] = {(cuuint32_t)kChunkK, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(map, dtype == RAG_DTYPE_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2,
                     const_cast<void*>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(RAG_ECUDA, "cuTensorMapEncodeTiled failed with CUresult %d (rows %lld cols %d box %d)", (int)r, rows, cols, box_rows);
    return RAG_OK;
}

// activations: NHWC [n][S][S][C] seen through an im2col window for a width x width 'same' filter:
// base points range over [-pad, S - 1 - pad] in both directions (lower corner -pad, upper corner
// -pad), every load brings kTileM points x 64 channels, off-board points are zero-filled
static int make_map_im2col(CUtensorMap* map, const void* ptr, int n, int S, int C, int width, int dtype) {
    static EncodeIm2colFn enc = (EncodeIm2colFn)driver_symbol("cuTensorMapEncodeIm2col");
    if (!enc) return fail(RAG_ECUDA, "cuTensorMapEncodeIm2col is not available from this driver");
    const int pad = width / 2;
    cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)S, (cuuint64_t)S, (cuuint64_t)n};
    cuuint64_t strides[3] = {(cuuint64_t)C * 2, (cuuint64_t)S * C * 2, (cuuint64_t)S * S * C * 2};
    int lower[2] = {-pad, -pad}, upper[2] = {-pad, -pad};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    CUresult r = enc(map, dtype == RAG_DTYPE_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 4,
                     const_cast<void*>(ptr), dims, strides, lower, upper, (cuuint32_t)kChunkK, (cuuint32_t)kTileM, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(RAG_ECUDA, "cuTensorMapEncodeIm2col failed with CUresult %d (n %d S %d C %d width %d)", (int)r, n, S, C, width);
    return RAG_OK;
}

static int round_up(int v, int m) { return (v + m - 1) / m * m; }

struct Layer {
    int width, cin, cout, cin_pad, n_pad, taps;
    // split operands: the first lo_in input channels come as hi + lo.  Input row / weight columns per tap:
    //   lo_in = 0        [hi]                          [W_hi]
    //   lo_in = cin_pad  [hi | lo]                     [W_hi | W_hi | W_lo]            (fp16x3; A chunks hi, lo, hi)
    //   lo_in = 64       [hi | lo 0..63]               [W_hi | W_hi 0..63 | W_lo 0..63] (A chunks hi.., lo, hi chunk 0)
    //   lo_in = 32       [hi | lo 0..31 | hi 0..31]    [W_hi | W_hi 0..31 | W_lo 0..31] (one extra chunk)
    int lo_in = 0;
    int cin_store = 0;            // channels of this layer's input rows
    int kdim = 0;                 // K extent of the weights per tap
    unsigned long long a_map = 0; // K chunk -> chunk of the input row, 4 bits each
    unsigned long long k_map = 0; // K chunk -> K=16 steps that hold real channels, 4 bits each
    std::vector<int> out_order;   // position j of this layer's output holds filter out_order[j] (energetic first)
    CUtensorMap tmB, tmB_pair;    // weight tiles, encoded by rag_net_load_weights
    struct ACache { const void* ptr; int n; CUtensorMap map; };
    std::vector<ACache> a_maps;   // im2col views of recently used input buffers
    int a_next = 0;
    std::vector<ACache> o_maps;   // store views of recently used output buffers
    int o_next = 0;
    uint16_t* d_w = nullptr;      // [taps][n_pad][kdim]
    float* d_bias = nullptr;      // [n_pad]
    float* d_bn_scale = nullptr;  // residual nets: BatchNormalization applied to this layer's INPUT, [S] each
    float* d_bn_shift = nullptr;
};

}  // namespace rag

using namespace rag;

struct rag_net {
    rag_net_desc desc;
    int device = 0, sms = 148;
    std::vector<Layer> layers;
    int max_c = 0;                 // widest padded channel count of any hidden activation
    float* d_head_w = nullptr;     // [n_pad of last layer]  1x1 conv weights
    float head_b = 0.f;
    float* d_pos_bias = nullptr;   // policy: Bias layer [S*S]
    float* d_dense1_w = nullptr;   // value: [S*S][D]
    float* d_dense1_b = nullptr;
    float* d_dense2_w = nullptr;   // [D]
    float* d_dense2_b = nullptr;   // [1]
    bool split = false;            // RAG_DTYPE_FP16X3: fp16 operands, every product formed as hi*hi + lo*hi + hi*lo
    int corr = 0;                  // RAG_DTYPE_FP16C32 / FP16C64: the same for the 32 / 64 most energetic input channels of layers 2..
    int max_pitch = 0;             // widest activation row any layer writes (channels)
    std::vector<float> in_density; // expected mean of every input plane (rag_net_set_input_planes), for the channel ranking
    int op_dtype = RAG_DTYPE_FP16; // what the tensor cores are fed
    bool loaded = false;
    bool profiling = false;        // record a CUDA event around every convolution launch
    std::vector<cudaEvent_t> events;   // layers + 1
};

namespace rag {

static size_t conv_smem_bytes(int N, int stages) {
    return 1024 + (size_t)stages * (kATileBytes + (size_t)N * 128) + 4 * kOutBoxBytes + (2 * stages + 4) * 8 + 16 + 2 * (size_t)N * 4;
}

// rag_set_conv_pair: layers whose filter count allows it can run on CTA pairs (cta_group::2): each CTA
// loads its own 128 rows of activations and HALF of the weight tile, one M=256 UMMA per pair reads both
// halves -- 30 % less operand traffic into shared memory per flop, 0.735 instead of 0.843 ms per fp16
// 3x3 layer.  OFF by default (RAG_CONV_PAIR=1 in the environment or rag_set_conv_pair(1) turns it on):
// beside the feature kernels of the pipeline the cluster launch sporadically never finishes on B200 /
// driver 580: after 1,000-4,000-step runs at every batch size had all completed, soaks of the pipelined
// path (tools/pair_soak.py, 1,200 steps per process) hung in 9 of 26 processes with this kernel -- in
// eager launches and in graph replays alike, more often when the pipeline is drained and restarted every
// 100 steps -- and in 0 of 22 with one CTA per tile.  The stall is not in any mbarrier wait of this kernel
// (those record where they stand and trap after ~4 s, rag_conv_stall_report; nothing was ever recorded),
// which leaves the co-scheduling of the CTA pair itself or tcgen05.alloc.  It needs pair CTAs and
// CTAs of another kernel on the same SM: with feature kernels that prefer a large L1 (RAG_FEATURE_CARVEOUT=0:
// an SM drains before it takes a convolution CTA) 20 of 20 processes completed -- at the throughput of one CTA
// per tile, because the two kinds of kernels then take turns -- with the all-shared or the 196 KB split 12 of 36
// hung.  Alone on the GPU the pair kernel never stalled.
static int g_conv_pair = getenv("RAG_CONV_PAIR") ? atoi(getenv("RAG_CONV_PAIR")) != 0 : 0;

static int pick_stages(int N) {
    for (int st = 6; st >= 2; st--)
        if (conv_smem_bytes(N, st) <= 225 * 1024) return st;
    return 0;
}

struct NetBuffers {
    uint16_t* lin[2];        // residual nets: unit outputs (the skip path)
    uint16_t* act[2];
    float* logits;
    uint16_t* in_rows;       // only for the NCHW entry
    size_t act_bytes, in_bytes, total;
};

// rows of a dense pixel-row buffer for n boards, rounded up to whole tiles
static long long dense_rows(int n, int S) {
    long long r = (long long)n * S * S;
    return (r + 2 * kTileM - 1) / (2 * kTileM) * (2 * kTileM);
}

static NetBuffers carve_net(const rag_net* net, int n, void* ws, bool with_input) {
    NetBuffers b;
    auto al = [](size_t v) { return (v + 1023) / 1024 * 1024; };
    const long long rows = dense_rows(n, net->desc.board);
    b.act_bytes = al((size_t)rows * net->max_pitch * 2);
    b.in_bytes = with_input ? al((size_t)rows * net->layers[0].cin_store * 2) : 0;
    size_t logits_bytes = al((size_t)n * net->desc.board * net->desc.board * 4);
    char* p = (char*)ws;
    b.act[0] = (uint16_t*)p; p += b.act_bytes;
    b.act[1] = (uint16_t*)p; p += b.act_bytes;
    b.logits = (float*)p; p += logits_bytes;
    b.in_rows = (uint16_t*)p; p += b.in_bytes;
    b.lin[0] = b.lin[1] = nullptr;
    size_t lin_bytes = 0;
    if (net->desc.resnet) {
        b.lin[0] = (uint16_t*)p; p += b.act_bytes;
        b.lin[1] = (uint16_t*)p; p += b.act_bytes;
        lin_bytes = 2 * b.act_bytes;
    }
    b.total = 2 * b.act_bytes + logits_bytes + b.in_bytes + lin_bytes;
    return b;
}

// the im2col view of an input buffer: encoded once per (buffer, batch size) and kept with the layer
static int input_map(rag_net* net, Layer& ly, const void* ptr, int n, const CUtensorMap** out) {
    for (auto& e : ly.a_maps)
        if (e.ptr == ptr && e.n == n) { *out = &e.map; return RAG_OK; }
    Layer::ACache e;
    e.ptr = ptr; e.n = n;
    if (int rc = make_map_im2col(&e.map, ptr, n, net->desc.board, ly.cin_store, ly.width, net->op_dtype)) return rc;
    if (ly.a_maps.size() < 16) { ly.a_maps.push_back(e); *out = &ly.a_maps.back().map; }
    else { ly.a_maps[ly.a_next] = e; *out = &ly.a_maps[ly.a_next].map; ly.a_next = (ly.a_next + 1) % 16; }
    return RAG_OK;
}

// the store view of an output buffer: [n*S*S rows][pitch channels] in boxes of 32 rows x 64 channels
static int output_map(rag_net* net, Layer& ly, const void* ptr, int n, int pitch, const CUtensorMap** out) {
    for (auto& e : ly.o_maps)
        if (e.ptr == ptr && e.n == n) { *out = &e.map; return RAG_OK; }
    Layer::ACache e;
    e.ptr = ptr; e.n = n;
    const int S = net->desc.board;
    if (int rc = make_map(&e.map, ptr, (long long)n * S * S, pitch, 32, net->op_dtype)) return rc;
    if (ly.o_maps.size() < 16) { ly.o_maps.push_back(e); *out = &ly.o_maps.back().map; }
    else { ly.o_maps[ly.o_next] = e; *out = &ly.o_maps[ly.o_next].map; ly.o_next = (ly.o_next + 1) % 16; }
    return RAG_OK;
}

static StallRecord* g_stall_host = nullptr;

// one record in portable host-mapped memory for the process; every device that runs a forward gets its address
// (the first forward on a device does this: make it before any stream capture, like every first call here)
static int stall_record_ready() {
    static bool done[64] = {false};
    int dev = 0;
    RAG_CUDA_CHECK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return fail(RAG_EUNSUPPORTED, "device ordinal %d not supported", dev);
    if (done[dev]) return RAG_OK;
    if (!g_stall_host) {
        RAG_CUDA_CHECK(cudaHostAlloc((void**)&g_stall_host, sizeof(StallRecord), cudaHostAllocMapped | cudaHostAllocPortable));
        memset(g_stall_host, 0, sizeof(StallRecord));
    }
    StallRecord* dev_view = nullptr;
    RAG_CUDA_CHECK(cudaHostGetDevicePointer((void**)&dev_view, g_stall_host, 0));
    RAG_CUDA_CHECK(cudaMemcpyToSymbol(g_stall_record, &dev_view, sizeof dev_view));
    done[dev] = true;
    return RAG_OK;
}

static int conv_smem_limit(bool pair, size_t smem) {
    static size_t set[2] = {0, 0};      // largest dynamic shared memory size already granted to each kernel
    if (smem <= set[pair]) return RAG_OK;
    if (pair) {
        RAG_CUDA_CHECK(cudaFuncSetAttribute(conv_umma_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        if (getenv("RAG_CONV_CARVEOUT"))      // experiment only: an explicit shared-memory / L1 preference on the cluster kernel
            RAG_CUDA_CHECK(cudaFuncSetAttribute(conv_umma_kernel<true>, cudaFuncAttributePreferredSharedMemoryCarveout, atoi(getenv("RAG_CONV_CARVEOUT"))));
    }
    else RAG_CUDA_CHECK(cudaFuncSetAttribute(conv_umma_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    set[pair] = smem;
    return RAG_OK;
}

static int run_layers(rag_net* net, const uint16_t* d_in, int n, float* d_out, const NetBuffers& b, cudaStream_t s) {
    if (int rc = stall_record_ready()) return rc;
    const int S = net->desc.board, L = (int)net->layers.size();
    int lin_cur = 0;
    const uint16_t* cur = d_in;
    for (int l = 0; l < L; l++) {
        Layer& ly = net->layers[l];
        const bool last = l == L - 1;
        uint16_t* dst = last ? nullptr : b.act[l & 1];

        ConvParams p;
        memset(&p, 0, sizeof p);
        p.rows = (long long)n * S * S;
        p.tiles = (int)((p.rows + kTileM - 1) / kTileM);
        p.width = ly.width; p.kchunks = ly.kdim / kChunkK; p.N = ly.n_pad;
        p.stages = pick_stages(ly.n_pad); p.dtype = net->op_dtype;
        p.a_map = ly.a_map; p.k_map = ly.k_map;
        static const int dbg = getenv("RAG_CONV_DEBUG") ? atoi(getenv("RAG_CONV_DEBUG")) : 0;
        static const int dbg_stages = getenv("RAG_CONV_STAGES") ? atoi(getenv("RAG_CONV_STAGES")) : 0;
        p.debug = dbg;
        if (dbg_stages >= 2 && dbg_stages < p.stages) p.stages = dbg_stages;
        if (!last) {
            const Layer& nx = net->layers[l + 1];
            p.out_pitch = nx.cin_store; p.lo_cols = nx.lo_in; p.dup_hi = nx.lo_in == 32;
        }
        p.S = S; p.n = n;
        p.mode = last ? 1 : 0; p.relu = 1;
        p.bias = ly.d_bias; p.head_w = net->d_head_w; p.head_b = net->head_b;
        p.out = last ? (void*)b.logits : (void*)dst;
        if (net->desc.resnet) {
            // conv l ends a unit (or is the first convolution): its linear output is the next skip path
            p.mode = last ? 3 : 2;
            const bool unit_end = net->desc.unit_end[l] != 0;
            if (l > 0 && unit_end) p.residual = b.lin[lin_cur];
            if (unit_end && !last) { p.linear_out = b.lin[l == 0 ? lin_cur : 1 - lin_cur]; if (l > 0) lin_cur = 1 - lin_cur; }
            if (!last) { p.bn_scale = net->layers[l + 1].d_bn_scale; p.bn_shift = net->layers[l + 1].d_bn_shift; }
        }

        // a CTA pair (cta_group::2) needs an even split of the weight tile into 8-row groups
        const bool pair = g_conv_pair && (ly.n_pad % 32 == 0) && net->sms >= 2;
        if (pair) {
            p.tiles = (int)((p.rows + 2 * kTileM - 1) / (2 * kTileM));
            p.stages = pick_stages(ly.n_pad / 2);
        }
        const CUtensorMap* tmA = nullptr;
        if (int rc = input_map(net, ly, cur, n, &tmA)) return rc;
        const CUtensorMap* tmO = tmA;                      // only read by layers that write plain pixel rows
        if (p.mode == 0)
            if (int rc = output_map(net, ly, dst, n, p.out_pitch, &tmO)) return rc;
        if (net->profiling) RAG_CUDA_CHECK(cudaEventRecord(net->events[l], s));
        if (pair) {
            const size_t smem = conv_smem_bytes(ly.n_pad / 2, p.stages) + (size_t)ly.n_pad * 4;      // bias/head tables are N wide
            if (int rc = conv_smem_limit(true, smem)) return rc;
            const int pairs = p.tiles < net->sms / 2 ? p.tiles : net->sms / 2;
            cudaLaunchConfig_t cfg;
            memset(&cfg, 0, sizeof cfg);
            cfg.gridDim = dim3(2 * pairs); cfg.blockDim = dim3(kConvThreads); cfg.dynamicSmemBytes = smem; cfg.stream = s;
            cudaLaunchAttribute attr[1];
            attr[0].id = cudaLaunchAttributeClusterDimension;
            attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
            cfg.attrs = attr; cfg.numAttrs = 1;
            RAG_CUDA_CHECK(cudaLaunchKernelEx(&cfg, conv_umma_kernel<true>, *tmA, ly.tmB_pair, *tmO, p));
        } else {
            const size_t smem = conv_smem_bytes(ly.n_pad, p.stages);
            if (int rc = conv_smem_limit(false, smem)) return rc;
            const int grid = p.tiles < net->sms ? p.tiles : net->sms;
            conv_umma_kernel<false><<<grid, kConvThreads, smem, s>>>(*tmA, ly.tmB, *tmO, p);
        }
        RAG_CUDA_CHECK(cudaGetLastError());
        if (net->profiling && last) RAG_CUDA_CHECK(cudaEventRecord(net->events[L], s));
        cur = dst;
    }
    const int P = S * S;
    if (net->desc.head == RAG_HEAD_POLICY) {
        softmax_head_kernel<<<(n + 3) / 4, 128, 0, s>>>(b.logits, net->d_pos_bias, n, P, d_out);
    } else {
        value_head_kernel<<<n, 256, 0, s>>>(b.logits, net->d_dense1_w, net->d_dense1_b, net->d_dense2_w, net->d_dense2_b,
                                            n, P, net->desc.dense_units, d_out);
    }
    RAG_CUDA_CHECK(cudaGetLastError());
    return RAG_OK;
}

static uint16_t to16(float f, int dtype) {
    if (dtype == RAG_DTYPE_BF16) { __nv_bfloat16 h = __float2bfloat16_rn(f); uint16_t u; memcpy(&u, &h, 2); return u; }
    __half h = __float2half_rn(f); uint16_t u; memcpy(&u, &h, 2); return u;
}

template <typename T>
static int upload(T** dptr, const std::vector<T>& host) {
    if (*dptr) { cudaFree(*dptr); *dptr = nullptr; }
    RAG_CUDA_CHECK(cudaMalloc((void**)dptr, host.size() * sizeof(T)));
    RAG_CUDA_CHECK(cudaMemcpy(*dptr, host.data(), host.size() * sizeof(T), cudaMemcpyHostToDevice));
    return RAG_OK;
}

}  // namespace rag

extern "C" {

int rag_net_create(const rag_net_desc* desc, rag_net** out)
{
    if (!desc || !out) return fail(RAG_EINVAL, "null argument");
    if (desc->board < 2 || desc->board > 19) return fail(RAG_EUNSUPPORTED, "board %d not in [2, 19]", desc->board);
    if (desc->layers < 1 || desc->layers > RAG_MAX_LAYERS) return fail(RAG_EINVAL, "layers %d not in [1, %d]", desc->layers, RAG_MAX_LAYERS);
    if (desc->in_planes < 1 || desc->in_planes > 128) return fail(RAG_EUNSUPPORTED, "in_planes %d not in [1, 128]", desc->in_planes);
    if (desc->head != RAG_HEAD_POLICY && desc->head != RAG_HEAD_VALUE) return fail(RAG_EINVAL, "head %d", desc->head);
    if (desc->dtype < RAG_DTYPE_BF16 || desc->dtype > RAG_DTYPE_FP16C64) return fail(RAG_EINVAL, "dtype %d", desc->dtype);
    if (desc->dtype >= RAG_DTYPE_FP16X3 && desc->resnet) return fail(RAG_EUNSUPPORTED, "split-operand precision is not available for residual networks");
    if (desc->dtype == RAG_DTYPE_FP16X3 && desc->in_planes > 64) return fail(RAG_EUNSUPPORTED, "split-operand precision takes at most 64 input planes (%d given)", desc->in_planes);
    if (desc->head == RAG_HEAD_VALUE && (desc->dense_units < 1 || desc->dense_units > 4096)) return fail(RAG_EINVAL, "dense_units %d", desc->dense_units);
    if (desc->resnet) {
        if (desc->head != RAG_HEAD_POLICY) return fail(RAG_EUNSUPPORTED, "residual networks are policy networks (policy.py:141)");
        for (int l = 1; l < desc->layers; l++)
            if (desc->filters[l] != desc->filters[0]) return fail(RAG_EINVAL, "a residual network needs the same filter count in every layer");
        if (!desc->unit_end[desc->layers - 1]) return fail(RAG_EINVAL, "the last convolution must close a residual unit");
    }
    rag_net* net = new rag_net();
    net->desc = *desc;
    net->split = desc->dtype == RAG_DTYPE_FP16X3;
    net->corr = desc->dtype == RAG_DTYPE_FP16C32 ? 32 : desc->dtype == RAG_DTYPE_FP16C64 ? 64 : 0;
    net->in_density.assign(desc->in_planes, 0.25f);
    net->op_dtype = desc->dtype == RAG_DTYPE_BF16 ? RAG_DTYPE_BF16 : RAG_DTYPE_FP16;
    int cin = desc->in_planes;
    for (int l = 0; l < desc->layers; l++) {
        Layer ly;
        ly.width = desc->widths[l]; ly.cin = cin; ly.cout = desc->filters[l];
        if (ly.width < 1 || ly.width > 7 || ly.width % 2 == 0) { delete net; return fail(RAG_EUNSUPPORTED, "filter_width_%d = %d (odd, <= 7 supported)", l + 1, ly.width); }
        if (ly.cout < 1 || ly.cout > 256) { delete net; return fail(RAG_EUNSUPPORTED, "filters_per_layer_%d = %d (<= 256 supported)", l + 1, ly.cout); }
        ly.cin_pad = round_up(cin, kChunkK); ly.n_pad = round_up(ly.cout, kChunkK); ly.taps = ly.width * ly.width;
        // the network's input planes are taken as exact in fp16 (feature planes are 0/1), so the first
        // layer of the FP16C modes has nothing to correct on its activation side
        ly.lo_in = net->split ? ly.cin_pad : (l > 0 ? net->corr : 0);
        const int nc = ly.cin_pad / kChunkK;
        unsigned long long map = 0;
        for (int c = 0; c < nc; c++) map |= (unsigned long long)c << (4 * c);
        if (ly.lo_in == 0) { ly.cin_store = ly.cin_pad; ly.kdim = ly.cin_pad; }
        else if (net->split) {
            ly.cin_store = 2 * ly.cin_pad; ly.kdim = 3 * ly.cin_pad;
            for (int c = 0; c < nc; c++) map |= (unsigned long long)(nc + c) << (4 * (nc + c)) | (unsigned long long)c << (4 * (2 * nc + c));
        } else {
            ly.cin_store = ly.cin_pad + 64; ly.kdim = ly.cin_pad + (ly.lo_in == 64 ? 128 : 64);
            map |= (unsigned long long)nc << (4 * nc);            // [lo 0..63] or [lo 0..31 | hi 0..31]; then chunk 0 again
        }
        ly.a_map = map;
        // K=16 steps per chunk: a chunk of plain channels holds min(64, cin - 64 c) real ones; the
        // corrected-channel chunks hold lo_in (and, for FP16C32, the copied hi channels after them)
        auto steps = [](int channels) { return (unsigned long long)((channels < 1 ? 1 : channels > 64 ? 64 : channels) + 15) / 16; };
        unsigned long long km = 0;
        const int total_chunks = ly.kdim / kChunkK;
        for (int c = 0; c < total_chunks; c++) {
            int real;
            if (net->split || ly.lo_in == 0 || c < nc) real = cin - 64 * (c % nc);
            else if (ly.lo_in == 64) real = cin < 64 ? cin : 64;
            else real = 64;                                            // [lo 0..31 | hi 0..31]: the hi half sits at 32..63
            km |= steps(real) << (4 * c);
        }
        ly.k_map = km;
        if (ly.kdim / kChunkK > 16 || ly.cin_store / kChunkK > 16) { delete net; return fail(RAG_EUNSUPPORTED, "layer %d has too many input channels for this precision mode", l + 1); }
        if (pick_stages(ly.n_pad) < 2) { delete net; return fail(RAG_EUNSUPPORTED, "layer %d does not fit shared memory", l + 1); }
        if (ly.n_pad > net->max_c) net->max_c = ly.n_pad;
        if (l > 0 && ly.cin_store > net->max_pitch) net->max_pitch = ly.cin_store;
        if (ly.n_pad > net->max_pitch) net->max_pitch = ly.n_pad;
        net->layers.push_back(ly);
        cin = ly.cout;
    }
    if (cudaGetDevice(&net->device) != cudaSuccess ||
        cudaDeviceGetAttribute(&net->sms, cudaDevAttrMultiProcessorCount, net->device) != cudaSuccess) {
        delete net; return fail(RAG_ECUDA, "no CUDA device");
    }
    int major = 0;
    cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, net->device);
    if (major != 10) { delete net; return fail(RAG_EUNSUPPORTED, "needs an sm_100a device (compute capability %d.x found)", major); }
    *out = net;
    return RAG_OK;
}

int rag_net_destroy(rag_net* net)
{
    if (!net) return RAG_OK;
    for (auto& ly : net->layers) { cudaFree(ly.d_w); cudaFree(ly.d_bias); cudaFree(ly.d_bn_scale); cudaFree(ly.d_bn_shift); }
    for (auto& e : net->events) cudaEventDestroy(e);
    cudaFree(net->d_head_w); cudaFree(net->d_pos_bias);
    cudaFree(net->d_dense1_w); cudaFree(net->d_dense1_b); cudaFree(net->d_dense2_w); cudaFree(net->d_dense2_b);
    delete net;
    return RAG_OK;
}

int rag_net_num_weight_arrays(const rag_net* net)
{
    if (!net) return 0;
    const int L = (int)net->layers.size();
    return 2 * L + (net->desc.resnet ? 4 * (L - 1) : 0) + 2 + (net->desc.head == RAG_HEAD_POLICY ? 1 : 4);
}

int rag_net_load_weights(rag_net* net, const float* const* h_arrays, int n_arrays)
{
    if (!net || !h_arrays) return fail(RAG_EINVAL, "null argument");
    if (n_arrays != rag_net_num_weight_arrays(net)) return fail(RAG_EINVAL, "expected %d weight arrays, got %d", rag_net_num_weight_arrays(net), n_arrays);
    for (int i = 0; i < n_arrays; i++) if (!h_arrays[i]) return fail(RAG_EINVAL, "weight array %d is null", i);
    RAG_CUDA_CHECK(cudaSetDevice(net->device));
    const int dtype = net->op_dtype, L = (int)net->layers.size(), P = net->desc.board * net->desc.board;
    int ai = 0;
    std::vector<float> mean_in = net->in_density;     // expected mean of every input channel, original order
    std::vector<int> in_order;                        // input position -> the producing layer's filter (empty: identity)
    for (int l = 0; l < L; l++) {
        Layer& ly = net->layers[l];
        if (net->desc.resnet && l > 0) {
            // BatchNormalization() in front of this convolution (policy.py:226-228), Keras-1 defaults:
            // mode 0, axis -1 (= the y coordinate in 'th' ordering), epsilon 1e-3; at inference
            // x -> gamma * (x - running_mean) / sqrt(running_std + eps) + beta  (running_std holds the variance)
            const float* gamma = h_arrays[ai++]; const float* beta = h_arrays[ai++];
            const float* mean = h_arrays[ai++];  const float* var = h_arrays[ai++];
            const int S = net->desc.board;
            std::vector<float> sc(S), sh(S);
            for (int y = 0; y < S; y++) { sc[y] = gamma[y] / sqrtf(var[y] + 1e-3f); sh[y] = beta[y] - mean[y] * sc[y]; }
            if (int rc = upload(&ly.d_bn_scale, sc)) return rc;
            if (int rc = upload(&ly.d_bn_shift, sh)) return rc;
        }
        const float* W = h_arrays[ai++];      // (cout, cin, kh, kw)  Keras 'th' kernel layout
        const float* bsrc = h_arrays[ai++];   // (cout,)
        const int k = ly.width;
        // Which output channels will carry the most energy?  ReLU activations are non-negative, so a
        // filter's mean response to the mean input tells: rank the filters by it and put the energetic
        // ones first when the next layer corrects its leading input channels (FP16C32 / FP16C64).
        std::vector<float> resp(ly.cout);
        for (int co = 0; co < ly.cout; co++) {
            double r = bsrc[co];
            for (int ci = 0; ci < ly.cin; ci++) {
                double sw = 0;
                for (int t = 0; t < k * k; t++) sw += W[((size_t)co * ly.cin + ci) * k * k + t];
                r += sw * mean_in[ci];
            }
            resp[co] = (float)r;
        }
        ly.out_order.resize(ly.cout);
        for (int co = 0; co < ly.cout; co++) ly.out_order[co] = co;
        if (l + 1 < L && net->corr)
            std::stable_sort(ly.out_order.begin(), ly.out_order.end(), [&](int a, int b) { return resp[a] > resp[b]; });
        std::vector<uint16_t> w((size_t)ly.taps * ly.n_pad * ly.kdim, 0);
        for (int i = 0; i < k; i++)
            for (int j = 0; j < k; j++) {
                // tap (i, j) reads in[x + i - pad, y + j - pad].  Theano's conv2d is a true
                // convolution, i.e. a correlation with the kernel rotated by 180 degrees.
                const int si = net->desc.kernel_flip ? k - 1 - i : i, sj = net->desc.kernel_flip ? k - 1 - j : j;
                for (int jo = 0; jo < ly.cout; jo++)
                    for (int ji = 0; ji < ly.cin; ji++) {
                        const int co = ly.out_order[jo], ci = in_order.empty() ? ji : in_order[ji];
                        const float v = W[(((size_t)co * ly.cin + ci) * k + si) * k + sj];
                        uint16_t* row = &w[((size_t)(i * k + j) * ly.n_pad + jo) * ly.kdim];
                        row[ji] = to16(v, dtype);
                        if (ji < ly.lo_in) {
                            // the extra K chunks pair lo(x) with W_hi and hi(x) with W_lo; the lo.lo term (2^-22 relative) is dropped
                            const uint16_t lo = to16(v - __half2float(__float2half_rn(v)), dtype);
                            if (net->split) { row[ly.cin_pad + ji] = row[ji]; row[2 * ly.cin_pad + ji] = lo; }
                            else if (ly.lo_in == 64) { row[ly.cin_pad + ji] = row[ji]; row[ly.cin_pad + 64 + ji] = lo; }
                            else { row[ly.cin_pad + ji] = row[ji]; row[ly.cin_pad + 32 + ji] = lo; }
                        }
                    }
            }
        std::vector<float> bias(ly.n_pad, 0.f);
        for (int jo = 0; jo < ly.cout; jo++) bias[jo] = bsrc[ly.out_order[jo]];
        if (int rc = upload(&ly.d_w, w)) return rc;
        if (int rc = upload(&ly.d_bias, bias)) return rc;
        if (int rc = make_map(&ly.tmB, ly.d_w, (long long)ly.taps * ly.n_pad, ly.kdim, ly.n_pad, dtype)) return rc;
        if (ly.n_pad % 32 == 0)
            if (int rc = make_map(&ly.tmB_pair, ly.d_w, (long long)ly.taps * ly.n_pad, ly.kdim, ly.n_pad / 2, dtype)) return rc;
        mean_in.assign(ly.cout, 0.f);
        for (int co = 0; co < ly.cout; co++) mean_in[co] = resp[co] > 0.f ? resp[co] : 0.f;
        in_order = ly.out_order;
    }
    {   // the 1x1 convolution to one plane (policy.py:125-130, value.py:23-25): W (1, cin, 1, 1), b (1,)
        const Layer& ly = net->layers[L - 1];
        const float* W = h_arrays[ai++];
        const float* bsrc = h_arrays[ai++];
        std::vector<float> hw(ly.n_pad, 0.f);
        for (int c = 0; c < ly.cout; c++) hw[c] = W[ly.out_order[c]];
        if (int rc = upload(&net->d_head_w, hw)) return rc;
        net->head_b = bsrc[0];
    }
    if (net->desc.head == RAG_HEAD_POLICY) {
        std::vector<float> pb(h_arrays[ai], h_arrays[ai] + P); ai++;
        if (int rc = upload(&net->d_pos_bias, pb)) return rc;
    } else {
        const int D = net->desc.dense_units;
        std::vector<float> w1(h_arrays[ai], h_arrays[ai] + (size_t)P * D); ai++;
        std::vector<float> b1(h_arrays[ai], h_arrays[ai] + D); ai++;
        std::vector<float> w2(h_arrays[ai], h_arrays[ai] + D); ai++;
        std::vector<float> b2(h_arrays[ai], h_arrays[ai] + 1); ai++;
        if (int rc = upload(&net->d_dense1_w, w1)) return rc;
        if (int rc = upload(&net->d_dense1_b, b1)) return rc;
        if (int rc = upload(&net->d_dense2_w, w2)) return rc;
        if (int rc = upload(&net->d_dense2_b, b2)) return rc;
    }
    net->loaded = true;
    return RAG_OK;
}

size_t rag_net_workspace_bytes(const rag_net* net, int n)
{
    if (!net || n <= 0) return 0;
    return carve_net(net, n, nullptr, true).total;
}

// Mean of every canonical feature plane (include/rocalphago_b200.h numbering) over random-playout
// 19x19 positions -- only used to rank channels by expected energy, never in the arithmetic.
static const float kPlaneDensity[RAG_NUM_PLANES] = {
    0.26f, 0.26f, 0.48f, 1.00f, 0.003f, 0.003f, 0.003f, 0.003f, 0.003f, 0.003f, 0.003f, 0.50f,
    0.052f, 0.071f, 0.085f, 0.074f, 0.048f, 0.040f, 0.029f, 0.117f,
    0.457f, 0.009f, 0.002f, 0.001f, 0.0005f, 0.0003f, 0.0002f, 0.0006f,
    0.013f, 0.002f, 0.001f, 0.0006f, 0.0004f, 0.0003f, 0.0002f, 0.0007f,
    0.018f, 0.047f, 0.107f, 0.155f, 0.041f, 0.041f, 0.020f, 0.041f,
    0.014f, 0.007f, 0.466f, 0.0f, 0.471f, 0.50f};

int rag_net_set_input_planes(rag_net* net, const int32_t* h_plane_src, int F)
{
    if (!net || !h_plane_src) return fail(RAG_EINVAL, "null argument");
    if (F != net->desc.in_planes) return fail(RAG_EINVAL, "network takes %d planes, %d listed", net->desc.in_planes, F);
    for (int f = 0; f < F; f++) {
        if (h_plane_src[f] < 0 || h_plane_src[f] >= RAG_NUM_PLANES) return fail(RAG_EINVAL, "plane %d is not a canonical plane", h_plane_src[f]);
        net->in_density[f] = kPlaneDensity[h_plane_src[f]];
    }
    return RAG_OK;
}

int rag_conv_stall_report(char* buf, int size)
{
    if (!buf || size < 1) return fail(RAG_EINVAL, "null argument");
    buf[0] = 0;
    if (!g_stall_host || g_stall_host->magic != 0x57A11u) return 0;
    static const char* what[] = {"?", "TMA producer waiting for a free stage", "MMA lane waiting for a drained accumulator",
                                 "MMA lane waiting for a loaded stage", "epilogue waiting for a finished accumulator"};
    const StallRecord& r = *g_stall_host;
    snprintf(buf, (size_t)size, "convolution kernel gave up after %lld cycles: %s (block %u, thread %u, tile %u, slot %u, parity %u)",
             kStallCycles, what[r.code < 5 ? r.code : 0], r.block, r.thread, r.tile, r.stage, r.parity);
    return 1;
}

int rag_set_conv_pair(int enable) { int old = g_conv_pair; g_conv_pair = enable != 0; return old; }

int rag_net_set_profiling(rag_net* net, int enable)
{
    if (!net) return fail(RAG_EINVAL, "null net");
    if (enable && net->events.empty()) {
        net->events.resize(net->layers.size() + 1);
        for (auto& e : net->events) RAG_CUDA_CHECK(cudaEventCreate(&e));
    }
    net->profiling = enable != 0;
    return RAG_OK;
}

int rag_net_layer_times(rag_net* net, float* h_ms, int n_layers)
{
    if (!net || !h_ms) return fail(RAG_EINVAL, "null argument");
    if (!net->profiling || n_layers != (int)net->layers.size()) return fail(RAG_EINVAL, "profiling is off or n_layers != %zu", net->layers.size());
    for (int l = 0; l < n_layers; l++) {
        RAG_CUDA_CHECK(cudaEventSynchronize(net->events[l + 1]));
        RAG_CUDA_CHECK(cudaEventElapsedTime(&h_ms[l], net->events[l], net->events[l + 1]));
    }
    return RAG_OK;
}

size_t rag_net_logits_offset(const rag_net* net, int n)
{
    if (!net || n <= 0) return 0;
    NetBuffers b = carve_net(net, n, nullptr, true);
    return (size_t)((char*)b.logits - (char*)nullptr);
}

long long rag_net_input_rows(const rag_net* net, int n) { return net ? dense_rows(n, net->desc.board) : 0; }
int rag_net_input_channels(const rag_net* net) { return net ? net->layers[0].cin_store : 0; }

static int net_ready(const rag_net* net, int n, const void* in, const void* out, const void* ws, size_t bytes) {
    if (!net) return fail(RAG_EINVAL, "null net");
    if (!net->loaded) return fail(RAG_EINVAL, "rag_net_load_weights has not been called");
    if (n < 0 || (n > 0 && (!in || !out || !ws))) return fail(RAG_EINVAL, "null argument");
    if (n > 0 && bytes < rag_net_workspace_bytes(net, n)) return fail(RAG_ENOMEM, "workspace has %zu bytes, %zu needed for n = %d", bytes, rag_net_workspace_bytes(net, n), n);
    return RAG_OK;
}

int rag_net_forward_rows(rag_net* net, const void* d_in_rows, int n, float* d_out, void* d_workspace, size_t workspace_bytes, void* stream)
{
    if (int rc = net_ready(net, n, d_in_rows, d_out, d_workspace, workspace_bytes)) return rc;
    if (n == 0) return RAG_OK;
    NetBuffers b = carve_net(net, n, d_workspace, true);
    return run_layers(net, (const uint16_t*)d_in_rows, n, d_out, b, (cudaStream_t)stream);
}

int rag_net_forward(rag_net* net, const float* d_nchw, int n, float* d_out, void* d_workspace, size_t workspace_bytes, void* stream)
{
    if (int rc = net_ready(net, n, d_nchw, d_out, d_workspace, workspace_bytes)) return rc;
    if (n == 0) return RAG_OK;
    cudaStream_t s = (cudaStream_t)stream;
    NetBuffers b = carve_net(net, n, d_workspace, true);
    const Layer& l0 = net->layers[0];
    const int S = net->desc.board;
    long long threads = (long long)n * S * S * (l0.cin_store / 8);
    nchw_to_rows_kernel<<<(int)((threads + 255) / 256), 256, 0, s>>>(d_nchw, n, net->desc.in_planes, S, l0.cin_store, net->op_dtype, net->split ? 1 : 0, b.in_rows);
    RAG_CUDA_CHECK(cudaGetLastError());
    return run_layers(net, b.in_rows, n, d_out, b, s);
}


size_t rag_eval_workspace_bytes(const rag_net* net, int n)
{
    if (!net || n <= 0) return 0;
    size_t f = (rag_features_workspace_bytes(n) + 1023) / 1024 * 1024;
    return f + rag_net_workspace_bytes(net, n);
}

int rag_eval_batch(rag_net* net, const uint32_t* d_states, int n, const int32_t* h_plane_src, int F, float* d_out,
                   void* d_workspace, size_t workspace_bytes, void* stream)
{
    if (!net) return fail(RAG_EINVAL, "null net");
    if (F != net->desc.in_planes) return fail(RAG_EINVAL, "network takes %d planes, %d listed", net->desc.in_planes, F);
    if (n > 0 && workspace_bytes < rag_eval_workspace_bytes(net, n))
        return fail(RAG_ENOMEM, "workspace has %zu bytes, %zu needed for n = %d", workspace_bytes, rag_eval_workspace_bytes(net, n), n);
    if (int rc = net_ready(net, n, d_states, d_out, d_workspace, (size_t)-1)) return rc;
    if (n == 0) return RAG_OK;
    const size_t fbytes = (rag_features_workspace_bytes(n) + 1023) / 1024 * 1024;
    char* ws = (char*)d_workspace;
    NetBuffers b = carve_net(net, n, ws + fbytes, true);
    const Layer& l0 = net->layers[0];
    // feature planes are 0/1, exact in 16 bits: in split mode the lo half of the rows is the zero padding
    if (int rc = rag_features(d_states, n, net->desc.board, h_plane_src, F, RAG_LAYOUT_CONV_IN, net->op_dtype,
                              l0.cin_store, b.in_rows, ws, fbytes, stream)) return rc;
    return run_layers(net, b.in_rows, n, d_out, b, (cudaStream_t)stream);
}

}  // extern "C"
