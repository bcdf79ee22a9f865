// Probe of TMA im2col-mode semantics on sm_100a: which pixels land in which smem rows?
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o im2col_probe im2col_probe.cu -lcuda
#include <cuda.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <vector>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e_), __LINE__); exit(1); } } while (0)

__global__ void probe(const __grid_constant__ CUtensorMap tm, int c, int w, int h, int n, int woff, int hoff, uint16_t* out, int rows)
{
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint32_t raw = (uint32_t)__cvta_generic_to_shared(smem_raw);
    uint32_t base = (raw + 1023u) & ~1023u;
    uint8_t* tile = smem_raw + (base - raw);
    __shared__ __align__(8) uint64_t bar;
    uint32_t b = (uint32_t)__cvta_generic_to_shared(&bar);
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(b));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    for (int i = threadIdx.x; i < rows * 64; i += blockDim.x) ((uint16_t*)tile)[i] = 0x7bff;   // sentinel
    __syncthreads();
    if (threadIdx.x == 0) {
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(b), "r"(rows * 128) : "memory");
        uint16_t wo = (uint16_t)woff, ho = (uint16_t)hoff;
        asm volatile(
            "cp.async.bulk.tensor.4d.shared::cluster.global.im2col.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2], {%7, %8};"
            ::"r"(base), "l"(&tm), "r"(b), "r"(c), "r"(w), "r"(h), "r"(n), "h"(wo), "h"(ho) : "memory");
    }
    asm volatile(
        "{\n\t.reg .pred p;\n\tW_%=:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n\t@p bra D_%=;\n\tbra W_%=;\n\tD_%=:\n\t}" ::"r"(b) : "memory");
    __syncthreads();
    for (int i = threadIdx.x; i < rows * 64; i += blockDim.x) out[i] = ((uint16_t*)tile)[i];
}

typedef CUresult (*EncIm2col)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                              const int*, const int*, cuuint32_t, cuuint32_t, const cuuint32_t*, CUtensorMapInterleave,
                              CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main()
{
    const int N = 3, H = 19, W = 19, C = 64, rows = 128;
    std::vector<__half> host((size_t)N * H * W * C);
    for (int n = 0; n < N; n++) for (int h = 0; h < H; h++) for (int w = 0; w < W; w++) for (int c = 0; c < C; c++)
        host[(((size_t)n * H + h) * W + w) * C + c] = __float2half(c == 0 ? (float)(n * 361 + h * 19 + w + 1) : (c == 1 ? (float)c : 0.f));
    __half* d; CK(cudaMalloc(&d, host.size() * 2)); CK(cudaMemcpy(d, host.data(), host.size() * 2, cudaMemcpyHostToDevice));
    uint16_t* dout; CK(cudaMalloc(&dout, rows * 128));
    void* sym = nullptr; cudaDriverEntryPointQueryResult q;
    CK(cudaGetDriverEntryPoint("cuTensorMapEncodeIm2col", &sym, cudaEnableDefault, &q));
    EncIm2col enc = (EncIm2col)sym;
    CUtensorMap tm;
    cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N};
    cuuint64_t strides[3] = {(cuuint64_t)C * 2, (cuuint64_t)W * C * 2, (cuuint64_t)H * W * C * 2};
    int lower[2] = {-1, -1}, upper[2] = {-1, -1};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    CUresult r = enc(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 4, d, dims, strides, lower, upper, 64, rows, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("encode rc=%d\n", (int)r);
    if (r != CUDA_SUCCESS) return 1;
    CK(cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, rows * 128 + 2048));
    struct Case { int w, h, n, wo, ho; } cases[] = {
        {-1, -1, 0, 0, 0}, {-1, -1, 0, 1, 1}, {-1, -1, 0, 2, 2}, {0, 0, 0, 1, 1}, {5, 3, 0, 1, 1}, {5, 3, 0, 0, 2},
        {10, 17, 0, 1, 1}, {-1, 16, 1, 1, 1}, {3, 17, 2, 1, 1}, {17, 17, 0, 2, 2}};
    std::vector<uint16_t> out(rows * 64);
    for (auto& cs : cases) {
        probe<<<1, 128, rows * 128 + 2048>>>(tm, 0, cs.w, cs.h, cs.n, cs.wo, cs.ho, dout, rows);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("case w=%d h=%d n=%d off=(%d,%d): launch error %s\n", cs.w, cs.h, cs.n, cs.wo, cs.ho, cudaGetErrorString(e)); return 1; }
        CK(cudaMemcpy(out.data(), dout, rows * 128, cudaMemcpyDeviceToHost));
        printf("case coords(w=%d,h=%d,n=%d) offsets(w=%d,h=%d): rows ->", cs.w, cs.h, cs.n, cs.wo, cs.ho);
        for (int rr = 0; rr < rows; rr++) {
            // un-swizzle: 16-byte chunk j of row rr sits at chunk j ^ (rr & 7); channel 0 is in chunk 0
            int chunk = 0 ^ (rr & 7);
            uint16_t v = out[rr * 64 + chunk * 8 + 0];
            float f = __half2float(*reinterpret_cast<__half*>(&v));
            if (rr < 44 || rr > 120) {
                if (v == 0x7bff) printf(" [%d:untouched]", rr);
                else if (f == 0.f) printf(" [%d:0]", rr);
                else { int id = (int)f - 1; printf(" [%d:n%d h%d w%d]", rr, id / 361, (id % 361) / 19, id % 19); }
            }
        }
        printf("\n");
    }
    return 0;
}
