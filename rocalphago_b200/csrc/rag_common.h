// rag_common.h -- error plumbing shared by the C-ABI translation units.
#pragma once
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdio.h>
#include "../../include/rocalphago_b200.h"

namespace rag {

char* last_error_buffer();                 // thread-local, 512 bytes (defined in features.cu)

inline int fail(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(last_error_buffer(), 512, fmt, ap);
    va_end(ap);
    return code;
}

#define RAG_CUDA_CHECK(expr)                                                                     \
    do {                                                                                         \
        cudaError_t _e = (expr);                                                                 \
        if (_e != cudaSuccess)                                                                   \
            return ::rag::fail(RAG_ECUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), \
                               __FILE__, __LINE__);                                              \
    } while (0)

// Geometry of the zero-padded pixel-row layout the convolutions read (see conv.cu):
// pixel (x, y) of board b lives in row  lead + b*Sp*Sp + x*Sp + y,  Sp = S + pad; every other
// row is zero and is never written, so a kxk tap is a plain row shift of dx*Sp + dy.
constexpr int kConvLeadRows = 128;
constexpr int kConvTileRows = 128;
inline long long conv_rows(int n, int S, int pad) {
    long long sp = S + pad, body = (long long)n * sp * sp;
    body = (body + kConvTileRows - 1) / kConvTileRows * kConvTileRows;
    return kConvLeadRows + body + kConvTileRows;
}

}  // namespace rag
