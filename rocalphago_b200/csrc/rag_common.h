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

// Rows of a dense pixel-row buffer (conv.cu): n*S*S, rounded up to whole 256-row tiles so that the
// last tile of a layer may be addressed in full.
inline long long conv_rows(int n, int S) {
    long long r = (long long)n * S * S;
    return (r + 255) / 256 * 256;
}

}  // namespace rag
