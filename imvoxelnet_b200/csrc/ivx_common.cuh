// Shared host/device helpers of libivx (error reporting, launch checks).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include <cstdarg>
#include <cstdio>
#include <cstdlib>

#include "ivx.h"

namespace ivx {

// Message of the last failure on this thread; read through ivx_last_error().
char *last_error_buffer();

int fail(int code, const char *fmt, ...);

inline int check_cuda(cudaError_t e, const char *what) {
    if (e == cudaSuccess) return IVX_OK;
    return fail(IVX_ERR_CUDA, "%s: %s", what, cudaGetErrorString(e));
}

#define IVX_REQUIRE(cond, ...)                                         \
    do {                                                               \
        if (!(cond)) return ::ivx::fail(IVX_ERR_INVALID_ARGUMENT, __VA_ARGS__); \
    } while (0)

#define IVX_CHECK_LAUNCH(what)                                        \
    do {                                                              \
        int rc__ = ::ivx::check_cuda(cudaGetLastError(), what);       \
        if (rc__ != IVX_OK) return rc__;                              \
    } while (0)

constexpr int kNumSMs = 148;  // B200

// n / d and n % d for 0 <= n < 2^31 as one multiply-high and a shift (Granlund-Montgomery with
// p = 31 + ceil(log2 d)); d == 1 is flagged by m == 0.  Tile decoding of the persistent kernels.
struct TileDiv {
    unsigned m;
    int s, d;
    __host__ __device__ __forceinline__ int div(int n) const {
#ifdef __CUDA_ARCH__
        return m ? (int)(__umulhi((unsigned)n, m) >> s) : n;
#else
        return n / d;
#endif
    }
    __host__ __device__ __forceinline__ int divmod(int n, int &rem) const {
        const int q = div(n);
        rem = n - q * d;
        return q;
    }
};

inline TileDiv make_tilediv(int d) {
    TileDiv f = {0u, 0, d < 1 ? 1 : d};
    if (d <= 1) return f;
    int lg = 0;
    while ((1ll << lg) < d) ++lg;
    const int p = 31 + lg;
    f.m = (unsigned)(((1ull << p) + (unsigned long long)d - 1) / (unsigned long long)d);
    f.s = p - 32;
    return f;
}

#define IVX_STR2(x) #x
#define IVX_STR(x) IVX_STR2(x)

}  // namespace ivx
