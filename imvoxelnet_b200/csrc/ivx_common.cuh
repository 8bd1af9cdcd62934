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

#define IVX_STR2(x) #x
#define IVX_STR(x) IVX_STR2(x)

}  // namespace ivx
