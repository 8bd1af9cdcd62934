// Library identification and per-thread error reporting of libivx.
#include <cstdlib>
#include <cstring>

#include "ivx_common.cuh"

namespace ivx {

static thread_local char g_last_error[512] = "";

char *last_error_buffer() { return g_last_error; }

int fail(int code, const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_last_error, sizeof(g_last_error), fmt, ap);
    va_end(ap);
    return code;
}

}  // namespace ivx

extern "C" const char *ivx_version(void) {
    return "libivx 0.1.0 (sm_100a, CUDA " IVX_STR(CUDART_VERSION) ")";
}

extern "C" const char *ivx_last_error(void) { return ivx::last_error_buffer(); }
