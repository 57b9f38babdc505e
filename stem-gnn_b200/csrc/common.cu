// common.cu -- error string, ABI version, device attribute cache.
#include <stdarg.h>

#include <mutex>

#include "common.cuh"

namespace stemgnn {
static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int sm_count() {
    static std::once_flag once;
    static int n = 148;
    std::call_once(once, [] {
        int dev = 0, v = 0;
        if (cudaGetDevice(&dev) == cudaSuccess &&
            cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && v > 0)
            n = v;
    });
    return n;
}
}  // namespace stemgnn

extern "C" const char* stemgnn_last_error(void) { return stemgnn::g_err; }
extern "C" int stemgnn_abi_version(void) { return 1; }
