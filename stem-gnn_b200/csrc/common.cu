// common.cu -- error string, ABI version, device attribute cache.
#include <stdarg.h>
#include <stdlib.h>

#include <atomic>
#include <mutex>

#include "common.cuh"

namespace stemgnn {
static thread_local char g_err[512] = "";
static std::atomic<uint64_t> g_launches{0};

void count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }
uint64_t launches() { return g_launches.load(std::memory_order_relaxed); }

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
static std::atomic<int> g_sm_reserve{0};
int persistent_sms() {
    const int n = sm_count() - g_sm_reserve.load(std::memory_order_relaxed);
    return n < 8 ? 8 : n;
}
int set_sm_reserve(int n) { return g_sm_reserve.exchange(n < 0 ? 0 : n, std::memory_order_relaxed); }

bool l2_pingpong() {
    static const bool on = [] { const char* e = getenv("STEMGNN_L2_PINGPONG"); return e ? atoi(e) != 0 : true; }();
    return on;
}
}  // namespace stemgnn

extern "C" const char* stemgnn_last_error(void) { return stemgnn::g_err; }
extern "C" int stemgnn_abi_version(void) { return 6; }
extern "C" int stemgnn_set_sm_reserve(int sms) { return stemgnn::set_sm_reserve(sms); }
extern "C" uint64_t stemgnn_launch_count(void) { return stemgnn::launches(); }
