// Error channel + version of the fastb200 C ABI.
#include "common.cuh"
#include <atomic>

namespace fb {

static std::atomic<long long> g_launches{0};      // kernel launches issued through this library (captured ones included)

thread_local char g_err[512] = {0};

int fail(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}

int check_launch(const char* what) {
    g_launches.fetch_add(1, std::memory_order_relaxed);        // every <<< >>> of the library is followed by exactly one check_launch
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) return 0;
    return fail((int)e, "%s: %s", what, cudaGetErrorString(e));
}

}  // namespace fb

extern "C" int fastb200_version(void) { return 100; }
extern "C" int fastb200_sizeof_model(void) { return (int)sizeof(fastb200_model); }
extern "C" const char* fastb200_last_error(void) { return fb::g_err; }
extern "C" int64_t fastb200_launch_count(void) { return fb::g_launches.load(std::memory_order_relaxed); }
