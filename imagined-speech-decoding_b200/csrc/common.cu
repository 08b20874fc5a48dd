// Error channel + version of the fastb200 C ABI.
#include "common.cuh"

namespace fb {

thread_local char g_err[512] = {0};

int fail(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}

int check_launch(const char* what) {
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) return 0;
    return fail((int)e, "%s: %s", what, cudaGetErrorString(e));
}

}  // namespace fb

extern "C" int fastb200_version(void) { return 100; }
extern "C" int fastb200_sizeof_model(void) { return (int)sizeof(fastb200_model); }
extern "C" const char* fastb200_last_error(void) { return fb::g_err; }
