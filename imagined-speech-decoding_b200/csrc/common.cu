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

// Strided host <-> device row copy on the caller's stream (one cudaMemcpy2DAsync): `rows` rows of `width_bytes` each, source
// rows `src_pitch` bytes apart, destination rows `dst_pitch` apart.  Lets the Python side ship only the samples of every
// trial row that the windows actually read (750 of 795: reference fast.py:247 never touches samples 750..794).
extern "C" int fastb200_copy_rows_async(void* dst, int64_t dst_pitch, const void* src, int64_t src_pitch, int64_t width_bytes,
                                        int64_t rows, void* stream) {
    FB_REQUIRE(width_bytes >= 0 && rows >= 0 && dst_pitch >= width_bytes && src_pitch >= width_bytes, "copy_rows: bad pitch / width");
    if (rows == 0 || width_bytes == 0) return 0;
    FB_REQUIRE(dst != nullptr && src != nullptr, "copy_rows: NULL buffer");
    cudaError_t e = cudaMemcpy2DAsync(dst, (size_t)dst_pitch, src, (size_t)src_pitch, (size_t)width_bytes, (size_t)rows,
                                      cudaMemcpyDefault, fb::as_stream(stream));
    if (e != cudaSuccess) return fb::fail((int)e, "cudaMemcpy2DAsync: %s", cudaGetErrorString(e));
    return 0;
}
