// Batched zero-phase IIR (sosfiltfilt) and polyphase resampling over trials x channels.
//
// north_star (c) of BASELINE.json; the reference repo has no implementation, the arithmetic
// follows scipy.signal.sosfiltfilt / resample_poly (scipy 1.18.1 _signaltools.py:5091-5190, 3884-4100).
//
// B200 design:
//   * one warp owns one row (trial, channel); the row is staged once through shared memory with coalesced loads,
//     the odd extension is built there (in the INPUT dtype, as scipy does) and promoted to fp64;
//   * the cascade of second-order sections is ONE linear recurrence with a 2*n_sections state.  The row is cut into
//     32 equal chunks, one per lane.  Pass A runs every chunk from a zero state to get its end state r_j; a 5-step
//     Hillis-Steele warp scan  E_j = M^(chunk) E_{j-1} + r_j  (the powers M^(chunk*2^k) are precomputed on the host in
//     fp64 and live in the kernel parameter block) gives every chunk its true start state; pass B re-runs the chunk
//     and writes the outputs in place.  Forward and backward directions use the same code with a mirrored index.
//   * fp64 state and data throughout: an fp32 recursion misses the 1e-5 bar.
#include "common.cuh"

namespace fb {

constexpr int FILT_MAX_SEC = 4;
constexpr int FILT_WARPS = 8;

template <int NS>
struct SosParams {
    double sos[NS][6];              // normalised: a0 == 1
    double zi[NS][2];               // sosfilt_zi
    double P[5][2 * NS][2 * NS];    // P[k] = M_chunk^(2^k), M_chunk = zero-input transition over `chunk` samples
};

template <int NS>
__device__ __forceinline__ double cascade_step(const SosParams<NS>& p, double (&z)[2 * NS], double v) {
#pragma unroll
    for (int s = 0; s < NS; ++s) {
        const double out = fma(p.sos[s][0], v, z[2 * s]);
        z[2 * s] = fma(p.sos[s][1], v, fma(-p.sos[s][4], out, z[2 * s + 1]));
        z[2 * s + 1] = fma(p.sos[s][2], v, -p.sos[s][5] * out);
        v = out;
    }
    return v;
}

// one direction of the filtfilt over buf[0..N): REV mirrors the index.  In place.
template <int NS, bool REV>
__device__ __forceinline__ void scan_direction(const SosParams<NS>& p, double* buf, int N, int chunk, int lane) {
    constexpr int SD = 2 * NS;
    const int p0 = lane * chunk;
    auto phys = [&](int q) { return REV ? N - 1 - q : q; };
    double z[SD], init[SD];
    const double first = buf[phys(0)];
#pragma unroll
    for (int s = 0; s < NS; ++s) {
        init[2 * s] = lane == 0 ? p.zi[s][0] * first : 0.0;
        init[2 * s + 1] = lane == 0 ? p.zi[s][1] * first : 0.0;
    }
#pragma unroll
    for (int i = 0; i < SD; ++i) z[i] = init[i];
    // pass A: end state of this chunk from (zero | initial-condition) state
    for (int q = 0; q < chunk; ++q) {
        const int pq = p0 + q;
        const double v = pq < N ? buf[phys(pq)] : 0.0;
        cascade_step<NS>(p, z, v);
    }
    // inclusive scan over chunks: E_j = sum_{i<=j} M^(j-i) r_i
#pragma unroll
    for (int k = 0; k < 5; ++k) {
        const int d = 1 << k;
        double up[SD];
#pragma unroll
        for (int i = 0; i < SD; ++i) up[i] = __shfl_up_sync(0xffffffffu, z[i], d);
        if (lane >= d) {
#pragma unroll
            for (int i = 0; i < SD; ++i) {
                double acc = z[i];
#pragma unroll
                for (int j = 0; j < SD; ++j) acc = fma(p.P[k][i][j], up[j], acc);
                z[i] = acc;
            }
        }
    }
    // start state of chunk j = end state of chunk j-1
#pragma unroll
    for (int i = 0; i < SD; ++i) {
        const double prev = __shfl_up_sync(0xffffffffu, z[i], 1);
        z[i] = lane == 0 ? init[i] : prev;
    }
    __syncwarp();
    // pass B: true outputs
    for (int q = 0; q < chunk; ++q) {
        const int pq = p0 + q;
        if (pq < N) {
            const int a = phys(pq);
            buf[a] = cascade_step<NS>(p, z, buf[a]);
        }
    }
    __syncwarp();
}

template <int NS, typename XT, typename YT>
__global__ void __launch_bounds__(FILT_WARPS * 32) sosfiltfilt_kernel(const __grid_constant__ SosParams<NS> p,
                                                                      const XT* __restrict__ x, YT* __restrict__ y,
                                                                      int64_t rows, int T, int edge, int chunk,
                                                                      int warps_per_cta) {
    extern __shared__ __align__(16) double fsm[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp >= warps_per_cta) return;
    const int N = T + 2 * edge;
    double* buf = fsm + (size_t)warp * (32 * chunk);
    for (int64_t row = (int64_t)blockIdx.x * warps_per_cta + warp; row < rows; row += (int64_t)gridDim.x * warps_per_cta) {
        const XT* xr = x + row * T;
        for (int t = lane; t < T; t += 32) buf[edge + t] = (double)xr[t];
        if (edge > 0) {
            // odd extension, computed in the input dtype (scipy builds it before promoting to float64)
            const XT x0 = xr[0], xl = xr[T - 1];
            for (int i = lane; i < edge; i += 32) {
                buf[i] = (double)(XT)((XT)2 * x0 - xr[edge - i]);
                buf[edge + T + i] = (double)(XT)((XT)2 * xl - xr[T - 2 - i]);
            }
        }
        __syncwarp();
        scan_direction<NS, false>(p, buf, N, chunk, lane);
        scan_direction<NS, true>(p, buf, N, chunk, lane);
        YT* yr = y + row * T;
        for (int t = lane; t < T; t += 32) yr[t] = (YT)buf[edge + t];
        __syncwarp();
    }
}

// y[j] = sum_i x[i] * h[(j + n_pre_remove) * down - i * up]
constexpr int RS_MAXT = 28;        // taps per polyphase branch held in registers (scipy's default FIR: 20 * max(up, down) + 1 taps)
template <typename XT, typename YT>
__global__ void __launch_bounds__(FILT_WARPS * 32) resample_poly_kernel(const XT* __restrict__ x, YT* __restrict__ y,
                                                                        int64_t rows, int T_in, int T_out,
                                                                        const double* __restrict__ h, int len_h, int up,
                                                                        int down, int n_pre_remove, int warps_per_cta) {
    extern __shared__ __align__(16) double fsm[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    double* hs = fsm;
    for (int i = threadIdx.x; i < len_h; i += blockDim.x) hs[i] = h[i];
    __syncthreads();
    if (warp >= warps_per_cta) return;
    double* xs = fsm + ((len_h + 1) & ~1) + (size_t)warp * T_in;
    // a lane's outputs j = lane, lane + 32, ... share the filter phase r = ((j + n_pre_remove) * down) % up when up | 32 * down
    const bool fast = (32 * down) % up == 0 && (len_h + up - 1) / up <= RS_MAXT;
    double cf[RS_MAXT];
    {
        const int r = ((lane + n_pre_remove) * down) % up;
#pragma unroll
        for (int q = 0; q < RS_MAXT; ++q) cf[q] = (fast && r + q * up < len_h) ? hs[r + q * up] : 0.0;
    }
    for (int64_t row = (int64_t)blockIdx.x * warps_per_cta + warp; row < rows; row += (int64_t)gridDim.x * warps_per_cta) {
        const XT* xr = x + row * T_in;
        for (int t = lane; t < T_in; t += 32) xs[t] = (double)xr[t];
        __syncwarp();
        YT* yr = y + row * T_out;
        if (fast) {
            // polyphase fast path: this lane's outputs all use the same sub-filter (phase r), held in registers, so a tap
            // costs one shared-memory read instead of two (the kernel is shared-memory-bandwidth bound on fp64 operands)
            for (int j = lane; j < T_out; j += 32) {
                const int pos = (j + n_pre_remove) * down;
                const int i_hi = pos / up;
                double acc = 0.0;
                if (i_hi <= T_in - 1) {
#pragma unroll
                    for (int q = RS_MAXT - 1; q >= 0; --q) {
                        const int i = i_hi - q;
                        if (i >= 0) acc = fma(xs[i], cf[q], acc);          // cf[q] == 0 past the end of the filter
                    }
                } else {                                                    // right edge: the input ran out before the filter did
                    int i_lo = pos - len_h + 1;
                    i_lo = i_lo <= 0 ? 0 : (i_lo + up - 1) / up;
                    for (int i = i_lo; i <= T_in - 1; ++i) acc = fma(xs[i], hs[pos - i * up], acc);
                }
                yr[j] = (YT)acc;
            }
            __syncwarp();
            continue;
        }
        for (int j = lane; j < T_out; j += 32) {
            const int pos = (j + n_pre_remove) * down;
            int i_hi = pos / up;
            if (i_hi > T_in - 1) i_hi = T_in - 1;
            int i_lo = pos - len_h + 1;
            i_lo = i_lo <= 0 ? 0 : (i_lo + up - 1) / up;
            double acc = 0.0;
            for (int i = i_lo; i <= i_hi; ++i) acc = fma(xs[i], hs[pos - i * up], acc);
            yr[j] = (YT)acc;
        }
        __syncwarp();
    }
}

// ---- host: transition-matrix powers -----------------------------------------------------------------------------
template <int NS>
static void fill_params(SosParams<NS>& p, const double* sos, int chunk) {
    constexpr int SD = 2 * NS;
    for (int s = 0; s < NS; ++s) {
        const double a0 = sos[s * 6 + 3];
        for (int k = 0; k < 6; ++k) p.sos[s][k] = sos[s * 6 + k] / a0;
    }
    // sosfilt_zi: per-section lfilter_zi scaled by the DC gain of the sections before it
    double scale = 1.0;
    for (int s = 0; s < NS; ++s) {
        const double b0 = p.sos[s][0], b1 = p.sos[s][1], b2 = p.sos[s][2], a1 = p.sos[s][4], a2 = p.sos[s][5];
        const double gain = (b0 + b1 + b2) / (1.0 + a1 + a2);
        const double d1 = b1 - gain * a1, d2 = b2 - gain * a2;
        p.zi[s][0] = scale * (d1 + d2);
        p.zi[s][1] = scale * d2;
        scale *= gain;
    }
    // M_chunk: zero-input response over `chunk` samples for each unit state
    double M[SD][SD];
    for (int c = 0; c < SD; ++c) {
        double z[SD];
        for (int i = 0; i < SD; ++i) z[i] = i == c ? 1.0 : 0.0;
        for (int q = 0; q < chunk; ++q) {
            double v = 0.0;
            for (int s = 0; s < NS; ++s) {
                const double out = p.sos[s][0] * v + z[2 * s];
                z[2 * s] = p.sos[s][1] * v - p.sos[s][4] * out + z[2 * s + 1];
                z[2 * s + 1] = p.sos[s][2] * v - p.sos[s][5] * out;
                v = out;
            }
        }
        for (int i = 0; i < SD; ++i) M[i][c] = z[i];
    }
    for (int k = 0; k < 5; ++k) {
        for (int i = 0; i < SD; ++i)
            for (int j = 0; j < SD; ++j) p.P[k][i][j] = M[i][j];
        double Q[SD][SD];
        for (int i = 0; i < SD; ++i)
            for (int j = 0; j < SD; ++j) {
                double s = 0.0;
                for (int l = 0; l < SD; ++l) s += M[i][l] * M[l][j];
                Q[i][j] = s;
            }
        for (int i = 0; i < SD; ++i)
            for (int j = 0; j < SD; ++j) M[i][j] = Q[i][j];
    }
}

template <int NS, typename XT, typename YT>
static int launch_sos(const void* x, void* y, int64_t rows, int T, const double* sos, int edge, cudaStream_t st) {
    const int N = T + 2 * edge;
    const int chunk = ceil_div(N, 32);
    const size_t per_warp = (size_t)32 * chunk * sizeof(double);
    int warps = (int)((220 * 1024) / per_warp);
    if (warps < 1) return fail(-2, "row of %d samples does not fit the shared-memory resident filter", T);
    if (warps > FILT_WARPS) warps = FILT_WARPS;
    // keep >= 2 CTAs per SM resident when rows are short
    if (warps == FILT_WARPS && per_warp * FILT_WARPS > 110 * 1024) warps = FILT_WARPS / 2;
    const size_t smem = per_warp * warps;
    SosParams<NS> p;
    fill_params<NS>(p, sos, chunk);
    auto kern = sosfiltfilt_kernel<NS, XT, YT>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return fail((int)e, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    int64_t grid = (rows + warps - 1) / warps;
    if (grid > NUM_SMS * 8) grid = NUM_SMS * 8;
    kern<<<(unsigned)grid, FILT_WARPS * 32, smem, st>>>(p, (const XT*)x, (YT*)y, rows, T, edge, chunk, warps);
    return check_launch("sosfiltfilt");
}

template <int NS>
static int dispatch_sos(const void* x, int xb, void* y, int yb, int64_t rows, int T, const double* sos, int edge,
                        cudaStream_t st) {
    if (xb == 4 && yb == 4) return launch_sos<NS, float, float>(x, y, rows, T, sos, edge, st);
    if (xb == 4 && yb == 8) return launch_sos<NS, float, double>(x, y, rows, T, sos, edge, st);
    if (xb == 8 && yb == 4) return launch_sos<NS, double, float>(x, y, rows, T, sos, edge, st);
    return launch_sos<NS, double, double>(x, y, rows, T, sos, edge, st);
}

template <typename XT, typename YT>
static int launch_resample(const void* x, void* y, int64_t rows, int T_in, int T_out, const double* h, int len_h,
                           int up, int down, int npr, cudaStream_t st) {
    const size_t hbytes = (size_t)((len_h + 1) & ~1) * sizeof(double);
    const size_t per_warp = (size_t)T_in * sizeof(double);
    int warps = (int)((100 * 1024 - hbytes) / per_warp);
    if (warps < 1) return fail(-2, "row of %d samples does not fit the shared-memory resident resampler", T_in);
    if (warps > FILT_WARPS) warps = FILT_WARPS;
    const size_t smem = hbytes + per_warp * warps;
    auto kern = resample_poly_kernel<XT, YT>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return fail((int)e, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    int64_t grid = (rows + warps - 1) / warps;
    if (grid > NUM_SMS * 8) grid = NUM_SMS * 8;
    kern<<<(unsigned)grid, FILT_WARPS * 32, smem, st>>>((const XT*)x, (YT*)y, rows, T_in, T_out, h, len_h, up, down, npr, warps);
    return check_launch("resample_poly");
}

}  // namespace fb

using namespace fb;

extern "C" int fastb200_sosfiltfilt(const void* x, int x_bytes, void* y, int y_bytes, int64_t rows, int T,
                                    const double* sos_host, int n_sections, int padlen, void* stream) {
    FB_REQUIRE(sos_host != nullptr, "sos is NULL");
    FB_REQUIRE(n_sections >= 1 && n_sections <= FILT_MAX_SEC, "n_sections=%d unsupported (1..%d)", n_sections, FILT_MAX_SEC);
    FB_REQUIRE((x_bytes == 4 || x_bytes == 8) && (y_bytes == 4 || y_bytes == 8), "x / y must be fp32 or fp64");
    FB_REQUIRE(rows >= 0 && T >= 1 && padlen >= 0, "bad rows / T / padlen");
    // scipy raises ValueError for x.shape[axis] <= padlen (_signaltools.py _validate_pad)
    FB_REQUIRE(T > padlen, "The length of the input vector x must be greater than padlen, which is %d.", padlen);
    for (int s = 0; s < n_sections; ++s) FB_REQUIRE(sos_host[s * 6 + 3] != 0.0, "sos[%d, 3] (a0) is zero", s);
    if (rows == 0) return 0;
    cudaStream_t st = as_stream(stream);
    switch (n_sections) {
        case 1: return dispatch_sos<1>(x, x_bytes, y, y_bytes, rows, T, sos_host, padlen, st);
        case 2: return dispatch_sos<2>(x, x_bytes, y, y_bytes, rows, T, sos_host, padlen, st);
        case 3: return dispatch_sos<3>(x, x_bytes, y, y_bytes, rows, T, sos_host, padlen, st);
        default: return dispatch_sos<4>(x, x_bytes, y, y_bytes, rows, T, sos_host, padlen, st);
    }
}

extern "C" int fastb200_resample_poly(const void* x, int x_bytes, void* y, int y_bytes, int64_t rows, int T_in,
                                      int T_out, const double* h_dev, int len_h, int up, int down, int n_pre_remove,
                                      void* stream) {
    FB_REQUIRE((x_bytes == 4 || x_bytes == 8) && (y_bytes == 4 || y_bytes == 8), "x / y must be fp32 or fp64");
    FB_REQUIRE(rows >= 0 && T_in >= 1 && T_out >= 1 && len_h >= 1 && up >= 1 && down >= 1 && n_pre_remove >= 0, "bad dims");
    FB_REQUIRE(len_h <= 4096, "FIR of %d taps is too long", len_h);
    if (rows == 0) return 0;
    cudaStream_t st = as_stream(stream);
    if (x_bytes == 4 && y_bytes == 4) return launch_resample<float, float>(x, y, rows, T_in, T_out, h_dev, len_h, up, down, n_pre_remove, st);
    if (x_bytes == 4 && y_bytes == 8) return launch_resample<float, double>(x, y, rows, T_in, T_out, h_dev, len_h, up, down, n_pre_remove, st);
    if (x_bytes == 8 && y_bytes == 4) return launch_resample<double, float>(x, y, rows, T_in, T_out, h_dev, len_h, up, down, n_pre_remove, st);
    return launch_resample<double, double>(x, y, rows, T_in, T_out, h_dev, len_h, up, down, n_pre_remove, st);
}
