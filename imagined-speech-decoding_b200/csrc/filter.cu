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
// impulse-response table of the cascade state: G[n] = state n steps after a unit sample entered a zero-state cascade.
// The end state of a chunk run from the zero state is then  r = sum_q x[q] * G[chunk-1-q]  -- 2*NS independent fmas per
// sample instead of the 5*NS dependent ones of the recursion (pass A is FP64-pipe bound).  Built once per CTA.
template <int NS>
__device__ __forceinline__ void build_impulse_table(const SosParams<NS>& p, double* G, int chunk) {
    constexpr int SD = 2 * NS;
    if (threadIdx.x == 0) {
        double z[SD];
#pragma unroll
        for (int i = 0; i < SD; ++i) z[i] = 0.0;
        for (int n = 0; n < chunk; ++n) {
            cascade_step<NS>(p, z, n == 0 ? 1.0 : 0.0);
#pragma unroll
            for (int i = 0; i < SD; ++i) G[n * SD + i] = z[i];
        }
    }
    __syncthreads();
}

template <int NS, bool REV>
__device__ __forceinline__ void scan_direction(const SosParams<NS>& p, const double* __restrict__ G, double* buf, int N, int chunk,
                                               int lane) {
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
    // pass A: end state of this chunk from the zero state (impulse-table dot products) ...
#pragma unroll
    for (int i = 0; i < SD; ++i) z[i] = 0.0;
    for (int q = 0; q < chunk; ++q) {
        const int pq = p0 + q;
        const double v = pq < N ? buf[phys(pq)] : 0.0;
        const double* g = G + (chunk - 1 - q) * SD;
#pragma unroll
        for (int i = 0; i < SD; ++i) z[i] = fma(v, g[i], z[i]);
    }
    // ... plus, in lane 0, the zero-input response of the steady-state initial condition: M_chunk * init
    // (the cascade is feed-forward: section s never sees the state of a later section, so M and its powers are block lower
    // triangular and the products skip the zero blocks)
#pragma unroll
    for (int i = 0; i < SD; ++i) {
        double acc = z[i];
#pragma unroll
        for (int j = 0; j < SD; ++j)
            if ((j >> 1) <= (i >> 1)) acc = fma(p.P[0][i][j], init[j], acc);
        z[i] = acc;
    }
    // inclusive scan over chunks: E_j = sum_{i<=j} M^(j-i) r_i
#pragma unroll
    for (int k = 0; k < 5; ++k) {
        const int d = 1 << k;
        double up[SD];
#pragma unroll
        for (int i = 0; i < SD; ++i) up[i] = __shfl_up_sync(0xffffffffu, z[i], d);
        // branch-free: lanes below d receive their own value from the shuffle and simply add nothing
#pragma unroll
        for (int j = 0; j < SD; ++j) up[j] = lane >= d ? up[j] : 0.0;
#pragma unroll
        for (int i = 0; i < SD; ++i) {
            double acc = z[i];
#pragma unroll
            for (int j = 0; j < SD; ++j)
                if ((j >> 1) <= (i >> 1)) acc = fma(p.P[k][i][j], up[j], acc);
            z[i] = acc;
        }
    }
    // start state of chunk j = end state of chunk j-1
#pragma unroll
    for (int i = 0; i < SD; ++i) {
        const double prev = __shfl_up_sync(0xffffffffu, z[i], 1);
        z[i] = lane == 0 ? init[i] : prev;
    }
    __syncwarp();
    // pass B: true outputs.  Branch-free body (only the store is predicated): inside a divergent region the compiler reloads
    // every filter coefficient from the constant bank on every sample (12 LDCU per 24 FP64 instructions, measured with ncu).
    // Lanes past the end of the row (the last lane's tail) run on clamped addresses; their state is never used.
    for (int q = 0; q < chunk; ++q) {
        const int pq = p0 + q;
        const int a = phys(pq < N ? pq : N - 1);
        const double out = cascade_step<NS>(p, z, buf[a]);
        if (pq < N) buf[a] = out;
    }
    __syncwarp();
}

template <int NS, typename XT, typename YT>
__global__ void __launch_bounds__(FILT_WARPS * 32, 4) sosfiltfilt_kernel(const __grid_constant__ SosParams<NS> p,
                                                                      const XT* __restrict__ x, YT* __restrict__ y,
                                                                      int64_t rows, int T, int edge, int chunk,
                                                                      int warps_per_cta) {
    extern __shared__ __align__(16) double fsm[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    double* G = fsm + (size_t)warps_per_cta * (32 * chunk);
    build_impulse_table<NS>(p, G, chunk);
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
        scan_direction<NS, false>(p, G, buf, N, chunk, lane);
        scan_direction<NS, true>(p, G, buf, N, chunk, lane);
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

// ---- phase-major polyphase resampler for a compile-time up/down (the 250 -> 200 Hz default) ------------------------------
// y[j] = sum_i x[i] * h0[j*DOWN + c - i*UP] with h0 = the FIR without its leading zeros, c = UP*cq + CR.  Output j = UP*m + p
// reads x[DOWN*m + cq + (p*DOWN+CR)/UP - k] * H[p][k], H[p][k] = h0[(p*DOWN+CR)%UP + k*UP].  One lane owns one m (UP
// consecutive outputs): the UP phases share one window of x values (NTAP + BMAX shared-memory reads for UP*NTAP fmas instead
// of one read per fma), lanes are DOWN doubles apart (conflict-free for odd DOWN), and every coefficient is a constant-bank
// operand of the DFMA (H lives in the kernel parameter block, all indices compile-time), so the LSU pipe carries x only.
template <int UP, int NTAP>
struct RsFir {
    double H[UP][NTAP];
};

// MB outputs-groups (m, m + 32, ...) per lane: every coefficient (one uniform constant-bank load each) feeds MB fmas
template <int UP, int DOWN, int NTAP, int CR, int MB, typename YT>
__device__ __forceinline__ void resample_group(const RsFir<UP, NTAP>& fir, const double* xs, YT* __restrict__ yr, int T_out, int cq,
                                               int n_m, int m0) {
    constexpr int BMAX = ((UP - 1) * DOWN + CR) / UP;
    const double* xb[MB];
    double acc[MB][UP];
#pragma unroll
    for (int u = 0; u < MB; ++u) {
        const int m = m0 + 32 * u < n_m ? m0 + 32 * u : m0;      // a lane without this m recomputes its first (not stored)
        xb[u] = xs + DOWN * m + cq + BMAX;                       // xs is readable (zeros) outside [0, T_in): pads at the callers
#pragma unroll
        for (int p = 0; p < UP; ++p) acc[u][p] = 0.0;
    }
#pragma unroll
    for (int t = 0; t < NTAP + BMAX; ++t) {
        // compiler fence every 6 taps: without it all NTAP + BMAX shared-memory reads are hoisted to the top (60 live doubles)
        if (t % 6 == 0) asm volatile("" ::: "memory");
        double xv[MB];
#pragma unroll
        for (int u = 0; u < MB; ++u) xv[u] = xb[u][-t];
#pragma unroll
        for (int p = 0; p < UP; ++p) {
            const int k = t - (BMAX - (p * DOWN + CR) / UP);
            if (k >= 0 && k < NTAP) {
#pragma unroll
                for (int u = 0; u < MB; ++u) acc[u][p] = fma(xv[u], fir.H[p][k], acc[u][p]);
            }
        }
    }
#pragma unroll
    for (int u = 0; u < MB; ++u) {
        const int m = m0 + 32 * u;
#pragma unroll
        for (int p = 0; p < UP; ++p)
            if (m < n_m && UP * m + p < T_out) yr[UP * m + p] = (YT)acc[u][p];
    }
}

template <int UP, int DOWN, int NTAP, int CR, typename YT>
__device__ __forceinline__ void resample_row_phases(const RsFir<UP, NTAP>& fir, const double* xs, YT* __restrict__ yr, int T_out,
                                                    int cq, int lane) {
    const int n_m = (T_out + UP - 1) / UP;
    // MB = 1: two groups per lane (MB = 2) halve the uniform coefficient loads but push the fused kernel over its 64-register
    // budget (4 CTAs per SM keep the FP64 pipe fed); measured slower
    for (int m0 = lane; m0 < n_m; m0 += 32) resample_group<UP, DOWN, NTAP, CR, 1, YT>(fir, xs, yr, T_out, cq, n_m, m0);
}

template <int UP, int DOWN, int NTAP, int CR, typename XT, typename YT>
__global__ void __launch_bounds__(FILT_WARPS * 32, 2) resample_phase_kernel(const __grid_constant__ RsFir<UP, NTAP> fir,
                                                                         const XT* __restrict__ x, YT* __restrict__ y,
                                                                         int64_t rows, int T_in, int T_out, int cq, int PL,
                                                                         int PR, int warps_per_cta) {
    extern __shared__ __align__(16) double fsm[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp >= warps_per_cta) return;
    const int span = PL + T_in + PR;
    double* xs = fsm + (size_t)warp * span + PL;
    for (int i = lane; i < PL; i += 32) xs[-1 - i] = 0.0;
    for (int i = lane; i < PR; i += 32) xs[T_in + i] = 0.0;
    for (int64_t row = (int64_t)blockIdx.x * warps_per_cta + warp; row < rows; row += (int64_t)gridDim.x * warps_per_cta) {
        const XT* xr = x + row * T_in;
        for (int t = lane; t < T_in; t += 32) xs[t] = (double)xr[t];
        __syncwarp();
        resample_row_phases<UP, DOWN, NTAP, CR, YT>(fir, xs, y + row * T_out, T_out, cq, lane);
        __syncwarp();
    }
}

// sosfiltfilt and the resampler in one pass over the row: the fp64 filtered row never leaves shared memory (the two-kernel
// chain writes and re-reads it through HBM: 16 of its 23 bytes per sample), and the resampler's shared-memory reads overlap
// the filter's FP64-pipe-bound recursion of the other warps.
template <int NS, int UP, int DOWN, int NTAP, int CR, typename XT, typename YT>
__global__ void __launch_bounds__(FILT_WARPS * 32, 4) sosfiltfilt_resample_kernel(const __grid_constant__ SosParams<NS> p,
                                                                               const __grid_constant__ RsFir<UP, NTAP> fir,
                                                                               const XT* __restrict__ x, YT* __restrict__ y,
                                                                               int64_t rows, int T, int edge, int chunk,
                                                                               int warps_per_cta, int T_out, int cq) {
    extern __shared__ __align__(16) double fsm[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    double* G = fsm + (size_t)warps_per_cta * (32 * chunk);
    build_impulse_table<NS>(p, G, chunk);
    if (warp >= warps_per_cta) return;
    const int N = T + 2 * edge;
    double* buf = fsm + (size_t)warp * (32 * chunk);
    for (int64_t row = (int64_t)blockIdx.x * warps_per_cta + warp; row < rows; row += (int64_t)gridDim.x * warps_per_cta) {
        const XT* xr = x + row * T;
        for (int t = lane; t < T; t += 32) buf[edge + t] = (double)xr[t];
        const XT x0 = xr[0], xl = xr[T - 1];
        for (int i = lane; i < edge; i += 32) {
            buf[i] = (double)(XT)((XT)2 * x0 - xr[edge - i]);
            buf[edge + T + i] = (double)(XT)((XT)2 * xl - xr[T - 2 - i]);
        }
        __syncwarp();
        scan_direction<NS, false>(p, G, buf, N, chunk, lane);
        scan_direction<NS, true>(p, G, buf, N, chunk, lane);
        // the resampler pads with zeros, not with the filtered extension
        for (int i = lane; i < edge; i += 32) {
            buf[i] = 0.0;
            buf[edge + T + i] = 0.0;
        }
        __syncwarp();
        resample_row_phases<UP, DOWN, NTAP, CR, YT>(fir, buf + edge, y + row * T_out, T_out, cq, lane);
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
    int warps = (int)((220 * 1024 - (size_t)chunk * 2 * NS * sizeof(double)) / per_warp);
    if (warps < 1) return fail(-2, "row of %d samples does not fit the shared-memory resident filter", T);
    if (warps > FILT_WARPS) warps = FILT_WARPS;
    // keep >= 2 CTAs per SM resident when rows are short
    if (warps == FILT_WARPS && per_warp * FILT_WARPS > 110 * 1024) warps = FILT_WARPS / 2;
    const size_t smem = per_warp * warps + (size_t)chunk * 2 * NS * sizeof(double);       // + the impulse table
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

// ---- host: the compile-time-ratio resampler (4/5, scipy's default FIR alignment) -------------------------------------------
constexpr int RS_UP = 4, RS_DOWN = 5, RS_NTAP = 26, RS_CR = 2;
struct RsPlan {
    RsFir<RS_UP, RS_NTAP> fir;
    int cq, PL, PR;
};

// true when (h, up, down, n_pre_remove) is the shape resample_phase_kernel is compiled for; fills the coefficient table
static bool plan_phases(const double* h, int len_h, int up, int down, int npr, int T_in, int T_out, RsPlan& pl) {
    if (h == nullptr || up != RS_UP || down != RS_DOWN) return false;
    int lead = 0, L = len_h;
    while (lead < len_h && h[lead] == 0.0) ++lead;
    while (L > lead && h[L - 1] == 0.0) --L;
    const int L0 = L - lead;
    const long long c = (long long)npr * down - lead;
    if (L0 < 1 || L0 > RS_UP * RS_NTAP || c < 0 || c % RS_UP != RS_CR || c / RS_UP > (1 << 20)) return false;
    pl.cq = (int)(c / RS_UP);
    for (int p = 0; p < RS_UP; ++p) {
        const int r = (p * RS_DOWN + RS_CR) % RS_UP;
        for (int k = 0; k < RS_NTAP; ++k) pl.fir.H[p][k] = r + k * RS_UP < L0 ? h[lead + r + k * RS_UP] : 0.0;
    }
    constexpr int BMAX = ((RS_UP - 1) * RS_DOWN + RS_CR) / RS_UP;
    const int n_m = (T_out + RS_UP - 1) / RS_UP;
    const int lo = pl.cq - (RS_NTAP - 1), hi = RS_DOWN * (n_m - 1) + pl.cq + BMAX;
    pl.PL = lo < 0 ? -lo : 0;
    pl.PR = hi > T_in - 1 ? hi - (T_in - 1) : 0;
    return pl.PL <= 4096 && pl.PR <= 4096;
}

template <typename XT, typename YT>
static int launch_resample_phases(const RsPlan& pl, const void* x, void* y, int64_t rows, int T_in, int T_out,
                                  cudaStream_t st) {
    const size_t per_warp = (size_t)(pl.PL + T_in + pl.PR) * sizeof(double);
    int warps = (int)((100 * 1024) / per_warp);
    if (warps < 1) return -2;
    if (warps > FILT_WARPS) warps = FILT_WARPS;
    const size_t smem = per_warp * warps;
    auto kern = resample_phase_kernel<RS_UP, RS_DOWN, RS_NTAP, RS_CR, XT, YT>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return fail((int)e, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    int64_t grid = (rows + warps - 1) / warps;
    if (grid > NUM_SMS * 8) grid = NUM_SMS * 8;
    kern<<<(unsigned)grid, FILT_WARPS * 32, smem, st>>>(pl.fir, (const XT*)x, (YT*)y, rows, T_in, T_out, pl.cq, pl.PL, pl.PR, warps);
    return check_launch("resample_poly (phase-major)");
}

// geometry of the fused kernel: the row buffer of the filter (T + 2*edge, rounded up to 32 chunks) must hold the resampler's pads
static bool fused_geometry(int T, int edge, const RsPlan& pl, int& chunk, int& warps) {
    const int N = T + 2 * edge;
    chunk = ceil_div(N, 32);
    if (pl.PL > edge || pl.PR > edge) return false;
    const size_t per_warp = (size_t)32 * chunk * sizeof(double);
    warps = (int)((220 * 1024 - (size_t)chunk * 2 * FILT_MAX_SEC * sizeof(double)) / per_warp);
    if (warps < 1) return false;
    if (warps > FILT_WARPS) warps = FILT_WARPS;
    if (warps == FILT_WARPS && per_warp * FILT_WARPS > 110 * 1024) warps = FILT_WARPS / 2;
    return true;
}

template <int NS, typename XT, typename YT>
static int launch_fused(const RsPlan& pl, const void* x, void* y, int64_t rows, int T, int T_out, const double* sos, int edge,
                        cudaStream_t st) {
    int chunk, warps;
    if (!fused_geometry(T, edge, pl, chunk, warps)) return fail(-3, "sosfiltfilt_resample: unsupported geometry");
    const size_t smem = (size_t)32 * chunk * sizeof(double) * warps + (size_t)chunk * 2 * NS * sizeof(double);
    SosParams<NS> p;
    fill_params<NS>(p, sos, chunk);
    auto kern = sosfiltfilt_resample_kernel<NS, RS_UP, RS_DOWN, RS_NTAP, RS_CR, XT, YT>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return fail((int)e, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    int64_t grid = (rows + warps - 1) / warps;
    if (grid > NUM_SMS * 8) grid = NUM_SMS * 8;
    kern<<<(unsigned)grid, FILT_WARPS * 32, smem, st>>>(p, pl.fir, (const XT*)x, (YT*)y, rows, T, edge, chunk, warps, T_out, pl.cq);
    return check_launch("sosfiltfilt_resample");
}

template <int NS>
static int dispatch_fused(const RsPlan& pl, const void* x, int xb, void* y, int yb, int64_t rows, int T, int T_out,
                          const double* sos, int edge, cudaStream_t st) {
    if (xb == 4 && yb == 4) return launch_fused<NS, float, float>(pl, x, y, rows, T, T_out, sos, edge, st);
    if (xb == 4 && yb == 8) return launch_fused<NS, float, double>(pl, x, y, rows, T, T_out, sos, edge, st);
    if (xb == 8 && yb == 4) return launch_fused<NS, double, float>(pl, x, y, rows, T, T_out, sos, edge, st);
    return launch_fused<NS, double, double>(pl, x, y, rows, T, T_out, sos, edge, st);
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
                                      const double* h_host, void* stream) {
    FB_REQUIRE((x_bytes == 4 || x_bytes == 8) && (y_bytes == 4 || y_bytes == 8), "x / y must be fp32 or fp64");
    FB_REQUIRE(rows >= 0 && T_in >= 1 && T_out >= 1 && len_h >= 1 && up >= 1 && down >= 1 && n_pre_remove >= 0, "bad dims");
    FB_REQUIRE(len_h <= 4096, "FIR of %d taps is too long", len_h);
    if (rows == 0) return 0;
    cudaStream_t st = as_stream(stream);
    RsPlan pl;
    if (plan_phases(h_host, len_h, up, down, n_pre_remove, T_in, T_out, pl)) {
        int rc;
        if (x_bytes == 4 && y_bytes == 4) rc = launch_resample_phases<float, float>(pl, x, y, rows, T_in, T_out, st);
        else if (x_bytes == 4 && y_bytes == 8) rc = launch_resample_phases<float, double>(pl, x, y, rows, T_in, T_out, st);
        else if (x_bytes == 8 && y_bytes == 4) rc = launch_resample_phases<double, float>(pl, x, y, rows, T_in, T_out, st);
        else rc = launch_resample_phases<double, double>(pl, x, y, rows, T_in, T_out, st);
        if (rc != -2) return rc;                                    // -2: row too long for this kernel's staging, use the generic one
    }
    FB_REQUIRE(h_dev != nullptr, "h_dev is NULL");
    if (x_bytes == 4 && y_bytes == 4) return launch_resample<float, float>(x, y, rows, T_in, T_out, h_dev, len_h, up, down, n_pre_remove, st);
    if (x_bytes == 4 && y_bytes == 8) return launch_resample<float, double>(x, y, rows, T_in, T_out, h_dev, len_h, up, down, n_pre_remove, st);
    if (x_bytes == 8 && y_bytes == 4) return launch_resample<double, float>(x, y, rows, T_in, T_out, h_dev, len_h, up, down, n_pre_remove, st);
    return launch_resample<double, double>(x, y, rows, T_in, T_out, h_dev, len_h, up, down, n_pre_remove, st);
}

extern "C" int fastb200_sosfiltfilt_resample_supported(int T, int n_sections, int padlen, int T_out, const double* h_host,
                                                       int len_h, int up, int down, int n_pre_remove) {
    if (T < 1 || T <= padlen || padlen < 0 || n_sections < 1 || n_sections > FILT_MAX_SEC || T_out < 1 || len_h < 1) return 0;
    RsPlan pl;
    if (!plan_phases(h_host, len_h, up, down, n_pre_remove, T, T_out, pl)) return 0;
    int chunk, warps;
    return fused_geometry(T, padlen, pl, chunk, warps) ? 1 : 0;
}

extern "C" int fastb200_sosfiltfilt_resample(const void* x, int x_bytes, void* y, int y_bytes, int64_t rows, int T,
                                             const double* sos_host, int n_sections, int padlen, int T_out,
                                             const double* h_host, int len_h, int up, int down, int n_pre_remove,
                                             void* stream) {
    FB_REQUIRE(sos_host != nullptr && h_host != nullptr, "sos / h is NULL");
    FB_REQUIRE(n_sections >= 1 && n_sections <= FILT_MAX_SEC, "n_sections=%d unsupported (1..%d)", n_sections, FILT_MAX_SEC);
    FB_REQUIRE((x_bytes == 4 || x_bytes == 8) && (y_bytes == 4 || y_bytes == 8), "x / y must be fp32 or fp64");
    FB_REQUIRE(rows >= 0 && T >= 1 && padlen >= 0 && T_out >= 1 && len_h >= 1, "bad rows / T / padlen / T_out / len_h");
    FB_REQUIRE(T > padlen, "The length of the input vector x must be greater than padlen, which is %d.", padlen);
    for (int s = 0; s < n_sections; ++s) FB_REQUIRE(sos_host[s * 6 + 3] != 0.0, "sos[%d, 3] (a0) is zero", s);
    RsPlan pl;
    if (!plan_phases(h_host, len_h, up, down, n_pre_remove, T, T_out, pl))
        return fail(-3, "sosfiltfilt_resample: up/down/FIR shape not supported by the fused kernel (use the two-call chain)");
    if (rows == 0) return 0;
    cudaStream_t st = as_stream(stream);
    switch (n_sections) {
        case 1: return dispatch_fused<1>(pl, x, x_bytes, y, y_bytes, rows, T, T_out, sos_host, padlen, st);
        case 2: return dispatch_fused<2>(pl, x, x_bytes, y, y_bytes, rows, T, T_out, sos_host, padlen, st);
        case 3: return dispatch_fused<3>(pl, x, x_bytes, y, y_bytes, rows, T, T_out, sos_host, padlen, st);
        default: return dispatch_fused<4>(pl, x, x_bytes, y, y_bytes, rows, T, T_out, sos_host, padlen, st);
    }
}
