// Eval-mode forward of the reference's ALTERNATIVE zone heads as one kernel per call.
//
// Reference: src/fast/models/fast.py -- CVBlock :32-100, EEGNet_Encoder :122-167, HeadConv_Paper_Version :170-196, used
// through Head.forward :208-210 inside FAST.forward_head :242-252.  In eval mode (BatchNorm on its running statistics,
// Dropout the identity) each of the three is a short chain of the SAME stage:
//
//     out[o][t'] = pool_{t in window t'} act( scale[o] * sum_{i in in(o)} sum_k w[o][i][k] * in[i][t + k - pad]  +  shift[o] )
//
// with in(o) = every input channel (dense conv; k = 1 is the spatial / point-wise mix) or the single channel
// o / (out_ch / in_ch) (depth-wise conv), zero padding, act in {none, ELU, GELU}, pool in {none, avg, max} (stride = size,
// floor), followed by one of three finals: temporal mean, temporal mean + Linear, flatten + Linear.  The host folds every
// BatchNorm into (scale, shift), moves the EEGNet spatial mix in front of the temporal conv (both linear; BN1 becomes part
// of the affine) and collapses the paper head's cnn1_t . cnn1_s into one conv -- see fast/models/alt_fused.py -- and passes
// the chain as a fastb200_alt_head descriptor.  Persistent CTAs, each serving one zone: the zone's stage weights (output
// channel fastest) and affines stay in shared memory, the window of a (trial, window) pair is gathered from x (unfold + zone
// gather fused, like the Conv4Layers kernels) with the first stage's zero halo, stages ping-pong between two shared-memory
// buffers, and a thread accumulates 8 consecutive conv positions of one channel in a register window.
// These heads cost ~0.3 MFLOP per item, so this is CUDA-core code; training mode (batch statistics) stays on stock ops.
#include "common.cuh"

namespace fb {
namespace alt {

constexpr int THREADS = 256;
constexpr int KC = 8;                                   // taps per register-window chunk

__device__ __forceinline__ float act_f(float v, int act) {
    if (act == FASTB200_ALT_ACT_ELU) return v > 0.f ? v : expm1f(v);
    if (act == FASTB200_ALT_ACT_GELU) return gelu_f(v);
    return v;
}

// Q consecutive conv positions t0 .. t0+Q-1 of output channel o, accumulated together so that an input sample and a
// weight are loaded once per Q multiply-adds (a register window of KC + Q - 1 samples per chunk of KC taps).  `in` rows
// carry the stage's zero padding as a halo (row index 0 = position -pad), so the loop has no bounds checks.
// ws: the stage's weights in shared memory, output channel fastest: dense [i][k][o], depth-wise [k][o].
constexpr int Q = 8;
__device__ __forceinline__ void conv_window(const fastb200_alt_stage& s, const float* ws, const float* in, int in_stride, int in_ch,
                                            int o, int t0, float (&acc)[Q]) {
#pragma unroll
    for (int q = 0; q < Q; ++q) acc[q] = 0.f;
    const int oc = s.out_ch, K = s.k;
    const bool dw = s.kind == FASTB200_ALT_DEPTHWISE;
    const int i0 = dw ? o / (oc / in_ch) : 0, i1 = dw ? i0 + 1 : in_ch;
    const float* xi = in + i0 * in_stride + t0;
    const float* wi = ws + (dw ? 0 : i0 * K * oc) + o;
    for (int i = i0; i < i1; ++i, xi += in_stride, wi += K * oc) {
        int k = 0;
        for (; k + KC <= K; k += KC) {
            float xw[KC + Q - 1];
#pragma unroll
            for (int j = 0; j < KC + Q - 1; ++j) xw[j] = xi[k + j];
#pragma unroll
            for (int kk = 0; kk < KC; ++kk) {
                const float w = wi[(k + kk) * oc];
#pragma unroll
                for (int q = 0; q < Q; ++q) acc[q] = fmaf(w, xw[kk + q], acc[q]);
            }
        }
        if (k < K) {                                     // K % KC taps (the k = 1 and k = 3 stages live here entirely)
            float xw[KC + Q - 2];
#pragma unroll
            for (int j = 0; j < KC + Q - 2; ++j) xw[j] = j < K - k + Q - 1 ? xi[k + j] : 0.f;
#pragma unroll
            for (int kk = 0; kk < KC - 1; ++kk) {
                if (k + kk < K) {
                    const float w = wi[(k + kk) * oc];
#pragma unroll
                    for (int q = 0; q < Q; ++q) acc[q] = fmaf(w, xw[kk + q], acc[q]);
                }
            }
        }
    }
}

__device__ __forceinline__ int stage_w_floats(const fastb200_alt_stage& s, int in_ch) {
    return s.kind == FASTB200_ALT_DEPTHWISE ? s.out_ch * s.k : s.out_ch * in_ch * s.k;
}

// Persistent CTAs: CTA c serves zone c % Z (the zone's stage weights and affines stay in shared memory) and walks the
// (trial, window) pairs c / Z, c / Z + gridDim / Z, ...
__global__ void __launch_bounds__(THREADS) alt_head_fwd_kernel(const __grid_constant__ fastb200_alt_head h, const float* __restrict__ packed,
                                                               const float* __restrict__ x, int BN, int C, int T, int n_win, int step,
                                                               int buf_floats, float* __restrict__ feat) {
    extern __shared__ __align__(16) float smem[];
    float* bufs[2] = {smem, smem + buf_floats};
    float* wsm = smem + 2 * buf_floats;
    const int Z = h.n_zones, z = blockIdx.x % Z;
    const int Cz = h.zone_size[z], W = h.window_len;
    const float* pz = packed + h.zone_base[z];
    // ---- the zone's weights: per stage [weights | scale | shift]
    int wbase[FASTB200_ALT_MAX_STAGES];
    {
        int at = 0, in_ch = Cz;
        for (int si = 0; si < h.n_stages; ++si) {
            const fastb200_alt_stage& s = h.stage[si];
            const int nw = stage_w_floats(s, in_ch);
            wbase[si] = at;
            for (int i = threadIdx.x; i < nw; i += THREADS) wsm[at + i] = __ldg(pz + s.w_off + i);
            for (int i = threadIdx.x; i < s.out_ch; i += THREADS) {
                wsm[at + nw + i] = s.scale_off >= 0 ? __ldg(pz + s.scale_off + i) : 1.f;
                wsm[at + nw + s.out_ch + i] = s.scale_off >= 0 ? __ldg(pz + s.shift_off + i) : 0.f;
            }
            at += nw + 2 * s.out_ch;
            in_ch = s.out_ch;
        }
    }
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int bn = blockIdx.x / Z; bn < BN; bn += gridDim.x / Z) {
        const int n = bn % n_win, b = bn / n_win;
        __syncthreads();                                  // weights staged / previous item's buffers free
        // window gather with the first stage's zero halo: x[b][zone_chan][n * step + t]
        const float* xb = x + (size_t)b * C * T + (size_t)n * step;
        int pad = h.stage[0].pad, stride = W + 2 * pad;
        for (int idx = threadIdx.x; idx < Cz * stride; idx += THREADS) {
            const int c = idx / stride, t = idx - c * stride - pad;
            bufs[0][idx] = (unsigned)t < (unsigned)W ? __ldg(xb + (size_t)h.zone_chan[h.zone_off[z] + c] * T + t) : 0.f;
        }
        __syncthreads();
        int cur = 0, in_ch = Cz, Lin = W;
        for (int si = 0; si < h.n_stages; ++si) {
            const fastb200_alt_stage& s = h.stage[si];
            const int P = s.pool_kind == FASTB200_ALT_POOL_NONE ? 1 : s.pool;
            const int Lout = (Lin + 2 * s.pad - s.k + 1) / P;
            const int pad_n = si + 1 < h.n_stages ? h.stage[si + 1].pad : 0, stride_n = Lout + 2 * pad_n;
            const float* in = bufs[cur];
            float* out = bufs[cur ^ 1];
            const float* ws = wsm + wbase[si];
            const float* aff = ws + stage_w_floats(s, in_ch);
            if (pad_n > 0)
                for (int idx = threadIdx.x; idx < s.out_ch * 2 * pad_n; idx += THREADS) {      // the next stage's zero halo
                    const int o = idx / (2 * pad_n), j = idx - o * 2 * pad_n;
                    out[o * stride_n + (j < pad_n ? j : Lout + j)] = 0.f;
                }
            const int G = Q / P, n_grp = (Lout + G - 1) / G;                                    // a thread: G pooled outputs of one channel
            for (int idx = threadIdx.x; idx < s.out_ch * n_grp; idx += THREADS) {               // output channel fastest
                const int grp = idx / s.out_ch, o = idx - grp * s.out_ch;
                const float scale = aff[o], shift = aff[s.out_ch + o];
                float acc[Q];
                conv_window(s, ws, in, stride, in_ch, o, grp * Q, acc);
#pragma unroll
                for (int q = 0; q < Q; ++q) acc[q] = act_f(fmaf(acc[q], scale, shift), s.act);
                float* orow = out + o * stride_n + pad_n + grp * G;
                const bool mx = s.pool_kind == FASTB200_ALT_POOL_MAX;
                if (P == 1) {
#pragma unroll
                    for (int q = 0; q < Q; ++q) if (grp * G + q < Lout) orow[q] = acc[q];
                } else if (P == 2) {
#pragma unroll
                    for (int q = 0; q < Q / 2; ++q)
                        if (grp * G + q < Lout) orow[q] = mx ? fmaxf(acc[2 * q], acc[2 * q + 1]) : 0.5f * (acc[2 * q] + acc[2 * q + 1]);
                } else if (P == 4) {
#pragma unroll
                    for (int q = 0; q < Q / 4; ++q)
                        if (grp * G + q < Lout)
                            orow[q] = mx ? fmaxf(fmaxf(acc[4 * q], acc[4 * q + 1]), fmaxf(acc[4 * q + 2], acc[4 * q + 3]))
                                         : 0.25f * ((acc[4 * q] + acc[4 * q + 1]) + (acc[4 * q + 2] + acc[4 * q + 3]));
                } else {                                                                        // host: P in {1, 2, 4, 8}
                    float r = acc[0];
#pragma unroll
                    for (int q = 1; q < Q; ++q) r = mx ? fmaxf(r, acc[q]) : r + acc[q];
                    orow[0] = mx ? r : r * 0.125f;
                }
            }
            __syncthreads();
            cur ^= 1; in_ch = s.out_ch; Lin = Lout; pad = pad_n; stride = stride_n;
        }
        const float* y = bufs[cur];                                          // [in_ch][Lin], no halo
        float* vec = bufs[cur ^ 1];
        float* dst = feat + ((size_t)bn * Z + z) * h.lin_out;
        if (h.final_kind != FASTB200_ALT_FINAL_FLATTEN_LINEAR) {
            for (int o = warp; o < in_ch; o += THREADS / 32) {               // temporal mean per channel
                float sum = 0.f;
                for (int t = lane; t < Lin; t += 32) sum += y[o * Lin + t];
                sum = warp_sum(sum);
                if (lane == 0) {
                    sum /= (float)Lin;
                    if (h.final_kind == FASTB200_ALT_FINAL_MEAN) dst[o] = sum; else vec[o] = sum;
                }
            }
            if (h.final_kind == FASTB200_ALT_FINAL_MEAN) continue;
            __syncthreads();
            y = vec;
        }
        const float* lw = pz + h.lin_w_off;
        for (int f = warp; f < h.lin_out; f += THREADS / 32) {               // Linear: one warp per output, weights from L1 / L2
            float sum = 0.f;
            for (int i = lane; i < h.lin_in; i += 32) sum = fmaf(__ldg(lw + (size_t)f * h.lin_in + i), y[i], sum);
            sum = warp_sum(sum);
            if (lane == 0) dst[f] = sum + __ldg(pz + h.lin_b_off + f);
        }
    }
}

}  // namespace alt
}  // namespace fb

using namespace fb;

// shape walk shared by the validation and the launch: activation buffer size and weight area (floats)
static int alt_walk(const fastb200_alt_head* h, int* buf_floats, int* w_floats) {
    FB_REQUIRE(h != nullptr, "alt-head descriptor is NULL");
    FB_REQUIRE(h->n_zones >= 1 && h->n_zones <= FASTB200_MAX_ZONES, "n_zones=%d out of range", h->n_zones);
    FB_REQUIRE(h->n_stages >= 1 && h->n_stages <= FASTB200_ALT_MAX_STAGES, "n_stages=%d out of range", h->n_stages);
    FB_REQUIRE(h->window_len >= 1, "window_len=%d", h->window_len);
    int cz_max = 0;
    for (int z = 0; z < h->n_zones; ++z) {
        FB_REQUIRE(h->zone_size[z] >= 1 && h->zone_off[z] >= 0 && h->zone_off[z] + h->zone_size[z] <= FASTB200_MAX_CHANS, "zone %d table out of range", z);
        cz_max = cz_max > h->zone_size[z] ? cz_max : h->zone_size[z];
    }
    int ch = cz_max, len = h->window_len, wf = 0;
    int64_t big = 0;
    for (int i = 0; i < h->n_stages; ++i) {
        const fastb200_alt_stage& s = h->stage[i];
        FB_REQUIRE(s.kind == FASTB200_ALT_DENSE || s.kind == FASTB200_ALT_DEPTHWISE, "stage %d: unknown kind %d", i, s.kind);
        FB_REQUIRE(s.out_ch >= 1 && s.k >= 1 && s.pad >= 0, "stage %d: bad out_ch/k/pad", i);
        FB_REQUIRE(i > 0 || s.kind == FASTB200_ALT_DENSE, "stage 0 reads the zone's electrodes and must be dense");
        FB_REQUIRE(s.kind == FASTB200_ALT_DENSE || s.out_ch % ch == 0, "stage %d: depth-wise out_ch %d not a multiple of in_ch %d", i, s.out_ch, ch);
        const int P = s.pool_kind == FASTB200_ALT_POOL_NONE ? 1 : s.pool;
        FB_REQUIRE(P == 1 || P == 2 || P == 4 || P == 8, "stage %d: pool size %d (instantiated: 1, 2, 4, 8)", i, P);
        big = big > (int64_t)ch * (len + 2 * s.pad) ? big : (int64_t)ch * (len + 2 * s.pad);          // this stage's input incl. halo
        wf += (s.kind == FASTB200_ALT_DEPTHWISE ? s.out_ch * s.k : s.out_ch * ch * s.k) + 2 * s.out_ch;
        const int lc = len + 2 * s.pad - s.k + 1;
        FB_REQUIRE(lc >= 1, "stage %d: kernel %d longer than its padded input %d", i, s.k, len + 2 * s.pad);
        len = lc / P;
        FB_REQUIRE(len >= 1, "stage %d: pooled length is 0", i);
        ch = s.out_ch;
    }
    big = big > (int64_t)ch * len ? big : (int64_t)ch * len;
    FB_REQUIRE(h->final_kind >= FASTB200_ALT_FINAL_MEAN && h->final_kind <= FASTB200_ALT_FINAL_FLATTEN_LINEAR, "unknown final %d", h->final_kind);
    if (h->final_kind == FASTB200_ALT_FINAL_MEAN) FB_REQUIRE(h->lin_out == ch, "final mean: lin_out %d != channels %d", h->lin_out, ch);
    if (h->final_kind == FASTB200_ALT_FINAL_MEAN_LINEAR) FB_REQUIRE(h->lin_in == ch && h->lin_out >= 1, "final mean+linear: lin_in %d != channels %d", h->lin_in, ch);
    if (h->final_kind == FASTB200_ALT_FINAL_FLATTEN_LINEAR)
        FB_REQUIRE(h->lin_in == ch * len && h->lin_out >= 1, "final flatten+linear: lin_in %d != %d x %d (the reference's Linear would reject this window length too)", h->lin_in, ch, len);
    FB_REQUIRE(big < (1 << 20), "alt head: activation buffer too large");
    *buf_floats = round_up((int)big, 4) + 4 * alt::Q;                  // + slack: the last group's window may run past the last row
    *w_floats = wf;
    return 0;
}

extern "C" int fastb200_sizeof_alt_head(void) { return (int)sizeof(fastb200_alt_head); }

extern "C" int fastb200_alt_head_fwd(const fastb200_alt_head* h, const float* packed, const float* x, int B, int C, int T, int n_win,
                                     int step, float* feat, void* stream) {
    int buf = 0, wf = 0;
    if (int r = alt_walk(h, &buf, &wf)) return r;
    FB_REQUIRE(B >= 0 && C >= 1 && n_win >= 1 && step >= 1, "bad B/C/n_win/step");
    FB_REQUIRE((n_win - 1) * (int64_t)step + h->window_len <= T, "windows run past the trial: n_win=%d step=%d W=%d T=%d", n_win, step, h->window_len, T);
    for (int z = 0; z < h->n_zones; ++z)
        for (int c = 0; c < h->zone_size[z]; ++c)
            FB_REQUIRE(h->zone_chan[h->zone_off[z] + c] >= 0 && h->zone_chan[h->zone_off[z] + c] < C, "zone %d electrode index out of range", z);
    FB_REQUIRE((int64_t)B * n_win * h->n_zones < (1ll << 31), "too many items for one launch");
    if (B == 0) return 0;
    FB_REQUIRE(packed && x && feat, "alt_head_fwd: NULL buffer");
    const size_t smem = (2 * (size_t)buf + wf) * sizeof(float);
    FB_REQUIRE(smem <= 200 * 1024, "alt head: %zu bytes of shared memory per CTA (window too long / too many channels)", smem);
    cudaError_t e = cudaFuncSetAttribute(alt::alt_head_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return fail((int)e, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    const int Z = h->n_zones, BN = B * n_win;
    int per_sm = (int)((220 * 1024) / (smem + 1024));
    per_sm = per_sm < 1 ? 1 : per_sm > 8 ? 8 : per_sm;
    int per_zone = NUM_SMS * per_sm / Z;                               // CTAs per zone: fill the GPU, never more than the items
    per_zone = per_zone < 1 ? 1 : per_zone > BN ? BN : per_zone;
    alt::alt_head_fwd_kernel<<<Z * per_zone, alt::THREADS, smem, as_stream(stream)>>>(*h, packed, x, BN, C, T, n_win, step, buf, feat);
    return check_launch("alt_head_fwd");
}
