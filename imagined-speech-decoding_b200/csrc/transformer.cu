// Token path of FAST: input_layer (+GELU), CLS token, positional embedding, L pre-LN AttentionBlocks and the
// classifier -- forward and hand-written backward.
//
// Reference: FAST.forward_transformer src/fast/models/fast.py:260-268, AttentionBlock :10-29 (nn.MultiheadAttention,
// LayerNorm eps 1e-5, exact GELU), 'train_head' branch :273-278.  This is 0.2 % of the model's arithmetic
// (SURVEY.md section 0.2), so the kernel is built for residency and launch count, not for the tensor pipe:
// one CTA owns a group of trials, keeps every activation of those trials in shared memory for the whole depth of
// the network, stages one block's weights at a time (33.7 KB at d_model 32) and never touches HBM in between.
// LayerNorm / softmax reductions are warp shuffles; dropout masks come from a counter-based generator so the
// backward kernel regenerates them instead of storing them.
#include "common.cuh"
#include "transformer.cuh"
#include <cstdlib>

namespace fb {

constexpr int TR_THREADS = 512;
constexpr int TR_WARPS = TR_THREADS / 32;

// ---- shared-memory resident building blocks (all threads of the CTA call these; R = live rows) --------------
// Asynchronous (cp.async, 4 bytes: the padded pitch breaks 16-byte alignment): a load -> store loop through registers
// serialises on the L2 latency of every element (the compiler cannot hoist the loads over the shared-memory stores), which
// cost ~12 us per block at one trial per CTA.  Callers run stage_wait() + __syncthreads() before the first use.
__device__ __noinline__ void stage_w(float* dst, const float* __restrict__ src, int rows, int cols) {
    const int ld = cols + 1;                                   // odd-ish pitch: conflict free both ways
    int r = threadIdx.x / cols, c = threadIdx.x - r * cols;
    const int dr = TR_THREADS / cols, dc = TR_THREADS - dr * cols;
    for (int i = threadIdx.x; i < rows * cols; i += TR_THREADS) {
        const unsigned d = (unsigned)__cvta_generic_to_shared(dst + r * ld + c);
        asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d), "l"(src + i) : "memory");
        r += dr; c += dc;
        if (c >= cols) { c -= cols; ++r; }
    }
}
__device__ __forceinline__ void stage_wait() { asm volatile("cp.async.wait_all;" ::: "memory"); }

// Y[r][j] = (RES ? Y[r][j] : 0) + post( act(b[j] + sum_i X[r][i] W[j][i]) )   -- RPT rows per thread.
// bias may live in shared memory (the per-block vectors staged by stage_block) or in global memory: plain loads.
template <int ACT, bool RES, int RPT, typename Post>
__device__ __forceinline__ void linear_fwd_rpt(const float* Xs, int ldx, const float* Ws, int Din, int Dout, const float* bias,
                                               float* Ys, int ldy, int R, float* pre_s, Post post) {
    const int ldw = Din + 1;
    const int RG = (R + RPT - 1) / RPT;
    for (int idx = threadIdx.x; idx < RG * Dout; idx += TR_THREADS) {
        const int j = idx % Dout, r0 = (idx / Dout) * RPT;
        const float b = bias ? bias[j] : 0.f;
        float acc[RPT];
        const float* xr[RPT];
#pragma unroll
        for (int q = 0; q < RPT; ++q) { acc[q] = b; xr[q] = Xs + min(r0 + q, R - 1) * ldx; }
        const float* wr = Ws + j * ldw;
#pragma unroll 8
        for (int i = 0; i < Din; ++i) {
            const float w = wr[i];
#pragma unroll
            for (int q = 0; q < RPT; ++q) acc[q] = fmaf(xr[q][i], w, acc[q]);
        }
#pragma unroll
        for (int q = 0; q < RPT; ++q) {
            const int r = r0 + q;
            if (r < R) {
                if (pre_s) pre_s[r * ldy + j] = acc[q];
                float v = ACT == ACT_GELU ? gelu_f(acc[q]) : acc[q];
                v = post(r, j, v);
                if (RES) Ys[r * ldy + j] += v; else Ys[r * ldy + j] = v;
            }
        }
    }
}
// Runtime epilogue (activation, residual, dropout site) and ONE out-of-line copy: the kernels call this from a dozen
// places; inlined and specialised per call site the backward kernel grew to 15 k instructions (~240 KB) and a third of
// its stall samples were instruction-cache misses (ncu: stall_no_inst).
// Few rows (one trial per CTA: the cross-validation batch sizes): one row per thread keeps all warps busy instead of 2 of 8.
struct Epi {
    int act, res, site, layer, trial0, tok;      // site < 0: no dropout
};
__device__ __noinline__ void linear_fwd(const float* Xs, int ldx, const float* Ws, int Din, int Dout, const float* bias,
                                        float* Ys, int ldy, int R, float* pre_s, const Epi e, const DropCtx& dc) {
    auto post = [&](int r, int j, float v) {
        if (e.act == ACT_GELU) v = gelu_f(v);
        if (e.site >= 0) v *= dc.mask(e.layer, e.site, e.trial0 + r / e.tok, (r % e.tok) * Dout + j);
        return e.res ? Ys[r * ldy + j] + v : v;
    };
    if (((R + 3) >> 2) * Dout >= TR_THREADS) linear_fwd_rpt<ACT_NONE, false, 4>(Xs, ldx, Ws, Din, Dout, bias, Ys, ldy, R, pre_s, post);
    else if (((R + 1) >> 1) * Dout >= TR_THREADS) linear_fwd_rpt<ACT_NONE, false, 2>(Xs, ldx, Ws, Din, Dout, bias, Ys, ldy, R, pre_s, post);
    else linear_fwd_rpt<ACT_NONE, false, 1>(Xs, ldx, Ws, Din, Dout, bias, Ys, ldy, R, pre_s, post);
}

// dX[r][i] (ACC ? += : =) sum_j dY[r][j] W[j][i]
template <bool ACC, int RPT>
__device__ __forceinline__ void linear_dgrad_rpt(const float* dYs, int ldy, const float* Ws, int Din, int Dout, float* dXs, int ldx,
                                                 int R) {
    const int ldw = Din + 1;
    const int RG = (R + RPT - 1) / RPT;
    for (int idx = threadIdx.x; idx < RG * Din; idx += TR_THREADS) {
        const int i = idx % Din, r0 = (idx / Din) * RPT;
        float acc[RPT];
        const float* yr[RPT];
#pragma unroll
        for (int q = 0; q < RPT; ++q) { acc[q] = 0.f; yr[q] = dYs + min(r0 + q, R - 1) * ldy; }
#pragma unroll 8
        for (int j = 0; j < Dout; ++j) {
            const float w = Ws[j * ldw + i];
#pragma unroll
            for (int q = 0; q < RPT; ++q) acc[q] = fmaf(yr[q][j], w, acc[q]);
        }
#pragma unroll
        for (int q = 0; q < RPT; ++q) {
            const int r = r0 + q;
            if (r < R) { if (ACC) dXs[r * ldx + i] += acc[q]; else dXs[r * ldx + i] = acc[q]; }
        }
    }
}
__device__ __noinline__ void linear_dgrad(const float* dYs, int ldy, const float* Ws, int Din, int Dout, float* dXs, int ldx,
                                          int R) {
    if (((R + 3) >> 2) * Din >= TR_THREADS) linear_dgrad_rpt<false, 4>(dYs, ldy, Ws, Din, Dout, dXs, ldx, R);
    else if (((R + 1) >> 1) * Din >= TR_THREADS) linear_dgrad_rpt<false, 2>(dYs, ldy, Ws, Din, Dout, dXs, ldx, R);
    else linear_dgrad_rpt<false, 1>(dYs, ldy, Ws, Din, Dout, dXs, ldx, R);
}

// gW[j][i] += sum_r dY[r][j] X[r][i] ;  gb[j] += sum_r dY[r][j]      (one atomic per element per CTA)
__device__ __noinline__ void linear_wgrad(const float* dYs, int ldy, const float* Xs, int ldx, int Din, int Dout,
                                             int R, float* gW, float* gb) {
    if (R <= 8 && TR_THREADS % Din == 0) {
        // few rows (one trial per CTA): a thread keeps its input column X[:, i] in registers and walks the output rows j --
        // no index division, R broadcast reads + R fmas + one reduction per element (the generic loop below spent 4x that
        // on index arithmetic: 25 % of the backward kernel at the cross-validation batch size)
        const int i = threadIdx.x % Din, j0 = threadIdx.x / Din, dj = TR_THREADS / Din;
        float x[8];
#pragma unroll
        for (int r = 0; r < 8; ++r) x[r] = r < R ? Xs[r * ldx + i] : 0.f;
        for (int j = j0; j < Dout; j += dj) {
            float s = 0.f;
#pragma unroll
            for (int r = 0; r < 8; ++r)
                if (r < R) s = fmaf(dYs[r * ldy + j], x[r], s);
            atomicAdd(gW + j * Din + i, s);
        }
    } else
    for (int idx = threadIdx.x; idx < Dout * Din; idx += TR_THREADS) {
        const int i = idx % Din, j = idx / Din;
        float s = 0.f;
        for (int r = 0; r < R; ++r) s = fmaf(dYs[r * ldy + j], Xs[r * ldx + i], s);
        atomicAdd(gW + idx, s);
    }
    if (gb)
        for (int j = threadIdx.x; j < Dout; j += TR_THREADS) {
            float s = 0.f;
            for (int r = 0; r < R; ++r) s += dYs[r * ldy + j];
            atomicAdd(gb + j, s);
        }
}

__device__ __noinline__ void ln_fwd(const float* Xs, float* Ys, const float* w, const float* b, int R, int D, float* stat) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int r = warp; r < R; r += TR_WARPS) {
        float s = 0.f;
        for (int i = lane; i < D; i += 32) s += Xs[r * D + i];
        const float mean = warp_sum(s) / (float)D;
        float v = 0.f;
        for (int i = lane; i < D; i += 32) { const float d = Xs[r * D + i] - mean; v = fmaf(d, d, v); }
        const float rstd = 1.0f / sqrtf(warp_sum(v) / (float)D + LN_EPS);
        for (int i = lane; i < D; i += 32) Ys[r * D + i] = (Xs[r * D + i] - mean) * rstd * w[i] + b[i];
        if (stat && lane == 0) { stat[2 * r] = mean; stat[2 * r + 1] = rstd; }
    }
}

// dXout[r][i] += rstd (g - mean(g) - xhat mean(g xhat)), g = dY w ;  gw += sum_r dY xhat ; gb += sum_r dY
__device__ __noinline__ void ln_bwd(const float* dYs, const float* Xs, const float* stat, const float* w, float* dXout, int R,
                                       int D, float* gw, float* gb) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int r = warp; r < R; r += TR_WARPS) {
        const float mean = stat[2 * r], rstd = stat[2 * r + 1];
        float a = 0.f, c = 0.f;
        for (int i = lane; i < D; i += 32) {
            const float g = dYs[r * D + i] * w[i], xh = (Xs[r * D + i] - mean) * rstd;
            a += g; c = fmaf(g, xh, c);
        }
        a = warp_sum(a) / (float)D; c = warp_sum(c) / (float)D;
        for (int i = lane; i < D; i += 32) {
            const float g = dYs[r * D + i] * w[i], xh = (Xs[r * D + i] - mean) * rstd;
            dXout[r * D + i] += rstd * (g - a - xh * c);
        }
    }
    for (int i = threadIdx.x; i < D; i += TR_THREADS) {
        float sw = 0.f, sb = 0.f;
        for (int r = 0; r < R; ++r) {
            const float dy = dYs[r * D + i];
            sw = fmaf(dy, (Xs[r * D + i] - stat[2 * r]) * stat[2 * r + 1], sw);
            sb += dy;
        }
        atomicAdd(gw + i, sw);
        atomicAdd(gb + i, sb);
    }
}

// softmax(q k^T / sqrt(hd)) per (trial, head, query) thread; P kept for the backward; O = dropout(P) v
__device__ __noinline__ void attention_fwd(const float* QKV, float* P, float* O, int G, int tok, int D, int H,
                                              const DropCtx& dc, int layer, int trial0) {
    const int hd = D / H;
    const float scale = 1.0f / sqrtf((float)hd);
    for (int idx = threadIdx.x; idx < G * H * tok; idx += TR_THREADS) {
        const int i = idx % tok, h = (idx / tok) % H, g = idx / (tok * H);
        const float* q = QKV + (g * tok + i) * 3 * D + h * hd;
        const float* kbase = QKV + (g * tok) * 3 * D + D + h * hd;
        float* prow = P + ((g * H + h) * tok + i) * tok;
        float mx = -INFINITY;
        for (int j = 0; j < tok; ++j) {
            float s = 0.f;
            for (int d = 0; d < hd; ++d) s = fmaf(q[d] * scale, kbase[j * 3 * D + d], s);
            prow[j] = s;
            mx = fmaxf(mx, s);
        }
        float sum = 0.f;
        for (int j = 0; j < tok; ++j) { const float e = expf(prow[j] - mx); prow[j] = e; sum += e; }
        const float inv = 1.0f / sum;
        for (int j = 0; j < tok; ++j) prow[j] *= inv;
        const float* vbase = kbase + D;
        if (hd <= 8) {                                  // one mask draw per key instead of one per (key, channel)
            float acc[8];
#pragma unroll
            for (int d = 0; d < 8; ++d) acc[d] = 0.f;
            for (int j = 0; j < tok; ++j) {
                const float pm = prow[j] * dc.mask(layer, SITE_ATTN, trial0 + g, (h * tok + i) * tok + j);
#pragma unroll
                for (int d = 0; d < 8; ++d)
                    if (d < hd) acc[d] = fmaf(pm, vbase[j * 3 * D + d], acc[d]);
            }
#pragma unroll
            for (int d = 0; d < 8; ++d)
                if (d < hd) O[(g * tok + i) * D + h * hd + d] = acc[d];
        } else {
            for (int d = 0; d < hd; ++d) {
                float acc = 0.f;
                for (int j = 0; j < tok; ++j)
                    acc = fmaf(prow[j] * dc.mask(layer, SITE_ATTN, trial0 + g, (h * tok + i) * tok + j), vbase[j * 3 * D + d], acc);
                O[(g * tok + i) * D + h * hd + d] = acc;
            }
        }
    }
}

// one AttentionBlock forward over the CTA's rows.  Buffers: X (in/out), Hn, QKV, P, O, U (hidden), Upre (or null).
struct BlockBufs { float *X, *Hn, *QKV, *P, *O, *U, *Upre, *stat1, *stat2, *Xmid, *Hn2; };

// the block's vectors (LayerNorm affine, biases) ride along in shared memory: every phase used to open with a dependent global load
struct BlockVec { const float *ln1_w, *ln1_b, *qkv_b, *out_b, *ln2_w, *ln2_b, *ff1_b, *ff2_b; };

__device__ __forceinline__ const float* stage_v(float*& dst, const float* __restrict__ src, int n) {
    float* d0 = dst;
    for (int i = threadIdx.x; i < n; i += TR_THREADS) {
        const unsigned d = (unsigned)__cvta_generic_to_shared(d0 + i);
        asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d), "l"(src + i) : "memory");
    }
    dst += n;
    return d0;
}

__device__ __forceinline__ void stage_block(const fastb200_model& m, const float* params, int l, float* Ws,
                                            float*& Wqkv, float*& Wo, float*& W1, float*& W2, BlockVec& bv) {
    const int D = m.dim_token, Hd = m.dim_hidden;
    Wqkv = Ws; Wo = Wqkv + 3 * D * (D + 1); W1 = Wo + D * (D + 1); W2 = W1 + Hd * (D + 1);
    stage_w(Wqkv, params + m.off_qkv_w[l], 3 * D, D);
    stage_w(Wo, params + m.off_out_w[l], D, D);
    stage_w(W1, params + m.off_ff1_w[l], Hd, D);
    stage_w(W2, params + m.off_ff2_w[l], D, Hd);
    float* v = W2 + D * (Hd + 1);
    bv.ln1_w = stage_v(v, params + m.off_ln1_w[l], D);
    bv.ln1_b = stage_v(v, params + m.off_ln1_b[l], D);
    bv.qkv_b = stage_v(v, params + m.off_qkv_b[l], 3 * D);
    bv.out_b = stage_v(v, params + m.off_out_b[l], D);
    bv.ln2_w = stage_v(v, params + m.off_ln2_w[l], D);
    bv.ln2_b = stage_v(v, params + m.off_ln2_b[l], D);
    bv.ff1_b = stage_v(v, params + m.off_ff1_b[l], Hd);
    bv.ff2_b = stage_v(v, params + m.off_ff2_b[l], D);
}

// KEEP = true (backward recompute): keeps Xmid, Hn2, Upre, LN statistics, and leaves X untouched.
template <bool KEEP>
__device__ __forceinline__ void block_fwd(const fastb200_model& m, const float* params, int l, const BlockBufs& bb,
                                          const float* Wqkv, const float* Wo, const float* W1, const float* W2,
                                          const BlockVec& bv, int G, int tok, const DropCtx& dc, int trial0) {
    const int D = m.dim_token, Hd = m.dim_hidden, R = G * tok;
    ln_fwd(bb.X, bb.Hn, bv.ln1_w, bv.ln1_b, R, D, bb.stat1);
    __syncthreads();
    linear_fwd(bb.Hn, D, Wqkv, D, 3 * D, bv.qkv_b, bb.QKV, 3 * D, R, nullptr, Epi{ACT_NONE, 0, -1, l, trial0, tok}, dc);
    __syncthreads();
    attention_fwd(bb.QKV, bb.P, bb.O, G, tok, D, m.n_heads, dc, l, trial0);
    __syncthreads();
    float* xmid = KEEP ? bb.Xmid : bb.X;
    if (KEEP) {
        for (int i = threadIdx.x; i < R * D; i += TR_THREADS) bb.Xmid[i] = bb.X[i];
        __syncthreads();
    }
    linear_fwd(bb.O, D, Wo, D, D, bv.out_b, xmid, D, R, nullptr, Epi{ACT_NONE, 1, -1, l, trial0, tok}, dc);
    __syncthreads();
    float* hn2 = KEEP ? bb.Hn2 : bb.Hn;
    ln_fwd(xmid, hn2, bv.ln2_w, bv.ln2_b, R, D, bb.stat2);
    __syncthreads();
    linear_fwd(hn2, D, W1, D, Hd, bv.ff1_b, bb.U, Hd, R, KEEP ? bb.Upre : nullptr, Epi{ACT_GELU, 0, SITE_FFN_HID, l, trial0, tok}, dc);
    __syncthreads();
    if (!KEEP) {
        linear_fwd(bb.U, Hd, W2, Hd, D, bv.ff2_b, bb.X, D, R, nullptr, Epi{ACT_NONE, 1, SITE_FFN_OUT, l, trial0, tok}, dc);
        __syncthreads();
    }
}

// ---------------------------------------------------------------------------------------------------------
// forward kernel
// ---------------------------------------------------------------------------------------------------------

__device__ __forceinline__ int tr_ws_floats(const fastb200_model& m) {
    const int D = m.dim_token, Hd = m.dim_hidden, ZF = m.n_zones * m.dim_cnn;
    const int blk = 3 * D * (D + 1) + D * (D + 1) + Hd * (D + 1) + D * (Hd + 1) + 9 * D + Hd;     // weights + vectors
    const int inp = D * (ZF + 1);
    return round_up(blk > inp ? blk : inp, 4);
}

// tokens of the input layer: T[g*n_win + n][D] = gelu(W_in feat + b)  (pre-activation optionally kept)
__device__ __forceinline__ void input_layer_fwd(const fastb200_model& m, const TrArgs& a, float* Ws, float* Fs,
                                                float* Ts, float* pre, int trial0, int G) {
    const int D = m.dim_token, ZF = m.n_zones * m.dim_cnn, Rin = G * a.n_win;
    stage_w(Ws, a.params + m.off_in_w, D, ZF);
    for (int i = threadIdx.x; i < Rin * ZF; i += TR_THREADS) Fs[i] = __ldg(a.feat + (size_t)trial0 * a.n_win * ZF + i);
    stage_wait();
    __syncthreads();
    DropCtx none{};
    linear_fwd(Fs, ZF, Ws, ZF, D, a.params + m.off_in_b, Ts, D, Rin, pre, Epi{ACT_GELU, 0, -1, 0, trial0, 1}, none);
    __syncthreads();
}

__global__ void __launch_bounds__(TR_THREADS) transformer_fwd_kernel(const __grid_constant__ fastb200_model m,
                                                                     const __grid_constant__ TrArgs a) {
    extern __shared__ __align__(16) float smem[];
    const int D = m.dim_token, Hd = m.dim_hidden, H = m.n_heads, ZF = m.n_zones * m.dim_cnn, NC = m.n_classes;
    const int tok = a.tok, n_win = a.n_win;
    const int trial0 = blockIdx.x * a.G;
    const int G = min(a.G, a.B - trial0);
    const int R = G * tok, Rmax = a.G * tok;
    DropCtx dc = a.dc;
    dc.resolve();
    float* Ws = smem;
    float* X = Ws + a.ws_floats;
    float* Hn = X + Rmax * D;
    float* QKV = Hn + Rmax * D;
    float* O = QKV + Rmax * 3 * D;
    float* U = O + Rmax * D;
    float* P = U + Rmax * Hd;
    float* Fs = P + a.G * H * tok * tok;         // (G*n_win, ZF)
    float* Ts = Fs + a.G * n_win * ZF;           // (G*n_win, D)

    input_layer_fwd(m, a, Ws, Fs, Ts, nullptr, trial0, G);

    if (a.mode == FASTB200_MODE_TRAIN_HEAD) {
        // logits = mean_n last_layer(tokens[n])      (fast.py:276-277)
        for (int idx = threadIdx.x; idx < G * NC; idx += TR_THREADS) {
            const int c = idx % NC, g = idx / NC;
            const float* wl = a.params + m.off_last_w + c * D;
            float s = 0.f;
            for (int n = 0; n < n_win; ++n)
                for (int i = 0; i < D; ++i) s = fmaf(Ts[(g * n_win + n) * D + i], __ldg(wl + i), s);
            a.logits[(size_t)(trial0 + g) * NC + c] = s / (float)n_win + __ldg(a.params + m.off_last_b + c);
        }
        return;
    }
    // X = cat(cls, tokens) + pos      (fast.py:263-265)
    for (int idx = threadIdx.x; idx < R * D; idx += TR_THREADS) {
        const int i = idx % D, r = idx / D, t = r % tok, g = r / tok;
        const float base = t == 0 ? __ldg(a.params + m.off_cls + i) : Ts[(g * n_win + t - 1) * D + i];
        X[idx] = base + __ldg(a.params + m.off_pos + t * D + i);
    }
    __syncthreads();
    BlockBufs bb{X, Hn, QKV, P, O, U, nullptr, nullptr, nullptr, nullptr, nullptr};
    for (int l = 0; l < m.n_layers; ++l) {
        float *Wqkv, *Wo, *W1, *W2;
        BlockVec bv;
        stage_block(m, a.params, l, Ws, Wqkv, Wo, W1, W2, bv);
        if (a.saved)
            for (int idx = threadIdx.x; idx < R * D; idx += TR_THREADS) {
                const int g = idx / (tok * D), rem = idx - g * tok * D;
                a.saved[((size_t)(trial0 + g) * (m.n_layers + 1) + l) * tok * D + rem] = X[idx];
            }
        stage_wait();
        __syncthreads();
        block_fwd<false>(m, a.params, l, bb, Wqkv, Wo, W1, W2, bv, G, tok, dc, trial0);
    }
    if (a.saved)
        for (int idx = threadIdx.x; idx < R * D; idx += TR_THREADS) {
            const int g = idx / (tok * D), rem = idx - g * tok * D;
            a.saved[((size_t)(trial0 + g) * (m.n_layers + 1) + m.n_layers) * tok * D + rem] = X[idx];
        }
    // logits = last_layer(dropout(tokens[:, 0]))     (fast.py:267)
    for (int idx = threadIdx.x; idx < G * NC; idx += TR_THREADS) {
        const int c = idx % NC, g = idx / NC;
        const float* wl = a.params + m.off_last_w + c * D;
        float s = __ldg(a.params + m.off_last_b + c);
        for (int i = 0; i < D; ++i)
            s = fmaf(X[(g * tok) * D + i] * dc.mask(0, SITE_FINAL, trial0 + g, i), __ldg(wl + i), s);
        a.logits[(size_t)(trial0 + g) * NC + c] = s;
    }
}

// ---------------------------------------------------------------------------------------------------------
// backward kernel
// ---------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(TR_THREADS) transformer_bwd_kernel(const __grid_constant__ fastb200_model m,
                                                                     const __grid_constant__ TrArgs a) {
    extern __shared__ __align__(16) float smem[];
    const int D = m.dim_token, Hd = m.dim_hidden, H = m.n_heads, ZF = m.n_zones * m.dim_cnn, NC = m.n_classes;
    const int hd = D / H;
    const int tok = a.tok, n_win = a.n_win, L = m.n_layers;
    const int trial0 = blockIdx.x * a.G;
    const int G = min(a.G, a.B - trial0);
    const int R = G * tok, Rmax = a.G * tok, Rin = G * n_win;
    const int big = 3 * D > Hd ? 3 * D : Hd;
    float* Ws = smem;
    float* X = Ws + a.ws_floats;
    float* Xmid = X + Rmax * D;
    float* Hn = Xmid + Rmax * D;
    float* Hn2 = Hn + Rmax * D;
    float* QKV = Hn2 + Rmax * D;
    float* O = QKV + Rmax * 3 * D;
    float* Upre = O + Rmax * D;
    float* U = Upre + Rmax * Hd;
    float* dX = U + Rmax * Hd;
    float* dT = dX + Rmax * D;
    float* dBig = dT + Rmax * D;
    float* P = dBig + Rmax * big;
    float* dS = P + a.G * H * tok * tok;
    float* stat1 = dS + a.G * H * tok * tok;
    float* stat2 = stat1 + 2 * Rmax;
    float* Fs = stat2 + 2 * Rmax;                // (G*n_win, ZF)
    float* g = a.grads;
    DropCtx dc = a.dc;
    dc.resolve();

    if (a.mode == FASTB200_MODE_DEFAULT) {
        // ---- classifier: logits = W_l drop(X_L[:,0]) + b_l
        const float* XL = a.saved;
        for (int idx = threadIdx.x; idx < R * D; idx += TR_THREADS) {
            const int i = idx % D, r = idx / D, t = r % tok, gg = r / tok;
            float v = 0.f;
            if (t == 0) {
                const float* dl = a.dlogits + (size_t)(trial0 + gg) * NC;
                for (int c = 0; c < NC; ++c) v = fmaf(__ldg(dl + c), __ldg(a.params + m.off_last_w + c * D + i), v);
                v *= dc.mask(0, SITE_FINAL, trial0 + gg, i);
            }
            dX[idx] = v;
        }
        for (int idx = threadIdx.x; idx < NC * D; idx += TR_THREADS) {
            const int i = idx % D, c = idx / D;
            float s = 0.f;
            for (int gg = 0; gg < G; ++gg) {
                const float x0 = __ldg(XL + ((size_t)(trial0 + gg) * (L + 1) + L) * tok * D + i) * dc.mask(0, SITE_FINAL, trial0 + gg, i);
                s = fmaf(__ldg(a.dlogits + (size_t)(trial0 + gg) * NC + c), x0, s);
            }
            atomicAdd(g + m.off_last_w + idx, s);
        }
        for (int c = threadIdx.x; c < NC; c += TR_THREADS) {
            float s = 0.f;
            for (int gg = 0; gg < G; ++gg) s += __ldg(a.dlogits + (size_t)(trial0 + gg) * NC + c);
            atomicAdd(g + m.off_last_b + c, s);
        }
        __syncthreads();

        BlockBufs bb{X, Hn, QKV, P, O, U, Upre, stat1, stat2, Xmid, Hn2};
        for (int l = L - 1; l >= 0; --l) {
            float *Wqkv, *Wo, *W1, *W2;
            BlockVec bv;
            stage_block(m, a.params, l, Ws, Wqkv, Wo, W1, W2, bv);
            for (int idx = threadIdx.x; idx < R * D; idx += TR_THREADS) {
                const int gg = idx / (tok * D), rem = idx - gg * tok * D;
                X[idx] = __ldg(a.saved + ((size_t)(trial0 + gg) * (L + 1) + l) * tok * D + rem);
            }
            stage_wait();
            __syncthreads();
            block_fwd<true>(m, a.params, l, bb, Wqkv, Wo, W1, W2, bv, G, tok, dc, trial0);   // recompute, X intact

            // ---- FFN: X_out = Xmid + drop(W2 U + b2),  U = drop(gelu(Upre)),  Upre = W1 Hn2 + b1
            for (int idx = threadIdx.x; idx < R * D; idx += TR_THREADS) {
                const int j = idx % D, r = idx / D;
                dT[idx] = dX[idx] * dc.mask(l, SITE_FFN_OUT, trial0 + r / tok, (r % tok) * D + j);
            }
            __syncthreads();
            linear_wgrad(dT, D, U, Hd, Hd, D, R, g + m.off_ff2_w[l], g + m.off_ff2_b[l]);
            linear_dgrad(dT, D, W2, Hd, D, dBig, Hd, R);
            __syncthreads();
            for (int idx = threadIdx.x; idx < R * Hd; idx += TR_THREADS) {
                const int j = idx % Hd, r = idx / Hd;
                dBig[idx] *= dc.mask(l, SITE_FFN_HID, trial0 + r / tok, (r % tok) * Hd + j) * gelu_grad_f(Upre[idx]);
            }
            __syncthreads();
            linear_wgrad(dBig, Hd, Hn2, D, D, Hd, R, g + m.off_ff1_w[l], g + m.off_ff1_b[l]);
            linear_dgrad(dBig, Hd, W1, D, Hd, dT, D, R);          // dT <- dHn2
            __syncthreads();
            ln_bwd(dT, Xmid, stat2, bv.ln2_w, dX, R, D, g + m.off_ln2_w[l], g + m.off_ln2_b[l]);
            __syncthreads();                                             // dX now holds dL/dXmid
            // ---- attention: Xmid = X + Wo O + bo
            linear_wgrad(dX, D, O, D, D, D, R, g + m.off_out_w[l], g + m.off_out_b[l]);
            linear_dgrad(dX, D, Wo, D, D, dT, D, R);              // dT <- dO
            __syncthreads();
            const float scale = 1.0f / sqrtf((float)hd);
            // dS[i][j] = P (dP - sum_j dP P), dP = mask * (dO_i . v_j)
            for (int idx = threadIdx.x; idx < G * H * tok; idx += TR_THREADS) {
                const int i = idx % tok, h = (idx / tok) % H, gg = idx / (tok * H);
                const float* dOi = dT + (gg * tok + i) * D + h * hd;
                const float* vbase = QKV + (gg * tok) * 3 * D + 2 * D + h * hd;
                const float* prow = P + ((gg * H + h) * tok + i) * tok;
                float* dsrow = dS + ((gg * H + h) * tok + i) * tok;
                float dot = 0.f;
                for (int j = 0; j < tok; ++j) {
                    float dp = 0.f;
                    for (int d = 0; d < hd; ++d) dp = fmaf(dOi[d], vbase[j * 3 * D + d], dp);
                    dp *= dc.mask(l, SITE_ATTN, trial0 + gg, (h * tok + i) * tok + j);
                    dsrow[j] = dp;
                    dot = fmaf(dp, prow[j], dot);
                }
                for (int j = 0; j < tok; ++j) dsrow[j] = prow[j] * (dsrow[j] - dot);
            }
            __syncthreads();
            // dQKV: per (trial, head, token, d)
            for (int idx = threadIdx.x; idx < R * D; idx += TR_THREADS) {
                const int c = idx % D, r = idx / D, t = r % tok, gg = r / tok, h = c / hd;
                const float* qk = QKV + (gg * tok) * 3 * D;
                const float* Pg = P + (gg * H + h) * tok * tok;
                const float* dSg = dS + (gg * H + h) * tok * tok;
                float dq = 0.f, dk = 0.f, dv = 0.f;
                for (int j = 0; j < tok; ++j) {
                    dq = fmaf(dSg[t * tok + j], qk[j * 3 * D + D + c], dq);              // sum_j dS[t][j] k_j
                    dk = fmaf(dSg[j * tok + t], qk[j * 3 * D + c], dk);                  // sum_i dS[i][t] q_i
                    dv = fmaf(Pg[j * tok + t] * dc.mask(l, SITE_ATTN, trial0 + gg, (h * tok + j) * tok + t),
                              dT[(gg * tok + j) * D + c], dv);                          // sum_i Pd[i][t] dO_i
                }
                dBig[r * 3 * D + c] = dq * scale;
                dBig[r * 3 * D + D + c] = dk * scale;
                dBig[r * 3 * D + 2 * D + c] = dv;
            }
            __syncthreads();
            linear_wgrad(dBig, 3 * D, Hn, D, D, 3 * D, R, g + m.off_qkv_w[l], g + m.off_qkv_b[l]);
            linear_dgrad(dBig, 3 * D, Wqkv, D, 3 * D, dT, D, R);   // dT <- dHn
            __syncthreads();
            ln_bwd(dT, X, stat1, bv.ln1_w, dX, R, D, g + m.off_ln1_w[l], g + m.off_ln1_b[l]);
            __syncthreads();                                             // dX now holds dL/dX_l
        }
        // ---- X_0 = cat(cls, T) + pos
        for (int idx = threadIdx.x; idx < tok * D; idx += TR_THREADS) {
            float s = 0.f;
            for (int gg = 0; gg < G; ++gg) s += dX[gg * tok * D + idx];
            atomicAdd(g + m.off_pos + idx, s);
            if (idx < D) atomicAdd(g + m.off_cls + idx, s);
        }
        // dT rows (g*n_win + n) <- dX rows (g*tok + 1 + n)
        for (int idx = threadIdx.x; idx < Rin * D; idx += TR_THREADS) {
            const int i = idx % D, r = idx / D, n = r % n_win, gg = r / n_win;
            dT[idx] = dX[(gg * tok + 1 + n) * D + i];
        }
        __syncthreads();
    }

    // ---- input layer (both modes): T = gelu(pre), pre = W_in feat + b_in.   Upre/U reused as pre / T.
    float* pre = Upre;
    float* Ts = U;
    input_layer_fwd(m, a, Ws, Fs, Ts, pre, trial0, G);
    if (a.mode == FASTB200_MODE_TRAIN_HEAD) {
        // logits = mean_n (W_l T_n + b_l)
        const float inv_n = 1.0f / (float)n_win;
        for (int idx = threadIdx.x; idx < Rin * D; idx += TR_THREADS) {
            const int i = idx % D, gg = (idx / D) / n_win;
            const float* dl = a.dlogits + (size_t)(trial0 + gg) * NC;
            float v = 0.f;
            for (int c = 0; c < NC; ++c) v = fmaf(__ldg(dl + c), __ldg(a.params + m.off_last_w + c * D + i), v);
            dT[idx] = v * inv_n;
        }
        for (int idx = threadIdx.x; idx < NC * D; idx += TR_THREADS) {
            const int i = idx % D, c = idx / D;
            float s = 0.f;
            for (int r = 0; r < Rin; ++r) s = fmaf(__ldg(a.dlogits + (size_t)(trial0 + r / n_win) * NC + c), Ts[r * D + i], s);
            atomicAdd(g + m.off_last_w + idx, s * inv_n);
        }
        for (int c = threadIdx.x; c < NC; c += TR_THREADS) {
            float s = 0.f;
            for (int gg = 0; gg < G; ++gg) s += __ldg(a.dlogits + (size_t)(trial0 + gg) * NC + c);
            atomicAdd(g + m.off_last_b + c, s);
        }
        __syncthreads();
    }
    for (int idx = threadIdx.x; idx < Rin * D; idx += TR_THREADS) dT[idx] *= gelu_grad_f(pre[idx]);
    __syncthreads();
    linear_wgrad(dT, D, Fs, ZF, ZF, D, Rin, g + m.off_in_w, g + m.off_in_b);
    __syncthreads();                                                     // every thread is done reading Fs
    if (a.dfeat) linear_dgrad(dT, D, Ws, ZF, D, Fs, ZF, Rin);     // Fs <- dfeat
    __syncthreads();
    if (a.dfeat)
        for (int i = threadIdx.x; i < Rin * ZF; i += TR_THREADS) a.dfeat[(size_t)trial0 * n_win * ZF + i] = Fs[i];
}

// ---------------------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------------------
static int host_ws_floats(const fastb200_model& m) {
    const int D = m.dim_token, Hd = m.dim_hidden, ZF = m.n_zones * m.dim_cnn;
    const int blk = 3 * D * (D + 1) + D * (D + 1) + Hd * (D + 1) + D * (Hd + 1) + 9 * D + Hd;     // weights + vectors
    const int inp = D * (ZF + 1);
    return round_up(blk > inp ? blk : inp, 4);
}
static size_t fwd_smem_floats(const fastb200_model& m, int G, int tok, int n_win) {
    const int D = m.dim_token, Hd = m.dim_hidden, ZF = m.n_zones * m.dim_cnn, R = G * tok;
    return (size_t)host_ws_floats(m) + (size_t)R * (D * 3 + 3 * D + Hd) + (size_t)G * m.n_heads * tok * tok +
           (size_t)G * n_win * (ZF + D);
}
static size_t bwd_smem_floats(const fastb200_model& m, int G, int tok, int n_win) {
    const int D = m.dim_token, Hd = m.dim_hidden, ZF = m.n_zones * m.dim_cnn, R = G * tok;
    const int big = 3 * D > Hd ? 3 * D : Hd;
    return (size_t)host_ws_floats(m) + (size_t)R * (D * 7 + 3 * D + 2 * Hd + big + 4) +
           (size_t)2 * G * m.n_heads * tok * tok + (size_t)G * n_win * ZF;
}

static int validate_tr(const fastb200_model* m, int B, int n_win, int mode, float p) {
    FB_REQUIRE(m != nullptr, "model descriptor is NULL");
    FB_REQUIRE(m->dim_token >= 1 && m->n_heads >= 1 && m->dim_token % m->n_heads == 0, "dim_token %d not divisible by heads %d", m->dim_token, m->n_heads);
    FB_REQUIRE(m->n_layers >= 0 && m->n_layers <= FASTB200_MAX_LAYERS, "n_layers=%d out of range", m->n_layers);
    FB_REQUIRE(mode == FASTB200_MODE_DEFAULT || mode == FASTB200_MODE_TRAIN_HEAD, "unknown mode %d", mode);
    FB_REQUIRE(B >= 0 && n_win >= 1, "bad B/n_win");
    FB_REQUIRE(mode != FASTB200_MODE_DEFAULT || n_win + 1 <= m->n_pos, "n_win+1=%d exceeds pos_embedding rows %d (fast.py:265 would fail to broadcast)", n_win + 1, m->n_pos);
    FB_REQUIRE(p >= 0.f && p < 1.f, "dropout p out of range");
    FB_REQUIRE(B < (1 << 24) && (n_win + 1) * m->dim_hidden * 1ll < (1 << 20) && (int64_t)m->n_heads * (n_win + 1) * (n_win + 1) < (1 << 20), "shape too large for the dropout counter layout");
    return 0;
}

// trials per CTA: as many as fit in `budget` bytes, but keep >= ~2 CTAs per SM worth of parallelism
template <typename SizeFn>
static int pick_group(const fastb200_model& m, int B, int tok, int n_win, SizeFn size_fn, size_t budget) {
    int G = 1;
    int want = ceil_div(B, 2 * NUM_SMS);
#ifdef FASTB200_DEBUG
    if (const char* e = getenv("FASTB200_TR_GROUP")) want = atoi(e) > 0 ? atoi(e) : want;     // developer builds: tuning probe
#endif
    while (G < want && G < 16 && size_fn(m, G + 1, tok, n_win) * sizeof(float) <= budget) ++G;
    return G;
}

// production shape -> one warp per trial (transformer_warp.cu); anything else -> the CTA-per-group kernels of this file
static bool use_warp_path(const fastb200_model& m, const float* params, const float* feat, int n_win, int mode, int backward = 0) {
#ifdef FASTB200_DEBUG
    if (const char* e = getenv("FASTB200_TR_GENERIC")) if (atoi(e) > 0) return false;        // developer builds: A/B probe
#endif
    return (reinterpret_cast<uintptr_t>(feat) & 15) == 0 && tw_supported(m, params, n_win, mode, backward) != 0;
}

}  // namespace fb

using namespace fb;

extern "C" int64_t fastb200_transformer_saved_floats(const fastb200_model* m, int B, int n_win) {
    if (m == nullptr) return -1;
    return (int64_t)B * (m->n_layers + 1) * (n_win + 1) * m->dim_token;
}

extern "C" int fastb200_transformer_fwd(const fastb200_model* m, const float* params, const float* feat, int B,
                                        int n_win, int mode, float dropout_p, uint64_t seed, uint64_t offset,
                                        const int32_t* dropout_step_dev, float* logits, float* saved, void* stream) {
    if (int r = validate_tr(m, B, n_win, mode, dropout_p)) return r;
    if (B == 0) return 0;
    TrArgs a{};
    a.params = params; a.feat = feat; a.logits = logits; a.saved = saved;
    a.B = B; a.n_win = n_win; a.tok = n_win + 1; a.mode = mode;
    a.dc.p = dropout_p; a.dc.scale = dropout_p > 0.f ? 1.0f / (1.0f - dropout_p) : 1.f; a.dc.seed = seed; a.dc.offset = offset; a.dc.step_dev = dropout_step_dev;
    if (use_warp_path(*m, params, feat, n_win, mode)) return tw_fwd_launch(*m, a, stream);
    a.ws_floats = host_ws_floats(*m);
    const size_t budget = 100 * 1024;
    FB_REQUIRE(fwd_smem_floats(*m, 1, a.tok, n_win) * sizeof(float) <= 227 * 1024, "sequence too long for the shared-memory resident transformer");
    a.G = pick_group(*m, B, a.tok, n_win, fwd_smem_floats, budget);
    const size_t smem = fwd_smem_floats(*m, a.G, a.tok, n_win) * sizeof(float);
    cudaError_t e = cudaFuncSetAttribute(transformer_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return fail((int)e, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    transformer_fwd_kernel<<<ceil_div(B, a.G), TR_THREADS, smem, as_stream(stream)>>>(*m, a);
    return check_launch("transformer_fwd");
}

extern "C" int fastb200_transformer_bwd(const fastb200_model* m, const float* params, const float* feat,
                                        const float* saved, const float* dlogits, int B, int n_win, int mode,
                                        float dropout_p, uint64_t seed, uint64_t offset,
                                        const int32_t* dropout_step_dev, float* grads, float* dfeat, void* stream) {
    if (int r = validate_tr(m, B, n_win, mode, dropout_p)) return r;
    FB_REQUIRE(mode != FASTB200_MODE_DEFAULT || saved != nullptr, "transformer_bwd needs the block inputs saved by transformer_fwd");
    if (B == 0) return 0;
    TrArgs a{};
    a.params = params; a.feat = feat; a.saved = const_cast<float*>(saved); a.dlogits = dlogits; a.grads = grads; a.dfeat = dfeat;
    a.B = B; a.n_win = n_win; a.tok = n_win + 1; a.mode = mode;
    a.dc.p = dropout_p; a.dc.scale = dropout_p > 0.f ? 1.0f / (1.0f - dropout_p) : 1.f; a.dc.seed = seed; a.dc.offset = offset; a.dc.step_dev = dropout_step_dev;
    if (((reinterpret_cast<uintptr_t>(grads) | reinterpret_cast<uintptr_t>(dfeat) | reinterpret_cast<uintptr_t>(saved)) & 15) == 0 &&
        use_warp_path(*m, params, feat, n_win, mode, 1))
        return tw_bwd_launch(*m, a, stream);
    a.ws_floats = host_ws_floats(*m);
    const size_t budget = 160 * 1024;
    FB_REQUIRE(bwd_smem_floats(*m, 1, a.tok, n_win) * sizeof(float) <= 227 * 1024, "sequence too long for the shared-memory resident transformer backward");
    a.G = pick_group(*m, B, a.tok, n_win, bwd_smem_floats, budget);
    const size_t smem = bwd_smem_floats(*m, a.G, a.tok, n_win) * sizeof(float);
    cudaError_t e = cudaFuncSetAttribute(transformer_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return fail((int)e, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    transformer_bwd_kernel<<<ceil_div(B, a.G), TR_THREADS, smem, as_stream(stream)>>>(*m, a);
    return check_launch("transformer_bwd");
}
