// Token path of FAST at the production shape, ONE WARP PER TRIAL.
//
// Reference: FAST.forward_transformer src/fast/models/fast.py:260-268, AttentionBlock :10-29, 'train_head' :273-278 -- the
// same function as transformer.cu, which stays the implementation for every other shape.  A trial is 6 tokens x 32
// features: transformer.cu spreads those 192 values over a 512-thread CTA and pays a CTA barrier between each of the ~12
// phases of a block (ncu, round 1: 25 % of the issue slots used, the rest barrier / dependency stalls).  Here lane j of a
// warp owns feature column j of every token of its trial:
//   * activations live in REGISTERS (X[t], q/k/v[t], the 6 x 6 attention probabilities of the lane's head ...);
//   * a linear layer is  acc[t] += x[t][i] * W[j][i]: the lane reads ITS weight row as float4 (rows are stored with their
//     16-byte chunks XOR-swizzled by (row & 7): conflict-free for row-per-lane float4 reads AND for the column-per-lane
//     scalar reads of the backward) and the token rows as warp-wide broadcast float4 from a 1.5 KB per-warp buffer;
//   * LayerNorm is a shuffle reduction per token, a head's 4-wide dot product is two xor-shuffles;
//   * ALL weights of the network (4 x 34 KB + the 32 KB input layer) are staged once per CTA by 16-byte cp.async, one
//     mbarrier per layer, so the first block starts while the later layers are still in flight; CTAs are persistent and
//     a warp walks trials blockIdx + gridDim * (slot + n_slots k) -- no CTA-wide barrier after the prologue.
//   * small batches (the cross-validation regime: 64 trials) cannot fill the GPU with one warp per trial and are bound by
//     the LATENCY of one warp walking the whole network, so the token rows of a trial can be SPLIT over S warps (S = 6:
//     one token each).  Everything but the attention is row-local; a block then costs one named barrier of the S warps,
//     around the exchange of the k / v rows through shared memory.  Per warp the instruction count drops ~5x, the total
//     grows 16 % (the weight-row reads are no longer amortised over 6 tokens), so large batches keep S = 1.
// Dropout masks use the same counter layout as transformer.cu (the two implementations draw identical masks).
#include "transformer.cuh"
#include <cstdlib>

namespace fb {
namespace tw {

constexpr int D = 32, HD = 64, H = 8;                 // dim_token, hidden, heads (head dim 4)
constexpr int MAX_WARPS = 8;
constexpr int LAYER_W = 3 * D * D + D * D + HD * D + D * HD;      // 8192 floats: Wqkv | Wo | W1 | W2
constexpr int O_WQKV = 0, O_WO = 3 * D * D, O_W1 = O_WO + D * D, O_W2 = O_W1 + HD * D;
enum { V_LN1W = 0, V_LN1B = 32, V_QKVB = 64, V_OUTB = 160, V_LN2W = 192, V_LN2B = 224, V_FF1B = 256, V_FF2B = 320, LAYER_V = 352 };
constexpr int LAYER_FLOATS = LAYER_W + LAYER_V;
constexpr int HDR_FLOATS = 32;                        // mbarriers: one per staging group (input layer + L blocks)

struct Lay {                                          // shared-memory map (float offsets)
    int win, in_b, cls, pos, last_w, last_b, layer0, slot0, per_slot, warp0, per_warp, xs_off, n_warps, total;
};

// n_warps = S * slots; a warp owns R = tok / S token rows: `fs_rows` staged feature rows + an [R][xs_cols] broadcast buffer;
// a slot (the S warps of one trial) owns `slot_floats` of exchange space.
__host__ __device__ inline Lay make_lay(const fastb200_model& m, int n_win, int S, int n_slots, int xs_cols, int slot_floats) {
    const int ZF = m.n_zones * m.dim_cnn, tok = n_win + 1, R = tok / S;
    Lay l;
    l.win = HDR_FLOATS;
    l.in_b = l.win + D * ZF;
    l.cls = l.in_b + D;
    l.pos = l.cls + D;
    l.last_w = l.pos + tok * D;
    l.last_b = l.last_w + m.n_classes * D;
    l.layer0 = round_up(l.last_b + m.n_classes, 4);
    l.slot0 = l.layer0 + m.n_layers * LAYER_FLOATS;
    l.per_slot = S > 1 ? slot_floats : 0;
    l.warp0 = l.slot0 + n_slots * l.per_slot;
    l.xs_off = (S == 1 ? R - 1 : R) * ZF;
    l.per_warp = l.xs_off + R * xs_cols;
    l.n_warps = S * n_slots;
    l.total = l.warp0 + l.n_warps * l.per_warp;
    return l;
}

__device__ __forceinline__ void cp16(float* dst, const float* src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp4(float* dst, const float* src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_wait_committed() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void cp_arrive(uint64_t* bar) {
    asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// (rows x cols) row-major -> rows of `cols` floats whose 16-byte chunk c sits at chunk c ^ (row & 7)
__device__ __forceinline__ void stage_swz(float* dst, const float* __restrict__ src, int rows, int cols, int tid, int nthreads) {
    const int cpr = cols >> 2;
    for (int i = tid; i < rows * cpr; i += nthreads) {
        const int j = i / cpr, c = i - j * cpr;
        cp16(dst + j * cols + ((c ^ (j & 7)) << 2), src + 4 * (size_t)i);
    }
}
__device__ __forceinline__ void stage_lin(float* dst, const float* __restrict__ src, int n, int tid, int nthreads) {
    if ((n & 3) == 0) { for (int i = tid; i < (n >> 2); i += nthreads) cp16(dst + 4 * i, src + 4 * i); }
    else { for (int i = tid; i < n; i += nthreads) cp4(dst + i, src + i); }
}
__device__ __forceinline__ float4 wrow4(const float* W, int cols, int j, int c) {
    return *reinterpret_cast<const float4*>(W + j * cols + ((c ^ (j & 7)) << 2));
}
__device__ __forceinline__ float wcol(const float* W, int cols, int j, int i) {          // W[j][i], lane-per-column readers
    return W[j * cols + ((((i >> 2) ^ (j & 7)) << 2) | (i & 3))];
}

// acc[o][t] += sum_i xs[t][i] * W[lane + 32 o][i],  i < 4 * CH  (ascending i, like the reference loop of transformer.cu)
template <int NR, int NO, int CH>
__device__ __forceinline__ void lin_rows(const float* W, const float* xs, int ldx, float (&acc)[NO][NR], int lane) {
#pragma unroll
    for (int c = 0; c < CH; ++c) {
        float4 w[NO];
#pragma unroll
        for (int o = 0; o < NO; ++o) w[o] = wrow4(W, 4 * CH, lane + 32 * o, c);
#pragma unroll
        for (int t = 0; t < NR; ++t) {
            const float4 x = *reinterpret_cast<const float4*>(xs + t * ldx + 4 * c);
#pragma unroll
            for (int o = 0; o < NO; ++o) {
                acc[o][t] = fmaf(x.x, w[o].x, acc[o][t]);
                acc[o][t] = fmaf(x.y, w[o].y, acc[o][t]);
                acc[o][t] = fmaf(x.z, w[o].z, acc[o][t]);
                acc[o][t] = fmaf(x.w, w[o].w, acc[o][t]);
            }
        }
    }
}

// The two schedules (S = 1 / S = 6) must give the same bits for the same trial (a trial's logits do not depend on its
// batch size): every multiply-add whose contraction into an FMA could depend on the surrounding code is written out.
__device__ __forceinline__ float gelu_nc(float x) { return __fmul_rn(0.5f * x, 1.0f + erff(x * 0.70710678118654752440f)); }

template <int R>
__device__ __forceinline__ void put_rows(float* xs, int ldx, const float (&v)[R], int lane, int col0 = 0) {
    __syncwarp();                                        // every lane is done reading the previous contents
#pragma unroll
    for (int r = 0; r < R; ++r) xs[r * ldx + col0 + lane] = v[r];
    __syncwarp();
}

template <int R>
__device__ __forceinline__ void layer_norm(const float (&x)[R], float (&y)[R], float w, float b, float* mean_o, float* rstd_o) {
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const float mean = warp_sum(x[r]) * (1.0f / D);
        const float d = x[r] - mean;
        const float rstd = 1.0f / sqrtf(warp_sum(d * d) * (1.0f / D) + LN_EPS);
        y[r] = fmaf(d * rstd, w, b);
        if (mean_o) { mean_o[r] = mean; rstd_o[r] = rstd; }
    }
}

// softmax(q k^T / 2) of the lane's head for the warp's R query rows (the 4 lanes of a head hold the same values)
template <int R, int TOK>
__device__ __forceinline__ void attn_probs(const float (&q)[R], const float (&k)[TOK], float (&p)[R][TOK]) {
#pragma unroll
    for (int i = 0; i < R; ++i) {
        float mx = -INFINITY;
        const float qi = q[i] * 0.5f;                    // 1 / sqrt(head dim 4)
#pragma unroll
        for (int j = 0; j < TOK; ++j) {
            float s = qi * k[j];
            s += __shfl_xor_sync(0xffffffffu, s, 1);
            s += __shfl_xor_sync(0xffffffffu, s, 2);
            p[i][j] = s;
            mx = fmaxf(mx, s);
        }
        float sum = 0.f;
#pragma unroll
        for (int j = 0; j < TOK; ++j) { p[i][j] = expf(p[i][j] - mx); sum += p[i][j]; }
        const float inv = 1.0f / sum;
#pragma unroll
        for (int j = 0; j < TOK; ++j) p[i][j] *= inv;
    }
}

// the S warps of one trial meet here (named barrier 1 + slot)
template <int S>
__device__ __forceinline__ void slot_sync(int slot) {
    if constexpr (S > 1) asm volatile("bar.sync %0, %1;" ::"r"(1 + slot), "n"(S * 32) : "memory");
}

// Who am I: warp `sub` of the S warps of slot `slot`; my row r is token sub + S r.
struct Who { int lane, sub, slot, trial; };

// prologue shared by the forward and backward kernels: barriers, this warp's first feature rows, all weights
template <int TOK, int S>
__device__ __forceinline__ void stage_feat(const TrArgs& a, float* fs, int ZF, int trial, int sub, int lane) {
    constexpr int R = TOK / S;
    // token t >= 1 reads feature row t - 1; warp `sub` keeps its rows in order (the CLS row of sub 0 has none)
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const int t = sub + S * r;
        if (t >= 1) stage_lin(fs + (r - (sub == 0 ? 1 : 0)) * ZF, a.feat + ((size_t)trial * (TOK - 1) + t - 1) * ZF, ZF, lane, 32);
    }
}
__device__ __forceinline__ void stage_weights(const fastb200_model& m, const TrArgs& a, const Lay& ly, float* smem, int tid, int nthreads) {
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem);
    const int ZF = m.n_zones * m.dim_cnn;
    stage_swz(smem + ly.win, a.params + m.off_in_w, D, ZF, tid, nthreads);
    stage_lin(smem + ly.in_b, a.params + m.off_in_b, D, tid, nthreads);
    stage_lin(smem + ly.cls, a.params + m.off_cls, D, tid, nthreads);
    stage_lin(smem + ly.pos, a.params + m.off_pos, a.tok * D, tid, nthreads);
    stage_lin(smem + ly.last_w, a.params + m.off_last_w, m.n_classes * D, tid, nthreads);
    stage_lin(smem + ly.last_b, a.params + m.off_last_b, m.n_classes, tid, nthreads);
    cp_arrive(&bars[0]);
    for (int l = 0; l < m.n_layers; ++l) {
        float* w = smem + ly.layer0 + l * LAYER_FLOATS;
        stage_swz(w + O_WQKV, a.params + m.off_qkv_w[l], 3 * D, D, tid, nthreads);
        stage_swz(w + O_WO, a.params + m.off_out_w[l], D, D, tid, nthreads);
        stage_swz(w + O_W1, a.params + m.off_ff1_w[l], HD, D, tid, nthreads);
        stage_swz(w + O_W2, a.params + m.off_ff2_w[l], D, HD, tid, nthreads);
        float* v = w + LAYER_W;
        stage_lin(v + V_LN1W, a.params + m.off_ln1_w[l], D, tid, nthreads);
        stage_lin(v + V_LN1B, a.params + m.off_ln1_b[l], D, tid, nthreads);
        stage_lin(v + V_QKVB, a.params + m.off_qkv_b[l], 3 * D, tid, nthreads);
        stage_lin(v + V_OUTB, a.params + m.off_out_b[l], D, tid, nthreads);
        stage_lin(v + V_LN2W, a.params + m.off_ln2_w[l], D, tid, nthreads);
        stage_lin(v + V_LN2B, a.params + m.off_ln2_b[l], D, tid, nthreads);
        stage_lin(v + V_FF1B, a.params + m.off_ff1_b[l], HD, tid, nthreads);
        stage_lin(v + V_FF2B, a.params + m.off_ff2_b[l], D, tid, nthreads);
        cp_arrive(&bars[1 + l]);
    }
}
__device__ __forceinline__ void init_bars(const fastb200_model& m, float* smem, int tid, int nthreads) {
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem);
    if (tid == 0) {
        for (int g = 0; g <= m.n_layers; ++g) mbar_init(&bars[g], nthreads);
        fence_mbar_init();
    }
    __syncthreads();
}

// pre[n] = b_in[lane] + sum_i fs[n][i] W_in[lane][i]   (input_layer.0, fast.py:231-233; GELU by the caller)
template <int NR>
__device__ __forceinline__ void input_rows(const float* win, const float* fs, int ZF, float b, float* pre, int lane) {
    if constexpr (NR > 0) {
        float acc[NR];
#pragma unroll
        for (int n = 0; n < NR; ++n) acc[n] = b;
#pragma unroll 4
        for (int c = 0; c < (ZF >> 2); ++c) {
            const float4 w = wrow4(win, ZF, lane, c);
#pragma unroll
            for (int n = 0; n < NR; ++n) {
                const float4 x = *reinterpret_cast<const float4*>(fs + n * ZF + 4 * c);
                acc[n] = fmaf(x.x, w.x, acc[n]);
                acc[n] = fmaf(x.y, w.y, acc[n]);
                acc[n] = fmaf(x.z, w.z, acc[n]);
                acc[n] = fmaf(x.w, w.w, acc[n]);
            }
        }
#pragma unroll
        for (int n = 0; n < NR; ++n) pre[n] = acc[n];
    }
}
// pre[r] for the warp's rows (pre[0] of sub 0 -- the CLS row -- is left alone)
template <int R>
__device__ __forceinline__ void input_layer(const float* win, const float* fs, int ZF, float b, float (&pre)[R], int sub, int lane) {
    if (sub == 0) input_rows<R - 1>(win, fs, ZF, b, pre + 1, lane);
    else input_rows<R>(win, fs, ZF, b, pre, lane);
}

// What the backward needs from the recomputed forward of one block (registers of the owning lane).
template <int R, int TOK>
struct Keep {
    float mean1[R], rstd1[R], q[R], kall[TOK], vall[TOK], p[R][TOK], xmid[R], mean2[R], rstd2[R], upre[2][R];
};

// x <- block_l(x) for the warp's R rows.  xs: the warp's [R][64] broadcast buffer; kv: the slot's exchange buffer
// (2 parities x [TOK][2][32]), par: running parity of the slot barrier.  KEEP: fill `kp` and leave in xs the rows the
// weight gradients multiply with, through `emit(kind, rows)` (see the backward kernel).
template <int TOK, int S, bool KEEP, typename Emit>
__device__ __forceinline__ void block_fwd(const float* w, float (&x)[TOK / S], float* xs, float* kv, int& par, const DropCtx& dc,
                                          int l, const Who& me, Keep<TOK / S, TOK>* kp, Emit emit) {
    constexpr int R = TOK / S;
    const int lane = me.lane;
    const float* v = w + LAYER_W;
    float hn[R];
    layer_norm<R>(x, hn, v[V_LN1W + lane], v[V_LN1B + lane], KEEP ? kp->mean1 : nullptr, KEEP ? kp->rstd1 : nullptr);
    put_rows<R>(xs, HD, hn, lane);
    emit(0, hn);
    float qkv[3][R];
#pragma unroll
    for (int o = 0; o < 3; ++o)
#pragma unroll
        for (int r = 0; r < R; ++r) qkv[o][r] = v[V_QKVB + 32 * o + lane];
    lin_rows<R, 3, 8>(w + O_WQKV, xs, HD, qkv, lane);
    float kall[TOK], vall[TOK];
    if constexpr (S == 1) {
#pragma unroll
        for (int t = 0; t < TOK; ++t) { kall[t] = qkv[1][t]; vall[t] = qkv[2][t]; }
    } else {
        float* buf = kv + par * (TOK * 2 * D);
        par ^= 1;
#pragma unroll
        for (int r = 0; r < R; ++r) {
            buf[((me.sub + S * r) * 2 + 0) * D + lane] = qkv[1][r];
            buf[((me.sub + S * r) * 2 + 1) * D + lane] = qkv[2][r];
        }
        slot_sync<S>(me.slot);
#pragma unroll
        for (int t = 0; t < TOK; ++t) { kall[t] = buf[(t * 2 + 0) * D + lane]; vall[t] = buf[(t * 2 + 1) * D + lane]; }
    }
    float p[R][TOK];
    attn_probs<R, TOK>(qkv[0], kall, p);
    const int h = lane >> 2;
    float o[R];
#pragma unroll
    for (int i = 0; i < R; ++i) {
        const int ti = me.sub + S * i;
        float acc = 0.f;
#pragma unroll
        for (int j = 0; j < TOK; ++j) acc = fmaf(p[i][j] * dc.mask(l, SITE_ATTN, me.trial, (h * TOK + ti) * TOK + j), vall[j], acc);
        o[i] = acc;
    }
    if constexpr (KEEP) {
#pragma unroll
        for (int i = 0; i < R; ++i) {
            kp->q[i] = qkv[0][i];
#pragma unroll
            for (int j = 0; j < TOK; ++j) kp->p[i][j] = p[i][j];
        }
#pragma unroll
        for (int t = 0; t < TOK; ++t) { kp->kall[t] = kall[t]; kp->vall[t] = vall[t]; }
    }
    put_rows<R>(xs, HD, o, lane);
    emit(1, o);
    float y[1][R];
#pragma unroll
    for (int r = 0; r < R; ++r) y[0][r] = v[V_OUTB + lane];
    lin_rows<R, 1, 8>(w + O_WO, xs, HD, y, lane);
#pragma unroll
    for (int r = 0; r < R; ++r) x[r] += y[0][r];                                 // x = Xmid
    if constexpr (KEEP) {
#pragma unroll
        for (int r = 0; r < R; ++r) kp->xmid[r] = x[r];
    }
    layer_norm<R>(x, hn, v[V_LN2W + lane], v[V_LN2B + lane], KEEP ? kp->mean2 : nullptr, KEEP ? kp->rstd2 : nullptr);
    put_rows<R>(xs, HD, hn, lane);
    emit(2, hn);
    float u[2][R];
#pragma unroll
    for (int o2 = 0; o2 < 2; ++o2)
#pragma unroll
        for (int r = 0; r < R; ++r) u[o2][r] = v[V_FF1B + 32 * o2 + lane];
    lin_rows<R, 2, 8>(w + O_W1, xs, HD, u, lane);
#pragma unroll
    for (int o2 = 0; o2 < 2; ++o2)
#pragma unroll
        for (int r = 0; r < R; ++r) {
            if constexpr (KEEP) kp->upre[o2][r] = u[o2][r];
            u[o2][r] = gelu_nc(u[o2][r]) * dc.mask(l, SITE_FFN_HID, me.trial, (me.sub + S * r) * HD + 32 * o2 + lane);
        }
    put_rows<R>(xs, HD, u[0], lane);
    put_rows<R>(xs, HD, u[1], lane, 32);
    emit(3, u[0]);
    emit(4, u[1]);
    if constexpr (!KEEP) {
#pragma unroll
        for (int r = 0; r < R; ++r) y[0][r] = v[V_FF2B + lane];
        lin_rows<R, 1, 16>(w + O_W2, xs, HD, y, lane);
#pragma unroll
        for (int r = 0; r < R; ++r) x[r] = fmaf(y[0][r], dc.mask(l, SITE_FFN_OUT, me.trial, (me.sub + S * r) * D + lane), x[r]);
    }
}

struct NoEmit { template <typename T> __device__ __forceinline__ void operator()(int, const T&) const {} };

// logits row of one trial from a 32-wide vector spread over the lanes: out[c] = bias[c] + scale * sum_i vec[i] W[c][i]
__device__ __forceinline__ void classify(const float* smem, const Lay& ly, int NC, float vec, float scale, float* out, int lane) {
    float mine = 0.f;
    for (int c = 0; c < NC; ++c) {
        const float s = warp_sum(vec * smem[ly.last_w + c * D + lane]);
        if (lane == c) mine = fmaf(s, scale, smem[ly.last_b + c]);
    }
    if (lane < NC) out[lane] = mine;
}

template <int TOK, int S>
__global__ void __launch_bounds__(S == 1 ? MAX_WARPS * 32 : 384, 1) tw_fwd_kernel(const __grid_constant__ fastb200_model m,
                                                                   const __grid_constant__ TrArgs a, const __grid_constant__ Lay ly) {
    extern __shared__ __align__(16) float smem[];
    constexpr int R = TOK / S;
    const int tid = threadIdx.x, warp = tid >> 5, nthreads = blockDim.x;
    const int L = m.n_layers, ZF = m.n_zones * m.dim_cnn, NC = m.n_classes;
    Who me{tid & 31, warp % S, warp / S, 0};
    const int lane = me.lane;
    float* fs = smem + ly.warp0 + warp * ly.per_warp;
    float* xs = fs + ly.xs_off;
    float* kv = smem + ly.slot0 + me.slot * ly.per_slot;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem);
    int trial = blockIdx.x + gridDim.x * me.slot;
    const int tstride = gridDim.x * (nthreads / (32 * S));
    init_bars(m, smem, tid, nthreads);
    if (trial < a.B) stage_feat<TOK, S>(a, fs, ZF, trial, me.sub, lane);
    cp_commit();
    stage_weights(m, a, ly, smem, tid, nthreads);
    DropCtx dc = a.dc;
    dc.resolve();
    bool first = true;
    int par = 0;
#pragma unroll 1
    for (; trial < a.B; trial += tstride) {
        me.trial = trial;
        cp_wait_committed();
        __syncwarp();
        if (first) mbar_wait(&bars[0], 0);
        float pre[R];
        input_layer<R>(smem + ly.win, fs, ZF, smem[ly.in_b + lane], pre, me.sub, lane);
        __syncwarp();                                    // the feature rows are consumed: fetch the next trial's behind the blocks
        if (trial + tstride < a.B) stage_feat<TOK, S>(a, fs, ZF, trial + tstride, me.sub, lane);
        cp_commit();
        if (a.mode == FASTB200_MODE_TRAIN_HEAD) {
            // logits = mean_n last_layer(tokens[n])      (fast.py:276-277)
            float ts = 0.f;
#pragma unroll
            for (int r = 0; r < R; ++r)
                if (me.sub + S * r >= 1) ts += gelu_nc(pre[r]);
            if constexpr (S > 1) {
                float* buf = kv + par * (TOK * 2 * D);
                par ^= 1;
                buf[me.sub * D + lane] = ts;
                slot_sync<S>(me.slot);
                ts = 0.f;
#pragma unroll
                for (int s2 = 0; s2 < S; ++s2) ts += buf[s2 * D + lane];
            }
            if (me.sub == 0) classify(smem, ly, NC, ts, 1.0f / (TOK - 1), a.logits + (size_t)trial * NC, lane);
            first = false;
            continue;
        }
        float x[R];                                      // X = cat(cls, tokens) + pos      (fast.py:263-265)
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const int t = me.sub + S * r;
            x[r] = (t == 0 ? smem[ly.cls + lane] : gelu_nc(pre[r])) + smem[ly.pos + t * D + lane];
        }
        for (int l = 0; l < L; ++l) {
            if (a.saved) {
#pragma unroll
                for (int r = 0; r < R; ++r) a.saved[(((size_t)trial * (L + 1) + l) * TOK + me.sub + S * r) * D + lane] = x[r];
            }
            if (first) mbar_wait(&bars[1 + l], 0);
            block_fwd<TOK, S, false>(smem + ly.layer0 + l * LAYER_FLOATS, x, xs, kv, par, dc, l, me, nullptr, NoEmit{});
        }
        if (a.saved) {
#pragma unroll
            for (int r = 0; r < R; ++r) a.saved[(((size_t)trial * (L + 1) + L) * TOK + me.sub + S * r) * D + lane] = x[r];
        }
        // logits = last_layer(dropout(tokens[:, 0]))     (fast.py:267)
        if (me.sub == 0) classify(smem, ly, NC, x[0] * dc.mask(0, SITE_FINAL, trial, lane), 1.0f, a.logits + (size_t)trial * NC, lane);
        first = false;
    }
    asm volatile("cp.async.wait_all;" ::: "memory");
}

static bool aligned4(int off) { return (off & 3) == 0; }
constexpr int SMEM_MAX = 227 * 1024;
constexpr int KV_FLOATS = 2 * 6 * 2 * D;               // forward exchange space of a slot: 2 parities x [6 tokens][k | v][32]

struct Plan { int S, slots, grid; Lay ly; };
// One warp per trial when the batch fills the GPU that way; below that the rows of a trial are split over 6 warps (the
// latency of one warp walking the network is what a small batch waits for).
static Plan make_plan(const fastb200_model& m, int B, int n_win, int xs_cols, int slot_floats, int max_warps_split) {
    Plan p;
    p.S = B <= 4 * NUM_SMS ? 6 : 1;
#ifdef FASTB200_DEBUG
    if (const char* e = getenv("FASTB200_TW_SPLIT")) if (atoi(e) == 1 || atoi(e) == 6) p.S = atoi(e);        // developer builds: tuning probe
#endif
    const int max_slots = p.S == 1 ? MAX_WARPS : max_warps_split / p.S;
    p.grid = B < NUM_SMS ? B : NUM_SMS;
    p.slots = ceil_div(B, p.grid);
    if (p.slots > max_slots) p.slots = max_slots;
    while (p.slots > 1 && (size_t)make_lay(m, n_win, p.S, p.slots, xs_cols, slot_floats).total * sizeof(float) > SMEM_MAX) --p.slots;
    p.ly = make_lay(m, n_win, p.S, p.slots, xs_cols, slot_floats);
    return p;
}

}  // namespace tw

using namespace tw;

int tw_supported(const fastb200_model& m, const float* params, int n_win, int mode) {
    (void)mode;
    const int ZF = m.n_zones * m.dim_cnn, tok = n_win + 1;
    if (m.dim_token != D || m.dim_hidden != HD || m.n_heads != H) return 0;
    if (tok != 6) return 0;                                          // instantiated token counts
    if (m.n_layers < 0 || m.n_layers > 4 || m.n_classes < 1 || m.n_classes > 32 || ZF % 32 != 0) return 0;
    if ((reinterpret_cast<uintptr_t>(params) & 15) != 0) return 0;
    bool ok = aligned4(m.off_in_w) && aligned4(m.off_in_b) && aligned4(m.off_cls) && aligned4(m.off_pos) && aligned4(m.off_last_w);
    for (int l = 0; l < m.n_layers; ++l)
        ok = ok && aligned4(m.off_ln1_w[l]) && aligned4(m.off_ln1_b[l]) && aligned4(m.off_qkv_w[l]) && aligned4(m.off_qkv_b[l]) &&
             aligned4(m.off_out_w[l]) && aligned4(m.off_out_b[l]) && aligned4(m.off_ln2_w[l]) && aligned4(m.off_ln2_b[l]) &&
             aligned4(m.off_ff1_w[l]) && aligned4(m.off_ff1_b[l]) && aligned4(m.off_ff2_w[l]) && aligned4(m.off_ff2_b[l]);
    if (!ok) return 0;
    return (size_t)make_lay(m, n_win, 1, 1, HD, KV_FLOATS).total * sizeof(float) <= SMEM_MAX &&
           (size_t)make_lay(m, n_win, 6, 1, HD, KV_FLOATS).total * sizeof(float) <= SMEM_MAX ? 1 : 0;
}

template <typename K>
static int launch_tw(K kernel, const Plan& p, const fastb200_model& m, const TrArgs& a, void* stream, const char* what) {
    const size_t smem = (size_t)p.ly.total * sizeof(float);
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return fail((int)e, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    kernel<<<p.grid, p.ly.n_warps * 32, smem, as_stream(stream)>>>(m, a, p.ly);
    return check_launch(what);
}

int tw_fwd_launch(const fastb200_model& m, const TrArgs& a, void* stream) {
    const Plan p = make_plan(m, a.B, a.n_win, HD, KV_FLOATS, 12);
    if (p.S == 1) return launch_tw(tw_fwd_kernel<6, 1>, p, m, a, stream, "transformer_fwd_warp");
    return launch_tw(tw_fwd_kernel<6, 6>, p, m, a, stream, "transformer_fwd_warp6");
}

}  // namespace fb
