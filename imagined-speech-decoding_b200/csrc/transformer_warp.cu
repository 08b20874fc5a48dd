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
constexpr int HDR_FLOATS = 32;                        // mbarriers: one per staging group (vectors, L blocks, input layer)

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
__device__ __forceinline__ void stage_block(const fastb200_model& m, const TrArgs& a, int dst_off, float* smem, int l, int tid, int nthreads) {
    float* w = smem + dst_off;
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
}
// Forward staging groups (one mbarrier each): 0 = the small vectors, L + 1 = the input-layer matrix, 1 + l = block l.  A
// group's barrier completes when every copy a thread issued BEFORE its arrive has landed, so the groups are issued in the
// order the kernel needs them.
__device__ __forceinline__ void stage_small(const fastb200_model& m, const TrArgs& a, const Lay& ly, float* smem, int tid, int nthreads) {
    stage_lin(smem + ly.in_b, a.params + m.off_in_b, D, tid, nthreads);
    stage_lin(smem + ly.cls, a.params + m.off_cls, D, tid, nthreads);
    stage_lin(smem + ly.pos, a.params + m.off_pos, a.tok * D, tid, nthreads);
    stage_lin(smem + ly.last_w, a.params + m.off_last_w, m.n_classes * D, tid, nthreads);
    stage_lin(smem + ly.last_b, a.params + m.off_last_b, m.n_classes, tid, nthreads);
}
__device__ __forceinline__ void stage_weights(const fastb200_model& m, const TrArgs& a, const Lay& ly, float* smem, int tid, int nthreads) {
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem);
    const int ZF = m.n_zones * m.dim_cnn, L = m.n_layers;
    stage_small(m, a, ly, smem, tid, nthreads);
    cp_arrive(&bars[0]);
    stage_swz(smem + ly.win, a.params + m.off_in_w, D, ZF, tid, nthreads);
    cp_arrive(&bars[L + 1]);
    for (int l = 0; l < L; ++l) { stage_block(m, a, ly.layer0 + l * LAYER_FLOATS, smem, l, tid, nthreads); cp_arrive(&bars[1 + l]); }
}
__device__ __forceinline__ void init_bars(int n_bars, float* smem, int tid, int nthreads) {
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem);
    if (tid == 0) {
        for (int g = 0; g < n_bars; ++g) mbar_init(&bars[g], nthreads);
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

// Deposit row of the backward (one per token row of a trial, DEP_ROW floats): the operands of the block's four weight
// gradients -- the inputs of the linear layers (written by the recomputed forward) and the gradients of their outputs.
enum { XC_HN1 = 0, XC_O = 32, XC_HN2 = 64, XC_U = 96, YC_QKV = 160, YC_XMID = 256, YC_UPRE = 288, YC_T = 352, DEP_ROW = 384 };

// What the backward needs from the recomputed forward of one block (registers of the owning lane).
template <int R, int TOK>
struct Keep {
    float mean1[R], rstd1[R], q[R], kall[TOK], vall[TOK], p[R][TOK], xmid[R], mean2[R], rstd2[R], upre[2][R];
    uint64_t abits;                                    // bit i * TOK + j: attention-dropout keep flag of (query row i, key j)
    uint32_t hbits;                                    // bit o * R + r: hidden-dropout keep flag
};

// x <- block_l(x) for the warp's R rows.  xs: the warp's [R][64] broadcast buffer; kv: the slot's exchange buffer
// (2 parities x [TOK][2][32]), par: running parity of the slot barrier.  KEEP (the backward's recomputation): fill `kp`,
// xs is the warp's deposit rows (LDX = DEP_ROW) and every linear layer's input stays in its own columns; the last linear
// layer (whose output only the next block needs) is skipped.
template <int TOK, int S, bool KEEP, int LDX>
__device__ __forceinline__ void block_fwd(const float* w, float (&x)[TOK / S], float* xs, float* kv, int& par, const DropCtx& dc,
                                          int l, const Who& me, Keep<TOK / S, TOK>* kp) {
    constexpr int R = TOK / S;
    constexpr int C_HN1 = KEEP ? XC_HN1 : 0, C_O = KEEP ? XC_O : 0, C_HN2 = KEEP ? XC_HN2 : 0, C_U = KEEP ? XC_U : 0;
    const int lane = me.lane;
    const float* v = w + LAYER_W;
    float hn[R];
    layer_norm<R>(x, hn, v[V_LN1W + lane], v[V_LN1B + lane], KEEP ? kp->mean1 : nullptr, KEEP ? kp->rstd1 : nullptr);
    put_rows<R>(xs + C_HN1, LDX, hn, lane);
    float qkv[3][R];
#pragma unroll
    for (int o = 0; o < 3; ++o)
#pragma unroll
        for (int r = 0; r < R; ++r) qkv[o][r] = v[V_QKVB + 32 * o + lane];
    lin_rows<R, 3, 8>(w + O_WQKV, xs + C_HN1, LDX, qkv, lane);
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
    if constexpr (KEEP) { kp->abits = 0; kp->hbits = 0; }
    float o[R];
#pragma unroll
    for (int i = 0; i < R; ++i) {
        const int ti = me.sub + S * i;
        float acc = 0.f;
#pragma unroll
        for (int j = 0; j < TOK; ++j) {
            const float mk = dc.mask(l, SITE_ATTN, me.trial, (h * TOK + ti) * TOK + j);
            if constexpr (KEEP) { if (mk != 0.f) kp->abits |= 1ull << (i * TOK + j); }
            acc = fmaf(p[i][j] * mk, vall[j], acc);
        }
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
    put_rows<R>(xs + C_O, LDX, o, lane);
    float y[1][R];
#pragma unroll
    for (int r = 0; r < R; ++r) y[0][r] = v[V_OUTB + lane];
    lin_rows<R, 1, 8>(w + O_WO, xs + C_O, LDX, y, lane);
#pragma unroll
    for (int r = 0; r < R; ++r) x[r] += y[0][r];                                 // x = Xmid
    if constexpr (KEEP) {
#pragma unroll
        for (int r = 0; r < R; ++r) kp->xmid[r] = x[r];
    }
    layer_norm<R>(x, hn, v[V_LN2W + lane], v[V_LN2B + lane], KEEP ? kp->mean2 : nullptr, KEEP ? kp->rstd2 : nullptr);
    put_rows<R>(xs + C_HN2, LDX, hn, lane);
    float u[2][R];
#pragma unroll
    for (int o2 = 0; o2 < 2; ++o2)
#pragma unroll
        for (int r = 0; r < R; ++r) u[o2][r] = v[V_FF1B + 32 * o2 + lane];
    lin_rows<R, 2, 8>(w + O_W1, xs + C_HN2, LDX, u, lane);
#pragma unroll
    for (int o2 = 0; o2 < 2; ++o2)
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const float mk = dc.mask(l, SITE_FFN_HID, me.trial, (me.sub + S * r) * HD + 32 * o2 + lane);
            if constexpr (KEEP) { kp->upre[o2][r] = u[o2][r]; if (mk != 0.f) kp->hbits |= 1u << (o2 * R + r); }
            u[o2][r] = gelu_nc(u[o2][r]) * mk;
        }
    put_rows<R>(xs + C_U, LDX, u[0], lane);
    put_rows<R>(xs + C_U, LDX, u[1], lane, 32);
    if constexpr (!KEEP) {
#pragma unroll
        for (int r = 0; r < R; ++r) y[0][r] = v[V_FF2B + lane];
        lin_rows<R, 1, 16>(w + O_W2, xs + C_U, LDX, y, lane);
#pragma unroll
        for (int r = 0; r < R; ++r) x[r] = fmaf(y[0][r], dc.mask(l, SITE_FFN_OUT, me.trial, (me.sub + S * r) * D + lane), x[r]);
    }
}

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
    init_bars(L + 2, smem, tid, nthreads);
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
        if (first) { mbar_wait(&bars[0], 0); mbar_wait(&bars[L + 1], 0); }
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
            block_fwd<TOK, S, false, HD>(smem + ly.layer0 + l * LAYER_FLOATS, x, xs, kv, par, dc, l, me, nullptr);
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

// =========================================================================================================================
// Backward (mode 'default': dlogits -> flat parameter gradients (+ dfeat)), same lane = feature column layout.
//
// Per block, in reverse order, a warp (a) reloads the block input X_l its forward saved, (b) re-runs the block keeping the
// intermediate values in registers (Keep) and every linear layer's INPUT rows in its deposit rows, (c) walks the block
// backwards: an input gradient dx[r][i] = sum_j dy[r][j] W[j][i] reads the lane's weight COLUMN (scalar, conflict-free under
// the XOR swizzle) and the dy rows as warp-wide broadcasts out of the deposit rows, where they stay for (d): the WEIGHT
// gradients are not formed per warp (one trial's 8,192 products per block would cost 8,192 atomics per trial) but by all
// threads of the CTA after a barrier, over every deposited row of the CTA at once: a thread owns a 4 x 8 tile of one
// weight matrix (256 tiles per block), reads 3 x 16 bytes per row for 32 multiply-adds, and adds its tile to the global
// gradient with 16-byte vector reductions -- one reduction per 4 elements per CTA and block.
// Weights: only two layer slots are resident (the block being processed and the next one in flight behind an mbarrier; the
// input-layer matrix uses a slot as the last stage), which leaves the room for the deposit rows.
// =========================================================================================================================
struct BLay {                                         // shared-memory map of the backward (float offsets)
    int in_b, cls, pos, last_w, last_b, wslot0, ex0, per_ex, warp0, per_warp, part_off, head_off, fs_off, n_warps, total;
};
constexpr int EX_FWD = 2 * 6 * 2 * D;                 // forward k / v exchange of a slot (S > 1): 2 parities x [6][k | v][32]
__host__ __device__ inline BLay make_blay(const fastb200_model& m, int n_win, int S, int n_slots) {
    const int ZF = m.n_zones * m.dim_cnn, tok = n_win + 1, R = tok / S;
    BLay l;
    l.in_b = HDR_FLOATS;
    l.cls = l.in_b + D;
    l.pos = l.cls + D;
    l.last_w = l.pos + tok * D;
    l.last_b = l.last_w + m.n_classes * D;
    l.wslot0 = round_up(l.last_b + m.n_classes, 4);
    l.ex0 = l.wslot0 + 2 * LAYER_FLOATS;
    l.per_ex = S > 1 ? EX_FWD + S * tok * 2 * D : 0;   // + backward exchange: [S writers][tok][dk | dv][32]
    l.warp0 = l.ex0 + n_slots * l.per_ex;
    l.part_off = R * DEP_ROW;                          // 4 x 32 LayerNorm weight / bias gradient partials of the warp's rows
    l.head_off = l.part_off + 4 * D;                   // x0 * mask (32) | dlogits (32) of the warp's trial
    l.fs_off = l.head_off + 2 * D;
    l.per_warp = l.fs_off + (S == 1 ? R - 1 : R) * ZF;
    l.n_warps = S * n_slots;
    l.total = l.warp0 + l.n_warps * l.per_warp;
    return l;
}

// dx[o][r] += sum_{j < DOUT} ys[r][j] * W[j][lane + 32 (o0 + o)]       (W: DOUT rows of `cols` floats, swizzled)
template <int NR, int NO, int DOUT>
__device__ __forceinline__ void dgrad_rows(const float* W, int cols, int o0, const float* ys, int ldy, float (&dx)[NO][NR], int lane) {
    static_assert(DOUT % 8 == 0, "rows are walked in groups of 8 (the period of the swizzle)");
    int lo[8];                                           // column lane + 32 o of row j sits at 32 o + lo[j & 7]
#pragma unroll
    for (int jj = 0; jj < 8; ++jj) lo[jj] = ((((lane >> 2) ^ jj) << 2) | (lane & 3)) + 32 * o0;
#pragma unroll 1
    for (int j8 = 0; j8 < DOUT / 8; ++j8) {
        const float* Wj = W + 8 * j8 * cols;
        float4 y[2][NR];
#pragma unroll
        for (int r = 0; r < NR; ++r) {
            y[0][r] = *reinterpret_cast<const float4*>(ys + r * ldy + 8 * j8);
            y[1][r] = *reinterpret_cast<const float4*>(ys + r * ldy + 8 * j8 + 4);
        }
#pragma unroll
        for (int jj = 0; jj < 8; ++jj) {
            float wv[NO];
#pragma unroll
            for (int o = 0; o < NO; ++o) wv[o] = Wj[jj * cols + 32 * o + lo[jj]];
#pragma unroll
            for (int r = 0; r < NR; ++r) {
                const float4 yq = y[jj >> 2][r];
                const float yv = (jj & 3) == 0 ? yq.x : (jj & 3) == 1 ? yq.y : (jj & 3) == 2 ? yq.z : yq.w;
#pragma unroll
                for (int o = 0; o < NO; ++o) dx[o][r] = fmaf(yv, wv[o], dx[o][r]);
            }
        }
    }
}

// LayerNorm backward of the warp's rows: dxout[r] += rstd (g - mean(g) - xhat mean(g xhat)), g = dy w; returns the lane's
// partial weight / bias gradients (sum over the rows).
template <int R>
__device__ __forceinline__ void ln_bwd_rows(const float (&dy)[R], const float (&x)[R], const float (&mean)[R], const float (&rstd)[R],
                                            float w, float (&dxout)[R], float& gw, float& gb) {
    gw = 0.f; gb = 0.f;
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const float g = dy[r] * w, xh = (x[r] - mean[r]) * rstd[r];
        const float a = warp_sum(g) * (1.0f / D), c = warp_sum(g * xh) * (1.0f / D);
        dxout[r] += rstd[r] * (g - a - xh * c);
        gw = fmaf(dy[r], xh, gw);
        gb += dy[r];
    }
}

// Backward of one block for the warp's R rows.  dX: in dL/dX_{l+1}, out dL/dX_l.  dep: the warp's deposit rows (inputs of
// the linear layers already there, see block_fwd<KEEP>); ex: the slot's backward exchange buffer (S > 1); part: the warp's
// 4 x 32 LayerNorm gradient partials.
template <int TOK, int S>
__device__ __forceinline__ void block_bwd(const float* w, const float (&xin)[TOK / S], float (&dX)[TOK / S], float* dep, float* ex,
                                          const DropCtx& dc, int l, const Who& me, const Keep<TOK / S, TOK>& kp, float* part) {
    constexpr int R = TOK / S;
    const int lane = me.lane;
    const float* v = w + LAYER_W;
    // ---- FFN: X_out = Xmid + drop(W2 U + b2),  U = drop(gelu(Upre)),  Upre = W1 Hn2 + b1
    float dT[R];
#pragma unroll
    for (int r = 0; r < R; ++r) dT[r] = dX[r] * dc.mask(l, SITE_FFN_OUT, me.trial, (me.sub + S * r) * D + lane);
    put_rows<R>(dep + YC_T, DEP_ROW, dT, lane);
    float dU[2][R];
#pragma unroll
    for (int o = 0; o < 2; ++o)
#pragma unroll
        for (int r = 0; r < R; ++r) dU[o][r] = 0.f;
    dgrad_rows<R, 2, D>(w + O_W2, HD, 0, dep + YC_T, DEP_ROW, dU, lane);
#pragma unroll
    for (int o = 0; o < 2; ++o)
#pragma unroll
        for (int r = 0; r < R; ++r)
            dU[o][r] = ((kp.hbits >> (o * R + r)) & 1u) ? dU[o][r] * dc.scale * gelu_grad_f(kp.upre[o][r]) : 0.f;
    put_rows<R>(dep + YC_UPRE, DEP_ROW, dU[0], lane);
    put_rows<R>(dep + YC_UPRE, DEP_ROW, dU[1], lane, 32);
    float dH[1][R];
#pragma unroll
    for (int r = 0; r < R; ++r) dH[0][r] = 0.f;
    dgrad_rows<R, 1, HD>(w + O_W1, D, 0, dep + YC_UPRE, DEP_ROW, dH, lane);
    float gw, gb;
    ln_bwd_rows<R>(dH[0], kp.xmid, kp.mean2, kp.rstd2, v[V_LN2W + lane], dX, gw, gb);      // dX = dL/dXmid
    part[2 * D + lane] = gw;
    part[3 * D + lane] = gb;
    // ---- attention: Xmid = X + Wo O + bo
    put_rows<R>(dep + YC_XMID, DEP_ROW, dX, lane);
    float dO[1][R];
#pragma unroll
    for (int r = 0; r < R; ++r) dO[0][r] = 0.f;
    dgrad_rows<R, 1, D>(w + O_WO, D, 0, dep + YC_XMID, DEP_ROW, dO, lane);
    float dq[R], dk[TOK], dv[TOK];                      // dk / dv: what the warp's query rows contribute to every key row
#pragma unroll
    for (int t = 0; t < TOK; ++t) { dk[t] = 0.f; dv[t] = 0.f; }
#pragma unroll
    for (int i = 0; i < R; ++i) {
        float dp[TOK], dot = 0.f;
#pragma unroll
        for (int j = 0; j < TOK; ++j) {
            float s = dO[0][i] * kp.vall[j];
            s += __shfl_xor_sync(0xffffffffu, s, 1);
            s += __shfl_xor_sync(0xffffffffu, s, 2);
            dp[j] = ((kp.abits >> (i * TOK + j)) & 1ull) ? s * dc.scale : 0.f;
            dot = fmaf(dp[j], kp.p[i][j], dot);
        }
        float acc = 0.f;
#pragma unroll
        for (int j = 0; j < TOK; ++j) {
            const float ds = kp.p[i][j] * (dp[j] - dot);
            acc = fmaf(ds, kp.kall[j], acc);
            dk[j] = fmaf(ds, kp.q[i], dk[j]);
            const float pm = ((kp.abits >> (i * TOK + j)) & 1ull) ? kp.p[i][j] * dc.scale : 0.f;
            dv[j] = fmaf(pm, dO[0][i], dv[j]);
        }
        dq[i] = acc * 0.5f;                              // 1 / sqrt(head dim 4)
    }
    float dkr[R], dvr[R];
    if constexpr (S == 1) {
#pragma unroll
        for (int t = 0; t < TOK; ++t) { dkr[t] = dk[t] * 0.5f; dvr[t] = dv[t]; }
    } else {
#pragma unroll
        for (int t = 0; t < TOK; ++t) {
            ex[((me.sub * TOK + t) * 2 + 0) * D + lane] = dk[t];
            ex[((me.sub * TOK + t) * 2 + 1) * D + lane] = dv[t];
        }
        slot_sync<S>(me.slot);
#pragma unroll
        for (int r = 0; r < R; ++r) {
            float sk = 0.f, sv = 0.f;
#pragma unroll
            for (int s2 = 0; s2 < S; ++s2) {
                sk += ex[((s2 * TOK + me.sub + S * r) * 2 + 0) * D + lane];
                sv += ex[((s2 * TOK + me.sub + S * r) * 2 + 1) * D + lane];
            }
            dkr[r] = sk * 0.5f; dvr[r] = sv;
        }
    }
    put_rows<R>(dep + YC_QKV, DEP_ROW, dq, lane);
    put_rows<R>(dep + YC_QKV, DEP_ROW, dkr, lane, 32);
    put_rows<R>(dep + YC_QKV, DEP_ROW, dvr, lane, 64);
#pragma unroll
    for (int r = 0; r < R; ++r) dH[0][r] = 0.f;
    dgrad_rows<R, 1, 3 * D>(w + O_WQKV, D, 0, dep + YC_QKV, DEP_ROW, dH, lane);
    ln_bwd_rows<R>(dH[0], xin, kp.mean1, kp.rstd1, v[V_LN1W + lane], dX, gw, gb);          // dX = dL/dX_l
    part[0 * D + lane] = gw;
    part[1 * D + lane] = gb;
}

__device__ __forceinline__ void red_add4(float* p, float a, float b, float c, float d) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

// deposit row q of the CTA (q < active slots x TOK): S = 1: warp = slot holds its TOK rows; S = TOK: one row per warp
template <int TOK, int S>
__device__ __forceinline__ const float* dep_row(const float* warp0, int per_warp, int q, int row_floats) {
    if constexpr (S == 1) return warp0 + (q / TOK) * per_warp + (q % TOK) * row_floats;
    else return warp0 + q * per_warp;
}

// One 4 (outputs j) x 8 (inputs i) tile of  gW[j][i] += sum_rows dy[row][j] x[row][i]  (+ the bias gradient when `gb`).
// row_y / row_x: functors row -> pointer to the 4 dy values / the 8 x values of that row.
template <typename FY, typename FX>
__device__ __forceinline__ void wgrad_tile(int n_rows, FY row_y, FX row_x, float* gW, int ld, float* gb) {
    float acc[4][8], bs[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        bs[j] = 0.f;
#pragma unroll
        for (int i = 0; i < 8; ++i) acc[j][i] = 0.f;
    }
#pragma unroll 2
    for (int q = 0; q < n_rows; ++q) {
        const float4 y = *reinterpret_cast<const float4*>(row_y(q));
        const float* xp = row_x(q);
        const float4 x0 = *reinterpret_cast<const float4*>(xp), x1 = *reinterpret_cast<const float4*>(xp + 4);
        const float yv[4] = {y.x, y.y, y.z, y.w};
        const float xv[8] = {x0.x, x0.y, x0.z, x0.w, x1.x, x1.y, x1.z, x1.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            bs[j] += yv[j];
#pragma unroll
            for (int i = 0; i < 8; ++i) acc[j][i] = fmaf(yv[j], xv[i], acc[j][i]);
        }
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        red_add4(gW + j * ld, acc[j][0], acc[j][1], acc[j][2], acc[j][3]);
        red_add4(gW + j * ld + 4, acc[j][4], acc[j][5], acc[j][6], acc[j][7]);
    }
    if (gb) red_add4(gb, bs[0], bs[1], bs[2], bs[3]);
}

// All threads: the four weight gradients, four bias gradients and the LayerNorm gradients of block l over the CTA's
// deposited rows (n_slots_act active slots).
template <int TOK, int S>
__device__ __forceinline__ void wgrad_block(const fastb200_model& m, float* g, int l, const float* warp0, const BLay& ly, int n_slots_act,
                                            int tid, int nthreads) {
    const int n_rows = n_slots_act * TOK;
    for (int tile = tid; tile < 256; tile += nthreads) {
        int yc, xc, ld, jg, ig, wo, bo;
        if (tile < 96) { jg = tile >> 2; ig = tile & 3; yc = YC_QKV; xc = XC_HN1; ld = D; wo = m.off_qkv_w[l]; bo = m.off_qkv_b[l]; }
        else if (tile < 128) { jg = (tile - 96) >> 2; ig = tile & 3; yc = YC_XMID; xc = XC_O; ld = D; wo = m.off_out_w[l]; bo = m.off_out_b[l]; }
        else if (tile < 192) { jg = (tile - 128) >> 2; ig = tile & 3; yc = YC_UPRE; xc = XC_HN2; ld = D; wo = m.off_ff1_w[l]; bo = m.off_ff1_b[l]; }
        else { jg = (tile - 192) >> 3; ig = tile & 7; yc = YC_T; xc = XC_U; ld = HD; wo = m.off_ff2_w[l]; bo = m.off_ff2_b[l]; }
        const int yo = yc + 4 * jg, xo = xc + 8 * ig, pw = ly.per_warp;
        wgrad_tile(n_rows,
                   [=](int q) { return dep_row<TOK, S>(warp0, pw, q, DEP_ROW) + yo; },
                   [=](int q) { return dep_row<TOK, S>(warp0, pw, q, DEP_ROW) + xo; },
                   g + wo + 4 * jg * ld + 8 * ig, ld, ig == 0 ? g + bo + 4 * jg : nullptr);
    }
    const int n_warps_act = n_slots_act * S;
    for (int idx = tid; idx < 4 * D; idx += nthreads) {
        float s = 0.f;
        for (int wq = 0; wq < n_warps_act; ++wq) s += warp0[wq * ly.per_warp + ly.part_off + idx];
        const int which = idx >> 5, c = idx & 31;
        const int off = which == 0 ? m.off_ln1_w[l] : which == 1 ? m.off_ln1_b[l] : which == 2 ? m.off_ln2_w[l] : m.off_ln2_b[l];
        atomicAdd(g + off + c, s);
    }
}

// All threads, last stage of a round: input-layer weight / bias gradients, positional / CLS gradients, classifier gradients.
// Rows of this stage (per warp, at the start of its deposit area): [dX_0 (32) | dL/dpre (32)] per token row.
constexpr int IN_ROW = 2 * D;
template <int TOK, int S>
__device__ __forceinline__ void wgrad_input(const fastb200_model& m, float* g, const float* warp0, const BLay& ly, int n_slots_act,
                                            int ZF, int NC, int tid, int nthreads) {
    const int n_rows = n_slots_act * (TOK - 1), pw = ly.per_warp, fo = ly.fs_off, tpr = ZF >> 3;
    // row q -> (slot, token t >= 1)
    auto row_y = [=](int q, int col) {
        const int s = q / (TOK - 1), t = 1 + q % (TOK - 1);
        return (S == 1 ? warp0 + s * pw + t * IN_ROW : warp0 + (s * S + t) * pw) + D + col;
    };
    auto row_x = [=](int q, int col) {
        const int s = q / (TOK - 1), t = 1 + q % (TOK - 1);
        return (S == 1 ? warp0 + s * pw + fo + (t - 1) * ZF : warp0 + (s * S + t) * pw + fo) + col;
    };
    for (int tile = tid; tile < 8 * tpr; tile += nthreads) {
        const int jg = tile / tpr, ig = tile - jg * tpr;
        wgrad_tile(n_rows, [=](int q) { return row_y(q, 4 * jg); }, [=](int q) { return row_x(q, 8 * ig); },
                   g + m.off_in_w + 4 * jg * ZF + 8 * ig, ZF, ig == 0 ? g + m.off_in_b + 4 * jg : nullptr);
    }
    for (int idx = tid; idx < TOK * D; idx += nthreads) {                 // X_0 = cat(cls, T) + pos
        const int t = idx >> 5, c = idx & 31;
        float s = 0.f;
        for (int sl = 0; sl < n_slots_act; ++sl) s += (S == 1 ? warp0 + sl * pw + t * IN_ROW : warp0 + (sl * S + t) * pw)[c];
        atomicAdd(g + m.off_pos + idx, s);
        if (t == 0) atomicAdd(g + m.off_cls + c, s);
    }
    for (int idx = tid; idx < NC * D + NC; idx += nthreads) {             // logits = W_l (x0 * mask) + b_l
        float s = 0.f;
        if (idx < NC * D) {
            const int c = idx >> 5, i = idx & 31;
            for (int sl = 0; sl < n_slots_act; ++sl) {
                const float* h = warp0 + sl * S * pw + ly.head_off;
                s = fmaf(h[D + c], h[i], s);
            }
            atomicAdd(g + m.off_last_w + idx, s);
        } else {
            const int c = idx - NC * D;
            for (int sl = 0; sl < n_slots_act; ++sl) s += warp0[sl * S * pw + ly.head_off + D + c];
            atomicAdd(g + m.off_last_b + c, s);
        }
    }
}

// stage k of a round: k < L -> block L-1-k, k == L -> the input-layer matrix (a slot holds LAYER_FLOATS >= 32 x ZF floats)
__device__ __forceinline__ void stage_bwd(const fastb200_model& m, const TrArgs& a, const BLay& ly, float* smem, int k, int slot,
                                          int tid, int nthreads) {
    const int L = m.n_layers, dst = ly.wslot0 + slot * LAYER_FLOATS;
    if (k < L) stage_block(m, a, dst, smem, L - 1 - k, tid, nthreads);
    else stage_swz(smem + dst, a.params + m.off_in_w, D, m.n_zones * m.dim_cnn, tid, nthreads);
}

template <int TOK, int S>
__global__ void __launch_bounds__(S == 1 ? MAX_WARPS * 32 : 384, 1) tw_bwd_kernel(const __grid_constant__ fastb200_model m,
                                                                   const __grid_constant__ TrArgs a, const __grid_constant__ BLay ly) {
    extern __shared__ __align__(16) float smem[];
    constexpr int R = TOK / S, ZFC = 256;                // feature row length of this instantiation (checked by the host)
    const int tid = threadIdx.x, warp = tid >> 5, nthreads = blockDim.x, n_slots = nthreads / (32 * S);
    const int L = m.n_layers, NC = m.n_classes;
    Who me{tid & 31, warp % S, warp / S, 0};
    const int lane = me.lane;
    float* warp0 = smem + ly.warp0;
    float* dep = warp0 + warp * ly.per_warp;
    float* part = dep + ly.part_off;
    float* head = dep + ly.head_off;
    float* fs = dep + ly.fs_off;
    float* kv = smem + ly.ex0 + me.slot * ly.per_ex;
    float* ex = kv + EX_FWD;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem);  // 0 / 1: the two weight slots, 2: the small vectors
    const int per_round = gridDim.x * n_slots;
    const int n_rounds = ceil_div(a.B - (int)blockIdx.x, per_round);
    const int n_stages = n_rounds * (L + 1);
    init_bars(3, smem, tid, nthreads);
    if ((int)blockIdx.x + (int)gridDim.x * me.slot < a.B) stage_feat<TOK, S>(a, fs, ZFC, blockIdx.x + gridDim.x * me.slot, me.sub, lane);
    cp_commit();
    stage_small(m, a, Lay{0, ly.in_b, ly.cls, ly.pos, ly.last_w, ly.last_b}, smem, tid, nthreads);
    cp_arrive(&bars[2]);
    for (int gs = 0; gs < 2 && gs < n_stages; ++gs) { stage_bwd(m, a, ly, smem, gs % (L + 1), gs & 1, tid, nthreads); cp_arrive(&bars[gs & 1]); }
    DropCtx dc = a.dc;
    dc.resolve();
    int par = 0, gs = 0;
    float* g = a.grads;
#pragma unroll 1
    for (int round = 0; round < n_rounds; ++round) {
        const int trial = blockIdx.x + gridDim.x * (me.slot + n_slots * round);
        const bool active = trial < a.B;
        const int left = a.B - (int)blockIdx.x - per_round * round;          // > 0
        const int n_act = min(n_slots, ceil_div(left, (int)gridDim.x));     // active slots of this round (a prefix)
        me.trial = trial;
        if (round == 0) mbar_wait(&bars[2], 0);
        // ---- classifier: logits = W_l drop(X_L[:, 0]) + b_l      (fast.py:267)
        float dX[R];
#pragma unroll
        for (int r = 0; r < R; ++r) dX[r] = 0.f;
        if (active && me.sub == 0) {
            const float mk = dc.mask(0, SITE_FINAL, trial, lane);
            const float x0 = __ldg(a.saved + (((size_t)trial * (L + 1) + L) * TOK) * D + lane) * mk;
            const float dl = lane < NC ? __ldg(a.dlogits + (size_t)trial * NC + lane) : 0.f;
            float vsum = 0.f;
            for (int c = 0; c < NC; ++c) vsum = fmaf(__shfl_sync(0xffffffffu, dl, c), smem[ly.last_w + c * D + lane], vsum);
            dX[0] = vsum * mk;
            head[lane] = x0;
            head[D + lane] = dl;
        }
#pragma unroll 1
        for (int k = 0; k < L; ++k, ++gs) {
            const int l = L - 1 - k, slot = gs & 1;
            const float* w = smem + ly.wslot0 + slot * LAYER_FLOATS;
            if (active) {
                mbar_wait(&bars[slot], (gs >> 1) & 1);
                float x[R], xin[R];
#pragma unroll
                for (int r = 0; r < R; ++r) xin[r] = x[r] = __ldg(a.saved + (((size_t)trial * (L + 1) + l) * TOK + me.sub + S * r) * D + lane);
                Keep<R, TOK> kp;
                block_fwd<TOK, S, true, DEP_ROW>(w, x, dep, kv, par, dc, l, me, &kp);
                block_bwd<TOK, S>(w, xin, dX, dep, ex, dc, l, me, kp, part);
            }
            __syncthreads();                             // every active warp's rows are deposited; the slot's weights are free
            if (gs + 2 < n_stages) { stage_bwd(m, a, ly, smem, (gs + 2) % (L + 1), slot, tid, nthreads); cp_arrive(&bars[slot]); }
            wgrad_block<TOK, S>(m, g, l, warp0, ly, n_act, tid, nthreads);
            __syncthreads();                             // deposit rows free again
        }
        // ---- X_0 = cat(cls, gelu(input_layer(feat))) + pos: rows [dX_0 | dL/dpre], dfeat
        {
            const int slot = gs & 1;
            const float* win = smem + ly.wslot0 + slot * LAYER_FLOATS;
            if (active) {
                mbar_wait(&bars[slot], (gs >> 1) & 1);
                cp_wait_committed();
                __syncwarp();
                float pre[R], dpre[R];
                input_layer<R>(win, fs, ZFC, smem[ly.in_b + lane], pre, me.sub, lane);
#pragma unroll
                for (int r = 0; r < R; ++r) dpre[r] = me.sub + S * r >= 1 ? dX[r] * gelu_grad_f(pre[r]) : 0.f;
                put_rows<R>(dep, IN_ROW, dX, lane);
                put_rows<R>(dep, IN_ROW, dpre, lane, D);
                if (a.dfeat) {
                    constexpr int NR = S == 1 ? R - 1 : R;                  // S == 1: rows 1 .. TOK-1; S > 1: the warp's row unless it is the CLS row
                    if (S == 1 || me.sub >= 1) {
                        const float* ys = dep + D + (S == 1 ? IN_ROW : 0);
#pragma unroll 1
                        for (int half = 0; half < 2; ++half) {
                            float df[4][NR];
#pragma unroll
                            for (int o = 0; o < 4; ++o)
#pragma unroll
                                for (int r = 0; r < NR; ++r) df[o][r] = 0.f;
                            dgrad_rows<NR, 4, D>(win, ZFC, 4 * half, ys, IN_ROW, df, lane);
#pragma unroll
                            for (int r = 0; r < NR; ++r) {
                                const int t = S == 1 ? r + 1 : me.sub;
#pragma unroll
                                for (int o = 0; o < 4; ++o)
                                    a.dfeat[((size_t)trial * (TOK - 1) + t - 1) * ZFC + lane + 32 * (4 * half + o)] = df[o][r];
                            }
                        }
                    }
                }
            }
            __syncthreads();
            if (gs + 2 < n_stages) { stage_bwd(m, a, ly, smem, (gs + 2) % (L + 1), slot, tid, nthreads); cp_arrive(&bars[slot]); }
            wgrad_input<TOK, S>(m, g, warp0, ly, n_act, ZFC, NC, tid, nthreads);
            __syncthreads();
            ++gs;
            const int next = trial + per_round;
            if (next < a.B) stage_feat<TOK, S>(a, fs, ZFC, next, me.sub, lane);
            cp_commit();
        }
    }
    asm volatile("cp.async.wait_all;" ::: "memory");
}

static bool aligned4(int off) { return (off & 3) == 0; }
constexpr int SMEM_MAX = 227 * 1024;
constexpr int KV_FLOATS = 2 * 6 * 2 * D;               // forward exchange space of a slot: 2 parities x [6 tokens][k | v][32]

struct Plan { int S, slots, grid; };
// One warp per trial when the batch fills the GPU that way; below that the rows of a trial are split over 6 warps (the
// latency of one warp walking the network is what a small batch waits for).
static Plan make_plan(int B, int max_warps_split) {
    Plan p;
    p.S = B <= 4 * NUM_SMS ? 6 : 1;
#ifdef FASTB200_DEBUG
    if (const char* e = getenv("FASTB200_TW_SPLIT")) if (atoi(e) == 1 || atoi(e) == 6) p.S = atoi(e);        // developer builds: tuning probe
#endif
    const int max_slots = p.S == 1 ? MAX_WARPS : max_warps_split / p.S;
    p.grid = B < NUM_SMS ? B : NUM_SMS;
    p.slots = ceil_div(B, p.grid);
    if (p.slots > max_slots) p.slots = max_slots;
    return p;
}

}  // namespace tw

using namespace tw;

int tw_supported(const fastb200_model& m, const float* params, int n_win, int mode, int backward) {
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
    if (backward) {
        // the backward is instantiated for mode 'default', 256-float feature rows (one weight slot holds the input-layer matrix)
        if (mode != FASTB200_MODE_DEFAULT || ZF != 256) return 0;
        return (size_t)make_blay(m, n_win, 1, MAX_WARPS).total * sizeof(float) <= SMEM_MAX &&
               (size_t)make_blay(m, n_win, 6, 2).total * sizeof(float) <= SMEM_MAX ? 1 : 0;
    }
    return (size_t)make_lay(m, n_win, 1, 1, HD, KV_FLOATS).total * sizeof(float) <= SMEM_MAX &&
           (size_t)make_lay(m, n_win, 6, 1, HD, KV_FLOATS).total * sizeof(float) <= SMEM_MAX ? 1 : 0;
}

template <typename K, typename LAY>
static int launch_tw(K kernel, int grid, const LAY& ly, const fastb200_model& m, const TrArgs& a, void* stream, const char* what) {
    const size_t smem = (size_t)ly.total * sizeof(float);
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return fail((int)e, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    kernel<<<grid, ly.n_warps * 32, smem, as_stream(stream)>>>(m, a, ly);
    return check_launch(what);
}

int tw_fwd_launch(const fastb200_model& m, const TrArgs& a, void* stream) {
    Plan p = make_plan(a.B, 12);
    while (p.slots > 1 && (size_t)make_lay(m, a.n_win, p.S, p.slots, HD, KV_FLOATS).total * sizeof(float) > SMEM_MAX) --p.slots;
    const Lay ly = make_lay(m, a.n_win, p.S, p.slots, HD, KV_FLOATS);
    if (p.S == 1) return launch_tw(tw_fwd_kernel<6, 1>, p.grid, ly, m, a, stream, "transformer_fwd_warp");
    return launch_tw(tw_fwd_kernel<6, 6>, p.grid, ly, m, a, stream, "transformer_fwd_warp6");
}

int tw_bwd_launch(const fastb200_model& m, const TrArgs& a, void* stream) {
    Plan p = make_plan(a.B, 12);
    if (p.S == 1) p.slots = MAX_WARPS;                   // the CTA-wide weight-gradient pass has one tile per thread of 8 warps
    while (p.slots > 1 && (size_t)make_blay(m, a.n_win, p.S, p.slots).total * sizeof(float) > SMEM_MAX) --p.slots;
    const BLay ly = make_blay(m, a.n_win, p.S, p.slots);
    if (p.S == 1) return launch_tw(tw_bwd_kernel<6, 1>, p.grid, ly, m, a, stream, "transformer_bwd_warp");
    return launch_tw(tw_bwd_kernel<6, 6>, p.grid, ly, m, a, stream, "transformer_bwd_warp6");
}

}  // namespace fb
