// Zone embedding head of FAST: window unfold + zone gather + Conv4Layers for every functional area.
//
// Reference path (kidusabe1/Imagined-Speech-Decoding): FAST.forward_head src/fast/models/fast.py:242-252,
// Head.forward :209-210, Conv4Layers.forward :111-119.  B200 design:
//   * one CTA per (trial, window, zone) item; the zone gather and the window slice are folded into the loader
//     (the reference's (B*N, 64, 250) unfold copy and (B*N, 32, Cz, 246) cnn1 intermediate never exist);
//   * cnn1 o cnn2 are one collapsed (F, Cz, 5) filter (no nonlinearity between them in the reference);
//   * cnn3 / cnn4 / GELU / mean run out of shared memory; weights arrive by 1-D bulk async copy (TMA engine);
//   * fp32 FFMA register tiles (8 channels x 4 samples per thread): tensor-core operand rounding (TF32/BF16)
//     misses the 1e-4 parity bar on head features, so the exact path is the CUDA-core one.
#include "common.cuh"

namespace fb {

constexpr int HEAD_THREADS = 256;
constexpr int HEAD_WARPS = HEAD_THREADS / 32;

struct HeadGeom {
    int C, W, Z, F;
    int cz[FASTB200_MAX_ZONES];
    int zoff[FASTB200_MAX_ZONES];
    int pk[FASTB200_MAX_ZONES];          // packed base offset (floats) of zone z
    int chan[FASTB200_MAX_CHANS];
};

static HeadGeom make_geom(const fastb200_model& m) {
    HeadGeom g;
    g.C = m.n_chan; g.W = m.window_len; g.Z = m.n_zones; g.F = m.dim_cnn;
    for (int z = 0; z < m.n_zones; ++z) {
        g.cz[z] = m.zone_size[z]; g.zoff[z] = m.zone_off[z]; g.pk[z] = packed_zone(m, z).weff;
    }
    int n = m.zone_off[m.n_zones - 1] + m.zone_size[m.n_zones - 1];
    for (int i = 0; i < n; ++i) g.chan[i] = m.zone_chan[i];
    return g;
}

// offsets inside one zone's packed block (mirrors packed_zone())
__device__ __forceinline__ int pz_beff(int Cz, int F) { return round_up(Cz * TAPS * F, 4); }
__device__ __forceinline__ int pz_w3(int Cz, int F) { return pz_beff(Cz, F) + round_up(F, 4); }

// ---------------------------------------------------------------------------------------------------------
// weight packing
// ---------------------------------------------------------------------------------------------------------
// `staged`: cnn2.weight of the zone is first copied to shared memory (row f at pitch F*Cz + 1).  The collapse loops read it
// with one f per lane, i.e. at a stride of F*Cz floats: from global memory that is one cache line per lane and a chain of
// dependent-latency loads (these tiny kernels were 15-20 us each, 10 % of a batch-64 training step).
__global__ void head_pack_kernel(fastb200_model m, const float* __restrict__ params, float* __restrict__ packed, int staged) {
    extern __shared__ __align__(16) float pk_s[];
    const int z = blockIdx.x;
    const int F = m.dim_cnn, Cz = m.zone_size[z];
    const PackedZone p = packed_zone(m, z);
    const float* w1 = params + m.off_cnn1_w[z];
    const float* b1 = params + m.off_cnn1_b[z];
    const float* w2 = params + m.off_cnn2_w[z];
    const float* w3 = params + m.off_cnn3_w[z];
    const float* w4 = params + m.off_cnn4_w[z];
    int fp = F * Cz;                                              // pitch of one f-slab of w2
    if (staged) {
        stage_rows_pitched(pk_s, w2, F, F * Cz);
        stage_rows_wait();
        __syncthreads();
        w2 = pk_s;
        fp = F * Cz + 1;
    }
    for (int i = threadIdx.x + blockDim.x * blockIdx.y; i < Cz * TAPS * F; i += blockDim.x * gridDim.y) {
        const int f = i % F, k = (i / F) % TAPS, c = i / (F * TAPS);
        float s = 0.f;
#pragma unroll 8
        for (int g = 0; g < F; ++g) s = fmaf(w2[f * fp + g * Cz + c], __ldg(w1 + g * TAPS + k), s);
        packed[p.weff + i] = s;
    }
    // beff[f] = sum_g b1[g] * sum_c w2[f,g,c]: one warp per f, lane = g (a serial 32 x Cz chain of dependent loads per thread
    // made this tiny kernel 40 us long)
    {
        const int lane = threadIdx.x & 31, wg = (threadIdx.x >> 5) + (blockDim.x >> 5) * blockIdx.y, nw = (blockDim.x >> 5) * gridDim.y;
        for (int f = wg; f < F; f += nw) {
            float v = 0.f;
            if (lane < F) {
                float r = 0.f;
                for (int c = 0; c < Cz; ++c) r += w2[f * fp + lane * Cz + c];
                v = r * b1[lane];
            }
            v = warp_sum(v);
            if (lane == 0) packed[p.beff + f] = v;
        }
    }
    for (int i = threadIdx.x + blockDim.x * blockIdx.y; i < F * TAPS * F; i += blockDim.x * gridDim.y) {
        const int inner = i % F, k = (i / F) % TAPS, outer = i / (F * TAPS);
        // forward layout [ci][k][co]
        packed[p.w3 + i] = w3[(inner * F + outer) * TAPS + k];
        packed[p.w4 + i] = w4[(inner * F + outer) * TAPS + k];
        // input-gradient layout [co][k][ci] with the taps flipped
        packed[p.w3t + i] = w3[(outer * F + inner) * TAPS + (TAPS - 1 - k)];
        packed[p.w4t + i] = w4[(outer * F + inner) * TAPS + (TAPS - 1 - k)];
    }
}

// gradient chain rule: packed-layout grads -> parameter-layout grads
__global__ void head_unpack_grad_kernel(fastb200_model m, const float* __restrict__ params,
                                        const float* __restrict__ gp, float* __restrict__ grads, int staged) {
    extern __shared__ __align__(16) float pk_s[];
    const int z = blockIdx.x;
    const int F = m.dim_cnn, Cz = m.zone_size[z];
    const PackedZone p = packed_zone(m, z);
    const float* w1 = params + m.off_cnn1_w[z];
    const float* b1 = params + m.off_cnn1_b[z];
    const float* w2 = params + m.off_cnn2_w[z];
    const float* gweff = gp + p.weff;   // [c][k][f]
    const float* gbeff = gp + p.beff;   // [f]
    int fp = F * Cz, cp = TAPS * F;                               // pitches of w2's f-slabs and of gweff's c-slabs
    if (staged) {                                                 // see head_pack_kernel
        float* gs = pk_s + F * (F * Cz + 1);
        stage_rows_pitched(pk_s, w2, F, F * Cz);
        stage_rows_pitched(gs, gweff, Cz, TAPS * F);
        stage_rows_wait();
        __syncthreads();
        w2 = pk_s; fp = F * Cz + 1;
        gweff = gs; cp = TAPS * F + 1;
    }
    // d cnn1.weight[g,k] = sum_{f,c} gweff[c,k,f] * w2[f,g,c]
    // (one warp per output, lane = f: the serial F x Cz chains of dependent loads dominated this kernel)
    const int lane = threadIdx.x & 31, wg = (threadIdx.x >> 5) + (blockDim.x >> 5) * blockIdx.y, nw = (blockDim.x >> 5) * gridDim.y;
    for (int i = wg; i < F * TAPS; i += nw) {
        const int g = i / TAPS, k = i % TAPS;
        float s = 0.f;
        if (lane < F)
            for (int c = 0; c < Cz; ++c) s = fmaf(gweff[c * cp + k * F + lane], w2[lane * fp + g * Cz + c], s);
        s = warp_sum(s);
        if (lane == 0) grads[m.off_cnn1_w[z] + i] = s;
    }
    // d cnn1.bias[g] = sum_{f,c} gbeff[f] * w2[f,g,c]
    for (int g = wg; g < F; g += nw) {
        float s = 0.f;
        if (lane < F) {
            float r = 0.f;
            for (int c = 0; c < Cz; ++c) r += w2[lane * fp + g * Cz + c];
            s = gbeff[lane] * r;
        }
        s = warp_sum(s);
        if (lane == 0) grads[m.off_cnn1_b[z] + g] = s;
    }
    // d cnn2.weight[f,g,c] = sum_k gweff[c,k,f] * w1[g,k] + gbeff[f] * b1[g]
    for (int i = threadIdx.x + blockDim.x * blockIdx.y; i < F * F * Cz; i += blockDim.x * gridDim.y) {
        const int c = i % Cz, g = (i / Cz) % F, f = i / (Cz * F);
        float s = gbeff[f] * b1[g];
#pragma unroll
        for (int k = 0; k < TAPS; ++k) s = fmaf(gweff[c * cp + k * F + f], __ldg(w1 + g * TAPS + k), s);
        grads[m.off_cnn2_w[z] + i] = s;
    }
    // cnn3 / cnn4: packed [ci][k][co] -> (co, ci, 1, k)
    for (int i = threadIdx.x + blockDim.x * blockIdx.y; i < F * F * TAPS; i += blockDim.x * gridDim.y) {
        const int k = i % TAPS, ci = (i / TAPS) % F, co = i / (TAPS * F);
        grads[m.off_cnn3_w[z] + i] = gp[p.w3 + (ci * TAPS + k) * F + co];
        grads[m.off_cnn4_w[z] + i] = gp[p.w4 + (ci * TAPS + k) * F + co];
    }
}

// ---------------------------------------------------------------------------------------------------------
// the temporal-convolution phase shared by forward and backward
//   out[co][t] = bias[co] + sum_{ci<nci} sum_{k<5} w[ci][k][co] * in[ci][t + k]        (columns of `in`)
// A warp owns 8 output channels x 128 samples; a lane owns 8 x 4 accumulators.  Per input channel a lane reads
// its 8-sample strip with two LDS.128 and the 5 x 8 weights as warp-broadcast LDS.128: 160 FFMA per 12 LDS.
// ---------------------------------------------------------------------------------------------------------
enum { EPI_STORE = 0, EPI_GELU_MEAN = 1 };

template <int F, int EPI>
__device__ __forceinline__ void conv_phase(const float* __restrict__ in, int nci, const float* __restrict__ w,
                                           const float* __restrict__ bias, float* __restrict__ out_s,
                                           float* __restrict__ save_g, float* __restrict__ red,
                                           int RS, int TS, int T_out, int n_tt) {
    constexpr int NCG = F / 8;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int task = warp; task < NCG * n_tt; task += HEAD_WARPS) {
        const int cg = task % NCG, tt = task / NCG;
        const int co0 = cg * 8, t0 = tt * 128 + lane * 4;
        float acc[8][4];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const float b = bias ? bias[co0 + i] : 0.f;
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[i][j] = b;
        }
        const float* ip = in + t0;
        const float* wp = w + co0;
#pragma unroll 2
        for (int ci = 0; ci < nci; ++ci) {
            const float4 a0 = *reinterpret_cast<const float4*>(ip + ci * RS);
            const float4 a1 = *reinterpret_cast<const float4*>(ip + ci * RS + 4);
            const float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
#pragma unroll
            for (int k = 0; k < TAPS; ++k) {
                const float4 w0 = *reinterpret_cast<const float4*>(wp + (ci * TAPS + k) * F);
                const float4 w1 = *reinterpret_cast<const float4*>(wp + (ci * TAPS + k) * F + 4);
                const float wv[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
#pragma unroll
                for (int i = 0; i < 8; ++i)
#pragma unroll
                    for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(wv[i], a[j + k], acc[i][j]);
            }
        }
        // mask samples past the valid length (they double as the zero padding of the next conv)
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j)
                if (t0 + j >= T_out) acc[i][j] = 0.f;
        if (save_g != nullptr && t0 < TS) {
#pragma unroll
            for (int i = 0; i < 8; ++i)
                *reinterpret_cast<float4*>(save_g + (co0 + i) * TS + t0) =
                    make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
        }
        if (EPI == EPI_STORE) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                float* o = out_s + (co0 + i) * RS + 2 + t0;
                *reinterpret_cast<float2*>(o) = make_float2(acc[i][0], acc[i][1]);
                *reinterpret_cast<float2*>(o + 2) = make_float2(acc[i][2], acc[i][3]);
            }
        } else {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                float s = 0.f;
#pragma unroll
                for (int j = 0; j < 4; ++j) s += (t0 + j < T_out) ? gelu_f(acc[i][j]) : 0.f;
                s = warp_sum(s);
                if (lane == 0) red[tt * F + co0 + i] = s;
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------------------
// forward
// ---------------------------------------------------------------------------------------------------------
struct HeadFwdArgs {
    HeadGeom g;
    const float* packed;
    const float* x;
    float* feat;
    float* saved;      // nullptr for inference
    int B, T, n_win, step;
    int RS, TS, T_out, n_tt;
    int off_xs, off_weff;    // float offsets inside dynamic smem (may alias buffer B)
};

template <int F>
__global__ void __launch_bounds__(HEAD_THREADS, 2) head_fwd_kernel(const __grid_constant__ HeadFwdArgs a) {
    extern __shared__ __align__(128) float smem[];
    const int tid = threadIdx.x;
    const int RS = a.RS;
    float* bufA = smem;
    float* bufB = bufA + F * RS;
    float* w3s = bufB + F * RS;
    float* w4s = w3s + F * TAPS * F;
    float* beff = w4s + F * TAPS * F;                 // round_up(F,4)
    float* red = beff + round_up(F, 4);               // n_tt * F
    uint64_t* bar = reinterpret_cast<uint64_t*>(red + round_up(a.n_tt * F, 4));
    float* xs = smem + a.off_xs;
    float* weff = smem + a.off_weff;

    const int item = blockIdx.x;
    const int z = item % a.g.Z, bn = item / a.g.Z;
    const int b = bn / a.n_win, n = bn % a.n_win;
    const int Cz = a.g.cz[z];
    const float* pz = a.packed + a.g.pk[z];

    if (tid == 0) { mbar_init(bar, 1); fence_mbar_init(); }
    __syncthreads();
    if (tid == 0) {
        const uint32_t b_weff = (uint32_t)round_up(Cz * TAPS * F, 4) * 4u, b_beff = (uint32_t)round_up(F, 4) * 4u;
        const uint32_t b_w = (uint32_t)(F * TAPS * F) * 4u;
        mbar_expect_tx(bar, b_weff + b_beff + 2 * b_w);
        bulk_g2s(weff, pz, b_weff, bar);
        bulk_g2s(beff, pz + pz_beff(Cz, F), b_beff, bar);
        bulk_g2s(w3s, pz + pz_w3(Cz, F), b_w, bar);
        bulk_g2s(w4s, pz + pz_w3(Cz, F) + F * TAPS * F, b_w, bar);
    }
    // zero pad columns of A; gather this zone's electrodes for this window (the fused unfold + index)
    for (int i = tid; i < F * 4; i += HEAD_THREADS) {
        const int r = i >> 2, c = i & 3;
        bufA[r * RS + (c < 2 ? c : RS - 4 + c)] = 0.f;
    }
    {
        const int* chan = a.g.chan + a.g.zoff[z];
        const float* xb = a.x + (size_t)b * a.g.C * a.T + (size_t)n * a.step;
        for (int i = tid; i < Cz * RS; i += HEAD_THREADS) {
            const int c = i / RS, t = i - c * RS;
            xs[i] = (t < a.g.W) ? __ldg(xb + (size_t)chan[c] * a.T + t) : 0.f;
        }
    }
    mbar_wait(bar, 0);
    __syncthreads();

    const size_t item_sz = (size_t)F * a.TS;
    float* sv = a.saved ? a.saved + (size_t)item * 3 * item_sz : nullptr;

    // P1: collapsed cnn1 o cnn2 (valid conv over the gathered electrodes) -> A
    conv_phase<F, EPI_STORE>(xs, Cz, weff, beff, bufA, sv, nullptr, RS, a.TS, a.T_out, a.n_tt);
    __syncthreads();
    // P2: cnn3 (zero padded at the window edges) -> B
    for (int i = tid; i < F * 4; i += HEAD_THREADS) {
        const int r = i >> 2, c = i & 3;
        bufB[r * RS + (c < 2 ? c : RS - 4 + c)] = 0.f;
    }
    conv_phase<F, EPI_STORE>(bufA, F, w3s, nullptr, bufB, sv ? sv + item_sz : nullptr, nullptr, RS, a.TS, a.T_out, a.n_tt);
    __syncthreads();
    // P3: cnn4 -> GELU -> mean over time
    conv_phase<F, EPI_GELU_MEAN>(bufB, F, w4s, nullptr, nullptr, sv ? sv + 2 * item_sz : nullptr, red, RS, a.TS, a.T_out, a.n_tt);
    __syncthreads();
    if (tid < F) {
        float s = 0.f;
        for (int tt = 0; tt < a.n_tt; ++tt) s += red[tt * F + tid];
        a.feat[(size_t)item * F + tid] = s / (float)a.T_out;
    }
}

// ---------------------------------------------------------------------------------------------------------
// backward
// ---------------------------------------------------------------------------------------------------------
// weight-gradient phase:  acc[i][k] += sum_t dY[co0+i][t] * X[lane][t + k - pad]
// dY rows are stored at column 2+t; X rows at column (t' + 4) for conv inputs / (t' + 2) for the raw electrodes,
// so that for a 4-column step both operands are 16-byte aligned: X columns needed are col4 + j + k.
template <int F>
__device__ __forceinline__ void wgrad_phase(const float* __restrict__ dY, const float* __restrict__ X, int nci,
                                            float (&acc)[4][TAPS], int RS, int T_out) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int co0 = warp * 4;
    if (co0 >= F || lane >= nci) return;
    const int col_end = 2 + T_out;              // first column with no gradient
    const float* xr = X + lane * RS;
#pragma unroll 2
    for (int col4 = 0; col4 < col_end; col4 += 4) {
        const float4 x0 = *reinterpret_cast<const float4*>(xr + col4);
        const float4 x1 = *reinterpret_cast<const float4*>(xr + col4 + 4);
        const float xa[8] = {x0.x, x0.y, x0.z, x0.w, x1.x, x1.y, x1.z, x1.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const float4 d = *reinterpret_cast<const float4*>(dY + (co0 + i) * RS + col4);
            const float dv[4] = {d.x, d.y, d.z, d.w};
#pragma unroll
            for (int k = 0; k < TAPS; ++k)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][k] = fmaf(dv[j], xa[j + k], acc[i][k]);
        }
    }
}

struct HeadBwdArgs {
    HeadGeom g;
    const float* packed;
    const float* x;
    const float* saved;
    const float* dfeat;
    float* gpacked;
    float* dx;          // nullptr unless the input gradient is wanted
    int B, T, n_win, step;
    int RS, TS, T_out, n_tt;
    int items_per_cta, n_items;
};

template <int F>
__global__ void __launch_bounds__(HEAD_THREADS, 1) head_bwd_kernel(const __grid_constant__ HeadBwdArgs a) {
    extern __shared__ __align__(128) float smem[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int RS = a.RS, T_out = a.T_out;
    float* bufA = smem;
    float* bufB = bufA + F * RS;
    float* bufC = bufB + 32 * RS;                    // B holds up to 32 gathered electrodes
    float* w3t = bufC + F * RS;
    float* w4t = w3t + F * TAPS * F;
    float* dfs = w4t + F * TAPS * F;                 // dfeat / T_out for this item
    uint64_t* bar = reinterpret_cast<uint64_t*>(dfs + round_up(F, 4));

    const int BN = a.B * a.n_win;
    const int first = blockIdx.x * a.items_per_cta;
    const int last = min(first + a.items_per_cta, a.n_items);
    if (first >= last) return;

    float acc4[4][TAPS], acc3[4][TAPS], accE[4][TAPS], accb[4];
    int cur_z = -1;
    uint32_t parity = 0;
    if (tid == 0) { mbar_init(bar, 1); fence_mbar_init(); }
    const int co0 = warp * 4;
    const float inv_T = 1.0f / (float)T_out;
    const size_t item_sz = (size_t)F * a.TS;

    auto flush = [&](int z) {
        const int Cz = a.g.cz[z];
        float* gz = a.gpacked + a.g.pk[z];
        if (co0 < F) {
            if (lane < F) {
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int k = 0; k < TAPS; ++k) {
                        atomicAdd(gz + pz_w3(Cz, F) + (lane * TAPS + k) * F + co0 + i, acc3[i][k]);
                        atomicAdd(gz + pz_w3(Cz, F) + F * TAPS * F + (lane * TAPS + k) * F + co0 + i, acc4[i][k]);
                    }
            }
            if (lane < Cz) {
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int k = 0; k < TAPS; ++k) atomicAdd(gz + (lane * TAPS + k) * F + co0 + i, accE[i][k]);
            }
            if (lane == 0) {
#pragma unroll
                for (int i = 0; i < 4; ++i) atomicAdd(gz + pz_beff(Cz, F) + co0 + i, accb[i]);
            }
        }
    };

    for (int idx = first; idx < last; ++idx) {
        const int z = idx / BN, bn = idx - z * BN;
        const int b = bn / a.n_win, n = bn - b * a.n_win;
        const int Cz = a.g.cz[z];
        const size_t item = (size_t)bn * a.g.Z + z;                   // forward item index
        if (z != cur_z) {
            if (cur_z >= 0) flush(cur_z);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                accb[i] = 0.f;
#pragma unroll
                for (int k = 0; k < TAPS; ++k) acc4[i][k] = acc3[i][k] = accE[i][k] = 0.f;
            }
            __syncthreads();                                          // everyone done with the old weights
            if (tid == 0) {
                const float* pz = a.packed + a.g.pk[z];
                const uint32_t b_w = (uint32_t)(F * TAPS * F) * 4u;
                fence_proxy_async();
                mbar_expect_tx(bar, 2 * b_w);
                bulk_g2s(w3t, pz + pz_w3(Cz, F) + 2 * F * TAPS * F, b_w, bar);
                bulk_g2s(w4t, pz + pz_w3(Cz, F) + 3 * F * TAPS * F, b_w, bar);
            }
            mbar_wait(bar, parity);
            parity ^= 1;
            cur_z = z;
        }
        const float* sv = a.saved + item * 3 * item_sz;
        if (tid < F) dfs[tid] = a.dfeat[item * F + tid] * inv_T;
        __syncthreads();
        // Q0: A <- dL/dy4 (column 2+t), B <- y3 (column 4+t), C pad columns <- 0
        for (int i = tid; i < F * RS; i += HEAD_THREADS) {
            const int f = i / RS, col = i - f * RS;
            const int t = col - 2, t2 = col - 4;
            float va = 0.f, vb = 0.f;
            if (t >= 0 && t < T_out) va = dfs[f] * gelu_grad_f(__ldg(sv + 2 * item_sz + f * a.TS + t));
            if (t2 >= 0 && t2 < T_out) vb = __ldg(sv + item_sz + f * a.TS + t2);
            bufA[i] = va;
            bufB[i] = vb;
            if (col < 2 || col >= 2 + a.n_tt * 128) bufC[i] = 0.f;
        }
        __syncthreads();
        // Q1: dW4 += dy4 (x) y3 ;  C <- dL/dy3 = conv(dy4, flipped cnn4^T)
        wgrad_phase<F>(bufA, bufB, F, acc4, RS, T_out);
        conv_phase<F, EPI_STORE>(bufA, F, w4t, nullptr, bufC, nullptr, nullptr, RS, a.TS, T_out, a.n_tt);
        __syncthreads();
        // Q2: B <- y2 (column 4+t)
        for (int i = tid; i < F * RS; i += HEAD_THREADS) {
            const int f = i / RS, t2 = i - f * RS - 4;
            bufB[i] = (t2 >= 0 && t2 < T_out) ? __ldg(sv + f * a.TS + t2) : 0.f;
        }
        __syncthreads();
        // Q3: dW3 += dy3 (x) y2 ;  A <- dL/dy2 = conv(dy3, flipped cnn3^T)
        wgrad_phase<F>(bufC, bufB, F, acc3, RS, T_out);
        conv_phase<F, EPI_STORE>(bufC, F, w3t, nullptr, bufA, nullptr, nullptr, RS, a.TS, T_out, a.n_tt);
        __syncthreads();
        // Q4: B <- gathered electrodes of this window (column 2+t')
        {
            const int* chan = a.g.chan + a.g.zoff[z];
            const float* xb = a.x + (size_t)b * a.g.C * a.T + (size_t)n * a.step;
            for (int i = tid; i < Cz * RS; i += HEAD_THREADS) {
                const int c = i / RS, t2 = i - c * RS - 2;
                bufB[i] = (t2 >= 0 && t2 < a.g.W) ? __ldg(xb + (size_t)chan[c] * a.T + t2) : 0.f;
            }
        }
        __syncthreads();
        // Q5: dWeff += dy2 (x) x ; dbeff += sum_t dy2 ; optional dx
        wgrad_phase<F>(bufA, bufB, Cz, accE, RS, T_out);
        if (co0 < F) {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                float s = 0.f;
                for (int col = 2 + lane; col < 2 + T_out; col += 32) s += bufA[(co0 + i) * RS + col];
                s = warp_sum(s);
                accb[i] += s;
            }
        }
        if (a.dx != nullptr) {
            const int* chan = a.g.chan + a.g.zoff[z];
            const float* weff = a.packed + a.g.pk[z];
            float* dxb = a.dx + (size_t)b * a.g.C * a.T + (size_t)n * a.step;
            for (int i = tid; i < Cz * a.g.W; i += HEAD_THREADS) {
                const int c = i / a.g.W, tp = i - c * a.g.W;
                float s = 0.f;
#pragma unroll
                for (int k = 0; k < TAPS; ++k) {
                    const int t = tp - k;
                    if (t >= 0 && t < T_out) {
                        const float* wk = weff + (c * TAPS + k) * F;
                        for (int f = 0; f < F; ++f) s = fmaf(__ldg(wk + f), bufA[f * RS + 2 + t], s);
                    }
                }
                atomicAdd(dxb + (size_t)chan[c] * a.T + tp, s);
            }
        }
        __syncthreads();
    }
    flush(cur_z);
}


// ---------------------------------------------------------------------------------------------------------
// backward, tensor-core split: everything except the cnn3 / cnn4 weight gradients (those are head_wgrad_tc.cu)
// ---------------------------------------------------------------------------------------------------------
// Same chain as head_bwd_kernel, restructured so that no global-memory latency is exposed inside an item:
//   * y4 of the NEXT item arrives by one bulk async copy (TMA engine) into its own buffer while this item computes;
//   * the zone's electrodes of this window are gathered with 4-byte cp.async (the trial rows are only 4-byte aligned)
//     issued at the top of the item and awaited just before the embed weight-gradient phase;
//   * dL/dy4 and dL/dy3 are written to `dys` ([F][TS] each) for the tensor-core weight-gradient kernel.
struct HeadDgradArgs {
    HeadGeom g;
    const float* packed;
    const float* x;
    const float* saved;
    const float* dfeat;
    float* gpacked;
    float* dx;
    float* dys;
    int embed_tc;       // 1: dL/dy2 is written to dys[2] and the collapsed-embed gradients are left to head_wgrad_tc.cu
    int B, T, n_win, step;
    int RS, TS, T_out, n_tt;
    int items_per_cta, n_items;
};

template <int F>
__global__ void __launch_bounds__(HEAD_THREADS, 1) head_dgrad_kernel(const __grid_constant__ HeadDgradArgs a) {
    extern __shared__ __align__(128) float smem[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int RS = a.RS, T_out = a.T_out, TS = a.TS;
    float* bufA = smem;
    float* bufB = bufA + F * RS;                     // gathered electrodes (up to 32 rows)
    float* bufC = bufB + 32 * RS;
    float* y4s = bufC + F * RS;                      // [F][TS] y4 of the current item (bulk copy target)
    float* w3t = y4s + F * TS;
    float* w4t = w3t + F * TAPS * F;
    float* dfs = w4t + F * TAPS * F;
    uint64_t* bar = reinterpret_cast<uint64_t*>(dfs + round_up(F, 4));     // [0] weights, [1] y4

    const int BN = a.B * a.n_win;
    const int first = blockIdx.x * a.items_per_cta;
    const int last = min(first + a.items_per_cta, a.n_items);
    if (first >= last) return;

    float accE[4][TAPS], accb[4];
    int cur_z = -1;
    uint32_t parity_w = 0, parity_y = 0;
    const int co0 = warp * 4;
    const float inv_T = 1.0f / (float)T_out;
    const size_t item_sz = (size_t)F * TS;
    const uint32_t y_bytes = (uint32_t)(item_sz * sizeof(float));

    auto fwd_item = [&](int idx) -> size_t {
        const int z = idx / BN, bn = idx - z * BN;
        return (size_t)bn * a.g.Z + z;
    };
    if (tid == 0) {
        mbar_init(&bar[0], 1);
        mbar_init(&bar[1], 1);
        fence_mbar_init();
        mbar_expect_tx(&bar[1], y_bytes);
        bulk_g2s(y4s, a.saved + fwd_item(first) * 3 * item_sz + 2 * item_sz, y_bytes, &bar[1]);
    }
    // pad columns of the conv operand buffers are zero for the whole kernel (the phases only write columns [2, 2 + 128 n_tt))
    for (int i = tid; i < F * RS; i += HEAD_THREADS) { bufA[i] = 0.f; bufC[i] = 0.f; }
    __syncthreads();

    auto flush = [&](int z) {
        const int Cz = a.g.cz[z];
        float* gz = a.gpacked + a.g.pk[z];
        if (co0 < F) {
            if (lane < Cz) {
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int k = 0; k < TAPS; ++k) atomicAdd(gz + (lane * TAPS + k) * F + co0 + i, accE[i][k]);
            }
            if (lane == 0) {
#pragma unroll
                for (int i = 0; i < 4; ++i) atomicAdd(gz + pz_beff(Cz, F) + co0 + i, accb[i]);
            }
        }
    };

    for (int idx = first; idx < last; ++idx) {
        const int z = idx / BN, bn = idx - z * BN;
        const int b = bn / a.n_win, n = bn - b * a.n_win;
        const int Cz = a.g.cz[z];
        const size_t item = (size_t)bn * a.g.Z + z;
        if (z != cur_z) {
            if (cur_z >= 0 && !a.embed_tc) flush(cur_z);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                accb[i] = 0.f;
#pragma unroll
                for (int k = 0; k < TAPS; ++k) accE[i][k] = 0.f;
            }
            __syncthreads();
            if (tid == 0) {
                const float* pz = a.packed + a.g.pk[z];
                const uint32_t b_w = (uint32_t)(F * TAPS * F) * 4u;
                fence_proxy_async();
                mbar_expect_tx(&bar[0], 2 * b_w);
                bulk_g2s(w3t, pz + pz_w3(Cz, F) + 2 * F * TAPS * F, b_w, &bar[0]);
                bulk_g2s(w4t, pz + pz_w3(Cz, F) + 3 * F * TAPS * F, b_w, &bar[0]);
            }
            mbar_wait(&bar[0], parity_w);
            parity_w ^= 1;
            cur_z = z;
        }
        // electrodes of this window -> B (column 2 + t'), asynchronously; awaited before Q5
        if (!a.embed_tc) {
            const int* chan = a.g.chan + a.g.zoff[z];
            const float* xb = a.x + (size_t)b * a.g.C * a.T + (size_t)n * a.step;
            const uint32_t dst0 = smem_u32(bufB);
            for (int i = tid; i < Cz * RS; i += HEAD_THREADS) {
                const int c = i / RS, t2 = i - c * RS - 2;
                const bool ok = t2 >= 0 && t2 < a.g.W;
                const float* src = ok ? xb + (size_t)chan[c] * a.T + t2 : a.x;
                asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(dst0 + 4u * i), "l"(src), "r"(ok ? 4u : 0u) : "memory");
            }
            asm volatile("cp.async.commit_group;" ::: "memory");
        }
        float* dy = a.dys + item * 3 * item_sz;
        if (tid < F) dfs[tid] = __ldg(a.dfeat + item * F + tid) * inv_T;
        mbar_wait(&bar[1], parity_y);
        parity_y ^= 1;
        __syncthreads();
        // Q0: A <- dL/dy4 = dfeat / T * gelu'(y4) (column 2 + t), also written to dys[0]
        for (int i = tid; i < F * (TS / 4); i += HEAD_THREADS) {
            const int f = i / (TS / 4), t = (i - f * (TS / 4)) * 4;
            const float4 y = *reinterpret_cast<const float4*>(y4s + f * TS + t);
            const float d = dfs[f];
            float4 g;
            g.x = t + 0 < T_out ? d * gelu_grad_f(y.x) : 0.f;
            g.y = t + 1 < T_out ? d * gelu_grad_f(y.y) : 0.f;
            g.z = t + 2 < T_out ? d * gelu_grad_f(y.z) : 0.f;
            g.w = t + 3 < T_out ? d * gelu_grad_f(y.w) : 0.f;
            *reinterpret_cast<float4*>(dy + f * TS + t) = g;
            float* o = bufA + f * RS + 2 + t;
            *reinterpret_cast<float2*>(o) = make_float2(g.x, g.y);
            *reinterpret_cast<float2*>(o + 2) = make_float2(g.z, g.w);
        }
        // columns [2 + TS, 2 + 128 n_tt) of A are rewritten by Q3 of the previous item with masked (zero) values
        __syncthreads();
        if (tid == 0 && idx + 1 < last) {              // y4 buffer is free: fetch the next item's
            fence_proxy_async();
            mbar_expect_tx(&bar[1], y_bytes);
            bulk_g2s(y4s, a.saved + fwd_item(idx + 1) * 3 * item_sz + 2 * item_sz, y_bytes, &bar[1]);
        }
        // Q1: C <- dL/dy3 = conv(dy4, flipped cnn4^T), also written to dys[1]
        conv_phase<F, EPI_STORE>(bufA, F, w4t, nullptr, bufC, dy + item_sz, nullptr, RS, TS, T_out, a.n_tt);
        __syncthreads();
        // Q3: A <- dL/dy2 = conv(dy3, flipped cnn3^T)
        conv_phase<F, EPI_STORE>(bufC, F, w3t, nullptr, bufA, a.embed_tc ? dy + 2 * item_sz : nullptr, nullptr, RS, TS, T_out, a.n_tt);
        if (a.embed_tc) { __syncthreads(); continue; }
        asm volatile("cp.async.wait_group 0;" ::: "memory");
        __syncthreads();
        // Q5: dWeff += dy2 (x) x ; dbeff += sum_t dy2 ; optional dx
        wgrad_phase<F>(bufA, bufB, Cz, accE, RS, T_out);
        if (co0 < F) {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                float s = 0.f;
                for (int col = 2 + lane; col < 2 + T_out; col += 32) s += bufA[(co0 + i) * RS + col];
                s = warp_sum(s);
                accb[i] += s;
            }
        }
        if (a.dx != nullptr) {
            const int* chan = a.g.chan + a.g.zoff[z];
            const float* weff = a.packed + a.g.pk[z];
            float* dxb = a.dx + (size_t)b * a.g.C * a.T + (size_t)n * a.step;
            for (int i = tid; i < Cz * a.g.W; i += HEAD_THREADS) {
                const int c = i / a.g.W, tp = i - c * a.g.W;
                float s = 0.f;
#pragma unroll
                for (int k = 0; k < TAPS; ++k) {
                    const int t = tp - k;
                    if (t >= 0 && t < T_out) {
                        const float* wk = weff + (c * TAPS + k) * F;
                        for (int f = 0; f < F; ++f) s = fmaf(__ldg(wk + f), bufA[f * RS + 2 + t], s);
                    }
                }
                atomicAdd(dxb + (size_t)chan[c] * a.T + tp, s);
            }
        }
        __syncthreads();
    }
    if (!a.embed_tc) flush(cur_z);
}

// ---------------------------------------------------------------------------------------------------------
// stand-alone index kernel (parity / diagnostics; the fused path never materialises this tensor)
// ---------------------------------------------------------------------------------------------------------
__global__ void window_gather_kernel(const float* __restrict__ x, const int32_t* __restrict__ chan,
                                     float* __restrict__ out, int C, int T, int W, int S, int n_win, int n_sel,
                                     size_t total) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const int t = (int)(i % W);
        const size_t r = i / W;
        const int j = (int)(r % n_sel);
        const size_t bn = r / n_sel;
        const int n = (int)(bn % n_win);
        const size_t b = bn / n_win;
        out[i] = __ldg(x + (b * C + chan[j]) * T + (size_t)n * S + t);
    }
}

// ---------------------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------------------
static int validate_model(const fastb200_model* m) {
    FB_REQUIRE(m != nullptr, "model descriptor is NULL");
    FB_REQUIRE(m->dim_cnn == 8 || m->dim_cnn == 16 || m->dim_cnn == 32, "dim_cnn=%d unsupported (8, 16 or 32)", m->dim_cnn);
    FB_REQUIRE(m->n_zones >= 1 && m->n_zones <= FASTB200_MAX_ZONES, "n_zones=%d out of range", m->n_zones);
    FB_REQUIRE(m->window_len >= 8, "window_len=%d too short", m->window_len);
    for (int z = 0; z < m->n_zones; ++z)
        FB_REQUIRE(m->zone_size[z] >= 1 && m->zone_size[z] <= 32, "zone %d has %d electrodes (1..32 supported)", z, m->zone_size[z]);
    FB_REQUIRE(m->zone_off[m->n_zones - 1] + m->zone_size[m->n_zones - 1] <= FASTB200_MAX_CHANS, "too many zone electrodes");
    return 0;
}

static int max_cz(const fastb200_model& m) {
    int v = 0;
    for (int z = 0; z < m.n_zones; ++z) v = m.zone_size[z] > v ? m.zone_size[z] : v;
    return v;
}

template <typename K>
static int set_smem(K kernel, size_t bytes) {
    if (bytes > 227 * 1024) return fail(-2, "window_len too long for the shared-memory resident head kernel (%zu B)", bytes);
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
    if (e != cudaSuccess) return fail((int)e, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    return 0;
}

}  // namespace fb

using namespace fb;

// Shared-memory footprints of the CUDA-core kernels (the window is resident in shared memory): the same expressions the
// launchers below use, so the support predicate and the launch can never disagree.
static size_t head_fwd_smem_floats(const fastb200_model& m) {
    const int F = m.dim_cnn, RS = head_row_stride(m.window_len), n_tt = head_n_tt(m.window_len), mcz = max_cz(m);
    int fl = 2 * F * RS + 2 * F * TAPS * F + round_up(F, 4) + round_up(n_tt * F, 4) + 4;
    const int need_alias = mcz * RS + round_up(mcz * TAPS * F, 4);
    if (need_alias > F * RS) fl += need_alias;
    return (size_t)fl;
}
static size_t head_bwd_smem_floats(const fastb200_model& m) {
    const int F = m.dim_cnn, RS = head_n_tt(m.window_len) * 128 + 36;
    return (size_t)((2 * F + 32) * RS + 2 * F * TAPS * F + round_up(F, 4) + 4);
}

// 1 when the fused Conv4Layers kernels cover this head shape (dim_cnn 8 / 16 / 32, zones of 1..32 electrodes, a window that
// fits the shared-memory resident CUDA-core kernels -- about 1,600 samples forward, 550 backward at dim_cnn 32), else 0.  The
// Python drop-in runs shapes outside it (e.g. Conv4Layers(dim=64), reference tests/test_model.py:278-284) as stock torch ops.
extern "C" int fastb200_head_supported(const fastb200_model* m, int need_backward) {
    if (validate_model(m)) return 0;
    if (head_fwd_smem_floats(*m) * sizeof(float) > 227 * 1024) return 0;
    if (need_backward && head_bwd_smem_floats(*m) * sizeof(float) > 227 * 1024) return 0;
    return 1;
}

extern "C" int64_t fastb200_head_packed_floats(const fastb200_model* m) {
    if (validate_model(m)) return -1;
    return packed_zone(*m, m->n_zones - 1).end;
}

extern "C" int64_t fastb200_head_saved_floats(const fastb200_model* m, int B, int n_win) {
    if (validate_model(m)) return -1;
    return (int64_t)B * n_win * m->n_zones * 3 * m->dim_cnn * head_saved_stride(m->window_len);
}

extern "C" int fastb200_head_pack(const fastb200_model* m, const float* params, float* packed, void* stream) {
    if (int r = validate_model(m)) return r;
    int czmax = 0;
    for (int z = 0; z < m->n_zones; ++z) czmax = czmax > m->zone_size[z] ? czmax : m->zone_size[z];
    size_t smem = (size_t)m->dim_cnn * (m->dim_cnn * czmax + 1) * sizeof(float);
    int staged = smem <= 200 * 1024;
    if (staged && smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(head_pack_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return fail((int)e, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    }
    head_pack_kernel<<<dim3(m->n_zones, 8), 256, staged ? smem : 0, as_stream(stream)>>>(*m, params, packed, staged);
    return check_launch("head_pack");
}

extern "C" int fastb200_head_unpack_grad(const fastb200_model* m, const float* params, const float* gpacked,
                                         float* grads, void* stream) {
    if (int r = validate_model(m)) return r;
    int czmax = 0;
    for (int z = 0; z < m->n_zones; ++z) czmax = czmax > m->zone_size[z] ? czmax : m->zone_size[z];
    size_t smem = ((size_t)m->dim_cnn * (m->dim_cnn * czmax + 1) + (size_t)czmax * (TAPS * m->dim_cnn + 1)) * sizeof(float);
    int staged = smem <= 200 * 1024;
    if (staged && smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(head_unpack_grad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return fail((int)e, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    }
    head_unpack_grad_kernel<<<dim3(m->n_zones, 8), 256, staged ? smem : 0, as_stream(stream)>>>(*m, params, gpacked, grads, staged);
    return check_launch("head_unpack_grad");
}

extern "C" int fastb200_head_fwd(const fastb200_model* m, const float* packed, const float* x, int B, int T,
                                 int n_win, int step, float* feat, float* saved, void* stream) {
    if (int r = validate_model(m)) return r;
    FB_REQUIRE(B >= 0 && n_win >= 1 && step >= 1, "bad B/n_win/step");
    FB_REQUIRE((n_win - 1) * (int64_t)step + m->window_len <= T, "windows run past the trial: n_win=%d step=%d W=%d T=%d", n_win, step, m->window_len, T);
    if (B == 0) return 0;
    const int F = m->dim_cnn, W = m->window_len;
    HeadFwdArgs a;
    a.g = make_geom(*m);
    a.packed = packed; a.x = x; a.feat = feat; a.saved = saved;
    a.B = B; a.T = T; a.n_win = n_win; a.step = step;
    a.RS = head_row_stride(W); a.TS = head_saved_stride(W); a.T_out = head_t_out(W); a.n_tt = head_n_tt(W);
    const int mcz = max_cz(*m);
    int fl = 2 * F * a.RS + 2 * F * TAPS * F + round_up(F, 4) + round_up(a.n_tt * F, 4) + 4;
    const int need_alias = mcz * a.RS + round_up(mcz * TAPS * F, 4);
    if (need_alias <= F * a.RS) {            // electrodes + collapsed filter live in buffer B until P2 overwrites it
        a.off_xs = F * a.RS;
        a.off_weff = a.off_xs + mcz * a.RS;
    } else {
        a.off_xs = fl;
        a.off_weff = a.off_xs + mcz * a.RS;
        fl += need_alias;
    }
    const size_t smem = (size_t)fl * sizeof(float);
    const int64_t items = (int64_t)B * n_win * m->n_zones;
    FB_REQUIRE(items < (1ll << 31), "too many items for one launch");
    int r = 0;
#define LAUNCH_FWD(FF)                                                                     \
    if ((r = set_smem(head_fwd_kernel<FF>, smem))) return r;                                \
    head_fwd_kernel<FF><<<(unsigned)items, HEAD_THREADS, smem, as_stream(stream)>>>(a);
    if (F == 32) { LAUNCH_FWD(32) } else if (F == 16) { LAUNCH_FWD(16) } else { LAUNCH_FWD(8) }
#undef LAUNCH_FWD
    return check_launch("head_fwd");
}

extern "C" int fastb200_head_bwd(const fastb200_model* m, const float* packed, const float* x, const float* saved,
                                 const float* dfeat, int B, int T, int n_win, int step, float* gpacked, float* dx,
                                 void* stream) {
    if (int r = validate_model(m)) return r;
    FB_REQUIRE(B >= 0 && n_win >= 1 && step >= 1, "bad B/n_win/step");
    FB_REQUIRE(saved != nullptr, "head_bwd needs the activations saved by head_fwd");
    if (B == 0) return 0;
    const int F = m->dim_cnn, W = m->window_len;
    HeadBwdArgs a;
    a.g = make_geom(*m);
    a.packed = packed; a.x = x; a.saved = saved; a.dfeat = dfeat; a.gpacked = gpacked; a.dx = dx;
    a.B = B; a.T = T; a.n_win = n_win; a.step = step;
    a.TS = head_saved_stride(W); a.T_out = head_t_out(W); a.n_tt = head_n_tt(W);
    a.RS = a.n_tt * 128 + 36;                       // room for the 8-column operand strips; RS % 32 == 4
    const int64_t items = (int64_t)B * n_win * m->n_zones;
    FB_REQUIRE(items < (1ll << 31), "too many items for one launch");
    a.n_items = (int)items;
    int grid = (int)(items < NUM_SMS ? items : NUM_SMS);
    a.items_per_cta = ceil_div((int)items, grid);
    grid = ceil_div((int)items, a.items_per_cta);
    const size_t smem = (size_t)((2 * F + 32) * a.RS + 2 * F * TAPS * F + round_up(F, 4) + 4) * sizeof(float);
    int r = 0;
#define LAUNCH_BWD(FF)                                                                     \
    if ((r = set_smem(head_bwd_kernel<FF>, smem))) return r;                                \
    head_bwd_kernel<FF><<<grid, HEAD_THREADS, smem, as_stream(stream)>>>(a);
    if (F == 32) { LAUNCH_BWD(32) } else if (F == 16) { LAUNCH_BWD(16) } else { LAUNCH_BWD(8) }
#undef LAUNCH_BWD
    return check_launch("head_bwd");
}

int fb::head_dgrad_launch(const fastb200_model* m, const float* packed, const float* x, const float* saved,
                          const float* dfeat, int B, int T, int n_win, int step, float* gpacked, float* dx, float* dys,
                          int embed_tc, void* stream) {
    if (int r = validate_model(m)) return r;
    FB_REQUIRE(B >= 0 && n_win >= 1 && step >= 1, "bad B/n_win/step");
    FB_REQUIRE(saved != nullptr && dys != nullptr, "head_dgrad needs the saved activations and the dys workspace");
    if (B == 0) return 0;
    const int F = m->dim_cnn, W = m->window_len;
    HeadDgradArgs a;
    a.g = make_geom(*m);
    a.packed = packed; a.x = x; a.saved = saved; a.dfeat = dfeat; a.gpacked = gpacked; a.dx = dx; a.dys = dys;
    a.embed_tc = (embed_tc && dx == nullptr) ? 1 : 0;
    a.B = B; a.T = T; a.n_win = n_win; a.step = step;
    a.TS = head_saved_stride(W); a.T_out = head_t_out(W); a.n_tt = head_n_tt(W);
    a.RS = a.n_tt * 128 + 36;
    const int64_t items = (int64_t)B * n_win * m->n_zones;
    FB_REQUIRE(items < (1ll << 31), "too many items for one launch");
    a.n_items = (int)items;
    int grid = (int)(items < NUM_SMS ? items : NUM_SMS);
    a.items_per_cta = ceil_div((int)items, grid);
    grid = ceil_div((int)items, a.items_per_cta);
    const size_t smem = (size_t)((2 * F + 32) * a.RS + F * a.TS + 2 * F * TAPS * F + round_up(F, 4) + 8) * sizeof(float);
    FB_REQUIRE(smem <= 227 * 1024, "window_len=%d too long for the shared-memory resident backward", W);
    int r = 0;
#define LAUNCH_DG(FF)                                                                      \
    if ((r = set_smem(head_dgrad_kernel<FF>, smem))) return r;                              \
    head_dgrad_kernel<FF><<<grid, HEAD_THREADS, smem, as_stream(stream)>>>(a);
    if (F == 32) { LAUNCH_DG(32) } else if (F == 16) { LAUNCH_DG(16) } else { LAUNCH_DG(8) }
#undef LAUNCH_DG
    return check_launch("head_dgrad");
}

extern "C" int fastb200_window_gather(const float* x, const int32_t* chan, float* out, int B, int C, int T, int W,
                                      int S, int n_win, int n_sel, void* stream) {
    FB_REQUIRE(B >= 0 && C > 0 && T > 0 && W > 0 && S > 0 && n_win > 0 && n_sel > 0, "bad dims");
    FB_REQUIRE((n_win - 1) * (int64_t)S + W <= T, "windows run past the trial");
    const size_t total = (size_t)B * n_win * n_sel * W;
    if (total == 0) return 0;
    size_t blocks = (total + 255) / 256;
    if (blocks > (size_t)NUM_SMS * 16) blocks = (size_t)NUM_SMS * 16;
    window_gather_kernel<<<(unsigned)blocks, 256, 0, as_stream(stream)>>>(x, chan, out, C, T, W, S, n_win, n_sel, total);
    return check_launch("window_gather");
}
