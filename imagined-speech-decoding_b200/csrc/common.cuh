// Shared helpers for the fastb200 kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>
#include "fastb200.h"

namespace fb {

// ---- error channel of the C ABI ------------------------------------------------------------------------
extern thread_local char g_err[512];
int fail(int code, const char* fmt, ...);
int check_launch(const char* what);          // cudaGetLastError() -> return code (+ message)

#define FB_REQUIRE(cond, ...) do { if (!(cond)) return fb::fail(-1, __VA_ARGS__); } while (0)

inline cudaStream_t as_stream(void* s) { return reinterpret_cast<cudaStream_t>(s); }

constexpr int TAPS = FASTB200_TAPS;
constexpr int NUM_SMS = 148;

__host__ __device__ inline int round_up(int a, int b) { return (a + b - 1) / b * b; }
__host__ __device__ inline int ceil_div(int a, int b) { return (a + b - 1) / b; }

// Time-tile geometry shared by host and device for the head kernels.
//  T_out = W - 4 samples survive the valid temporal conv; a warp covers 128 of them (4 per lane).
//  Shared-memory rows are RS floats: 2 zero columns, the samples, zero fill; RS % 32 == 4 keeps the
//  lane-per-row float4 reads of the weight-gradient phase bank-conflict free.
__host__ __device__ inline int head_t_out(int W) { return W - 4; }
__host__ __device__ inline int head_n_tt(int W) { return ceil_div(W - 4, 128); }
__host__ __device__ inline int head_row_stride(int W) { return head_n_tt(W) * 128 + 4; }
__host__ __device__ inline int head_saved_stride(int W) { return round_up(W - 4, 4); }

// ---- packed head weight layout (floats), shared by pack / fwd / bwd / unpack -------------------------------
// per zone z (offset pk_zone_off(z)):  weff [Cz][5][F] | beff [F] (padded to 4) | w3 [F][5][F] | w4 [F][5][F]
//                                      | w3t [F][5][F] | w4t [F][5][F]
// weff[c][k][f]  = sum_g cnn2[f,g,c] * cnn1[g,k]           beff[f] = sum_{g,c} cnn2[f,g,c] * b1[g]
// w3[ci][k][co]  = cnn3[co,ci,k]                           w3t[co][k][ci] = cnn3[co,ci,4-k]
struct PackedZone { int weff, beff, w3, w4, w3t, w4t, end; };
__host__ __device__ inline PackedZone packed_zone(const fastb200_model& m, int z) {
    const int F = m.dim_cnn;
    int base = 0;
    for (int i = 0; i < z; ++i) base += round_up(m.zone_size[i] * TAPS * F, 4) + round_up(F, 4) + 4 * F * TAPS * F;
    PackedZone p;
    p.weff = base;
    p.beff = p.weff + round_up(m.zone_size[z] * TAPS * F, 4);
    p.w3 = p.beff + round_up(F, 4);
    p.w4 = p.w3 + F * TAPS * F;
    p.w3t = p.w4 + F * TAPS * F;
    p.w4t = p.w3t + F * TAPS * F;
    p.end = p.w4t + F * TAPS * F;
    return p;
}

// CUDA-core part of the tensor-core-split head backward (head.cu): input gradients of cnn4 / cnn3, collapsed-embed weight
// and bias gradients, optional dx; dL/dy4, dL/dy3 (and dL/dy2) are WRITTEN to dys (per item 3 x [F][TS]) for head_wgrad_tc.cu.
// embed_tc != 0: the collapsed-embed weight / bias gradients are left to that kernel too (needs dx == nullptr).
int head_dgrad_launch(const fastb200_model* m, const float* packed, const float* x, const float* saved, const float* dfeat,
                      int B, int T, int n_win, int step, float* gpacked, float* dx, float* dys, int embed_tc, void* stream);

// Tensor-core input-gradient chain (head_dgrad_tc.cu): writes dL/dy4, dL/dy3, dL/dy2 to dys; fpacked = head_pack output.
int head_dgrad_tc_supported(const fastb200_model* m);
int64_t head_dgrad_tc_packed_bytes(const fastb200_model* m, int n_terms);
// write_dy4 = 0: dL/dy4 is not written (head_wgrad_tc recomputes it from the saved y4 and dfeat in its converter).
int head_dgrad_tc_launch(const fastb200_model* m, const float* fpacked, const float* saved, const float* dfeat, int B,
                         int n_win, float* dys, void* packed_ws, int n_terms, int write_dy4, void* stream);

// ---- device helpers -------------------------------------------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// exact GELU (nn.GELU default / F.gelu, erf form) and its derivative
__device__ __forceinline__ float gelu_f(float x) { return 0.5f * x * (1.0f + erff(x * 0.70710678118654752440f)); }
__device__ __forceinline__ float gelu_grad_f(float x) {
    const float cdf = 0.5f * (1.0f + erff(x * 0.70710678118654752440f));
    const float pdf = 0.39894228040143267794f * __expf(-0.5f * x * x);
    return cdf + x * pdf;
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }

// ---- 1-D bulk async copy (TMA engine, no tensor map) global -> shared, completion on an mbarrier --------------
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst_smem)), "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
#ifndef FB_MBAR_SUSPEND_NS
#define FB_MBAR_SUSPEND_NS 500u
#endif
// Bounded wait: a byte-count mistake must surface as a trap (launch error), never as a hung GPU box.
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    uint32_t done = 0;
    #pragma unroll 1
    for (uint32_t spin = 0; spin < (1u << 24); ++spin) {
        // the suspend-time hint (ns) lets the hardware park the warp instead of re-issuing the poll
#if FB_MBAR_SUSPEND_NS > 0
        asm volatile("{\n\t.reg .pred p;\n\t"
                     "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
                     "selp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done) : "r"(smem_u32(bar)), "r"(parity), "r"(FB_MBAR_SUSPEND_NS) : "memory");
#else
        asm volatile("{\n\t.reg .pred p;\n\t"
                     "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
                     "selp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done) : "r"(smem_u32(bar)), "r"(parity) : "memory");
#endif
        if (done) return;
    }
    __trap();
}

// Cooperative copy of a [n_rows][row_len] global array into shared memory with pitch row_len + 1 (all loads in flight at
// once, coalesced; the odd pitch keeps "one row per lane" readers conflict-free).  Caller runs stage_rows_wait() and __syncthreads().
__device__ __forceinline__ void stage_rows_pitched(float* dst, const float* __restrict__ src, int n_rows, int row_len) {
    int r = threadIdx.x / row_len, c = threadIdx.x - r * row_len;
    const int dr = blockDim.x / row_len, dcol = blockDim.x - dr * row_len;
    for (int i = threadIdx.x; i < n_rows * row_len; i += blockDim.x) {
        // cp.async, not load + store: through registers every element waits out its own L2 latency (the compiler cannot
        // hoist the loads over the shared-memory stores)
        const unsigned d = (unsigned)__cvta_generic_to_shared(dst + r * (row_len + 1) + c);
        asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d), "l"(src + i) : "memory");
        r += dr; c += dcol;
        if (c >= row_len) { c -= row_len; ++r; }
    }
}
__device__ __forceinline__ void stage_rows_wait() { asm volatile("cp.async.wait_all;" ::: "memory"); }

// counter-based uniform in [0,1): (seed, offset, index) -> float.  Used for dropout masks that must be
// reproduced bit-for-bit by the backward kernel.  (splitmix64-style finaliser; not torch's Philox stream.)
__device__ __forceinline__ float uniform01(uint64_t seed, uint64_t offset, uint64_t idx) {
    uint64_t z = seed + 0x9E3779B97F4A7C15ull * (idx + 1) + (offset << 32 | offset >> 32);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    z = z ^ (z >> 31);
    return (float)(z >> 40) * (1.0f / 16777216.0f);
}

}  // namespace fb
