// tcgen05 / TMEM PTX wrappers, bf16 operand splitting and packed-fp32x2 helpers shared by the tensor-core head kernels
// (head_tc.cu: inference forward; head_wgrad_tc.cu: weight gradients of cnn3 / cnn4).
#pragma once
#include "common.cuh"
#include <cuda_bf16.h>

namespace fb {
namespace tc {

// ---- PTX wrappers ---------------------------------------------------------------------------------------------------
__host__ __device__ constexpr uint32_t idesc_bf16(int N) {      // D fp32, A/B bf16, K-major both, M = 128
    return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
}
// The MMA warp runs CONVERGED (every lane computes the same warp-uniform descriptors) and one elected lane issues:
// a divergent `if (lane == 0)` makes ptxas wrap every UTCHMMA in a per-lane uniformisation loop (measured: 70-105
// cycles per MMA of pure issue latency instead of the 47-cycle hardware rate).
// low 32 bits of a K-major SWIZZLE_NONE descriptor: start address and LBO (16-byte units); the high word is constant.
__device__ __forceinline__ uint32_t desc_lo(uint32_t saddr, uint32_t lbo_bytes) { return ((saddr & 0x3FFFFu) >> 4) | (lbo_bytes >> 4) << 16; }
constexpr uint32_t DESC_HI = (128u >> 4) | (1u << 14);      // SBO = 128 B, descriptor version 1
__device__ __forceinline__ void umma(uint32_t d_tmem, uint32_t a_lo, uint32_t b_lo, uint32_t idesc, uint32_t acc) {
    asm volatile("{\n\t.reg .pred p, q;\n\t.reg .b64 da, db;\n\t"
                 "elect.sync _|q, 0xffffffff;\n\t"
                 "setp.ne.b32 p, %4, 0;\n\t"
                 "mov.b64 da, {%1, %5};\n\t"
                 "mov.b64 db, {%2, %5};\n\t"
                 "@q tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %3, p;\n\t}"
                 ::"r"(d_tmem), "r"(a_lo), "r"(b_lo), "r"(idesc), "r"(acc), "r"(DESC_HI) : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
    asm volatile("{\n\t.reg .pred q;\n\telect.sync _|q, 0xffffffff;\n\t"
                 "@q tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n\t}" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tmem_ld32(uint32_t addr, float (&v)[32]) {
    uint32_t r[32];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,"
                 "%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
                   "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
                   "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
                   "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
                 : "r"(addr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
}

__device__ __forceinline__ void tmem_ld16(uint32_t addr, float (&v)[16]) {
    uint32_t r[16];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
                   "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
                 : "r"(addr));
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// ---- bf16 operand splitting ------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t pack_bf16x2(float even, float odd) {       // even -> low half (lower K index)
    uint32_t r;
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(odd), "f"(even));
    return r;
}
// eight fp32 -> NT uint4 of bf16 terms (term 0 = rounded value, term i = rounded residual after terms < i)
template <int NT>
__device__ __forceinline__ void split8(const float (&x)[8], uint4 (&out)[NT]) {
    float r[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) r[i] = x[i];
#pragma unroll
    for (int t = 0; t < NT; ++t) {
        uint32_t w[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            w[i] = pack_bf16x2(r[2 * i], r[2 * i + 1]);
            if (t + 1 < NT) {                               // residual = value - bf16 term (exact in fp32), two lanes per FADD2
                const float lo = __uint_as_float(w[i] << 16), hi = __uint_as_float(w[i] & 0xffff0000u);
                unsigned long long d;
                asm("{\n\t.reg .b64 a, b;\n\tmov.b64 a, {%1, %2};\n\tmov.b64 b, {%3, %4};\n\t"
                    "sub.rn.f32x2 %0, a, b;\n\t}" : "=l"(d) : "f"(r[2 * i]), "f"(r[2 * i + 1]), "f"(lo), "f"(hi));
                asm("mov.b64 {%0, %1}, %2;" : "=f"(r[2 * i]), "=f"(r[2 * i + 1]) : "l"(d));
            }
        }
        out[t] = make_uint4(w[0], w[1], w[2], w[3]);
    }
}

// ---- packed fp32x2 arithmetic (sm_100: FFMA2 / FMUL2 / FADD2 -- one issue slot for two lanes of math) ---------------------
// The epilogue of this kernel is bound by instruction ISSUE, not by a math pipe (ncu: issue slots ~50 % busy with every
// warp latency-stalled; 41 % of all issued instructions were the scalar GELU), so halving the instruction count pays.
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pk(float a, float b) { f32x2 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ void unpk(f32x2 v, float& a, float& b) { asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); }
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) { f32x2 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b) { f32x2 r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ f32x2 add2(f32x2 a, f32x2 b) { f32x2 r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ float rcp_approx(float x) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ float ex2_approx(float x) { float r; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }

// Exact-form GELU x * Phi(x) on two values.  Phi(x) = 0.5 erfc(-x / sqrt 2) with erfc from Abramowitz & Stegun 7.1.26
// (|erf error| <= 1.5e-7):  h = 0.5 erfc(|x| / sqrt 2) = t (c1 + t (c2 + t (c3 + t (c4 + t c5)))) exp(-x^2 / 2),
// t = 1 / (1 + p |x| / sqrt 2);  gelu(x) = x (1 - h) for x >= 0 and x h for x < 0, i.e.  max(x, 0) - |x h|  for both.
__device__ __forceinline__ void gelu2(float& x0, float& x1) {
    const float t0 = rcp_approx(fmaf(fabsf(x0), 0.3275911f * 0.70710678118654752440f, 1.0f));
    const float t1 = rcp_approx(fmaf(fabsf(x1), 0.3275911f * 0.70710678118654752440f, 1.0f));
    const f32x2 t = pk(t0, t1), x = pk(x0, x1);
    f32x2 q = fma2(t, pk(0.5f * 1.061405429f, 0.5f * 1.061405429f), pk(0.5f * -1.453152027f, 0.5f * -1.453152027f));
    q = fma2(t, q, pk(0.5f * 1.421413741f, 0.5f * 1.421413741f));
    q = fma2(t, q, pk(0.5f * -0.284496736f, 0.5f * -0.284496736f));
    q = fma2(t, q, pk(0.5f * 0.254829592f, 0.5f * 0.254829592f));
    q = mul2(q, t);
    const f32x2 arg = mul2(mul2(x, x), pk(-0.72134752044448170368f, -0.72134752044448170368f));   // -x^2/2 * log2(e)
    float a0, a1;
    unpk(arg, a0, a1);
    const f32x2 r = mul2(x, mul2(q, pk(ex2_approx(a0), ex2_approx(a1))));                            // x h
    float r0, r1;
    unpk(r, r0, r1);
    x0 = fmaxf(x0, 0.f) - fabsf(r0);
    x1 = fmaxf(x1, 0.f) - fabsf(r1);
}

// d/dx [x Phi(x)] = Phi(x) + x phi(x), with Phi from the same A&S 7.1.26 erfc as the forward's fused GELU (|error| <= 1.5e-7)
__device__ __forceinline__ float gelu_grad_fast(float x) {
    const float ax = fabsf(x);
    const float t = rcp_approx(fmaf(ax, 0.3275911f * 0.70710678118654752440f, 1.0f));
    float q = fmaf(t, 0.5f * 1.061405429f, 0.5f * -1.453152027f);
    q = fmaf(t, q, 0.5f * 1.421413741f);
    q = fmaf(t, q, 0.5f * -0.284496736f);
    q = fmaf(t, q, 0.5f * 0.254829592f);
    const float e = ex2_approx(x * x * -0.72134752044448170368f);       // exp(-x^2 / 2)
    const float h = q * t * e;                                          // 0.5 erfc(|x| / sqrt 2)
    const float cdf = x >= 0.f ? 1.0f - h : h;
    return fmaf(x * 0.39894228040143267794f, e, cdf);
}

}  // namespace tc
}  // namespace fb
