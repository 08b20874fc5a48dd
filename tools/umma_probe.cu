// Probe (dev tool, not product): validates on a B200 the tcgen05 operand tricks the tensor-core head kernel relies on:
//  (1) K-major SWIZZLE_NONE descriptors with SBO = 128 B are linear in the row index, so a temporal tap shift is a
//      +16 B start-address offset; (2) B stacked along N ([b0|b1|b2], N = 96); (3) SS-mode MMA issue rate vs N.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o umma_probe tools/umma_probe.cu
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cmath>
#include <vector>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr & 0x3FFFF) >> 4);
    d |= (uint64_t)(lbo_bytes >> 4) << 16;
    d |= (uint64_t)(sbo_bytes >> 4) << 32;
    d |= (uint64_t)1 << 46;                       // version = 1 (Blackwell)
    return d;                                     // layout_type = 0 (no swizzle), base_offset = 0
}
__host__ __device__ constexpr uint32_t make_idesc_bf16(int M, int N) {
    return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
__device__ __forceinline__ void mma_bf16(uint32_t d_tmem, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                 ::"r"(d_tmem), "l"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ void commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t c) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(c) : "memory");
}
__device__ __forceinline__ bool mbar_wait(uint64_t* bar, uint32_t parity) {
    uint32_t done = 0;
    for (uint32_t i = 0; i < (1u << 22); ++i) {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done) : "r"(smem_u32(bar)), "r"(parity) : "memory");
        if (done) return true;
    }
    return false;
}

constexpr int R = 136;        // rows of the activation tile (128 + taps)
constexpr int KC = 4;         // 16-byte K chunks per row (32 bf16)
constexpr int NB = 96;        // stacked B rows

// A layout [chunk][row][8 bf16]; B layout [chunk][n][8 bf16]
__global__ void __launch_bounds__(128, 1) probe_kernel(const __nv_bfloat16* A, const __nv_bfloat16* B, float* D, int shift, int N,
                                                       int iters, long long* cycles) {
    extern __shared__ __align__(128) uint8_t smem[];
    __nv_bfloat16* As = reinterpret_cast<__nv_bfloat16*>(smem);
    __nv_bfloat16* Bs = As + KC * R * 8;
    uint64_t* bar = reinterpret_cast<uint64_t*>(Bs + KC * NB * 8);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar + 1);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    for (int i = tid; i < KC * R * 8; i += 128) As[i] = A[i];
    for (int i = tid; i < KC * NB * 8; i += 128) Bs[i] = B[i];
    if (tid == 0) { mbar_init(bar, 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "n"(128) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = *tmem_slot;
    const uint32_t idesc = make_idesc_bf16(128, N);
    uint32_t parity = 0;
    if (tid == 0) {
        // correctness pass: D = A[shift : shift+128, 0:32] x B[0:N, 0:32]^T   (two K16 steps)
        for (int s = 0; s < 2; ++s) {
            uint64_t a = make_desc(smem_u32(As) + (2 * s) * R * 16 + shift * 16, R * 16, 128);
            uint64_t b = make_desc(smem_u32(Bs) + (2 * s) * NB * 16, NB * 16, 128);
            mma_bf16(tmem, a, b, idesc, s > 0);
        }
        commit(bar);
    }
    bool ok = mbar_wait(bar, parity); parity ^= 1;
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    if (!ok) { if (tid == 0) printf("timeout waiting for MMA\n"); }
    for (int c0 = 0; c0 < N; c0 += 32) {
        uint32_t v[32];
        const uint32_t addr = tmem + ((uint32_t)(warp * 32) << 16) + c0;
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
                     : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
                       "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]),
                       "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]),
                       "=r"(v[30]), "=r"(v[31]) : "r"(addr));
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        for (int j = 0; j < 32; ++j) D[(warp * 32 + lane) * NB + c0 + j] = __uint_as_float(v[j]);
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    // throughput pass: `iters` groups of 10 MMAs (5 taps x 2 K16 steps), one commit at the end
    if (tid == 0) {
        long long t0 = clock64();
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int k = 0; k < 5; ++k)
#pragma unroll
                for (int s = 0; s < 2; ++s) {
                    uint64_t a = make_desc(smem_u32(As) + (2 * s) * R * 16 + k * 16, R * 16, 128);
                    uint64_t b = make_desc(smem_u32(Bs) + (2 * s) * NB * 16, NB * 16, 128);
                    mma_bf16(tmem, a, b, idesc, 1);
                }
        }
        commit(bar);
        mbar_wait(bar, parity);
        long long t1 = clock64();
        cycles[0] = t1 - t0;
    }
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(128) : "memory");
}

static float bf(float x) { return __bfloat162float(__float2bfloat16(x)); }

int main() {
    std::vector<float> Af(R * 32), Bf(NB * 32);
    srand(1);
    for (auto& v : Af) v = bf((rand() % 2001 - 1000) / 500.0f);
    for (auto& v : Bf) v = bf((rand() % 2001 - 1000) / 1000.0f);
    std::vector<__nv_bfloat16> Ah(KC * R * 8), Bh(KC * NB * 8);
    for (int r = 0; r < R; ++r) for (int k = 0; k < 32; ++k) Ah[((k / 8) * R + r) * 8 + k % 8] = __float2bfloat16(Af[r * 32 + k]);
    for (int n = 0; n < NB; ++n) for (int k = 0; k < 32; ++k) Bh[((k / 8) * NB + n) * 8 + k % 8] = __float2bfloat16(Bf[n * 32 + k]);
    __nv_bfloat16 *Ad, *Bd; float* Dd; long long* cyc;
    cudaMalloc(&Ad, Ah.size() * 2); cudaMalloc(&Bd, Bh.size() * 2); cudaMalloc(&Dd, 128 * NB * 4); cudaMalloc(&cyc, 8);
    cudaMemcpy(Ad, Ah.data(), Ah.size() * 2, cudaMemcpyHostToDevice);
    cudaMemcpy(Bd, Bh.data(), Bh.size() * 2, cudaMemcpyHostToDevice);
    const size_t smem = (KC * R * 8 + KC * NB * 8) * 2 + 64;
    cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    int bad_total = 0;
    for (int N : {32, 64, 96}) {
        for (int shift : {0, 1, 3, 4, 7}) {
            cudaMemset(Dd, 0, 128 * NB * 4);
            probe_kernel<<<1, 128, smem>>>(Ad, Bd, Dd, shift, N, 200, cyc);
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("CUDA error: %s\n", cudaGetErrorString(e)); return 1; }
            std::vector<float> Dh(128 * NB); long long c;
            cudaMemcpy(Dh.data(), Dd, Dh.size() * 4, cudaMemcpyDeviceToHost);
            cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
            double maxerr = 0; int bad = 0;
            for (int m = 0; m < 128; ++m) for (int n = 0; n < N; ++n) {
                double ref = 0;
                for (int k = 0; k < 32; ++k) ref += (double)Af[(m + shift) * 32 + k] * Bf[n * 32 + k];
                double err = fabs(ref - Dh[m * NB + n]);
                if (err > maxerr) maxerr = err;
                if (err > 1e-3) ++bad;
            }
            bad_total += bad;
            printf("N=%d shift=%d  max|err|=%.3e bad=%d   cycles/MMA=%.1f (2000 MMAs M128 K16)\n", N, shift, maxerr, bad, c / 2000.0);
        }
    }
    printf(bad_total ? "PROBE FAILED\n" : "PROBE OK\n");
    return bad_total != 0;
}
