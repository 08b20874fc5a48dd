// Probe (dev tool): SS-mode tcgen05.mma issue rate vs N, M, operand layout (swizzle) and kind.  Garbage operands.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo, uint32_t layout) {
    return (uint64_t)((saddr & 0x3FFFF) >> 4) | (uint64_t)(lbo >> 4) << 16 | (uint64_t)(sbo >> 4) << 32 | (uint64_t)1 << 46 |
           (uint64_t)layout << 61;
}
template <int KIND>   // 0 = f16 (bf16 operands), 1 = tf32
__device__ __forceinline__ void mma(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc) {
    if (KIND == 0)
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, 1, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d), "l"(a), "l"(b), "r"(idesc) : "memory");
    else
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, 1, 0;\n\ttcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(d), "l"(a), "l"(b), "r"(idesc) : "memory");
}
template <int KIND>
__global__ void __launch_bounds__(128, 1) tput(int M, int N, uint32_t lboA, uint32_t sboA, uint32_t lboB, uint32_t sboB, uint32_t layout,
                                               int iters, long long* cycles) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t bar; __shared__ uint32_t slot;
    const int tid = threadIdx.x, warp = tid >> 5;
    for (int i = tid; i < 160 * 1024 / 4; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0;
    if (tid == 0) { asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar))); asm volatile("fence.mbarrier_init.release.cluster;"); }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&slot)), "n"(256) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = slot;
    const uint32_t fmt = KIND == 0 ? 1u : 2u;
    const uint32_t idesc = (1u << 4) | (fmt << 7) | (fmt << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
    if (tid == 0) {
        const uint32_t a0 = smem_u32(smem), b0 = smem_u32(smem) + 96 * 1024;
        long long t0 = clock64();
        for (int it = 0; it < iters; ++it)
#pragma unroll
            for (int k = 0; k < 10; ++k)
                mma<KIND>(tmem, make_desc(a0 + (k & 1) * 32 + (k >> 1) * 16 * (layout == 0), lboA, sboA, layout), make_desc(b0 + (k & 1) * 32, lboB, sboB, layout), idesc);
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
        uint32_t done = 0;
        while (!done) asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(done) : "r"(smem_u32(&bar)) : "memory");
        cycles[0] = clock64() - t0;
    }
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(256) : "memory");
}
int main() {
    long long* cyc; cudaMalloc(&cyc, 8);
    const size_t smem = 160 * 1024;
    cudaFuncSetAttribute(tput<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaFuncSetAttribute(tput<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    struct L { const char* name; uint32_t layout, lboA, sboA, lboB, sboB; };
    // no swizzle: [chunk][row][16B] rows contiguous (SBO 128, LBO = 136 rows * 16); swizzle 128B/64B/32B K-major: SBO = 8 rows * row bytes
    L ls[] = {{"none  ", 0, 136 * 16, 128, 256 * 16, 128}, {"sw128 ", 2, 16, 1024, 16, 1024}, {"sw64  ", 4, 16, 512, 16, 512}, {"sw32  ", 6, 16, 256, 16, 256}};
    for (int kind = 0; kind < 2; ++kind)
        for (auto& l : ls)
            for (int M : {128, 64})
                for (int N : {32, 64, 96, 128, 256}) {
                    if (kind == 0) tput<0><<<1, 128, smem>>>(M, N, l.lboA, l.sboA, l.lboB, l.sboB, l.layout, 100, cyc);
                    else tput<1><<<1, 128, smem>>>(M, N, l.lboA, l.sboA, l.lboB, l.sboB, l.layout, 100, cyc);
                    cudaError_t e = cudaDeviceSynchronize();
                    if (e != cudaSuccess) { printf("CUDA error %s\n", cudaGetErrorString(e)); return 1; }
                    long long c; cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
                    printf("%s layout=%s M=%3d N=%3d  cycles/MMA=%.1f\n", kind ? "tf32" : "bf16", l.name, M, N, c / 1000.0);
                }
    return 0;
}
