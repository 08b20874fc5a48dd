// Probe (dev tool): TMEM->register load rate (tcgen05.ld 32x32b.x32), alone and while MMAs run; MMA rate with mixed N.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
    return (uint64_t)((saddr & 0x3FFFF) >> 4) | (uint64_t)(lbo >> 4) << 16 | (uint64_t)(sbo >> 4) << 32 | (uint64_t)1 << 46;
}
__device__ __forceinline__ void mma(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, 1, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d), "l"(a), "l"(b), "r"(idesc) : "memory");
}
__device__ __forceinline__ uint32_t idesc(int N) { return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | (8u << 24); }
__device__ __forceinline__ float ld32(uint32_t addr) {
    uint32_t r[32];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,"
                 "%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
                   "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
                   "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
                   "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31]) : "r"(addr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    float s = 0;
#pragma unroll
    for (int i = 0; i < 32; ++i) s += __uint_as_float(r[i]);
    return s;
}
// mode bits: 1 = run LDTM loops (n_ld_warps warps), 2 = run MMA loop; pattern: 0 = N64 only, 1 = alternate 64/32 same D,
// 2 = ten N64 then ten N32 same D, 3 = alternate 64/32 into disjoint D columns
__global__ void __launch_bounds__(384, 1) k(int mode, int n_ld_warps, int pattern, int iters, long long* out, float* sink, int fill = 1) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t bar; __shared__ uint32_t slot;
    const int tid = threadIdx.x, warp = tid >> 5;
    for (int i = tid; i < 96 * 1024 / 4; i += blockDim.x) { uint32_t h = (uint32_t)i * 2654435761u; reinterpret_cast<uint32_t*>(smem)[i] = fill == 0 ? 0u : fill == 1 ? 0x3f803f80u + (i & 7) : fill == 2 ? ((h & 0x807f807fu) | 0x3f003f00u) : 0x3f803f80u; }
    if (tid == 0) { asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar))); asm volatile("fence.mbarrier_init.release.cluster;"); }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&slot)), "n"(512) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = slot;
    if (warp < n_ld_warps && (mode & 1)) {
        float s = 0;
        long long t0 = clock64();
        for (int i = 0; i < iters; ++i) s += ld32(tmem + ((uint32_t)((warp & 3) * 32) << 16) + 256 + ((i & 3) * 32));
        long long t1 = clock64();
        if ((tid & 31) == 0) out[warp] = t1 - t0;
        sink[tid] = s;
    }
    if (warp == 8 && (tid & 31) == 0 && (mode & 2)) {
        const uint32_t a0 = smem_u32(smem), b0 = smem_u32(smem) + 40 * 1024;
        long long t0 = clock64();
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int j = 0; j < 20; ++j) {
                int N = 64; uint32_t d = tmem;
                if (pattern == 1) N = (j & 1) ? 32 : 64;
                if (pattern == 2) N = j < 10 ? 64 : 32;
                if (pattern == 3) { N = (j & 1) ? 32 : 64; d = tmem + ((j & 1) ? 64 : 0); }
                if (pattern >= 4) {        // faithful conv stage of head_fwd_tc (NT = 2): j = (tap k, half h, term t)
                    const int t = j & 1, hh = (j >> 1) & 1, kk = (j >> 2);
                    N = t ? 32 : 64;
                    const uint32_t ry = pattern == 4 ? 132 : 136, lbB = pattern == 6 ? 1536 : 1024;
                    mma(d, make_desc(a0 + t * (4 * ry * 16) + 2 * hh * (ry * 16) + kk * 16, ry * 16, 128),
                        make_desc(b0 + (4 * kk + 2 * hh) * lbB, lbB, 128), idesc(N));
                    continue;
                }
                mma(d, make_desc(a0 + (j % 5) * 16, 132 * 16, 128), make_desc(b0, 64 * 16, 128), idesc(N));
            }
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
        uint32_t done = 0;
        while (!done) asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(done) : "r"(smem_u32(&bar)) : "memory");
        out[8] = clock64() - t0;
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(512) : "memory");
}
int main() {
    long long* out; float* sink; cudaMalloc(&out, 16 * 8); cudaMalloc(&sink, 384 * 4);
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024);
    long long h[16];
    const int iters = 200;
    for (int nld : {1, 4, 8}) {
        cudaMemset(out, 0, 128);
        k<<<1, 384, 96 * 1024>>>(1, nld, 0, iters, out, sink); cudaDeviceSynchronize();
        cudaMemcpy(h, out, 128, cudaMemcpyDeviceToHost);
        printf("LDTM only, %d warps: %.1f cycles per ld.x32 per warp (4 KB per warp-instr)\n", nld, h[0] / (double)iters);
    }
    for (int pat = 0; pat < 7; ++pat) {
        cudaMemset(out, 0, 128);
        k<<<1, 384, 96 * 1024>>>(2, 0, pat, iters, out, sink); cudaDeviceSynchronize();
        cudaMemcpy(h, out, 128, cudaMemcpyDeviceToHost);
        printf("MMA only pattern %d: %.1f cycles/MMA\n", pat, h[8] / (20.0 * iters));
    }
    for (int fill = 0; fill < 4; ++fill) {
        cudaMemset(out, 0, 128);
        k<<<1, 384, 96 * 1024>>>(2, 0, 4, iters, out, sink, fill); cudaDeviceSynchronize();
        cudaMemcpy(h, out, 128, cudaMemcpyDeviceToHost);
        printf("fill %d (0 zeros, 1 ~1.0 varied, 2 random +-[0.5,1), 3 all exactly 1.0): %.1f cycles/MMA\n", fill, h[8] / (20.0 * iters));
    }
    for (int nld : {4, 8})
        for (int pat : {0, 1}) {
            cudaMemset(out, 0, 128);
            k<<<1, 384, 96 * 1024>>>(3, nld, pat, iters * 10, out, sink); cudaDeviceSynchronize();
            cudaMemcpy(h, out, 128, cudaMemcpyDeviceToHost);
            printf("both, %d LDTM warps, MMA pattern %d: %.1f cycles/MMA, %.1f cycles per ld.x32 (the LDTM loop is 20x shorter: only its rate while overlapped is meaningful if it ended later)\n",
                   nld, pat, h[8] / (20.0 * iters * 10), h[0] / (double)(iters * 10));
        }
    cudaError_t e = cudaGetLastError();
    printf("%s\n", cudaGetErrorString(e));
    return 0;
}
