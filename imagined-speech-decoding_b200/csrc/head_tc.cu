// Tensor-core (tcgen05 / TMEM) zone embedding head of FAST -- inference forward.
//
// Same arithmetic as head.cu (reference: FAST.forward_head src/fast/models/fast.py:242-252, Head.forward :209-210,
// Conv4Layers.forward :111-119), mapped onto the 5th-generation tensor cores:
//
//   * each temporal convolution is an implicit GEMM  D[t][co] = sum_{tap,ci} A_tap[t][ci] * W[tap,ci][co]  with
//     M = 128 time samples, N = 32 output channels, K = 5 taps x 32 channels.  Activations live in shared memory as
//     [ci/8][row][8 x bf16] (K-major, SWIZZLE_NONE, SBO = 128 B), which is LINEAR in the row index, so the operand
//     of tap k is the same tile with its descriptor start address advanced by k rows (16 B each): no im2col copy;
//   * fp32 accuracy from bf16 tensor cores by operand splitting: x = x0 + x1 (+ x2) with bf16 terms, products
//     a_i * b_j kept for i + j < NT.  The weight terms are stacked along N ([b0|b1|b2], one descriptor), so the pass
//     over activation term a_i is ONE MMA of N = 32 * (NT - i) accumulating into TMEM columns [0, N): the epilogue
//     adds the NT column groups.  (Measured on B200, tools/umma_probe.cu: an SS-mode M128 K16 MMA costs
//     max(46.6, bytes/127) cycles -- issue/shared-memory bound, not tensor bound -- so stacking along N is free.)
//   * a 246-sample window is two independent 128-row streams (rows [0,128) and [118,246)); each stream recomputes
//     the few rows whose receptive field crosses the cut, so the streams never exchange data and ping-pong on the
//     tensor pipe: while one stream's accumulator is drained (TMEM -> registers -> bias/mask -> bf16 split -> next
//     layer's operand tile in shared memory), the other stream's MMAs run;
//   * persistent CTAs, one zone per CTA (its ~50-77 KB of split weights stay resident in shared memory, fetched once
//     by a bulk async copy); the 148 CTAs are apportioned to zones by cost so every zone walks the trial list at the
//     same rate and the 8 zone passes over a trial hit L2, not HBM.
//
// Warp roles (768 threads): warps 0-7 / 8-15 epilogue of stream A / B (TMEM lane quadrant = warp % 4, column half =
// (warp / 4) % 2), warps 16 / 17 MMA issuer of stream A / B (one elected lane each; while one issuer sits in a barrier
// wait the other keeps the tensor pipe fed), warps 18 / 19 electrode copier of stream A / B
// (fused unfold + zone gather, cp.async ring), warps 20-23 converters (staged fp32 -> split-bf16 operand tile).
#include "common.cuh"
#include "tc_ptx.cuh"
#include <cuda_bf16.h>
#include <stdlib.h>

namespace fb {
namespace tc {

constexpr int F = 32;                   // dim_cnn handled by this kernel
constexpr int RY = 132;                 // rows of a conv operand tile: 2 zero rows + 128 + 2 zero rows
constexpr int RX = 136;                 // rows of an electrode tile: 128 + 4 taps + 1 (tap-pair trick) + spare
constexpr int EPI_WARPS = 16, MMA_WARP = 16, MMA_WARPS = 2, COPY_WARP0 = 18, CONV_WARP0 = 20, LOAD_WARPS = 6;
// 24 warps launch with 80 registers each; the four epilogue warpgroups then grow to EPI_REGS and the two service
// warpgroups (issuers + copiers, converters) shrink to SVC_REGS (setmaxnreg): the CTA register pool is what it launched with.
constexpr int EPI_REGS = 96, SVC_REGS = 48;      // pool = 768 x 80: 16 warps x (96 - 80) == 8 warps x (80 - 48)
constexpr int CW = 16;                  // accumulator columns per epilogue warp (two warps share a TMEM lane quadrant)
constexpr int THREADS = (EPI_WARPS + MMA_WARPS + LOAD_WARPS) * 32;
constexpr int EMB_CHUNKS = 10;          // K chunks (8 bf16) of the collapsed embed filter: 5 taps x 16 electrodes
constexpr int CONV_CHUNKS = 20;         // 5 taps x 32 channels / 8

__host__ __device__ constexpr int lb(int NT) { return 32 * NT * 16; }                       // bytes of one weight K chunk
__host__ __device__ constexpr int zone_bytes(int NT) { return (EMB_CHUNKS + 2 * CONV_CHUNKS) * lb(NT) + 128; }
__host__ __device__ constexpr int x_bytes(int NT) { return NT * 2 * RX * 16; }              // one stream's electrode tile
__host__ __device__ constexpr int y_bytes(int NT) { return NT * 4 * RY * 16; }              // one stream's conv operand

// ---- weight packing --------------------------------------------------------------------------------------------------
// per zone (zone_bytes(NT) bytes):  embed [10 chunks][32*NT rows][8 bf16] | cnn3 [20][32*NT][8] | cnn4 [20][32*NT][8] | beff fp32[32]
//   row n = term * 32 + co.   K index of the embed filter: Czp = 16 -> kk = tap * 16 + c;  Czp = 8 -> kk = tap * 8 + c
//   (taps padded to 6 so that one K16 MMA covers two taps);  conv K index kk = tap * 32 + ci.
__device__ __forceinline__ void store_split(uint16_t* blk, int NT, int chunk, int n_co, int e, float v) {
    float r = v;
    for (int t = 0; t < NT; ++t) {
        const __nv_bfloat16 b = __float2bfloat16_rn(r);
        blk[((size_t)chunk * 32 * NT + t * 32 + n_co) * 8 + e] = __bfloat16_as_ushort(b);
        r -= __bfloat162float(b);
    }
}

__global__ void head_pack_tc_kernel(fastb200_model m, const float* __restrict__ params, uint8_t* __restrict__ packed, int NT) {
    extern __shared__ __align__(16) float pk_s[];                // cnn2.weight of the zone, f-slabs at pitch F*Cz + 1 (see head_pack_kernel)
    const int z = blockIdx.x;
    const int Cz = m.zone_size[z];
    const int Czp = Cz <= 8 ? 8 : 16;
    uint8_t* zb = packed + (size_t)z * (NT == 2 ? zone_bytes(2) : zone_bytes(3));
    uint16_t* emb = reinterpret_cast<uint16_t*>(zb);
    uint16_t* c3 = emb + EMB_CHUNKS * 32 * NT * 8;
    uint16_t* c4 = c3 + CONV_CHUNKS * 32 * NT * 8;
    float* beff = reinterpret_cast<float*>(c4 + CONV_CHUNKS * 32 * NT * 8);
    const float* w1 = params + m.off_cnn1_w[z];
    const float* b1 = params + m.off_cnn1_b[z];
    const float* w2 = params + m.off_cnn2_w[z];
    const float* w3 = params + m.off_cnn3_w[z];
    const float* w4 = params + m.off_cnn4_w[z];
    stage_rows_pitched(pk_s, w2, F, F * Cz);
    stage_rows_wait();
    __syncthreads();
    w2 = pk_s;
    const int fp = F * Cz + 1;
    // collapsed cnn1 o cnn2 (same summation order as head_pack_kernel in head.cu)
    for (int i = threadIdx.x + blockDim.x * blockIdx.y; i < EMB_CHUNKS * 8 * F; i += blockDim.x * gridDim.y) {
        const int f = i % F, kk = i / F;
        const int k = kk / Czp, c = kk % Czp;
        float s = 0.f;
        if (k < TAPS && c < Cz) {
#pragma unroll 8
            for (int g = 0; g < F; ++g) s = fmaf(w2[f * fp + g * Cz + c], __ldg(w1 + g * TAPS + k), s);
        }
        store_split(emb, NT, kk / 8, f, kk % 8, s);
    }
    {   // beff[f] = sum_g b1[g] * sum_c w2[f,g,c]: one warp per f, lane = g (as head_pack_kernel)
        const int lane = threadIdx.x & 31, wg = (threadIdx.x >> 5) + (blockDim.x >> 5) * blockIdx.y, nw = (blockDim.x >> 5) * gridDim.y;
        for (int f = wg; f < F; f += nw) {
            float r = 0.f;
            if (lane < F) {
                for (int c = 0; c < Cz; ++c) r += w2[f * fp + lane * Cz + c];
                r *= b1[lane];
            }
            const float v = warp_sum(r);
            if (lane == 0) beff[f] = v;
        }
    }
    for (int i = threadIdx.x + blockDim.x * blockIdx.y; i < CONV_CHUNKS * 8 * F; i += blockDim.x * gridDim.y) {
        const int co = i % F, kk = i / F;
        const int k = kk / F, ci = kk % F;
        store_split(c3, NT, kk / 8, co, kk % 8, w3[(co * F + ci) * TAPS + k]);
        store_split(c4, NT, kk / 8, co, kk % 8, w4[(co * F + ci) * TAPS + k]);
    }
}

// ---- kernel arguments ------------------------------------------------------------------------------------------------
struct FwdArgs {
    const uint8_t* packed;
    const float* x;
    float* feat;
    float* saved;                              // SAVE builds (training forward): per item 3 slots of F * TS floats: y2, y3 as split-bf16
                                               // operand images, y4 as fp32 [F][TS] (the input-gradient kernel needs gelu'(y4))
    int TS;
    long long* prof;                           // PROF builds only: per-CTA wait-cycle counters [grid][16]
    int dbg;                                   // PROF builds only: bit0 skip operand stores, bit1 skip GELU, bit2 skip electrode loads
    int B, T, n_win, step, C, W, Z;
    int T_out, split, tb1;                     // stream B starts at row tb1 = T_out - 128; stream A owns t < split
    int cz[FASTB200_MAX_ZONES];
    int zoff[FASTB200_MAX_ZONES];
    int chan[FASTB200_MAX_CHANS];
    uint8_t cta_zone[NUM_SMS];                 // zone of CTA c
    uint8_t cta_rank[NUM_SMS];                 // its rank among the CTAs of that zone
    uint8_t zone_ctas[FASTB200_MAX_ZONES];     // CTAs apportioned to each zone
};

struct Smem {                                  // fixed part at the start of dynamic shared memory
    uint64_t w_full;
    uint64_t x_full[2], x_empty[2], de_full[2], ye_ready[2], yc_ready[2], d_free[2];
    uint64_t dm_full[2][2];                    // per stream, per conv layer (cnn3, cnn4): one completion per item each
    uint64_t stg_full[2][2], stg_empty[2][2];
    uint64_t red_free[2];                      // partial-sum slot of item parity p may be reused (its feature was written)
    uint32_t tmem_base;
    int counter[2];
    float red[2][EPI_WARPS][CW];
};
constexpr int STG_BYTES = 16 * RX * 4;               // one staged electrode tile: raw fp32 [16 channels][RX rows]
__host__ __device__ constexpr int stg_tiles(int NT) { return NT == 2 ? 2 : 1; }   // staging ring depth PER STREAM (shared-memory budget)
constexpr int SMEM_HDR = 3072;                 // bytes reserved for Smem in front of the operand tiles
static_assert(sizeof(Smem) <= SMEM_HDR, "Smem header overflows its reservation");

// ---- forward kernel --------------------------------------------------------------------------------------------------
template <int NT, bool PROF, bool SAVE>
__global__ void __launch_bounds__(THREADS, 1) head_fwd_tc_kernel(const __grid_constant__ FwdArgs a) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    Smem& sm = *reinterpret_cast<Smem*>(smem_raw);
    uint8_t* wz = smem_raw + SMEM_HDR;                              // this zone's split weights (zone_bytes(NT))
    uint8_t* xs = wz + zone_bytes(NT);                              // [2 streams] electrode tiles
    uint8_t* ys = xs + 2 * x_bytes(NT);                             // [2 streams][2 item parities] conv operand tiles
    constexpr int LB = lb(NT);
    constexpr int STG = stg_tiles(NT);
    constexpr int DCOLS = 32 * NT;                                  // accumulator columns per stream
    constexpr int TMEM_COLS = 4 * DCOLS <= 256 ? 256 : 512;         // conv accumulators [0, 2 DCOLS), embed accumulators after

    // warp index through a shuffle: tells the compiler it is warp-uniform, so role-dependent values (stream, descriptors)
    // live in uniform registers -- the form UTCHMMA wants -- instead of being re-uniformised per MMA
    const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0), lane = tid & 31;
    const int z = a.cta_zone[blockIdx.x];
    const int rank = a.cta_rank[blockIdx.x], nz = a.zone_ctas[z];
    const int BN = a.B * a.n_win;
    const int Cz = a.cz[z];
    const int n_xchunk = Cz <= 8 ? 1 : 2;

    // ---- one-time setup: barriers, TMEM, weights (bulk async copy), zero rows of the operand tiles
    if (tid == 0) {
        mbar_init(&sm.w_full, 1);
        for (int s = 0; s < 2; ++s) {
            mbar_init(&sm.x_full[s], 64);
            for (int i = 0; i < 2; ++i) { mbar_init(&sm.stg_full[s][i], 32); mbar_init(&sm.stg_empty[s][i], 64); }
            mbar_init(&sm.x_empty[s], 1);
            mbar_init(&sm.de_full[s], 1);
            mbar_init(&sm.dm_full[s][0], 1);
            mbar_init(&sm.dm_full[s][1], 1);
            // every epilogue thread arrives for itself, as soon as ITS stores are fenced: one arrival per warp behind a
            // __syncwarp was measured 9 % slower (the operand hand-over sits on the critical path of the stream)
            mbar_init(&sm.ye_ready[s], EPI_WARPS * 16);
            mbar_init(&sm.yc_ready[s], EPI_WARPS * 16);
            mbar_init(&sm.d_free[s], EPI_WARPS * 16);
        }
        mbar_init(&sm.red_free[0], 1);
        mbar_init(&sm.red_free[1], 1);
        sm.counter[0] = sm.counter[1] = 0;
        fence_mbar_init();
    }
    for (int i = tid; i < 4 * y_bytes(NT) / 16; i += THREADS) reinterpret_cast<uint4*>(ys)[i] = make_uint4(0, 0, 0, 0);
    if (warp == MMA_WARP) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&sm.tmem_base)), "n"(TMEM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = __shfl_sync(0xffffffffu, sm.tmem_base, 0);
    if (rank >= BN) {                                               // nothing to do for this CTA (tiny batches)
        __syncthreads();
        if (warp == MMA_WARP) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(TMEM_COLS) : "memory");
        return;
    }
    const int n_it = (BN - rank + nz - 1) / nz;                     // items rank, rank + nz, ...

    if (warp < EPI_WARPS) asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(EPI_REGS));
    else asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(SVC_REGS));

    if (warp >= MMA_WARP && warp < MMA_WARP + MMA_WARPS) {
        // ===================================== MMA issuer of stream s = warp - 16 =====================================
        // Software pipeline across items (per stream s):   ... C3(i) | E(i+1) | C4(i) | C3(i+1) ...
        // The embed of the NEXT item is issued between cnn3 and cnn4 of the current one (own accumulator, operand tiles
        // double-buffered by item parity), so every epilogue has at least one full conv stage of the other stream to
        // hide behind and the tensor pipe never waits for the short embed stage's epilogue.
        const int s = warp - MMA_WARP;
        if (s == 0 && lane == 0) {
            mbar_expect_tx(&sm.w_full, (uint32_t)zone_bytes(NT));
            bulk_g2s(wz, a.packed + (size_t)z * zone_bytes(NT), (uint32_t)zone_bytes(NT), &sm.w_full);
        }
        __syncwarp();
        mbar_wait(&sm.w_full, 0);
        const uint32_t w_emb = smem_u32(wz), w_c3 = w_emb + EMB_CHUNKS * LB, w_c4 = w_c3 + CONV_CHUNKS * LB;
        long long pw[4] = {0, 0, 0, 0};
        const long long t_start = PROF ? clock64() : 0;

        auto issue_embed = [&](int it) {                     // collapsed cnn1 o cnn2 on the electrode tile
            long long c0 = 0;
            if (PROF) c0 = clock64();
            mbar_wait(&sm.x_full[s], it & 1);
            if (PROF) pw[0] += clock64() - c0;
            tc_fence_after();
            const uint32_t d = tmem + (2 + s) * DCOLS;
            // descriptors = one base per operand + compile-time offsets in 16-byte units (keeps the issuer's live state tiny)
            const uint32_t xb = smem_u32(xs) + s * x_bytes(NT);
            const uint32_t wd = desc_lo(w_emb, LB);
            uint32_t acc = 0;
            if (n_xchunk == 2) {
                const uint32_t xd = desc_lo(xb, RX * 16);
#pragma unroll
                for (int k = 0; k < TAPS; ++k)
#pragma unroll
                    for (int t = 0; t < NT; ++t) {
                        umma(d, xd + (t * (2 * RX * 16) + k * 16) / 16, wd + (2 * k * LB) / 16, idesc_bf16(32 * (NT - t)), acc);
                        acc = 1;
                    }
            } else {                                                // 8 electrodes: one K16 step = taps (2p, 2p+1), LBO = one row
                const uint32_t xd = desc_lo(xb, 16);
#pragma unroll
                for (int p = 0; p < 3; ++p)
#pragma unroll
                    for (int t = 0; t < NT; ++t) {
                        umma(d, xd + (t * (2 * RX * 16) + 2 * p * 16) / 16, wd + (2 * p * LB) / 16, idesc_bf16(32 * (NT - t)), acc);
                        acc = 1;
                    }
            }
            umma_commit(&sm.de_full[s]);
            umma_commit(&sm.x_empty[s]);
        };
        auto issue_conv = [&](int layer, int it) {           // cnn3 (layer 0) / cnn4 (layer 1) on the conv operand tile
            long long c0 = 0, c1 = 0;
            if (PROF) c0 = clock64();
            if (layer == 0) {
                mbar_wait(&sm.ye_ready[s], it & 1);
                if (PROF) { c1 = clock64(); pw[2] += c1 - c0; }
                if (it > 0) mbar_wait(&sm.d_free[s], (it - 1) & 1);  // GELU epilogue of the previous item drained D
                if (PROF) pw[1] += clock64() - c1;
            } else {
                mbar_wait(&sm.yc_ready[s], it & 1);
                if (PROF) pw[3] += clock64() - c0;
            }
            tc_fence_after();
            const uint32_t wl = layer == 0 ? w_c3 : w_c4;
            const uint32_t d = tmem + s * DCOLS;
            const uint32_t yd = desc_lo(smem_u32(ys) + (2 * s + (it & 1)) * y_bytes(NT), RY * 16);
            const uint32_t wd = desc_lo(wl, LB);
            uint32_t acc = 0;
#pragma unroll
            for (int k = 0; k < TAPS; ++k)
#pragma unroll
                for (int h = 0; h < 2; ++h)
#pragma unroll
                    for (int t = 0; t < NT; ++t) {
                        umma(d, yd + (t * (4 * RY * 16) + 2 * h * (RY * 16) + k * 16) / 16, wd + ((4 * k + 2 * h) * LB) / 16,
                             idesc_bf16(32 * (NT - t)), acc);
                        acc = 1;
                    }
            umma_commit(&sm.dm_full[s][layer]);
        };
        issue_embed(0);
        for (int it = 0; it < n_it; ++it) {
            issue_conv(0, it);
            if (it + 1 < n_it) issue_embed(it + 1);
            issue_conv(1, it);
        }
        if (PROF && lane == 0 && s == 0) {
            long long* o = a.prof + (size_t)blockIdx.x * 16;
            o[0] = clock64() - t_start; o[1] = pw[0]; o[2] = pw[1]; o[3] = pw[2]; o[4] = pw[3]; o[5] = n_it;
        }
    } else if (warp >= COPY_WARP0) {
        // ============================ electrode loaders: unfold + zone gather + split ============================
        // Copier warp of stream s (warps 18, 19) streams item after item of its stream's electrode tile global -> shared
        // staging with 4-byte cp.async (no alignment requirement on the trial rows; zero fill outside the zone / window),
        // a ring of STG tiles deep, so the global latency costs neither registers nor issue slots.  Two converter warps
        // per stream (warps 20-23) turn a landed tile into the split-bf16 K-major operand tile.  Copy and conversion are
        // different warps because the fence.proxy.async after the converter's stores would otherwise wait for the
        // copier's in-flight prefetches and serialise them.
        const int* chan = a.chan + a.zoff[z];
        const int n_ch = 8 * n_xchunk;
        if (warp < CONV_WARP0) {
            const int s = warp - COPY_WARP0;
            const uint32_t stg_u32 = smem_u32(ys + 4 * y_bytes(NT)) + s * STG * STG_BYTES;
            const int tb = s == 0 ? 0 : a.tb1;
#pragma unroll 1
            for (int it = 0; it < n_it; ++it) {
                const int slot = it % STG;
                if (it >= STG) mbar_wait(&sm.stg_empty[s][slot], ((it / STG) - 1) & 1);
                const int item = rank + it * nz;
                const int b = item / a.n_win, n = item - b * a.n_win;
                const float* xb = a.x + (size_t)b * a.C * a.T + (size_t)n * a.step + tb + lane;
                const uint32_t dst0 = stg_u32 + slot * STG_BYTES + lane * 4;
#pragma unroll 1
                for (int c = 0; c < n_ch; ++c) {
                    const float* p = xb + (size_t)chan[min(c, Cz - 1)] * a.T;
                    const bool ch_ok = c < Cz;
#pragma unroll
                    for (int k = 0; k < (RX + 31) / 32; ++k) {
                        const int r = lane + 32 * k;
                        if (r < RX) {
                            const uint32_t nbytes = (ch_ok && tb + r < a.W) ? 4u : 0u;
                            asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(dst0 + (c * RX + 32 * k) * 4),
                                         "l"(nbytes ? p + 32 * k : a.x), "r"(nbytes) : "memory");
                        }
                    }
                }
                asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(smem_u32(&sm.stg_full[s][slot])) : "memory");
            }
            asm volatile("cp.async.wait_all;" ::: "memory");
        } else {
            const int s = (warp - CONV_WARP0) >> 1, half = (warp - CONV_WARP0) & 1;
            const float* stg = reinterpret_cast<const float*>(ys + 4 * y_bytes(NT)) + s * STG * (STG_BYTES / 4);
            uint8_t* xt = xs + s * x_bytes(NT);
            long long cw[3] = {0, 0, 0};
#pragma unroll 1
            for (int it = 0; it < n_it; ++it) {
                const int slot = it % STG;
                long long q0 = 0, q1 = 0, q2 = 0;
                if (PROF) q0 = clock64();
#if !defined(FB_NO_GROUP_WAIT) && !defined(FB_NO_CONV_GROUP_WAIT)
                if (half == 0) {                                    // one poller per stream, its partner parks in a named barrier
                    mbar_wait(&sm.stg_full[s][slot], (it / STG) & 1);
                    if (PROF) q1 = clock64();
                    if (it > 0) mbar_wait(&sm.x_empty[s], (it - 1) & 1);
                }
                asm volatile("bar.sync %0, 64;" ::"r"(3 + s) : "memory");
                if (PROF) { q2 = clock64(); cw[0] += q1 - q0; cw[1] += q2 - q1; }
#else
                mbar_wait(&sm.stg_full[s][slot], (it / STG) & 1);
                if (it > 0) mbar_wait(&sm.x_empty[s], (it - 1) & 1);
#endif
                const float* sg = stg + slot * (STG_BYTES / 4);
                for (int j = 0; j < n_xchunk; ++j)
#pragma unroll
                    for (int k = 0; k < 3; ++k) {
                        const int r = lane + 32 * (2 * k + half);
                        if (r < RX) {
                            float v[8];
#pragma unroll
                            for (int e = 0; e < 8; ++e) v[e] = sg[(8 * j + e) * RX + r];
                            uint4 o[NT];
                            split8<NT>(v, o);
#pragma unroll
                            for (int tm = 0; tm < NT; ++tm) *reinterpret_cast<uint4*>(xt + tm * (2 * RX * 16) + j * (RX * 16) + r * 16) = o[tm];
                        }
                    }
                fence_proxy_async();
                mbar_arrive(&sm.x_full[s]);
                mbar_arrive(&sm.stg_empty[s][slot]);
                if (PROF) cw[2] += clock64() - q2;
            }
            if (PROF && warp == CONV_WARP0 && lane == 0) {
                long long* o = a.prof + (size_t)blockIdx.x * 16;
                o[13] = cw[0]; o[14] = cw[1]; o[15] = cw[2];
            }
        }
    } else {
        // ============================ epilogue warps (stream s, lane quadrant q, column half h) ============================
        const int s = warp >> 3, h = (warp >> 2) & 1, q = warp & 3;
        const int m = 32 * q + lane;                                // accumulator row = TMEM lane
        const int t = (s == 0 ? 0 : a.tb1) + m;                     // time sample of this row
        const uint32_t lane_addr = tmem + ((uint32_t)(32 * q) << 16) + CW * h;
        uint8_t* yt0 = ys + 2 * s * y_bytes(NT) + (m + 2) * 16 + 2 * h * (RY * 16);
        const float* beff = reinterpret_cast<const float*>(wz + (EMB_CHUNKS + 2 * CONV_CHUNKS) * LB) + CW * h;
        const bool live = t < a.T_out;                              // rows past the window are the next conv's zero padding
        const bool own = live && (s == 0 ? t < a.split : t >= a.split);
        // Waits of the stream's 8 epilogue warps: ONE warp polls the mbarrier, the other seven park in a named hardware
        // barrier.  A polling warp re-issues ~7 instructions every ~30 cycles; with all 16 epilogue warps spinning (ncu, round 2:
        // 3,000 SYNCS + 1,500 NANOSLEEP of the 20,900 instructions per item, and every SYNCS is a shared-memory wavefront
        // competing with the MMA operand fetch) the spinners took as many issue slots as the warps doing the epilogue
        // arithmetic they were waiting for.  The pollers sit on sub-partitions 2 / 3, away from the two MMA issuers.
        const bool poller = (warp & 7) == 2 + s;
        auto group_wait = [&](uint64_t* bar, uint32_t parity) {
#ifndef FB_NO_GROUP_WAIT
            if (poller) mbar_wait(bar, parity);
            asm volatile("bar.sync %0, 256;" ::"r"(1 + s) : "memory");
#else
            mbar_wait(bar, parity);
#endif
        };
        group_wait(&sm.w_full, 0);                                  // beff visible
        // this warp's 16 collapsed-bias values live in registers for NT = 2 (16 shared-memory loads per thread and item on the
        // critical path between the TMEM drain and the operand-tile store otherwise: -16 % kernel time, round 2)
        // (NT = 3 keeps them in shared memory: its drain holds 48 accumulator values and the extra 16 registers cost more
        // than the loads -- 10.23 vs 10.39 ms per 8192 trials.)
        float be_reg[NT == 2 ? CW : 1];
        if (NT == 2) {
#pragma unroll
            for (int i = 0; i < CW; ++i) be_reg[NT == 2 ? i : 0] = beff[i];
        }
        const float* be = NT == 2 ? be_reg : beff;
        long long ew = 0, ep[5] = {0, 0, 0, 0, 0};

        // accumulator (NT column groups) -> v[CW]
        auto drain = [&](uint32_t d_addr, float (&v)[CW]) {
            float u[NT - 1][CW];
            tmem_ld16(d_addr, v);
#pragma unroll
            for (int tm = 1; tm < NT; ++tm) tmem_ld16(d_addr + 32 * tm, u[tm - 1]);
            tmem_ld_wait();
#pragma unroll
            for (int tm = 1; tm < NT; ++tm)
#pragma unroll
                for (int i = 0; i < CW; i += 2) {
                    const f32x2 sum = add2(pk(v[i], v[i + 1]), pk(u[tm - 1][i], u[tm - 1][i + 1]));
                    unpk(sum, v[i], v[i + 1]);
                }
        };
        // v (+ bias) -> masked -> bf16 terms -> this item's operand tile; signals `ready`
        // training forward: row t of stage `stage` (0: y2, 1: y3, 2: y4) of this item, columns [CW h, CW h + CW) -- each row is
        // written by the stream that owns it (both streams hold valid values there, see the halo argument in the header)
        auto save_rows = [&](const float (&v)[CW], bool add_bias, int it, int stage) {
            if (!own) return;
            const size_t item = (size_t)(rank + it * nz) * a.Z + z;
            float* dst = a.saved + (item * 3 + stage) * (size_t)(F * a.TS) + (size_t)(CW * h) * a.TS + t;
#pragma unroll
            for (int i = 0; i < CW; ++i) dst[(size_t)i * a.TS] = add_bias ? v[i] + be[i] : v[i];
        };
        // SAVE (training forward), stage 0 / 1 (y2 / y3): the first two bf16 terms just computed ARE the operand image the
        // weight-gradient kernel needs (head_wgrad_tc.cu: A operand, plane = term * 4 + channel group, one 16-byte row per time
        // sample), so they are stored as such -- [term 2][channel group 4][t < T_out][8 x bf16] inside the item's stage slot
        // (128 T_out bytes of its 128 TS) -- instead of as fp32: same bytes, four 16-byte stores per thread instead of sixteen
        // scalar ones, and no fp32 -> bf16 conversion pass in the backward.  Each row is written by the stream that owns it.
        auto to_operand = [&](float (&v)[CW], bool add_bias, int it, uint64_t* ready, int stage) {
            uint8_t* yt = yt0 + (it & 1) * y_bytes(NT);
            uint8_t* img = nullptr;
            if (SAVE && own) img = reinterpret_cast<uint8_t*>(a.saved + (((size_t)(rank + it * nz) * a.Z + z) * 3 + stage) * (size_t)(F * a.TS));
            if (!(PROF && (a.dbg & 1)))
#pragma unroll
            for (int j = 0; j < CW / 8; ++j) {
                float x8[8];
#pragma unroll
                for (int e = 0; e < 8; ++e) {
                    float val = v[8 * j + e];
                    if (add_bias) val += be[8 * j + e];
                    x8[e] = live ? val : 0.f;
                }
                uint4 o[NT];
                split8<NT>(x8, o);
#pragma unroll
                for (int tm = 0; tm < NT; ++tm) *reinterpret_cast<uint4*>(yt + tm * (4 * RY * 16) + j * (RY * 16)) = o[tm];
                if (SAVE && img != nullptr) {
#pragma unroll
                    for (int tm = 0; tm < 2; ++tm)
                        *reinterpret_cast<uint4*>(img + ((size_t)(tm * 4 + 2 * h + j) * a.T_out + t) * 16) = o[tm];
                }
            }
            fence_proxy_async();
            tc_fence_before();
            mbar_arrive(ready);
        };
        auto epi_embed = [&](int it) {
            long long c0 = 0;
            if (PROF) c0 = clock64();
            group_wait(&sm.de_full[s], it & 1);
            if (PROF) ew += clock64() - c0;
            tc_fence_after();
            float v[CW];
            drain(lane_addr + (2 + s) * DCOLS, v);
            to_operand(v, true, it, &sm.ye_ready[s], 0);
            if (PROF) ep[0] += clock64() - c0;
        };
        auto epi_cnn3 = [&](int it) {
            long long c0 = 0;
            if (PROF) c0 = clock64();
            group_wait(&sm.dm_full[s][0], it & 1);
            if (PROF) ew += clock64() - c0;
            tc_fence_after();
            float v[CW];
            drain(lane_addr + s * DCOLS, v);
            to_operand(v, false, it, &sm.yc_ready[s], 1);
            if (PROF) ep[1] += clock64() - c0;
        };
        auto epi_cnn4 = [&](int it) {
            long long c0 = 0, c1 = 0;
            if (PROF) c0 = clock64();
            group_wait(&sm.dm_full[s][1], it & 1);
            if (PROF) { c1 = clock64(); ew += c1 - c0; }
            tc_fence_after();
            float v[CW];
            drain(lane_addr + s * DCOLS, v);
            tc_fence_before();
            mbar_arrive(&sm.d_free[s]);                             // accumulator drained: next item's cnn3 may overwrite it
            if (SAVE) save_rows(v, false, it, 2);
#pragma unroll
            for (int i = 0; i < CW; i += 2) {
                v[i] = own ? v[i] : 0.f;                            // gelu(0) == 0: rows this stream does not own add nothing
                v[i + 1] = own ? v[i + 1] : 0.f;
                if (!(PROF && (a.dbg & 2))) gelu2(v[i], v[i + 1]);
            }
            // sum over the warp's 32 rows by recursive halving (16 shuffles): lanes 2c, 2c+1 end up with column c
#pragma unroll
            for (int st = 16; st >= 2; st >>= 1) {
                const bool upper = (lane & st) != 0;
#pragma unroll
                for (int i = 0; i < st / 2; ++i) {
                    const float keep = upper ? v[i + st / 2] : v[i];
                    const float send = upper ? v[i] : v[i + st / 2];
                    v[i] = keep + __shfl_xor_sync(0xffffffffu, send, st);
                }
            }
            v[0] += __shfl_xor_sync(0xffffffffu, v[0], 1);
            const int par = it & 1;
            long long c2 = 0;
            if (PROF) { c2 = clock64(); ep[3] += c2 - c1; }         // drain + GELU + row reduction
            // the two streams have independent issuers and may drift apart: do not touch the slot of item `it` before the
            // feature of item it - 2 (same slot) has been written
            if (it >= 2) group_wait(&sm.red_free[par], ((it >> 1) - 1) & 1);
            if (PROF) ep[4] += clock64() - c2;                      // wait for the partial-sum slot
            if ((lane & 1) == 0) sm.red[par][warp][lane >> 1] = v[0];
            __syncwarp();
            int old = 0;
            if (lane == 0) {
                __threadfence_block();
                old = atomicAdd(&sm.counter[par], 1);
            }
            old = __shfl_sync(0xffffffffu, old, 0);
            if (old == EPI_WARPS - 1) {                             // last of the 16 warps: fixed-order sum -> deterministic
                __threadfence_block();
                float tot = 0.f;                                    // lane = output channel co = 16 h' + c
#pragma unroll
                for (int ss = 0; ss < 2; ++ss)
#pragma unroll
                    for (int qq = 0; qq < 4; ++qq) tot += sm.red[par][ss * 8 + (lane >> 4) * 4 + qq][lane & 15];
                const int item = rank + it * nz;
                a.feat[((size_t)item * a.Z + z) * F + lane] = tot / (float)a.T_out;
                __syncwarp();
                if (lane == 0) {
                    sm.counter[par] = 0;
                    mbar_arrive(&sm.red_free[par]);
                }
            }
            if (PROF) ep[2] += clock64() - c1;
        };
        epi_embed(0);
        for (int it = 0; it < n_it; ++it) {
            epi_cnn3(it);
            if (it + 1 < n_it) epi_embed(it + 1);
            epi_cnn4(it);
        }
        if (PROF && tid == 0) { long long* o = a.prof + (size_t)blockIdx.x * 16; o[6] = ew; o[8] = ep[0]; o[9] = ep[1]; o[10] = ep[2]; o[11] = ep[3]; o[12] = ep[4]; }
        if (PROF && tid == 256) a.prof[(size_t)blockIdx.x * 16 + 7] = ew;
    }
    tc_fence_before();
    __syncthreads();
    if (warp == MMA_WARP) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(TMEM_COLS) : "memory");
}

}  // namespace tc
}  // namespace fb

// ---- host side -------------------------------------------------------------------------------------------------------
using namespace fb;

static int tc_supported(const fastb200_model* m, int n_terms) {
    FB_REQUIRE(m != nullptr, "model descriptor is NULL");
    FB_REQUIRE(n_terms == 2 || n_terms == 3, "n_terms=%d: the split-bf16 head supports 2 or 3 terms", n_terms);
    FB_REQUIRE(m->dim_cnn == tc::F, "tensor-core head needs dim_cnn == 32 (got %d)", m->dim_cnn);
    FB_REQUIRE(m->n_zones >= 1 && m->n_zones <= FASTB200_MAX_ZONES, "n_zones=%d out of range", m->n_zones);
    const int T_out = m->window_len - 4;
    FB_REQUIRE(T_out >= 128 && T_out <= 248, "tensor-core head needs 132 <= window_len <= 252 (got %d)", m->window_len);
    for (int z = 0; z < m->n_zones; ++z)
        FB_REQUIRE(m->zone_size[z] >= 1 && m->zone_size[z] <= 16, "zone %d has %d electrodes (tensor-core head: 1..16)", z, m->zone_size[z]);
    return 0;
}

extern "C" int fastb200_head_tc_supported(const fastb200_model* m) { return tc_supported(m, 2) == 0 ? 1 : 0; }

extern "C" int64_t fastb200_head_packed_tc_bytes(const fastb200_model* m, int n_terms) {
    if (tc_supported(m, n_terms)) return -1;
    return (int64_t)m->n_zones * tc::zone_bytes(n_terms);
}

extern "C" int fastb200_head_pack_tc(const fastb200_model* m, const float* params, void* packed, int n_terms, void* stream) {
    if (int r = tc_supported(m, n_terms)) return r;
    // the tensor-core path requires F = 32 and zones of <= 16 electrodes: the staged cnn2.weight is at most 65.7 KB
    const size_t smem = (size_t)tc::F * (tc::F * 16 + 1) * sizeof(float);
    cudaError_t e = cudaFuncSetAttribute(tc::head_pack_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return fb::fail((int)e, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    tc::head_pack_tc_kernel<<<dim3(m->n_zones, 8), 256, smem, as_stream(stream)>>>(*m, params, static_cast<uint8_t*>(packed), n_terms);
    return check_launch("head_pack_tc");
}

#ifdef FASTB200_PROFILE
// Developer builds only (nvcc -DFASTB200_PROFILE): diagnostic buffer for the next fastb200_head_fwd_tc calls of this thread
// (148 x 16 int64 wait-cycle counters per launch); NULL = uninstrumented kernel.  The shipped library neither exports this
// symbol nor instantiates the instrumented kernels nor reads the environment.
static thread_local long long* g_tc_prof = nullptr;
extern "C" void fastb200_debug_set_tc_profile(long long* dev_buf) { g_tc_prof = dev_buf; }
#endif

static int head_fwd_tc_launch(const fastb200_model* m, const void* packed, const float* x, int B, int T, int n_win,
                              int step, float* feat, float* saved, int n_terms, void* stream) {
    if (int r = tc_supported(m, n_terms)) return r;
    FB_REQUIRE(B >= 0 && n_win >= 1 && step >= 1, "bad B/n_win/step");
    FB_REQUIRE((n_win - 1) * (int64_t)step + m->window_len <= T, "windows run past the trial: n_win=%d step=%d W=%d T=%d", n_win, step, m->window_len, T);
    FB_REQUIRE((int64_t)B * n_win < (1ll << 31) / FASTB200_MAX_ZONES, "too many items for one launch");
    if (B == 0) return 0;
    tc::FwdArgs a;
    a.packed = static_cast<const uint8_t*>(packed); a.x = x; a.feat = feat; a.prof = nullptr; a.dbg = 0;
    a.saved = saved; a.TS = head_saved_stride(m->window_len);
#ifdef FASTB200_PROFILE
    a.prof = g_tc_prof;
    a.dbg = (g_tc_prof && getenv("FASTB200_TC_DBG")) ? atoi(getenv("FASTB200_TC_DBG")) : 0;
#endif
    a.B = B; a.T = T; a.n_win = n_win; a.step = step; a.C = m->n_chan; a.W = m->window_len; a.Z = m->n_zones;
    a.T_out = m->window_len - 4; a.split = a.T_out / 2; a.tb1 = a.T_out - 128;
    int n_sel = 0;
    for (int z = 0; z < m->n_zones; ++z) { a.cz[z] = m->zone_size[z]; a.zoff[z] = m->zone_off[z]; n_sel += m->zone_size[z]; }
    for (int i = 0; i < n_sel; ++i) a.chan[i] = m->zone_chan[i];
    // apportion the CTAs to zones in proportion to per-item MMA cost (largest remainder), at least one each
    const int Z = m->n_zones;
    const int grid = NUM_SMS >= Z ? NUM_SMS : Z;
    double cost[FASTB200_MAX_ZONES], total = 0;
    for (int z = 0; z < Z; ++z) { cost[z] = 20.0 + (m->zone_size[z] <= 8 ? 3.0 : 5.0); total += cost[z]; }
    int n_cta[FASTB200_MAX_ZONES], used = 0;
    double frac[FASTB200_MAX_ZONES];
    for (int z = 0; z < Z; ++z) {
        const double share = cost[z] / total * grid;
        n_cta[z] = (int)share < 1 ? 1 : (int)share;
        frac[z] = share - n_cta[z];
        used += n_cta[z];
    }
    while (used < grid) {
        int best = 0;
        for (int z = 1; z < Z; ++z) if (frac[z] > frac[best]) best = z;
        n_cta[best]++; frac[best] -= 1.0; used++;
    }
    while (used > grid) {                                    // only when the minimum of one CTA per zone overshoots
        int best = 0;
        for (int z = 1; z < Z; ++z) if (n_cta[z] > n_cta[best]) best = z;
        n_cta[best]--; used--;
    }
    // interleave the zones over the CTA index so that co-resident waves mix zones
    int given[FASTB200_MAX_ZONES] = {0};
    int c = 0;
    while (c < grid)
        for (int z = 0; z < Z && c < grid; ++z)
            if (given[z] < n_cta[z]) { a.cta_zone[c] = (uint8_t)z; a.cta_rank[c] = (uint8_t)given[z]++; ++c; }
    for (int z = 0; z < Z; ++z) a.zone_ctas[z] = (uint8_t)n_cta[z];
    const size_t smem = tc::SMEM_HDR + tc::zone_bytes(n_terms) + 2 * tc::x_bytes(n_terms) + 4 * tc::y_bytes(n_terms) +
                        2 * tc::stg_tiles(n_terms) * tc::STG_BYTES;
    cudaError_t e;
#define LAUNCH_TC(NTT, PROFF, SAVEE)                                                                               \
    e = cudaFuncSetAttribute(tc::head_fwd_tc_kernel<NTT, PROFF, SAVEE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
    if (e != cudaSuccess) return fail((int)e, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));                 \
    tc::head_fwd_tc_kernel<NTT, PROFF, SAVEE><<<grid, tc::THREADS, smem, as_stream(stream)>>>(a);
    if (saved)       { if (n_terms == 2) { LAUNCH_TC(2, false, true) } else { LAUNCH_TC(3, false, true) } }
#ifdef FASTB200_PROFILE
    else if (a.prof) { if (n_terms == 2) { LAUNCH_TC(2, true, false) } else { LAUNCH_TC(3, true, false) } }
#endif
    else             { if (n_terms == 2) { LAUNCH_TC(2, false, false) } else { LAUNCH_TC(3, false, false) } }
#undef LAUNCH_TC
    return check_launch("head_fwd_tc");
}

extern "C" int fastb200_head_fwd_tc(const fastb200_model* m, const void* packed, const float* x, int B, int T, int n_win,
                                    int step, float* feat, int n_terms, void* stream) {
    return head_fwd_tc_launch(m, packed, x, B, T, n_win, step, feat, nullptr, n_terms, stream);
}

extern "C" int fastb200_head_fwd_tc_train(const fastb200_model* m, const void* packed, const float* x, int B, int T, int n_win,
                                          int step, float* feat, float* saved, int n_terms, void* stream) {
    FB_REQUIRE(saved != nullptr || B == 0, "head_fwd_tc_train needs the activation workspace (fastb200_head_saved_floats)");
    return head_fwd_tc_launch(m, packed, x, B, T, n_win, step, feat, saved, n_terms, stream);
}
