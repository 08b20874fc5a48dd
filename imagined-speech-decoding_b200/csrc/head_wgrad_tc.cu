// Tensor-core (tcgen05 / TMEM) weight gradients of cnn3 / cnn4 of the FAST zone embedding head.
//
// Reference: autograd of Conv4Layers.forward (src/fast/models/fast.py:115-116) -- for one (trial, window, zone) item
//     dW[ci][tap][co] = sum_t  ypad[ci][t + tap] * dY[co][t]        (ypad = layer input, zero padded by 2 at the window edges)
// i.e. per tap a (32 x T) x (T x 32) contraction over TIME, accumulated over every item of the zone.  The FFMA backward
// (head.cu) spends most of its time here (20 accumulators per thread fed by broadcast shared-memory loads); on the tensor
// pipe it is 80 MMAs per item:
//
//   * both operands are "MN-major": the shared-memory tile [plane = term*4 + c/8][row = time][8 channels x bf16] keeps
//     time on the 16-byte row index, so a K16 step is 16 consecutive rows (LBO = 128 B between the two 8-row core
//     matrices, SBO = one plane) and the temporal tap is a +16 B start-address offset of the A descriptor -- the same tile
//     serves all five taps (tools/umma_wgrad_probe.cu checks the descriptor semantics on the hardware);
//   * fp32 accuracy from bf16 MMAs by operand splitting: A = [a0; a1] stacked along M (lanes 0-31 / 32-63; lanes 64-127 of
//     the M = 128 tile read whatever follows the A tile and are never drained), B = [b0 | b1] stacked along N (64 columns):
//     ONE MMA forms all four products a_i b_j; the drain adds the four 32 x 32 blocks.  Dropped terms are O(2^-17) relative
//     and of random sign over the 246 x items products of a sum;
//   * the accumulators (5 taps x 64 columns = 320 TMEM columns) stay in TMEM for the CTA's whole item list and are added to
//     the packed gradient buffer once, at the end;
//   * persistent CTAs, one (zone, layer) pair each; the fp32 activations / gradients ([32][TS] per item, written by the
//     forward and by head_bwd_kernel) arrive by 1-D bulk async copies into a 2-slot staging ring, eight converter warps turn
//     them into split-bf16 tiles (double buffered), one elected thread issues the MMAs.
#include "common.cuh"
#include "tc_ptx.cuh"
#include <stdlib.h>

namespace fb {
namespace wg {

using namespace fb::tc;

constexpr int F = 32;
constexpr int R = 260;                          // tile rows: 2 zero rows + up to 256 time samples (16 K16 steps) + 2
constexpr int PLANE = R * 16;                   // bytes of one [row][8 x bf16] plane
constexpr int TILE = 8 * PLANE;                 // 2 terms x 4 channel groups
constexpr int CONV_WARPS = 8, THREADS = (2 + CONV_WARPS) * 32;
constexpr int NCOL = 64;                        // accumulator columns per tap: [b0 | b1]
constexpr int TMEM_COLS = 512;                  // 5 x 64 = 320 used
constexpr int HDR = 1024;

// D fp32, A/B bf16, both MN-major (bits 15 / 16), M = 128, N = 64
constexpr uint32_t IDESC = (1u << 4) | (1u << 7) | (1u << 10) | (1u << 15) | (1u << 16) | ((uint32_t)(NCOL >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
constexpr uint32_t DESC_HI_MN = (uint32_t)(PLANE >> 4) | (1u << 14);          // SBO = one plane, descriptor version 1
__device__ __forceinline__ uint32_t desc_lo_mn(uint32_t saddr) { return ((saddr & 0x3FFFFu) >> 4) | ((128u >> 4) << 16); }   // LBO = 128 B

__device__ __forceinline__ void umma_mn(uint32_t d_tmem, uint32_t a_lo, uint32_t b_lo, uint32_t acc, uint32_t desc_hi) {
    asm volatile("{\n\t.reg .pred p, q;\n\t.reg .b64 da, db;\n\t"
                 "elect.sync _|q, 0xffffffff;\n\t"
                 "setp.ne.b32 p, %4, 0;\n\t"
                 "mov.b64 da, {%1, %5};\n\t"
                 "mov.b64 db, {%2, %5};\n\t"
                 "@q tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %3, p;\n\t}"
                 ::"r"(d_tmem), "r"(a_lo), "r"(b_lo), "r"(IDESC), "r"(acc), "r"(desc_hi) : "memory");
}

struct Args {
    const float* saved;                        // per item 3 x [F][TS]: y2, y3, y4 (head_fwd)
    const float* dys;                          // per item 2 x [F][TS]: dL/dy4, dL/dy3 (head_bwd_kernel)
    float* gpacked;
    int BN, Z, TS, T_out, ksteps;
    uint32_t desc_hi, lbo16;                   // descriptor high word / LBO in 16-byte units (debug: FASTB200_WG_SWAP swaps LBO and SBO)
    int goff[FASTB200_MAX_ZONES][2];           // float offset of the packed cnn3 / cnn4 gradient block of zone z
    uint8_t cta_zone[NUM_SMS], cta_layer[NUM_SMS], cta_rank[NUM_SMS], cta_n[NUM_SMS];
};

struct Smem {
    uint64_t stg_full[2], stg_empty[2], tile_full[2], tile_empty[2], done;
    uint32_t tmem_base;
};

__global__ void __launch_bounds__(THREADS, 1) head_wgrad_tc_kernel(const __grid_constant__ Args a) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    Smem& sm = *reinterpret_cast<Smem*>(smem_raw);
    uint8_t* tiles = smem_raw + HDR;                                // [buf][A | B]
    uint8_t* stg = tiles + 4 * TILE;                                // [2 slots][F][TS] fp32
    const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0), lane = tid & 31;
    const int z = a.cta_zone[blockIdx.x], layer = a.cta_layer[blockIdx.x];        // layer 0: cnn3 (A = y2, B = dy3); 1: cnn4 (y3, dy4)
    const int rank = a.cta_rank[blockIdx.x], nc = a.cta_n[blockIdx.x];
    const int n_it = rank < a.BN ? (a.BN - rank + nc - 1) / nc : 0;
    const uint32_t tensor_bytes = (uint32_t)(F * a.TS * 4);

    if (tid == 0) {
        for (int i = 0; i < 2; ++i) {
            mbar_init(&sm.stg_full[i], 1);
            mbar_init(&sm.stg_empty[i], CONV_WARPS * 32);
            mbar_init(&sm.tile_full[i], 2 * CONV_WARPS * 32);
            mbar_init(&sm.tile_empty[i], 1);
        }
        mbar_init(&sm.done, 1);
        fence_mbar_init();
    }
    for (int i = tid; i < 4 * TILE / 16; i += THREADS) reinterpret_cast<uint4*>(tiles)[i] = make_uint4(0, 0, 0, 0);
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&sm.tmem_base)), "n"(TMEM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = __shfl_sync(0xffffffffu, sm.tmem_base, 0);
    const size_t item_sz = (size_t)F * a.TS;

    if (n_it > 0) {
        if (warp == 0) {
            // ---------------- producer: tensor loads A(0), B(0), A(1), B(1), ... into the 2-slot staging ring
            if (lane == 0) {
                for (int L = 0; L < 2 * n_it; ++L) {
                    const int slot = L & 1, it = L >> 1;
                    if (L >= 2) mbar_wait(&sm.stg_empty[slot], ((L >> 1) - 1) & 1);
                    const size_t item = (size_t)(rank + it * nc) * a.Z + z;
                    const float* src = (L & 1) == 0 ? a.saved + item * 3 * item_sz + (size_t)layer * item_sz        // y2 / y3
                                                    : a.dys + item * 2 * item_sz + (size_t)(1 - layer) * item_sz;   // dy3 / dy4
                    mbar_expect_tx(&sm.stg_full[slot], tensor_bytes);
                    bulk_g2s(stg + (size_t)slot * tensor_bytes, src, tensor_bytes, &sm.stg_full[slot]);
                }
            }
            __syncwarp();
        } else if (warp == 1) {
            // ---------------- MMA issuer (converged warp, one elected lane inside umma_mn)
#pragma unroll 1
            for (int it = 0; it < n_it; ++it) {
                const int buf = it & 1;
                mbar_wait(&sm.tile_full[buf], (it >> 1) & 1);
                tc_fence_after();
                const uint32_t ad = (((smem_u32(tiles) + buf * 2 * TILE) & 0x3FFFFu) >> 4) | (a.lbo16 << 16);
                const uint32_t bd = (((smem_u32(tiles) + buf * 2 * TILE + TILE) & 0x3FFFFu) >> 4) | (a.lbo16 << 16);
#pragma unroll 1
                for (int j = 0; j < a.ksteps; ++j) {
                    const uint32_t acc = (it > 0 || j > 0) ? 1u : 0u;
#pragma unroll
                    for (int k = 0; k < TAPS; ++k) umma_mn(tmem + k * NCOL, ad + 16 * j + k, bd + 16 * j, acc, a.desc_hi);
                }
                umma_commit(&sm.tile_empty[buf]);
            }
            umma_commit(&sm.done);
        } else {
            // ---------------- converters: staged fp32 [c][t] -> split-bf16 tile [term][c/8][row][8]
            const int ct = tid - 64;                                 // 0 .. 255
#pragma unroll 1
            for (int it = 0; it < n_it; ++it) {
                const int buf = it & 1;
                if (it >= 2) mbar_wait(&sm.tile_empty[buf], ((it >> 1) - 1) & 1);
#pragma unroll 1
                for (int half = 0; half < 2; ++half) {
                    const int L = 2 * it + half, slot = L & 1;
                    mbar_wait(&sm.stg_full[slot], (L >> 1) & 1);
                    const float* sg = reinterpret_cast<const float*>(stg + (size_t)slot * tensor_bytes);
                    uint8_t* tile = tiles + buf * 2 * TILE + half * TILE + (half == 0 ? 2 * 16 : 0);     // A rows are shifted by the 2 pad rows
                    for (int t = ct; t < a.T_out; t += CONV_WARPS * 32) {
#pragma unroll
                        for (int p = 0; p < 4; ++p) {
                            float v[8];
#pragma unroll
                            for (int e = 0; e < 8; ++e) v[e] = sg[(8 * p + e) * a.TS + t];
                            uint4 o[2];
                            split8<2>(v, o);
                            *reinterpret_cast<uint4*>(tile + p * PLANE + t * 16) = o[0];
                            *reinterpret_cast<uint4*>(tile + (4 + p) * PLANE + t * 16) = o[1];
                        }
                    }
                    fence_proxy_async();
                    mbar_arrive(&sm.stg_empty[slot]);
                    mbar_arrive(&sm.tile_full[buf]);
                }
            }
        }
        // ---------------- drain: lanes 0-31 hold the a0 rows, 32-63 the a1 rows; columns [b0 | b1] per tap
        if (warp < 2) {
            mbar_wait(&sm.done, 0);
            tc_fence_after();
            float* g = a.gpacked + a.goff[z][layer];
#pragma unroll 1
            for (int k = 0; k < TAPS; ++k) {
                float v0[32], v1[32];
                const uint32_t addr = tmem + ((uint32_t)(32 * warp) << 16) + k * NCOL;
                tmem_ld32(addr, v0);
                tmem_ld32(addr + 32, v1);
#pragma unroll
                for (int co = 0; co < 32; ++co) atomicAdd(g + (lane * TAPS + k) * F + co, v0[co] + v1[co]);
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(TMEM_COLS) : "memory");
}

}  // namespace wg
}  // namespace fb

using namespace fb;

static int wgrad_tc_supported(const fastb200_model* m) {
    FB_REQUIRE(m != nullptr, "model descriptor is NULL");
    FB_REQUIRE(m->dim_cnn == wg::F, "tensor-core weight gradient needs dim_cnn == 32 (got %d)", m->dim_cnn);
    FB_REQUIRE(m->n_zones >= 1 && m->n_zones <= FASTB200_MAX_ZONES, "n_zones=%d out of range", m->n_zones);
    const int T_out = m->window_len - 4;
    FB_REQUIRE(T_out >= 16 && T_out <= 256, "tensor-core weight gradient needs 20 <= window_len <= 260 (got %d)", m->window_len);
    return 0;
}

extern "C" int fastb200_head_bwd_tc_supported(const fastb200_model* m) { return wgrad_tc_supported(m) == 0 ? 1 : 0; }

extern "C" int64_t fastb200_head_bwd_tc_work_floats(const fastb200_model* m, int B, int n_win) {
    if (wgrad_tc_supported(m)) return -1;
    return (int64_t)B * n_win * m->n_zones * 2 * m->dim_cnn * head_saved_stride(m->window_len);
}

extern "C" int fastb200_head_bwd_tc(const fastb200_model* m, const float* packed, const float* x, const float* saved,
                                    const float* dfeat, int B, int T, int n_win, int step, float* gpacked, float* dx,
                                    float* work, void* stream) {
    if (int r = wgrad_tc_supported(m)) return r;
    if (B == 0) return 0;
    FB_REQUIRE(work != nullptr, "head_bwd_tc needs its workspace (fastb200_head_bwd_tc_work_floats)");
    // input gradients of cnn4 / cnn3, the collapsed-embed weight / bias gradients and the optional dx: CUDA-core kernel
    if (int r = head_dgrad_launch(m, packed, x, saved, dfeat, B, T, n_win, step, gpacked, dx, work, stream)) return r;
    wg::Args a;
    a.saved = saved; a.dys = work; a.gpacked = gpacked;
    a.BN = B * n_win; a.Z = m->n_zones; a.TS = head_saved_stride(m->window_len); a.T_out = m->window_len - 4;
    a.ksteps = ceil_div(a.T_out, 16);
    a.desc_hi = wg::DESC_HI_MN; a.lbo16 = 128u >> 4;
    if (const char* sw = getenv("FASTB200_WG_SWAP")) if (atoi(sw)) { a.desc_hi = (128u >> 4) | (1u << 14); a.lbo16 = (uint32_t)(wg::PLANE >> 4); }
    const int F = m->dim_cnn;
    for (int z = 0; z < m->n_zones; ++z) {
        const PackedZone p = packed_zone(*m, z);
        a.goff[z][0] = p.w3; a.goff[z][1] = p.w4;
    }
    (void)F;
    const int pairs = 2 * m->n_zones;
    const int grid = NUM_SMS >= pairs ? NUM_SMS : pairs;
    int c = 0;
    for (int p = 0; p < pairs; ++p) {
        const int n = grid / pairs + (p < grid % pairs ? 1 : 0);
        for (int r = 0; r < n; ++r, ++c) {
            a.cta_zone[c] = (uint8_t)(p >> 1); a.cta_layer[c] = (uint8_t)(p & 1); a.cta_rank[c] = (uint8_t)r; a.cta_n[c] = (uint8_t)n;
        }
    }
    const size_t smem = wg::HDR + 4 * wg::TILE + 2 * (size_t)wg::F * a.TS * 4;
    cudaError_t e = cudaFuncSetAttribute(wg::head_wgrad_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return fail((int)e, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    wg::head_wgrad_tc_kernel<<<grid, wg::THREADS, smem, as_stream(stream)>>>(a);
    return check_launch("head_wgrad_tc");
}
