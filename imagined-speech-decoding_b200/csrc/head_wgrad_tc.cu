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
    const float* dys;                          // per item 3 x [F][TS]: dL/dy4, dL/dy3, dL/dy2 (head_dgrad_kernel)
    const float* x;                            // (B, C, T) trials: A operand of the collapsed-embed layer (layer 2)
    const float* dfeat;                        // (B, n_win, Z, F): with dy4_from_y4, dL/dy4 = dfeat / T_out * gelu'(y4) is formed in the converter
    int dy4_from_y4;                           // 1: the cnn4 layer's B operand is recomputed from the saved y4 (dL/dy4 never went to HBM)
    float* gpacked;
    int BN, Z, TS, T_out, ksteps;
    int n_win, step, C, T, W;
    int cz[FASTB200_MAX_ZONES], zoff[FASTB200_MAX_ZONES], boff[FASTB200_MAX_ZONES];   // zone size / offset into chan / packed beff offset
    int chan[FASTB200_MAX_CHANS];
    uint32_t desc_hi, lbo16;                   // descriptor high word / LBO in 16-byte units
    int goff[FASTB200_MAX_ZONES][3];           // float offset of the packed cnn3 / cnn4 / collapsed-embed gradient block of zone z
    uint8_t cta_zone[NUM_SMS], cta_layer[NUM_SMS], cta_rank[NUM_SMS], cta_n[NUM_SMS];
    int dbg;                                   // developer builds: bit 0 no operand loads, bit 1 no MMAs, bit 2 no embed CTAs, bit 3 no conv CTAs (pipeline probes; wrong results)
};

// =====================================================================================================================
// Image-fed kernel (round 2, the shipped one).
//
// The forward (head_tc.cu, SAVE) and the input-gradient kernel (head_dgrad_tc.cu) now leave y2, y3 and dL/dy4, dL/dy3,
// dL/dy2 in HBM as the split-bf16 OPERAND IMAGES this kernel multiplies -- [term 2][channel group 4][t < T_out][8 x bf16],
// i.e. plane p = term * 4 + group of the tiles above, one 16-byte row per time sample -- so there is no fp32 staging ring
// and no conversion pass here: a producer lane fills the tiles of item i + 1 with bulk async copies (one per plane) while
// the MMA warp works on item i.  That also makes the second tap free: the A planes are copied twice, the second time one
// row earlier and directly behind the first copy (A2[p][r] = A[p][r + 1]), so the planes 8-15 that the M = 128 descriptor
// reads anyway hold the same operand at tap + 1 -- lanes 0-63 accumulate tap 2j, lanes 64-127 tap 2j + 1, three MMAs per
// K16 step instead of five (48 instead of 80 per item and layer; FB_WG_ONE_TAP keeps the five).
// Only the collapsed-embed layer still converts: its A operand is the zone's raw electrodes (fp32 trials), gathered and
// split by the eight converter warps (one time sample per thread, next item's 16 loads in flight), which also form the
// collapsed-bias gradient sum_t dL/dy2 from the landed B planes.
namespace im {

constexpr int R = 264;                          // tile rows: 2 zero rows + up to 256 time samples + 5 tap rows (+ spare)
constexpr int PLANE_I = R * 16;
constexpr int TILE_I = 8 * PLANE_I;
#ifdef FB_WG_ONE_TAP
constexpr bool TWO_TAP = false;
#else
constexpr bool TWO_TAP = true;
#endif
constexpr int NTILE = TWO_TAP ? 3 : 2;          // [A | (A one row earlier) | B]
constexpr int BUF_I = NTILE * TILE_I;
constexpr int NBUF = TWO_TAP ? 2 : 3;           // 202,752 bytes of tiles either way
constexpr int NACC = TWO_TAP ? 3 : TAPS;        // accumulators: tap pairs (0,1) (2,3) (4,-), or the five taps
constexpr int TMEM_COLS_I = TWO_TAP ? 256 : 512;
constexpr int CONV_T = 256, THREADS_I = CONV_T + 64;      // warp 0 producer, warp 1 MMA issuer, warps 2-9 converters / drain
constexpr uint32_t DESC_HI_I = (uint32_t)(PLANE_I >> 4) | (1u << 14);

struct SmemI {
    uint64_t full[NBUF], empty[NBUF], bfull[NBUF], done;
    uint32_t tmem_base;
    float bsum[F];
};

__global__ void __launch_bounds__(THREADS_I, 1) head_wgrad_img_kernel(const __grid_constant__ Args a) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    SmemI& sm = *reinterpret_cast<SmemI*>(smem_raw);
    uint8_t* tiles = smem_raw + HDR;                                // [buf][A | A2 | B]
    const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0), lane = tid & 31;
    const int z = a.cta_zone[blockIdx.x], layer = a.cta_layer[blockIdx.x];
    const bool emb = layer == 2;
    // Embed layer, zones of <= 15 electrodes: electrode slot 15 of the A tile carries the constant 1, so row 15 of the tap-0
    // accumulator IS sum_t dL/dy2 -- the collapsed-bias gradient comes out of the MMAs the layer issues anyway (1.0 is exact in
    // bf16) and the converter warps no longer read the landed dY2 planes to sum them (40 % of their instructions: the embed CTAs
    // were converter-bound and the kernel's long pole).
    const bool ones = emb && a.cz[z] <= 15;
    const int Cz = a.cz[z];
    const int rank = a.cta_rank[blockIdx.x], nc = a.cta_n[blockIdx.x];
    const int n_it = ((a.dbg & 4) && emb) || ((a.dbg & 8) && !emb) ? 0 : rank < a.BN ? (a.BN - rank + nc - 1) / nc : 0;   // dbg bits 2 / 3: skip a layer kind
    const size_t item_sz = (size_t)F * a.TS;                        // floats per stage slot
    const uint32_t plane_bytes = (uint32_t)a.T_out * 16u;           // one plane of an image

    if (tid == 0) {
        for (int i = 0; i < NBUF; ++i) {
            mbar_init(&sm.full[i], emb ? CONV_T : 1);
            mbar_init(&sm.empty[i], 1);
            mbar_init(&sm.bfull[i], 1);
        }
        mbar_init(&sm.done, 1);
        fence_mbar_init();
    }
    for (int i = tid; i < NBUF * BUF_I / 16; i += THREADS_I) reinterpret_cast<uint4*>(tiles)[i] = make_uint4(0, 0, 0, 0);
    if (tid < F) sm.bsum[tid] = 0.f;
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&sm.tmem_base)), "n"(TMEM_COLS_I) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = __shfl_sync(0xffffffffu, sm.tmem_base, 0);

    if (n_it > 0) {
        if (warp == 0) {
            // ---------------- producer: images -> tiles, one bulk async copy per plane
            if (lane == 0) {
                for (int it = 0; it < n_it; ++it) {
                    const int buf = it % NBUF, use = it / NBUF;
                    if (use >= 1) mbar_wait(&sm.empty[buf], (use - 1) & 1);
                    const size_t item = (size_t)(rank + it * nc) * a.Z + z;
                    uint8_t* A1 = tiles + buf * BUF_I;
                    uint8_t* Bt = A1 + (NTILE - 1) * TILE_I;
                    const uint8_t* srcB = reinterpret_cast<const uint8_t*>(a.dys + item * 3 * item_sz + (size_t)(emb ? 2 : 1 - layer) * item_sz);
                    if (a.dbg & 1) {
                        if (emb) mbar_arrive(&sm.bfull[buf]); else mbar_arrive(&sm.full[buf]);
                    } else if (emb) {
                        mbar_expect_tx(&sm.bfull[buf], 8 * plane_bytes);
                        for (int p = 0; p < 8; ++p) bulk_g2s(Bt + p * PLANE_I, srcB + (size_t)p * plane_bytes, plane_bytes, &sm.bfull[buf]);
                    } else {
                        const uint8_t* srcA = reinterpret_cast<const uint8_t*>(a.saved + item * 3 * item_sz + (size_t)layer * item_sz);
                        mbar_expect_tx(&sm.full[buf], (TWO_TAP ? 24u : 16u) * plane_bytes);
                        for (int p = 0; p < 8; ++p) {
                            bulk_g2s(A1 + p * PLANE_I + 2 * 16, srcA + (size_t)p * plane_bytes, plane_bytes, &sm.full[buf]);        // ypad: 2 zero rows in front
                            if (TWO_TAP) bulk_g2s(A1 + TILE_I + p * PLANE_I + 1 * 16, srcA + (size_t)p * plane_bytes, plane_bytes, &sm.full[buf]);
                            bulk_g2s(Bt + p * PLANE_I, srcB + (size_t)p * plane_bytes, plane_bytes, &sm.full[buf]);
                        }
                    }
                }
            }
            __syncwarp();
        } else if (warp == 1) {
            // ---------------- MMA issuer (converged warp, one elected lane inside umma_mn)
#pragma unroll 1
            for (int it = 0; it < n_it; ++it) {
                const int buf = it % NBUF, use = it / NBUF;
                mbar_wait(&sm.full[buf], use & 1);
                if (emb) mbar_wait(&sm.bfull[buf], use & 1);
                tc_fence_after();
                const uint32_t ad = (((smem_u32(tiles) + buf * BUF_I) & 0x3FFFFu) >> 4) | ((128u >> 4) << 16);
                const uint32_t bd = (((smem_u32(tiles) + buf * BUF_I + (NTILE - 1) * TILE_I) & 0x3FFFFu) >> 4) | ((128u >> 4) << 16);
#pragma unroll 1
                for (int j = 0; j < a.ksteps; ++j) {
                    const uint32_t acc = (it > 0 || j > 0) ? 1u : 0u;
#pragma unroll
                    for (int k = 0; k < NACC; ++k) if (!(a.dbg & 2)) umma_mn(tmem + k * NCOL, ad + 16 * j + (TWO_TAP ? 2 : 1) * k, bd + 16 * j, acc, DESC_HI_I);
                }
                umma_commit(&sm.empty[buf]);
            }
            umma_commit(&sm.done);
        } else {
            const int ct = tid - 64;                                 // 0 .. 255
            if (emb) {
                // ---------------- embed layer: electrodes (fp32) -> split-bf16 A tile(s); collapsed-bias gradient from the B planes
                const int* chan = a.chan + a.zoff[z];
                float bs[F];
#pragma unroll
                for (int c = 0; c < F; ++c) bs[c] = 0.f;
                float cur[16], nxt[16], nx2[16];                     // this item's samples and the next two items' (in flight)
                auto load = [&](int it, float (&v)[16]) {            // unfold + zone gather: sample ct of the window, every electrode of the zone
                    const int bn = rank + it * nc;
                    const int b = bn / a.n_win, n = bn - b * a.n_win;
                    const float* xb = a.x + (size_t)b * a.C * a.T + (size_t)n * a.step + ct;
#pragma unroll
                    for (int c = 0; c < 16; ++c) v[c] = (c < Cz && ct < a.W) ? __ldg(xb + (size_t)chan[c < Cz ? c : 0] * a.T) : 0.f;
                    if (ones) v[15] = ct < a.W ? 1.0f : 0.f;
                };
                load(0, cur);
                if (n_it > 1) load(1, nxt);
#pragma unroll 1
                for (int it = 0; it < n_it; ++it) {
                    const int buf = it % NBUF, use = it / NBUF;
                    // two items ahead: one iteration is shorter than a DRAM round trip under load (with a prefetch distance of one
                    // item the embed CTAs ran at the load latency and were the kernel's long pole; three items buy nothing more)
                    if (it + 2 < n_it) load(it + 2, nx2);
                    if (use >= 1) {
                        if (warp == 2) mbar_wait(&sm.empty[buf], (use - 1) & 1);
                        asm volatile("bar.sync 2, %0;" ::"n"(CONV_T) : "memory");
                    }
                    uint8_t* A1 = tiles + buf * BUF_I;
                    if (ct < a.W) {
#pragma unroll
                        for (int p = 0; p < 2; ++p) {                // 16 channels: planes 0, 1 hold term 0, planes 2, 3 term 1
                            float v[8];
#pragma unroll
                            for (int e = 0; e < 8; ++e) v[e] = cur[8 * p + e];
                            uint4 o[2];
                            split8<2>(v, o);
                            *reinterpret_cast<uint4*>(A1 + p * PLANE_I + ct * 16) = o[0];
                            *reinterpret_cast<uint4*>(A1 + (2 + p) * PLANE_I + ct * 16) = o[1];
                            if (TWO_TAP && ct > 0) {
                                *reinterpret_cast<uint4*>(A1 + TILE_I + p * PLANE_I + (ct - 1) * 16) = o[0];
                                *reinterpret_cast<uint4*>(A1 + TILE_I + (2 + p) * PLANE_I + (ct - 1) * 16) = o[1];
                            }
                        }
                    }
                    // sum_t dL/dy2: this thread's time sample of all 32 channels, both bf16 terms, from the landed B planes -- BEFORE the
                    // arrival below, so that the MMA (and with it the producer's refill of this buffer) cannot overtake the reads
                    if (!ones) {
                    if (warp == 2) mbar_wait(&sm.bfull[buf], use & 1);
                    asm volatile("bar.sync 2, %0;" ::"n"(CONV_T) : "memory");
                    if (ct < a.T_out) {
                        const uint8_t* Bt = A1 + (NTILE - 1) * TILE_I + ct * 16;
#pragma unroll
                        for (int p = 0; p < 8; ++p) {
                            const uint4 w = *reinterpret_cast<const uint4*>(Bt + p * PLANE_I);
                            const uint32_t ww[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
                            for (int i = 0; i < 4; ++i) {
                                bs[8 * (p & 3) + 2 * i] += __uint_as_float(ww[i] << 16);
                                bs[8 * (p & 3) + 2 * i + 1] += __uint_as_float(ww[i] & 0xffff0000u);
                            }
                        }
                    }
                    }
                    fence_proxy_async();
                    mbar_arrive(&sm.full[buf]);
#pragma unroll
                    for (int c = 0; c < 16; ++c) { cur[c] = nxt[c]; nxt[c] = nx2[c]; }
                }
                if (!ones) {
#pragma unroll
                    for (int c = 0; c < F; ++c) {
                        const float r = warp_sum(bs[c]);
                        if (lane == 0) atomicAdd(&sm.bsum[c], r);
                    }
                    asm volatile("bar.sync 1, %0;" ::"n"(CONV_T) : "memory");
                    if (ct < F) atomicAdd(a.gpacked + a.boff[z] + ct, sm.bsum[ct]);
                }
            }
            // ---------------- drain by warps 4-7 = TMEM lane quadrants 0-3 (a warp may only read lanes 32 (warp % 4) .. + 31):
            // TWO_TAP: quadrant 0 / 1 = a0 / a1 rows of the even tap of a pair, 2 / 3 of the odd tap; else quadrants 0 / 1 only.
            // Embed layer: lanes 0-15 a0 rows, 16-31 a1 rows of quadrant 0 (and 2).
            if (warp >= 4 && warp < 8) {
                const int q4 = warp & 3;
                mbar_wait(&sm.done, 0);
                tc_fence_after();
                float* g = a.gpacked + a.goff[z][layer];
                const int odd = q4 >> 1;
                const bool use = TWO_TAP ? (emb ? (q4 == 0 || q4 == 2) : true) : (emb ? q4 == 0 : q4 < 2);
                const int row = emb ? (lane & 15) : lane;
                const bool ok = use && (!emb || row < Cz);
#pragma unroll 1
                for (int kk = 0; kk < NACC; ++kk) {
                    const int k = TWO_TAP ? 2 * kk + odd : kk;
                    float v0[32], v1[32];
                    const uint32_t addr = tmem + ((uint32_t)(32 * q4) << 16) + kk * NCOL;
                    tmem_ld32(addr, v0);
                    tmem_ld32(addr + 32, v1);
                    if (ok && k < TAPS) {
#pragma unroll
                        for (int co = 0; co < 32; ++co) atomicAdd(g + (row * TAPS + k) * F + co, v0[co] + v1[co]);
                    }
                    if (ones && q4 == 0 && kk == 0 && lane == 15) {  // the ones row (a0 part; its a1 part is 0): collapsed-bias gradient
#pragma unroll
                        for (int co = 0; co < 32; ++co) atomicAdd(a.gpacked + a.boff[z] + co, v0[co] + v1[co]);
                    }
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(TMEM_COLS_I) : "memory");
}

}  // namespace im

}  // namespace wg
}  // namespace fb

using namespace fb;

static int wgrad_tc_supported(const fastb200_model* m) {
    FB_REQUIRE(m != nullptr, "model descriptor is NULL");
    FB_REQUIRE(m->dim_cnn == wg::F, "tensor-core backward needs dim_cnn == 32 (got %d)", m->dim_cnn);
    FB_REQUIRE(m->n_zones >= 1 && m->n_zones <= FASTB200_MAX_ZONES, "n_zones=%d out of range", m->n_zones);
    // the backward consumes the operand images the tensor-core training forward writes: same shape envelope as that kernel
    FB_REQUIRE(head_dgrad_tc_supported(m), "tensor-core backward needs 132 <= window_len <= 252 (got %d)", m->window_len);
    for (int z = 0; z < m->n_zones; ++z)
        FB_REQUIRE(m->zone_size[z] >= 1 && m->zone_size[z] <= 16, "zone %d has %d electrodes (tensor-core backward: 1..16)", z, m->zone_size[z]);
    return 0;
}

extern "C" int fastb200_head_bwd_tc_supported(const fastb200_model* m) { return wgrad_tc_supported(m) == 0 ? 1 : 0; }

extern "C" int64_t fastb200_head_bwd_tc_work_floats(const fastb200_model* m, int B, int n_win) {
    if (wgrad_tc_supported(m)) return -1;
    // operand images of dL/dy4, dL/dy3, dL/dy2 per item + the split-bf16 transposed cnn3 / cnn4 weights of the input-gradient kernel
    return (int64_t)B * n_win * m->n_zones * 3 * m->dim_cnn * head_saved_stride(m->window_len) + head_dgrad_tc_packed_bytes(m, 3) / 4;
}

extern "C" int fastb200_head_bwd_tc(const fastb200_model* m, const float* packed, const float* x, const float* saved,
                                    const float* dfeat, int B, int T, int n_win, int step, float* gpacked, float* dx,
                                    float* work, void* stream) {
    if (int r = wgrad_tc_supported(m)) return r;
    FB_REQUIRE(B >= 0 && n_win >= 1 && step >= 1, "bad B/n_win/step");
    FB_REQUIRE((n_win - 1) * (int64_t)step + m->window_len <= T, "windows run past the trial: n_win=%d step=%d W=%d T=%d", n_win, step, m->window_len, T);
    FB_REQUIRE((int64_t)B * n_win < (1ll << 31) / FASTB200_MAX_ZONES, "too many items for one launch");
    FB_REQUIRE(dx == nullptr, "head_bwd_tc computes no input gradient: run fastb200_head_fwd + fastb200_head_bwd (CUDA cores) when dx is wanted");
    if (B == 0) return 0;
    FB_REQUIRE(packed && x && saved && dfeat && gpacked, "head_bwd_tc: NULL buffer");
    FB_REQUIRE(work != nullptr, "head_bwd_tc needs its workspace (fastb200_head_bwd_tc_work_floats)");
    int dg_terms = 2;
#ifdef FASTB200_DEBUG
    if (const char* dg_env = getenv("FASTB200_DGRAD")) if (atoi(dg_env) == 3) dg_terms = 3;        // developer builds: 3-term input-gradient chain
#endif
    float* pk = work + (int64_t)B * n_win * m->n_zones * 3 * m->dim_cnn * head_saved_stride(m->window_len);
    if (int r = head_dgrad_tc_launch(m, packed, saved, dfeat, B, n_win, work, pk, dg_terms, /*write_dy4=*/1, stream)) return r;
    wg::Args a;
    a.saved = saved; a.dys = work; a.gpacked = gpacked; a.x = x; a.dfeat = dfeat; a.dy4_from_y4 = 0;
    a.BN = B * n_win; a.Z = m->n_zones; a.TS = head_saved_stride(m->window_len); a.T_out = m->window_len - 4;
    a.ksteps = ceil_div(a.T_out, 16);
    a.n_win = n_win; a.step = step; a.C = m->n_chan; a.T = T; a.W = m->window_len;
    a.desc_hi = wg::im::DESC_HI_I; a.lbo16 = 128u >> 4;
    a.dbg = 0;
#ifdef FASTB200_DEBUG
    if (const char* e = getenv("FASTB200_WG_DBG")) a.dbg = atoi(e);                                        // developer builds: pipeline probes
#endif
    int n_sel = 0;
    for (int z = 0; z < m->n_zones; ++z) {
        const PackedZone p = packed_zone(*m, z);
        a.goff[z][0] = p.w3; a.goff[z][1] = p.w4; a.goff[z][2] = p.weff; a.boff[z] = p.beff;
        a.cz[z] = m->zone_size[z]; a.zoff[z] = m->zone_off[z]; n_sel += m->zone_size[z];
    }
    for (int i = 0; i < n_sel; ++i) a.chan[i] = m->zone_chan[i];
    const int layers = 3;
    const int groups = layers * m->n_zones;                  // one (zone, layer) group per CTA; equal cost per item
    const int grid = NUM_SMS >= groups ? NUM_SMS : groups;
    int c = 0;
    for (int p = 0; p < groups; ++p) {
        const int n = grid / groups + (p < grid % groups ? 1 : 0);
        for (int r = 0; r < n; ++r, ++c) {
            a.cta_zone[c] = (uint8_t)(p / layers); a.cta_layer[c] = (uint8_t)(p % layers); a.cta_rank[c] = (uint8_t)r; a.cta_n[c] = (uint8_t)n;
        }
    }
    const size_t smem = wg::HDR + (size_t)wg::im::NBUF * wg::im::BUF_I;
    cudaError_t e = cudaFuncSetAttribute(wg::im::head_wgrad_img_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return fail((int)e, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    wg::im::head_wgrad_img_kernel<<<grid, wg::im::THREADS_I, smem, as_stream(stream)>>>(a);
    return check_launch("head_wgrad_img");
}

// ---- allocation-free form of fastb200_head_bwd_tc ------------------------------------------------------------------------
// The workspace of the call above lives in a file-scope device array instead of a caller allocation, for callers that may not
// allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).  Every GPU that loads the library holds the
// array until exit.  In host code `g_bwd_tc_static_work` is only a stand-in for the symbol: its device address is looked up once
// per device with cudaGetSymbolAddress (with that device current) and that pointer is what the kernels receive.
constexpr int64_t STATIC_WORK_FLOATS = (int64_t)FASTB200_HEAD_BWD_TC_STATIC_BYTES / 4;
static __device__ __align__(256) float g_bwd_tc_static_work[STATIC_WORK_FLOATS];

constexpr int STATIC_MAX_DEVICES = 64;
static float* g_static_work_ptr[STATIC_MAX_DEVICES];                 // per device: address of the array on that device (0 = not looked up yet)

static int static_work_address(float** out) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return fb::fail((int)e, "cudaGetDevice: %s", cudaGetErrorString(e));
    if (dev < 0 || dev >= STATIC_MAX_DEVICES) return fb::fail(-1, "device ordinal %d outside the static-workspace table (%d)", dev, STATIC_MAX_DEVICES);
    float* p = __atomic_load_n(&g_static_work_ptr[dev], __ATOMIC_ACQUIRE);
    if (p == nullptr) {
        void* sym = nullptr;
        e = cudaGetSymbolAddress(&sym, g_bwd_tc_static_work);        // the address on the CURRENT device
        if (e != cudaSuccess) return fb::fail((int)e, "cudaGetSymbolAddress(static workspace): %s", cudaGetErrorString(e));
        if (sym == nullptr) return fb::fail(-1, "cudaGetSymbolAddress returned a null address for the static workspace");
        p = static_cast<float*>(sym);
        __atomic_store_n(&g_static_work_ptr[dev], p, __ATOMIC_RELEASE);   // racing threads look up the same value
    }
    *out = p;
    return 0;
}

extern "C" int64_t fastb200_head_bwd_tc_static_max_floats(void) { return STATIC_WORK_FLOATS; }

extern "C" int fastb200_head_bwd_tc_static(const fastb200_model* m, const float* packed, const float* x, const float* saved,
                                           const float* dfeat, int B, int T, int n_win, int step, float* gpacked, void* stream) {
    using namespace fb;
    FB_REQUIRE(m != nullptr, "model descriptor is NULL");
    FB_REQUIRE(B >= 0 && n_win >= 1, "bad B/n_win");
    const int64_t need = fastb200_head_bwd_tc_work_floats(m, B, n_win);
    if (need < 0) return -1;                                          // unsupported shape: message already set
    FB_REQUIRE(need <= STATIC_WORK_FLOATS,
               "head_bwd_tc_static: B=%d x n_win=%d needs %lld workspace floats, the static workspace holds %lld (%d MiB): "
               "use fastb200_head_bwd_tc with a caller workspace, or split the batch",
               B, n_win, (long long)need, (long long)STATIC_WORK_FLOATS, (int)(FASTB200_HEAD_BWD_TC_STATIC_BYTES >> 20));
    if (B == 0) return 0;
    float* work = nullptr;
    if (int r = static_work_address(&work)) return r;
    // nothing in the workspace has to start at zero: the input-gradient kernel writes every image row the weight-gradient
    // kernel reads and the transposed weights are re-packed per call, so the previous call's contents are simply overwritten
    return fastb200_head_bwd_tc(m, packed, x, saved, dfeat, B, T, n_win, step, gpacked, /*dx=*/nullptr, work, stream);
}
