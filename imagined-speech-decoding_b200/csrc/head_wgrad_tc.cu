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
};

struct Smem {
    uint64_t stg_full[2], stg_empty[2], tile_full[2], tile_empty[2], done;
    uint32_t tmem_base;
    float bsum[F];                             // layer 2: sum_t dL/dy2 over the CTA's items (gradient of the collapsed bias)
};

__global__ void __launch_bounds__(THREADS, 1) head_wgrad_tc_kernel(const __grid_constant__ Args a) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    Smem& sm = *reinterpret_cast<Smem*>(smem_raw);
    uint8_t* tiles = smem_raw + HDR;                                // [buf][A | B]
    uint8_t* stg = tiles + 4 * TILE;                                // [2 slots][F][TS] fp32
    const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0), lane = tid & 31;
    // layer 0: cnn3 (A = y2, B = dy3); 1: cnn4 (A = y3, B = dy4); 2: collapsed cnn1 o cnn2 (A = the zone's electrodes, B = dy2)
    const int z = a.cta_zone[blockIdx.x], layer = a.cta_layer[blockIdx.x];
    const bool emb = layer == 2;
    const int Cz = a.cz[z];
    const int TSX = a.TS + 4;                                       // staging row stride of the electrode rows (W = T_out + 4 samples)
    const int rank = a.cta_rank[blockIdx.x], nc = a.cta_n[blockIdx.x];
    const int n_it = rank < a.BN ? (a.BN - rank + nc - 1) / nc : 0;
    const uint32_t tensor_bytes = (uint32_t)(F * a.TS * 4);
    const uint32_t slot_bytes = (uint32_t)(2 * 16 * (a.TS + 4) * 4);   // staging slot: one [F][TS] tensor, or two [16][TS + 4] electrode tiles

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
    if (emb) {                                                      // electrode rows >= Cz of staging slot 0 stay zero
        for (int i = tid; i < 2 * 16 * TSX; i += THREADS) reinterpret_cast<float*>(stg)[i] = 0.f;
        if (tid < F) sm.bsum[tid] = 0.f;
    }
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
            // (embed layer: only B = dy2 -- the electrode rows are gathered by the converter warps themselves)
            if (lane == 0) {
                for (int L = emb ? 1 : 0; L < 2 * n_it; L += emb ? 2 : 1) {
                    const int slot = L & 1, it = L >> 1;
                    if (L >= 2) mbar_wait(&sm.stg_empty[slot], ((L >> 1) - 1) & 1);
                    const size_t item = (size_t)(rank + it * nc) * a.Z + z;
                    const bool recompute = slot == 1 && layer == 1 && a.dy4_from_y4;
                    const float* src = slot == 0 ? a.saved + item * 3 * item_sz + (size_t)layer * item_sz              // y2 / y3
                                     : recompute ? a.saved + item * 3 * item_sz + 2 * item_sz                          // y4 (-> dy4 in the converter)
                                                 : a.dys + item * 3 * item_sz + (size_t)(emb ? 2 : 1 - layer) * item_sz;   // dy3 / dy4 / dy2
                    mbar_expect_tx(&sm.stg_full[slot], tensor_bytes + (recompute ? (uint32_t)(F * 4) : 0u));
                    bulk_g2s(stg + (size_t)slot * slot_bytes, src, tensor_bytes, &sm.stg_full[slot]);
                    // the item's dfeat row rides behind the tensor in the slot's spare bytes (slot_bytes - tensor_bytes = 512)
                    if (recompute) bulk_g2s(stg + (size_t)slot * slot_bytes + tensor_bytes, a.dfeat + item * F, F * 4, &sm.stg_full[slot]);
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
            float bs[4][8];
#pragma unroll
            for (int p = 0; p < 4; ++p)
#pragma unroll
                for (int e = 0; e < 8; ++e) bs[p][e] = 0.f;
#pragma unroll 1
            // embed layer: fused unfold + zone gather of item `it` into electrode tile it & 1 of staging slot 0, by 4-byte
            // cp.async (trial rows are only 4-byte aligned) spread over the 256 converter threads, one item ahead
            auto gather = [&](int it) {
                const int bn = rank + it * nc;
                const int b = bn / a.n_win, n = bn - b * a.n_win;
                const float* xb = a.x + (size_t)b * a.C * a.T + (size_t)n * a.step;
                const int* chan = a.chan + a.zoff[z];
                const uint32_t dst0 = smem_u32(stg) + (uint32_t)(it & 1) * 16u * TSX * 4u;
                for (int i = ct; i < Cz * a.W; i += CONV_WARPS * 32) {
                    const int c = i / a.W, t = i - c * a.W;
                    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst0 + 4u * (c * TSX + t)), "l"(xb + (size_t)chan[c] * a.T + t) : "memory");
                }
                asm volatile("cp.async.commit_group;" ::: "memory");
            };
            if (emb) gather(0);
#pragma unroll 1
            for (int it = 0; it < n_it; ++it) {
                const int buf = it & 1;
                if (emb) {
                    asm volatile("cp.async.wait_group 0;" ::: "memory");                      // this thread's part of gather(it) landed
                    asm volatile("bar.sync 1, %0;" ::"n"(CONV_WARPS * 32) : "memory");        // ... and everybody's; all are done with item it - 1
                    if (it + 1 < n_it) gather(it + 1);                                        // tile (it + 1) & 1 was last read for item it - 1
                }
                // one polling warp (warp 2), the other seven converter warps park in named barrier 2: eight warps spinning on
                // an mbarrier take the issue slots and shared-memory cycles the converting warps need (see head_tc.cu)
#ifndef FB_NO_GROUP_WAIT_WG
                if (it >= 2) {
                    if (warp == 2) mbar_wait(&sm.tile_empty[buf], ((it >> 1) - 1) & 1);
                    asm volatile("bar.sync 2, %0;" ::"n"(CONV_WARPS * 32) : "memory");
                }
#else
                if (it >= 2) mbar_wait(&sm.tile_empty[buf], ((it >> 1) - 1) & 1);
#endif
#pragma unroll 1
                for (int half = 0; half < 2; ++half) {
                    const int L = 2 * it + half, slot = L & 1;
#ifndef FB_NO_GROUP_WAIT_WG
                    if (!(emb && half == 0)) {
                        if (warp == 2) mbar_wait(&sm.stg_full[slot], (L >> 1) & 1);
                        asm volatile("bar.sync 2, %0;" ::"n"(CONV_WARPS * 32) : "memory");
                    }
#else
                    if (!(emb && half == 0)) mbar_wait(&sm.stg_full[slot], (L >> 1) & 1);
#endif
                    const float* sg = (emb && half == 0) ? reinterpret_cast<const float*>(stg) + (it & 1) * 16 * TSX
                                                         : reinterpret_cast<const float*>(stg + (size_t)slot * slot_bytes);
                    if (emb && half == 0) {
                        // electrodes: 16 channels (2 planes per term: a0 -> M rows 0-15, a1 -> 16-31), rows = samples 0 .. W-1
                        uint8_t* tile = tiles + buf * 2 * TILE;
                        for (int t = ct; t < a.W; t += CONV_WARPS * 32) {
#pragma unroll
                            for (int p = 0; p < 2; ++p) {
                                float v[8];
#pragma unroll
                                for (int e = 0; e < 8; ++e) v[e] = sg[(8 * p + e) * TSX + t];
                                uint4 o[2];
                                split8<2>(v, o);
                                *reinterpret_cast<uint4*>(tile + p * PLANE + t * 16) = o[0];
                                *reinterpret_cast<uint4*>(tile + (2 + p) * PLANE + t * 16) = o[1];
                            }
                        }
                    } else {
                        uint8_t* tile = tiles + buf * 2 * TILE + half * TILE + (half == 0 ? 2 * 16 : 0);     // A rows are shifted by the 2 pad rows
                        const bool recompute = half == 1 && layer == 1 && a.dy4_from_y4;
                        const float* df = sg + F * a.TS;                                  // dfeat row behind the staged tensor
                        const float inv_T = 1.0f / (float)a.T_out;
                        for (int t = ct; t < a.T_out; t += CONV_WARPS * 32) {
#pragma unroll
                            for (int p = 0; p < 4; ++p) {
                                float v[8];
#pragma unroll
                                for (int e = 0; e < 8; ++e) v[e] = sg[(8 * p + e) * a.TS + t];
                                if (recompute) {                    // same expression as head_dgrad_tc's first stage: bit-identical dL/dy4
#pragma unroll
                                    for (int e = 0; e < 8; ++e) v[e] = df[8 * p + e] * inv_T * gelu_grad_fast(v[e]);
                                }
                                if (emb) {
#pragma unroll
                                    for (int e = 0; e < 8; ++e) bs[p][e] += v[e];
                                }
                                uint4 o[2];
                                split8<2>(v, o);
                                *reinterpret_cast<uint4*>(tile + p * PLANE + t * 16) = o[0];
                                *reinterpret_cast<uint4*>(tile + (4 + p) * PLANE + t * 16) = o[1];
                            }
                        }
                    }
                    fence_proxy_async();
                    if (!(emb && half == 0)) mbar_arrive(&sm.stg_empty[slot]);
                    mbar_arrive(&sm.tile_full[buf]);
                }
            }
            if (emb) {
#pragma unroll
                for (int p = 0; p < 4; ++p)
#pragma unroll
                    for (int e = 0; e < 8; ++e) {
                        const float r = warp_sum(bs[p][e]);
                        if (lane == 0) atomicAdd(&sm.bsum[8 * p + e], r);
                    }
                asm volatile("bar.sync 1, %0;" ::"n"(CONV_WARPS * 32) : "memory");
                if (ct < F) atomicAdd(a.gpacked + a.boff[z] + ct, sm.bsum[ct]);
            }
        }
        // ---------------- drain: lanes 0-31 hold the a0 rows, 32-63 the a1 rows; columns [b0 | b1] per tap
        if (warp < 2) {
            mbar_wait(&sm.done, 0);
            tc_fence_after();
            float* g = a.gpacked + a.goff[z][layer];
            if (!emb || warp == 0) {
                const int row = emb ? (lane & 15) : lane;           // embed: lanes 0-15 = a0 rows, 16-31 = a1 rows (channel = lane & 15)
                const bool ok = !emb || row < Cz;
#pragma unroll 1
                for (int k = 0; k < TAPS; ++k) {
                    float v0[32], v1[32];
                    const uint32_t addr = tmem + ((uint32_t)(32 * warp) << 16) + k * NCOL;
                    tmem_ld32(addr, v0);
                    tmem_ld32(addr + 32, v1);
                    if (ok) {
#pragma unroll
                        for (int co = 0; co < 32; ++co) atomicAdd(g + (row * TAPS + k) * F + co, v0[co] + v1[co]);
                    }
                }
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
    // dL/dy4, dL/dy3, dL/dy2 per item + the split-bf16 transposed cnn3 / cnn4 weights of the tensor-core input-gradient kernel
    return (int64_t)B * n_win * m->n_zones * 3 * m->dim_cnn * head_saved_stride(m->window_len) + head_dgrad_tc_packed_bytes(m, 3) / 4;
}

extern "C" int fastb200_head_bwd_tc(const fastb200_model* m, const float* packed, const float* x, const float* saved,
                                    const float* dfeat, int B, int T, int n_win, int step, float* gpacked, float* dx,
                                    float* work, void* stream) {
    if (int r = wgrad_tc_supported(m)) return r;
    FB_REQUIRE(B >= 0 && n_win >= 1 && step >= 1, "bad B/n_win/step");
    FB_REQUIRE((n_win - 1) * (int64_t)step + m->window_len <= T, "windows run past the trial: n_win=%d step=%d W=%d T=%d", n_win, step, m->window_len, T);
    FB_REQUIRE((int64_t)B * n_win < (1ll << 31) / FASTB200_MAX_ZONES, "too many items for one launch");
    if (B == 0) return 0;
    FB_REQUIRE(packed && x && saved && dfeat && gpacked, "head_bwd_tc: NULL buffer");
    FB_REQUIRE(work != nullptr, "head_bwd_tc needs its workspace (fastb200_head_bwd_tc_work_floats)");
    // the collapsed-embed weight / bias gradients join the tensor-core kernel when no input gradient is wanted (training) and
    // every zone fits the 16-channel electrode tile; otherwise the CUDA-core kernel keeps them (and dx)
    bool embed_tc = dx == nullptr;
    for (int z = 0; z < m->n_zones; ++z) embed_tc = embed_tc && m->zone_size[z] <= 16;
    int dg_terms = 2, dy4_from_y4 = 0;
    bool dg_ffma = false;
#ifdef FASTB200_DEBUG
    if (const char* dg_env = getenv("FASTB200_DGRAD")) {                              // developer builds: "ffma" forces the CUDA-core chain, "3" = 3 terms
        dg_ffma = dg_env[0] == 'f';
        if (atoi(dg_env) == 3) dg_terms = 3;
    }
#endif
    if (embed_tc && head_dgrad_tc_supported(m) && !dg_ffma) {
        float* pk = work + (int64_t)B * n_win * m->n_zones * 3 * m->dim_cnn * head_saved_stride(m->window_len);
#ifndef FB_WRITE_DY4
        dy4_from_y4 = 1;        // dL/dy4 is element-wise in y4: the weight-gradient kernel re-forms it, the 1.3 GB per 1024 trials stay off HBM
#endif
        if (int r = head_dgrad_tc_launch(m, packed, saved, dfeat, B, n_win, work, pk, dg_terms, /*write_dy4=*/!dy4_from_y4, stream)) return r;
    } else if (int r = head_dgrad_launch(m, packed, x, saved, dfeat, B, T, n_win, step, gpacked, dx, work, embed_tc ? 1 : 0, stream)) return r;
    wg::Args a;
    a.saved = saved; a.dys = work; a.gpacked = gpacked; a.x = x; a.dfeat = dfeat; a.dy4_from_y4 = dy4_from_y4;
    a.BN = B * n_win; a.Z = m->n_zones; a.TS = head_saved_stride(m->window_len); a.T_out = m->window_len - 4;
    a.ksteps = ceil_div(a.T_out, 16);
    a.n_win = n_win; a.step = step; a.C = m->n_chan; a.T = T; a.W = m->window_len;
    a.desc_hi = wg::DESC_HI_MN; a.lbo16 = 128u >> 4;
    int n_sel = 0;
    for (int z = 0; z < m->n_zones; ++z) {
        const PackedZone p = packed_zone(*m, z);
        a.goff[z][0] = p.w3; a.goff[z][1] = p.w4; a.goff[z][2] = p.weff; a.boff[z] = p.beff;
        a.cz[z] = m->zone_size[z]; a.zoff[z] = m->zone_off[z]; n_sel += m->zone_size[z];
    }
    for (int i = 0; i < n_sel; ++i) a.chan[i] = m->zone_chan[i];
    const int layers = embed_tc ? 3 : 2;
    const int groups = layers * m->n_zones;                  // one (zone, layer) group per CTA; equal cost per item
    const int grid = NUM_SMS >= groups ? NUM_SMS : groups;
    int c = 0;
    for (int p = 0; p < groups; ++p) {
        const int n = grid / groups + (p < grid % groups ? 1 : 0);
        for (int r = 0; r < n; ++r, ++c) {
            a.cta_zone[c] = (uint8_t)(p / layers); a.cta_layer[c] = (uint8_t)(p % layers); a.cta_rank[c] = (uint8_t)r; a.cta_n[c] = (uint8_t)n;
        }
    }
    const size_t smem = wg::HDR + 4 * wg::TILE + 2 * (size_t)(2 * 16 * (a.TS + 4) * 4);
    cudaError_t e = cudaFuncSetAttribute(wg::head_wgrad_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return fail((int)e, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    wg::head_wgrad_tc_kernel<<<grid, wg::THREADS, smem, as_stream(stream)>>>(a);
    return check_launch("head_wgrad_tc");
}
