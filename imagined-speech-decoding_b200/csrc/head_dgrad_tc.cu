// Tensor-core (tcgen05 / TMEM) input-gradient chain of the FAST zone embedding head:
//     dL/dy4 = dfeat / T * gelu'(y4)   ->   dL/dy3 = conv(dL/dy4, flipped cnn4^T)   ->   dL/dy2 = conv(dL/dy3, flipped cnn3^T)
// (autograd of Conv4Layers.forward, src/fast/models/fast.py:115-118).  The transpose of a zero-padded "same" temporal conv is
// again a zero-padded "same" conv with the taps flipped and the channel roles swapped, so this is the forward kernel's
// implicit GEMM (head_tc.cu: M = 128 time samples, N = 32 channels, K = 5 taps x 32 channels, row-shifted K-major
// descriptors, split-bf16 operands with the weight terms stacked along N, two independent 128-row streams per window that
// ping-pong on the tensor pipe) with a different first stage and different epilogues:
//   * stage 0 is element-wise: copier warps stream the saved y4 rows of the item (16-byte cp.async ring), converter warps turn
//     them into dL/dy4 (fused GELU derivative), write it to `dys[0]` and, split into bf16 terms, into the operand tile;
//   * the two conv epilogues write dL/dy3 / dL/dy2 (fp32, [F][TS] per item) to `dys[1]` / `dys[2]` for the tensor-core
//     weight-gradient kernel (head_wgrad_tc.cu), which owns every weight gradient of the head.
// Each row is written by the stream that owns it (rows near the cut are computed by both; the 5-row overlap margin covers
// the two 2-row halos, exactly as in the forward).
#include "common.cuh"
#include "tc_ptx.cuh"

namespace fb {
namespace dg {

using namespace fb::tc;

constexpr int F = 32;
constexpr int RY = 132;                 // rows of a conv operand tile: 2 zero rows + 128 + 2 zero rows
constexpr int SR = 132;                 // staged y4 samples per channel: 128 + alignment slack (16-byte cp.async)
constexpr int EPI_WARPS = 16, MMA_WARP = 16, MMA_WARPS = 2, COPY_WARP0 = 18, CONV_WARP0 = 20, LOAD_WARPS = 6;
constexpr int EPI_REGS = 96, SVC_REGS = 48;
constexpr int CW = 16;
constexpr int THREADS = (EPI_WARPS + MMA_WARPS + LOAD_WARPS) * 32;
constexpr int CONV_CHUNKS = 20;
constexpr int STG_BYTES = F * SR * 4 + F * 4;        // staged y4 rows + the item's dfeat row
constexpr int SMEM_HDR = 2048;

__host__ __device__ constexpr int lb(int NT) { return 32 * NT * 16; }
__host__ __device__ constexpr int zone_bytes(int NT) { return 2 * CONV_CHUNKS * lb(NT); }          // [cnn4^T | cnn3^T]
__host__ __device__ constexpr int y_bytes(int NT) { return NT * 4 * RY * 16; }
__host__ __device__ constexpr int stg_tiles(int NT) { return NT == 2 ? 2 : 1; }

// per zone: cnn4^T [20 chunks][32*NT rows][8 bf16] | cnn3^T [20][32*NT][8].  Row n = term * 32 + ci (the conv's OUTPUT channel
// here), K index kk = tap' * 32 + co (its input channel), value w[co][ci][4 - tap'] (flipped taps).
// Source: the fp32 packed head weights of head_pack_kernel, whose w3t / w4t blocks are already [co][4 - tap][ci].
__global__ void head_pack_dgrad_tc_kernel(fastb200_model m, const float* __restrict__ fpacked, uint8_t* __restrict__ packed, int NT) {
    const int z = blockIdx.x;
    uint16_t* c4 = reinterpret_cast<uint16_t*>(packed + (size_t)z * (NT == 2 ? zone_bytes(2) : zone_bytes(3)));
    uint16_t* c3 = c4 + CONV_CHUNKS * 32 * NT * 8;
    const PackedZone pz = packed_zone(m, z);
    const float* w3t = fpacked + pz.w3t;
    const float* w4t = fpacked + pz.w4t;
    for (int i = threadIdx.x + blockDim.x * blockIdx.y; i < CONV_CHUNKS * 8 * F; i += blockDim.x * gridDim.y) {
        const int ci = i % F, kk = i / F;
        const int k = kk / F, co = kk % F;
        float r4 = w4t[(co * TAPS + k) * F + ci], r3 = w3t[(co * TAPS + k) * F + ci];
        for (int t = 0; t < NT; ++t) {
            const __nv_bfloat16 b4 = __float2bfloat16_rn(r4), b3 = __float2bfloat16_rn(r3);
            const size_t o = ((size_t)(kk / 8) * 32 * NT + t * 32 + ci) * 8 + kk % 8;
            c4[o] = __bfloat16_as_ushort(b4);
            c3[o] = __bfloat16_as_ushort(b3);
            r4 -= __bfloat162float(b4);
            r3 -= __bfloat162float(b3);
        }
    }
}

struct Args {
    const uint8_t* packed;                     // head_pack_dgrad_tc_kernel output
    const float* saved;                        // per item 3 x [F][TS]: y2, y3, y4
    const float* dfeat;                        // (B, n_win, Z, F)
    float* dys;                                // per item 3 slots of F * TS floats: split-bf16 operand images of dL/dy4, dL/dy3, dL/dy2 (written)
    int write_dy4;                             // 0: no dL/dy4 image (a weight-gradient kernel that re-forms it from y4 follows)
    int BN, Z, TS;
    int T_out, split, tb1;
    uint8_t cta_zone[NUM_SMS], cta_rank[NUM_SMS], zone_ctas[FASTB200_MAX_ZONES];
};

struct Smem {
    uint64_t w_full;
    uint64_t stg_full[2][2], stg_empty[2][2];
    uint64_t ya_ready[2][2], tile_free[2][2];  // per stream, per item parity
    uint64_t yb_ready[2], dm_full[2][2], d_free[2];    // dm_full[stream][layer]: one completion per item each
    uint32_t tmem_base;
};
static_assert(sizeof(Smem) <= SMEM_HDR, "Smem header overflows its reservation");

template <int NT>
__global__ void __launch_bounds__(THREADS, 1) head_dgrad_tc_kernel(const __grid_constant__ Args a) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    Smem& sm = *reinterpret_cast<Smem*>(smem_raw);
    uint8_t* wz = smem_raw + SMEM_HDR;                              // this zone's split weights [cnn4^T | cnn3^T]
    uint8_t* ys = wz + zone_bytes(NT);                              // [2 streams][2 item parities] conv operand tiles
    uint8_t* stg = ys + 4 * y_bytes(NT);                            // [2 streams][STG slots] staged y4 rows + dfeat
    constexpr int LB = lb(NT);
    constexpr int STG = stg_tiles(NT);
    constexpr int DCOLS = 32 * NT;
    constexpr int TMEM_COLS = 4 * DCOLS <= 256 ? 256 : 512;         // per stream one accumulator PER LAYER: columns (2 layer + s) DCOLS

    const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0), lane = tid & 31;
    const int z = a.cta_zone[blockIdx.x];
    const int rank = a.cta_rank[blockIdx.x], nz = a.zone_ctas[z];
    const size_t item_sz = (size_t)F * a.TS;

    if (tid == 0) {
        mbar_init(&sm.w_full, 1);
        for (int s = 0; s < 2; ++s) {
            for (int i = 0; i < 2; ++i) {
                mbar_init(&sm.stg_full[s][i], 32);
                mbar_init(&sm.stg_empty[s][i], 64);
                mbar_init(&sm.ya_ready[s][i], 64);
                mbar_init(&sm.tile_free[s][i], 1);
            }
            mbar_init(&sm.yb_ready[s], EPI_WARPS * 16);
            mbar_init(&sm.dm_full[s][0], 1);
            mbar_init(&sm.dm_full[s][1], 1);
            mbar_init(&sm.d_free[s], EPI_WARPS * 16);
        }
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
    if (rank >= a.BN) {
        __syncthreads();
        if (warp == MMA_WARP) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(TMEM_COLS) : "memory");
        return;
    }
    const int n_it = (a.BN - rank + nz - 1) / nz;

    if (warp < EPI_WARPS) asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(EPI_REGS));
    else asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(SVC_REGS));

    if (warp >= MMA_WARP && warp < MMA_WARP + MMA_WARPS) {
        // ===================================== MMA issuer of stream s =====================================
        const int s = warp - MMA_WARP;
        if (s == 0 && lane == 0) {
            mbar_expect_tx(&sm.w_full, (uint32_t)zone_bytes(NT));
            bulk_g2s(wz, a.packed + (size_t)z * zone_bytes(NT), (uint32_t)zone_bytes(NT), &sm.w_full);
        }
        __syncwarp();
        mbar_wait(&sm.w_full, 0);
        const uint32_t w_c4 = smem_u32(wz), w_c3 = w_c4 + CONV_CHUNKS * LB;
        auto issue_conv = [&](int layer, int it) {           // layer 0: cnn4^T on dL/dy4, layer 1: cnn3^T on dL/dy3
            if (layer == 0) {
                // its accumulator was drained by the dL/dy3 epilogue of item it - 1, which the issuer has already waited for
                // (yb_ready in front of layer 1 of item it - 1)
                mbar_wait(&sm.ya_ready[s][it & 1], (it >> 1) & 1);
            } else {
                mbar_wait(&sm.yb_ready[s], it & 1);
                if (it > 0) mbar_wait(&sm.d_free[s], (it - 1) & 1);  // the dL/dy2 epilogue of item it - 1 drained this layer's accumulator
            }
            tc_fence_after();
            const uint32_t d = tmem + (2 * layer + s) * DCOLS;
            const uint32_t yd = desc_lo(smem_u32(ys) + (2 * s + (it & 1)) * y_bytes(NT), RY * 16);
            const uint32_t wd = desc_lo(layer == 0 ? w_c4 : w_c3, LB);
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
            if (layer == 1) umma_commit(&sm.tile_free[s][it & 1]);       // the operand tile of this parity may be rewritten
        };
        // Software pipeline across items: cnn3^T of item it and cnn4^T of item it + 1 are issued back to back (separate
        // accumulators), so the tensor pipe has 40 MMAs of this stream queued while the epilogue warps turn dL/dy2(it) and then
        // dL/dy3(it + 1) around -- with one accumulator the chain cnn4^T -> epilogue -> cnn3^T -> drain was strictly serial per item.
        issue_conv(0, 0);
        for (int it = 0; it < n_it; ++it) {
            issue_conv(1, it);
            if (it + 1 < n_it) issue_conv(0, it + 1);
        }
    } else if (warp >= COPY_WARP0 && warp < CONV_WARP0) {
        // ============================ copier of stream s: saved y4 rows + dfeat row -> staging ring ============================
        const int s = warp - COPY_WARP0;
        const int tb = s == 0 ? 0 : a.tb1;
        const int a0 = tb & ~3;                                     // 16-byte aligned first sample
        const uint32_t stg_u32 = smem_u32(stg) + s * STG * STG_BYTES;
#pragma unroll 1
        for (int it = 0; it < n_it; ++it) {
            const int slot = it % STG;
            if (it >= STG) mbar_wait(&sm.stg_empty[s][slot], ((it / STG) - 1) & 1);
            const size_t item = (size_t)(rank + it * nz) * a.Z + z;
            const float* y4 = a.saved + item * 3 * item_sz + 2 * item_sz + a0;
            const uint32_t dst0 = stg_u32 + slot * STG_BYTES;
            for (int i = lane; i < F * (SR / 4); i += 32) {          // 33 16-byte pieces per channel
                const int c = i / (SR / 4), j = i - c * (SR / 4);
                const uint32_t nbytes = (a0 + 4 * j + 4 <= a.TS) ? 16u : 0u;
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst0 + (c * SR + 4 * j) * 4),
                             "l"(nbytes ? y4 + (size_t)c * a.TS + 4 * j : a.saved), "r"(nbytes) : "memory");
            }
            asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst0 + F * SR * 4 + lane * 4), "l"(a.dfeat + item * F + lane) : "memory");
            asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(smem_u32(&sm.stg_full[s][slot])) : "memory");
        }
        asm volatile("cp.async.wait_all;" ::: "memory");
    } else if (warp >= CONV_WARP0) {
        // ============================ converters: y4 -> dL/dy4 (global + split-bf16 operand tile) ============================
        const int s = (warp - CONV_WARP0) >> 1, half = (warp - CONV_WARP0) & 1;
        const int tb = s == 0 ? 0 : a.tb1;
        const int off = tb & 3;                                     // staged sample index of row m is off + m
        const float inv_T = 1.0f / (float)a.T_out;
#pragma unroll 1
        for (int it = 0; it < n_it; ++it) {
            const int slot = it % STG;
#ifndef FB_NO_GROUP_WAIT_DG
            if (half == 0) {                                        // one poller per stream, its partner parks in a named barrier (see head_tc.cu)
                mbar_wait(&sm.stg_full[s][slot], (it / STG) & 1);
                if (it >= 2) mbar_wait(&sm.tile_free[s][it & 1], ((it >> 1) - 1) & 1);
            }
            asm volatile("bar.sync %0, 64;" ::"r"(3 + s) : "memory");
#else
            mbar_wait(&sm.stg_full[s][slot], (it / STG) & 1);
            if (it >= 2) mbar_wait(&sm.tile_free[s][it & 1], ((it >> 1) - 1) & 1);
#endif
            const float* sg = reinterpret_cast<const float*>(stg + (size_t)(s * STG + slot) * STG_BYTES);
            const float* df = sg + F * SR;
            const size_t item = (size_t)(rank + it * nz) * a.Z + z;
            float* dy4 = a.dys + item * 3 * item_sz;
            uint8_t* yt = ys + (2 * s + (it & 1)) * y_bytes(NT);
#pragma unroll 1
            for (int u = 0; u < 8; ++u) {                           // 4 channel groups x 128 rows over the stream's 64 converter threads
                const int j = u >> 1;                               // channel group (8 channels)
                const int m = 64 * (u & 1) + 32 * half + lane;      // row of the stream's tile
                const int t = tb + m;
                const bool live = t < a.T_out;
                const bool own = live && (s == 0 ? t < a.split : t >= a.split);
                float v[8];
#pragma unroll
                for (int e = 0; e < 8; ++e) {
                    const float y = sg[(8 * j + e) * SR + off + m];
                    const float g = df[8 * j + e] * inv_T * gelu_grad_fast(y);
                    v[e] = live ? g : 0.f;
                }
                uint4 o[NT];
                split8<NT>(v, o);
                if (own && a.write_dy4) {                           // operand image for the weight-gradient kernel (see below)
#pragma unroll
                    for (int tm = 0; tm < 2; ++tm)
                        *reinterpret_cast<uint4*>(reinterpret_cast<uint8_t*>(dy4) + ((size_t)(tm * 4 + j) * a.T_out + t) * 16) = o[tm];
                }
#pragma unroll
                for (int tm = 0; tm < NT; ++tm) *reinterpret_cast<uint4*>(yt + tm * (4 * RY * 16) + j * (RY * 16) + (m + 2) * 16) = o[tm];
            }
            fence_proxy_async();
            mbar_arrive(&sm.ya_ready[s][it & 1]);
            mbar_arrive(&sm.stg_empty[s][slot]);
        }
    } else {
        // ============================ epilogue warps (stream s, lane quadrant q, column half h) ============================
        const int s = warp >> 3, h = (warp >> 2) & 1, q = warp & 3;
        const int m = 32 * q + lane;
        const int t = (s == 0 ? 0 : a.tb1) + m;
        const uint32_t lane_addr = tmem + ((uint32_t)(32 * q) << 16) + CW * h + s * (32 * NT);
        uint8_t* yt0 = ys + 2 * s * y_bytes(NT) + (m + 2) * 16 + 2 * h * (RY * 16);
        const bool live = t < a.T_out;
        const bool own = live && (s == 0 ? t < a.split : t >= a.split);
        // one polling warp per stream; the other seven epilogue warps park in a named hardware barrier (head_tc.cu explains why)
        const bool poller = (warp & 7) == 2 + s;
        auto group_wait = [&](uint64_t* bar, uint32_t parity) {
#ifndef FB_NO_GROUP_WAIT_DG
            if (poller) mbar_wait(bar, parity);
            asm volatile("bar.sync %0, 256;" ::"r"(1 + s) : "memory");
#else
            mbar_wait(bar, parity);
#endif
        };
        auto drain = [&](int layer, float (&v)[CW]) {
            float u[NT - 1][CW];
            const uint32_t addr = lane_addr + layer * (2 * 32 * NT);
            tmem_ld16(addr, v);
#pragma unroll
            for (int tm = 1; tm < NT; ++tm) tmem_ld16(addr + 32 * tm, u[tm - 1]);
            tmem_ld_wait();
#pragma unroll
            for (int tm = 1; tm < NT; ++tm)
#pragma unroll
                for (int i = 0; i < CW; i += 2) {
                    const f32x2 sum = add2(pk(v[i], v[i + 1]), pk(u[tm - 1][i], u[tm - 1][i + 1]));
                    unpk(sum, v[i], v[i + 1]);
                }
        };
        // dL/dy3 and dL/dy2 leave as split-bf16 OPERAND IMAGES -- [term 2][channel group 4][t < T_out][8 x bf16] inside the item's
        // stage slot, exactly the B tile of head_wgrad_tc.cu -- not as fp32: the weight-gradient kernel then needs no conversion
        // pass, its tiles are filled by bulk copies.  Each row is written by the stream that owns it.
        auto save_image = [&](const uint4 (&o)[NT], int j, int it, int stage) {
            if (!own) return;
            const size_t item = (size_t)(rank + it * nz) * a.Z + z;
            uint8_t* img = reinterpret_cast<uint8_t*>(a.dys + (item * 3 + stage) * item_sz);
#pragma unroll
            for (int tm = 0; tm < 2; ++tm) *reinterpret_cast<uint4*>(img + ((size_t)(tm * 4 + 2 * h + j) * a.T_out + t) * 16) = o[tm];
        };
#pragma unroll 1
        for (int it = 0; it < n_it; ++it) {
            // ---- dL/dy3: drain -> masked -> operand tile of this parity (overwrites dL/dy4, already consumed) + global
            group_wait(&sm.dm_full[s][0], it & 1);
            tc_fence_after();
            float v[CW];
            drain(0, v);
            {
                uint8_t* yt = yt0 + (it & 1) * y_bytes(NT);
#pragma unroll
                for (int j = 0; j < CW / 8; ++j) {
                    float x8[8];
#pragma unroll
                    for (int e = 0; e < 8; ++e) x8[e] = live ? v[8 * j + e] : 0.f;
                    uint4 o[NT];
                    split8<NT>(x8, o);
#pragma unroll
                    for (int tm = 0; tm < NT; ++tm) *reinterpret_cast<uint4*>(yt + tm * (4 * RY * 16) + j * (RY * 16)) = o[tm];
                    save_image(o, j, it, 1);
                }
                fence_proxy_async();
                tc_fence_before();
                mbar_arrive(&sm.yb_ready[s]);
            }
            // ---- dL/dy2: drain -> global
            group_wait(&sm.dm_full[s][1], it & 1);
            tc_fence_after();
            drain(1, v);
            tc_fence_before();
            mbar_arrive(&sm.d_free[s]);
#pragma unroll
            for (int j = 0; j < CW / 8; ++j) {
                float x8[8];
#pragma unroll
                for (int e = 0; e < 8; ++e) x8[e] = v[8 * j + e];
                uint4 o[NT];
                split8<NT>(x8, o);
                save_image(o, j, it, 2);
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == MMA_WARP) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(TMEM_COLS) : "memory");
}

}  // namespace dg
}  // namespace fb

using namespace fb;

int fb::head_dgrad_tc_supported(const fastb200_model* m) {
    if (m == nullptr || m->dim_cnn != dg::F) return 0;
    const int T_out = m->window_len - 4;
    return T_out >= 128 && T_out <= 248;
}

int64_t fb::head_dgrad_tc_packed_bytes(const fastb200_model* m, int n_terms) { return (int64_t)m->n_zones * dg::zone_bytes(n_terms); }

int fb::head_dgrad_tc_launch(const fastb200_model* m, const float* fpacked, const float* saved, const float* dfeat, int B,
                             int n_win, float* dys, void* packed_ws, int n_terms, int write_dy4, void* stream) {
    FB_REQUIRE(head_dgrad_tc_supported(m), "shape not supported by the tensor-core input-gradient kernel");
    FB_REQUIRE(n_terms == 2 || n_terms == 3, "n_terms must be 2 or 3");
    FB_REQUIRE((int64_t)B * n_win < (1ll << 31) / FASTB200_MAX_ZONES, "too many items for one launch");
    dg::head_pack_dgrad_tc_kernel<<<dim3(m->n_zones, 8), 256, 0, as_stream(stream)>>>(*m, fpacked, static_cast<uint8_t*>(packed_ws), n_terms);
    if (int r = check_launch("head_pack_dgrad_tc")) return r;
    dg::Args a;
    a.packed = static_cast<const uint8_t*>(packed_ws); a.saved = saved; a.dfeat = dfeat; a.dys = dys; a.write_dy4 = write_dy4;
    a.BN = B * n_win; a.Z = m->n_zones; a.TS = head_saved_stride(m->window_len);
    a.T_out = m->window_len - 4; a.split = a.T_out / 2; a.tb1 = a.T_out - 128;
    const int Z = m->n_zones;
    const int grid = NUM_SMS >= Z ? NUM_SMS : Z;
    int given[FASTB200_MAX_ZONES] = {0};
    for (int c = 0; c < grid; ++c) { const int z = c % Z; a.cta_zone[c] = (uint8_t)z; a.cta_rank[c] = (uint8_t)given[z]++; }
    for (int z = 0; z < Z; ++z) a.zone_ctas[z] = (uint8_t)given[z];
    const size_t smem = dg::SMEM_HDR + dg::zone_bytes(n_terms) + 4 * dg::y_bytes(n_terms) + 2 * dg::stg_tiles(n_terms) * dg::STG_BYTES;
    cudaError_t e;
    if (n_terms == 2) {
        e = cudaFuncSetAttribute(dg::head_dgrad_tc_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return fail((int)e, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
        dg::head_dgrad_tc_kernel<2><<<grid, dg::THREADS, smem, as_stream(stream)>>>(a);
    } else {
        e = cudaFuncSetAttribute(dg::head_dgrad_tc_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return fail((int)e, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
        dg::head_dgrad_tc_kernel<3><<<grid, dg::THREADS, smem, as_stream(stream)>>>(a);
    }
    return check_launch("head_dgrad_tc");
}
