// Shared by the two token-path implementations: transformer.cu (any shape, one CTA per group of trials) and
// transformer_warp.cu (production shape, one warp per trial).
#pragma once
#include "common.cuh"

namespace fb {

constexpr float LN_EPS = 1e-5f;

enum { ACT_NONE = 0, ACT_GELU = 1 };
enum { SITE_ATTN = 0, SITE_FFN_HID = 1, SITE_FFN_OUT = 2, SITE_FINAL = 3 };

struct DropCtx {
    float p, scale;          // p == 0 -> identity
    uint64_t seed, offset;
    const int32_t* step_dev; // optional device counter added to offset (lets a replayed CUDA graph draw fresh masks)
    __device__ __forceinline__ void resolve() {
        if (p > 0.f && step_dev != nullptr) offset += (uint64_t)(uint32_t)__ldg(step_dev);
    }
    __device__ __forceinline__ float mask(int layer, int site, int trial, int elem) const {
        if (p <= 0.f) return 1.f;
        const uint64_t idx = ((uint64_t)(layer * 4 + site) << 44) + ((uint64_t)trial << 20) + (uint64_t)elem;
        return uniform01(seed, offset, idx) >= p ? scale : 0.f;
    }
};

struct TrArgs {
    const float* params;
    const float* feat;      // (B, n_win, ZF)
    float* logits;          // (B, NC)
    float* saved;           // (B, L+1, tok, D) or null
    const float* dlogits;   // backward only
    float* grads;           // backward only (flat, accumulated)
    float* dfeat;           // backward only, may be null
    int B, n_win, tok, G, mode;
    DropCtx dc;
    int ws_floats;          // weight staging area
};

// ---- warp-per-trial kernels (transformer_warp.cu) -----------------------------------------------------------------------
// 1 when the shape is covered: dim_token 32, 8 heads, hidden 64, n_win + 1 tokens in [2, 8], 16-byte aligned parameter blocks
// and a shared-memory footprint that fits.
int tw_supported(const fastb200_model& m, const float* params, int n_win, int mode, int backward);
int tw_fwd_launch(const fastb200_model& m, const TrArgs& a, void* stream);
// backward: mode 'default' only, 16-byte aligned grads / dfeat; gradients are ACCUMULATED into a.grads like transformer_bwd
int tw_bwd_launch(const fastb200_model& m, const TrArgs& a, void* stream);

}  // namespace fb
