// Cross-entropy (+ argmax / accuracy counter), flat AdamW, and the .mat-layout -> trial-layout transpose.
//
// Reference: nn.CrossEntropyLoss + optim.AdamW(lr=5e-4) in src/fast/train/trainer.py:37,49,59; argmax in
// inference_on_loader trainer.py:89; transpose/cast/edge-pad in src/fast/data/preprocess.py:57-62.
#include "common.cuh"

namespace fb {

// One CTA; thread per trial, block-level tree reduction for the mean (deterministic).
// Labels follow nn.CrossEntropyLoss (trainer.py:37,59): a target of -100 (torch's default ignore_index; int64 labels only)
// takes no part in the loss, the mean or the gradient; any other label outside [0, NC) is an error that torch reports with a
// device-side assert -- here the loss and that row's gradient become NaN (loud at the job's single host read, and no
// out-of-bounds access), since the launch may sit inside a captured graph where nothing can be raised.
template <typename LabelT>
__global__ void ce_kernel(const float* __restrict__ logits, const LabelT* __restrict__ labels, int B, int NC,
                          float grad_scale, float* __restrict__ loss, float* __restrict__ dlogits,
                          int64_t* __restrict__ pred, int32_t* __restrict__ n_correct) {
    __shared__ float s_loss[32];
    __shared__ int s_hit[32];
    __shared__ int s_cnt;
    float my_loss = 0.f;
    int my_hit = 0, my_cnt = 0;
    if (threadIdx.x == 0) s_cnt = 0;
    __syncthreads();
    for (int r = threadIdx.x; r < B; r += blockDim.x) my_cnt += ((long long)labels[r] != -100ll);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) my_cnt += __shfl_xor_sync(0xffffffffu, my_cnt, o);
    if ((threadIdx.x & 31) == 0 && my_cnt) atomicAdd(&s_cnt, my_cnt);          // integer: order-independent
    __syncthreads();
    const int n_used = s_cnt;                                  // torch divides by the number of non-ignored targets
    const float gs = grad_scale / (float)n_used;
    const float qnan = __int_as_float(0x7fc00000);
    for (int r = threadIdx.x; r < B; r += blockDim.x) {
        const float* l = logits + (size_t)r * NC;
        float mx = l[0];
        int am = 0;
        for (int c = 1; c < NC; ++c)
            if (l[c] > mx) { mx = l[c]; am = c; }
        float se = 0.f;
        for (int c = 0; c < NC; ++c) se += expf(l[c] - mx);
        const float lse = mx + logf(se);
        const long long yl = (long long)labels[r];
        const bool ignored = yl == -100ll;
        const bool bad = !ignored && (yl < 0 || yl >= NC);
        const int y = (ignored || bad) ? 0 : (int)yl;
        if (!ignored) {
            my_loss += bad ? qnan : lse - l[y];
            my_hit += (!bad && am == y);
        }
        if (pred) pred[r] = am;
        if (dlogits) {
            const float inv = 1.0f / se;
            for (int c = 0; c < NC; ++c) {
                const float d = (expf(l[c] - mx) * inv - (c == y ? 1.f : 0.f)) * gs;
                dlogits[(size_t)r * NC + c] = ignored ? 0.f : (bad ? qnan : d);
            }
        }
    }
    my_loss = warp_sum(my_loss);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) my_hit += __shfl_xor_sync(0xffffffffu, my_hit, o);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0) { s_loss[warp] = my_loss; s_hit[warp] = my_hit; }
    __syncthreads();
    if (warp == 0) {
        const int nw = blockDim.x >> 5;
        float v = lane < nw ? s_loss[lane] : 0.f;
        int h = lane < nw ? s_hit[lane] : 0;
        v = warp_sum(v);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) h += __shfl_xor_sync(0xffffffffu, h, o);
        if (lane == 0) {
            if (loss) *loss = v / (float)n_used;               // 0 / 0 = NaN when every target is ignored, as torch
            if (n_correct) *n_correct = h;
        }
    }
}

// Scalars follow torch's single-tensor AdamW: Python-double hyper-parameter arithmetic, rounded to fp32 once
// where torch hands them to an fp32 tensor op.
// PEERS > 0: `g` is ignored and the gradient of element i is the sum over `peers.ptr[r][i]`, r = 0 .. n_peers-1 in rank
// order (so every rank forms bit-identical sums) -- a one-shot all-reduce over NVLink peer memory fused into the update.
struct PeerGrads { const float* ptr[FASTB200_MAX_PEERS]; int n; };

template <bool PEERS>
__global__ void adamw_kernel(float* __restrict__ p, const float* __restrict__ g, PeerGrads peers, float* __restrict__ m,
                             float* __restrict__ v, int64_t n, const float* __restrict__ lr_dev, int lr_by_step,
                             double beta1, double beta2, double eps_d, double wd,
                             const int32_t* __restrict__ step_dev) {
    __shared__ float s_c[3];
    if (threadIdx.x == 0) {
        const int done = *step_dev;                  // updates already applied
        const int step = done + 1;
        const double bc1 = 1.0 - pow(beta1, (double)step);
        const double bc2 = 1.0 - pow(beta2, (double)step);
        const double lr = (double)lr_dev[lr_by_step ? done : 0];
        s_c[0] = (float)(lr / bc1);                  // step size
        s_c[1] = (float)sqrt(bc2);
        s_c[2] = (float)(1.0 - lr * wd);
    }
    __syncthreads();
    const float step_size = s_c[0], bc2_sqrt = s_c[1], decay = s_c[2];
    const float w1 = (float)(1.0 - beta1), b2 = (float)beta2, w2 = (float)(1.0 - beta2), eps = (float)eps_d;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        float gi;
        if constexpr (PEERS) {
            gi = __ldcv(peers.ptr[0] + i);                    // peer memory: never serve from a stale L1 line
            for (int r = 1; r < peers.n; ++r) gi += __ldcv(peers.ptr[r] + i);
        } else {
            gi = g[i];
        }
        float pi = p[i] * decay;
        float mi = m[i];
        mi = mi + (gi - mi) * w1;                             // exp_avg.lerp_(grad, 1 - beta1)
        const float vi = v[i] * b2 + w2 * gi * gi;            // exp_avg_sq.mul_(beta2).addcmul_(g, g, 1 - beta2)
        const float denom = sqrtf(vi) / bc2_sqrt + eps;
        pi -= step_size * (mi / denom);
        p[i] = pi; m[i] = mi; v[i] = vi;
    }
}
__global__ void bump_step_kernel(int32_t* step_dev) { *step_dev += 1; }

template <typename SrcT>
__global__ void trials_from_mat_kernel(const SrcT* __restrict__ src, float* __restrict__ dst, int T, int C, int N,
                                       int T_pad) {
    __shared__ float tile[32][33];
    const int c = blockIdx.z;
    const int t0 = blockIdx.x * 32, n0 = blockIdx.y * 32;
#pragma unroll
    for (int j = 0; j < 32; j += 8) {
        const int t = t0 + threadIdx.y + j, n = n0 + threadIdx.x;
        if (t < T_pad && n < N) {
            const int ts = t < T ? t : T - 1;                       // np.pad(..., 'edge')
            tile[threadIdx.y + j][threadIdx.x] = (float)src[((size_t)ts * C + c) * N + n];
        }
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < 32; j += 8) {
        const int n = n0 + threadIdx.y + j, t = t0 + threadIdx.x;
        if (t < T_pad && n < N) dst[((size_t)n * C + c) * T_pad + t] = tile[threadIdx.x][threadIdx.y + j];
    }
}

}  // namespace fb

using namespace fb;

extern "C" int fastb200_ce_fwd_bwd(const float* logits, const void* labels, int label_bytes, int B, int n_classes,
                                   float grad_scale, float* loss, float* dlogits, int64_t* pred,
                                   int32_t* n_correct, void* stream) {
    FB_REQUIRE(B >= 1 && n_classes >= 1, "bad B / n_classes");
    FB_REQUIRE(label_bytes == 1 || label_bytes == 8, "labels must be uint8 or int64");
    const int threads = B >= 256 ? 256 : round_up(B, 32);
    if (label_bytes == 1)
        ce_kernel<uint8_t><<<1, threads, 0, as_stream(stream)>>>(logits, (const uint8_t*)labels, B, n_classes, grad_scale, loss, dlogits, pred, n_correct);
    else
        ce_kernel<int64_t><<<1, threads, 0, as_stream(stream)>>>(logits, (const int64_t*)labels, B, n_classes, grad_scale, loss, dlogits, pred, n_correct);
    return check_launch("ce_fwd_bwd");
}

extern "C" int fastb200_adamw_flat(float* p, const float* g, float* m, float* v, int64_t n, const float* lr_dev,
                                   int lr_by_step, double beta1, double beta2, double eps, double weight_decay,
                                   int32_t* step_dev, int bump_step, void* stream) {
    FB_REQUIRE(n >= 0, "bad n");
    if (n == 0) return 0;
    int64_t blocks = (n + 255) / 256;
    if (blocks > NUM_SMS * 8) blocks = NUM_SMS * 8;
    adamw_kernel<false><<<(unsigned)blocks, 256, 0, as_stream(stream)>>>(p, g, PeerGrads{}, m, v, n, lr_dev, lr_by_step, beta1, beta2, eps, weight_decay, step_dev);
    if (int r = check_launch("adamw_flat")) return r;
    if (bump_step) {
        bump_step_kernel<<<1, 1, 0, as_stream(stream)>>>(step_dev);
        return check_launch("adamw_bump_step");
    }
    return 0;
}

extern "C" int fastb200_adamw_flat_peers(float* p, const float* const* peer_grads_host, int n_peers, float* m, float* v,
                                         int64_t n, const float* lr_dev, int lr_by_step, double beta1, double beta2,
                                         double eps, double weight_decay, int32_t* step_dev, int bump_step,
                                         void* stream) {
    FB_REQUIRE(n >= 0, "bad n");
    FB_REQUIRE(n_peers >= 1 && n_peers <= FASTB200_MAX_PEERS, "n_peers must be 1..FASTB200_MAX_PEERS");
    FB_REQUIRE(peer_grads_host != nullptr, "peer_grads is NULL");
    PeerGrads pg{};
    pg.n = n_peers;
    for (int r = 0; r < n_peers; ++r) {
        FB_REQUIRE(peer_grads_host[r] != nullptr, "a peer gradient pointer is NULL");
        pg.ptr[r] = peer_grads_host[r];
    }
    if (n == 0) return 0;
    int64_t blocks = (n + 255) / 256;
    if (blocks > NUM_SMS * 8) blocks = NUM_SMS * 8;
    adamw_kernel<true><<<(unsigned)blocks, 256, 0, as_stream(stream)>>>(p, nullptr, pg, m, v, n, lr_dev, lr_by_step, beta1, beta2, eps, weight_decay, step_dev);
    if (int r = check_launch("adamw_flat_peers")) return r;
    if (bump_step) {
        bump_step_kernel<<<1, 1, 0, as_stream(stream)>>>(step_dev);
        return check_launch("adamw_bump_step");
    }
    return 0;
}

extern "C" int fastb200_trials_from_mat(const void* src, int src_bytes, float* dst, int T, int C, int N, int T_pad,
                                        void* stream) {
    FB_REQUIRE(T >= 1 && C >= 1 && N >= 0 && T_pad >= T, "bad dims");
    FB_REQUIRE(src_bytes == 4 || src_bytes == 8, "src must be fp32 or fp64");
    FB_REQUIRE(C <= 65535, "too many channels");
    if (N == 0) return 0;
    dim3 grid(ceil_div(T_pad, 32), ceil_div(N, 32), C), block(32, 8);
    if (src_bytes == 8)
        trials_from_mat_kernel<double><<<grid, block, 0, as_stream(stream)>>>((const double*)src, dst, T, C, N, T_pad);
    else
        trials_from_mat_kernel<float><<<grid, block, 0, as_stream(stream)>>>((const float*)src, dst, T, C, N, T_pad);
    return check_launch("trials_from_mat");
}
