/*
 * fastb200.h -- C ABI of the B200-native (sm_100a) FAST hot path.
 *
 * Every entry point is what a binding of the reference's hot path would call; the reference symbol each one
 * replaces is cited as file:line relative to kidusabe1/Imagined-Speech-Decoding.  The library is plain CUDA
 * (no torch types in any signature): pointers are device pointers unless stated, sizes are element counts,
 * `stream` is a cudaStream_t passed as void* (NULL = legacy default stream).
 *
 * Conventions
 *   - return 0 on success, <0 for an invalid argument / unsupported shape, >0 = cudaError_t of the launch.
 *     fastb200_last_error() returns a thread-local, human readable message for the last non-zero return.
 *   - no entry point allocates, frees, retains a pointer past the call, or synchronises the device.
 *   - all kernels launch on the caller's stream; entry points are re-entrant given disjoint buffers.
 *   - all floating point tensors are fp32 unless stated; all layouts are dense row-major.
 */
#ifndef FASTB200_H
#define FASTB200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FASTB200_MAX_ZONES   16
#define FASTB200_MAX_CHANS   256
#define FASTB200_MAX_LAYERS  16
#define FASTB200_TAPS        5      /* Conv4Layers temporal kernel length (fast.py:106-109) */

/* Forward modes of FAST.forward (src/fast/models/fast.py:270-284). */
#define FASTB200_MODE_DEFAULT     0
#define FASTB200_MODE_TRAIN_HEAD  1

/*
 * Static description of one FAST model: the 12 config attributes FAST.__init__ reads (fast.py:219-230), the
 * zone gather table Head builds (fast.py:205-206), and the element offsets of every parameter inside ONE flat
 * fp32 buffer holding the parameters in state_dict order (SURVEY.md section 8b "state_dict wire format").
 * The host fills it once per model; it is passed by pointer to host memory and copied by value into launches.
 */
typedef struct fastb200_model {
    int32_t n_chan;        /* C   electrodes per trial                           */
    int32_t window_len;    /* W   config.window_len                              */
    int32_t slide_step;    /* S   config.slide_step                              */
    int32_t n_zones;       /* Z   len(zone_dict)            <= FASTB200_MAX_ZONES */
    int32_t dim_cnn;       /* F   config.dim_cnn            in {8,16,32}          */
    int32_t dim_token;     /* D   config.dim_token          <= 128, D % heads == 0 */
    int32_t n_heads;       /* H   config.num_heads                                */
    int32_t n_layers;      /* L   config.num_layers         <= FASTB200_MAX_LAYERS */
    int32_t dim_hidden;    /* 2*D AttentionBlock hidden_dim (fast.py:235)         */
    int32_t n_classes;     /*     config.n_classes          <= 32                 */
    int32_t n_pos;         /*     rows of pos_embedding = n_tokens + 1 (fast.py:236) */
    int32_t n_params;      /*     total fp32 elements of the flat parameter buffer */
    int32_t zone_size[FASTB200_MAX_ZONES];   /* Cz                                              */
    int32_t zone_off[FASTB200_MAX_ZONES];    /* prefix sum of Cz into zone_chan                 */
    int32_t zone_chan[FASTB200_MAX_CHANS];   /* electrode row per (zone, position) -- Head.index_dict */
    /* parameter offsets (fp32 elements) into the flat parameter / gradient buffers */
    int32_t off_pos, off_cls;
    int32_t off_cnn1_w[FASTB200_MAX_ZONES];  /* (F,1,1,5)  */
    int32_t off_cnn1_b[FASTB200_MAX_ZONES];  /* (F)        */
    int32_t off_cnn2_w[FASTB200_MAX_ZONES];  /* (F,F,Cz,1) */
    int32_t off_cnn3_w[FASTB200_MAX_ZONES];  /* (F,F,1,5)  */
    int32_t off_cnn4_w[FASTB200_MAX_ZONES];  /* (F,F,1,5)  */
    int32_t off_in_w, off_in_b;              /* input_layer.0 (D, Z*F), (D) */
    int32_t off_ln1_w[FASTB200_MAX_LAYERS], off_ln1_b[FASTB200_MAX_LAYERS];
    int32_t off_qkv_w[FASTB200_MAX_LAYERS], off_qkv_b[FASTB200_MAX_LAYERS];   /* attn.in_proj (3D,D),(3D) */
    int32_t off_out_w[FASTB200_MAX_LAYERS], off_out_b[FASTB200_MAX_LAYERS];   /* attn.out_proj (D,D),(D)  */
    int32_t off_ln2_w[FASTB200_MAX_LAYERS], off_ln2_b[FASTB200_MAX_LAYERS];
    int32_t off_ff1_w[FASTB200_MAX_LAYERS], off_ff1_b[FASTB200_MAX_LAYERS];   /* linear.0 (2D,D),(2D)     */
    int32_t off_ff2_w[FASTB200_MAX_LAYERS], off_ff2_b[FASTB200_MAX_LAYERS];   /* linear.3 (D,2D),(D)      */
    int32_t off_last_w, off_last_b;          /* last_layer (n_classes, D), (n_classes) */
} fastb200_model;

/* ---- library ------------------------------------------------------------------------------------------- */
int         fastb200_version(void);
int         fastb200_sizeof_model(void);      /* sizeof(fastb200_model): lets a binding verify its struct mirror */
const char* fastb200_last_error(void);
/* Number of kernel launches this process has issued through the library so far (a launch recorded into a CUDA graph
 * under stream capture counts once, its replays do not pass through the library).  New: measurement aid for bench.py. */
int64_t     fastb200_launch_count(void);
/* Row-strided async copy (host<->device, one cudaMemcpy2DAsync on `stream`): `rows` rows of `width_bytes`, pitches in bytes.
 * Used by FAST.predict to upload only the samples the windows read (replaces the full-row x.cuda() of
 * inference_on_loader, src/fast/train/trainer.py:89).  Pointers may be pinned host or device memory. */
int         fastb200_copy_rows_async(void* dst, int64_t dst_pitch, const void* src, int64_t src_pitch, int64_t width_bytes,
                                     int64_t rows, void* stream);

/* ---- index kernels: window unfold + zone gather -------------------------------------------------------- */
/*
 * Replaces  x.unfold(-1, W, S) -> rearrange 'B C N T -> (B N) C T' (fast.py:247-249) followed by the per-zone
 * gathers x[:, index_dict[area]] (fast.py:210), all zones concatenated in zone order.
 *   x   (B, C, T) fp32          chan (n_sel) int32 electrode rows, device pointer
 *   out (B*n_win, n_sel, W) fp32, out[b*n_win+n, j, t] = x[b, chan[j], n*S + t]       -- bit exact
 */
int fastb200_window_gather(const float* x, const int32_t* chan, float* out,
                           int B, int C, int T, int W, int S, int n_win, int n_sel, void* stream);

/* ---- zone embedding head (Conv4Layers x Z) -------------------------------------------------------------- */
/* 1 when the fused Conv4Layers kernels cover this head shape (dim_cnn 8/16/32, zones of 1..32 electrodes, window resident in
 * shared memory; need_backward also checks the backward kernel's footprint), else 0: the Python drop-in then runs the head
 * as stock torch CUDA ops (reference Conv4Layers accepts any dim / length, src/fast/models/fast.py:103-119). */
int         fastb200_head_supported(const fastb200_model* m, int need_backward);
/* Size in floats of the packed head weights / packed head gradients for this model. */
int64_t fastb200_head_packed_floats(const fastb200_model* m);
/* Size in floats of the activations head_fwd saves for head_bwd: B*n_win*Z items x 3 x F x roundup4(W-4). */
int64_t fastb200_head_saved_floats(const fastb200_model* m, int B, int n_win);
/*
 * Re-lays the head parameters for the kernels (run once per optimiser step): collapses cnn1 o cnn2
 * (fast.py:113-114: no nonlinearity between them) into one (F, Cz, 5) filter + bias per zone and transposes
 * cnn3 / cnn4 into [ci][tap][co] (forward) and [co][4-tap][ci] (input-gradient) order.
 */
int fastb200_head_pack(const fastb200_model* m, const float* params, float* packed, void* stream);
/*
 * Replaces FAST.forward_head (fast.py:242-252): unfold, zone gather, Conv4Layers for every zone, stack.
 *   x (B, C, T)   feat (B, n_win, Z, F)   step = slide step actually used (step_override or config.slide_step)
 *   saved: NULL for inference; else workspace of fastb200_head_saved_floats floats (y2, y3, y4 per item).
 */
int fastb200_head_fwd(const fastb200_model* m, const float* packed, const float* x, int B, int T, int n_win,
                      int step, float* feat, float* saved, void* stream);
/*
 * Backward of the above (autograd of fast.py:111-119, 210, 247-251).
 *   dfeat (B, n_win, Z, F)    gpacked: packed-layout weight gradients, ACCUMULATED into (caller zeroes)
 *   dx: NULL, or (B, C, T) input gradient, ACCUMULATED into (caller zeroes)  [SHAP path, explain_fast.py:88-91]
 */
int fastb200_head_bwd(const fastb200_model* m, const float* packed, const float* x, const float* saved,
                      const float* dfeat, int B, int T, int n_win, int step, float* gpacked, float* dx,
                      void* stream);
/* Chain rule from packed-layout gradients to d(cnn1.weight, cnn1.bias, cnn2.weight, cnn3.weight, cnn4.weight),
 * WRITTEN (not accumulated) at the parameter offsets of `grads` (flat, same layout as params). */
int fastb200_head_unpack_grad(const fastb200_model* m, const float* params, const float* gpacked,
                              float* grads, void* stream);

/* ---- zone embedding head on the tensor cores (tcgen05 / TMEM), inference forward ------------------------------ */
/*
 * Same function as fastb200_head_fwd with saved == NULL (FAST.forward_head, fast.py:242-252, under no_grad), computed
 * by bf16 tensor-core MMAs with fp32 accumulation and split operands: every fp32 operand is the sum of n_terms bf16
 * terms (n_terms = 3: 24 mantissa bits, fp32-equivalent; n_terms = 2: 16 bits, still ~1e-6 on the features because
 * the feature is a 246-sample mean).  Shapes supported: dim_cnn == 32, 132 <= window_len <= 252, zones of <= 16
 * electrodes -- fastb200_head_tc_supported() returns 1 for them; the FFMA kernel above covers the rest.
 */
int     fastb200_head_tc_supported(const fastb200_model* m);
int64_t fastb200_head_packed_tc_bytes(const fastb200_model* m, int n_terms);
/* Collapses cnn1 o cnn2, splits every head weight into n_terms bf16 terms and lays them out as K-major UMMA operands. */
int fastb200_head_pack_tc(const fastb200_model* m, const float* params, void* packed, int n_terms, void* stream);
/*   packed: output of fastb200_head_pack_tc with the same n_terms.   x (B, C, T)   feat (B, n_win, Z, F) */
int fastb200_head_fwd_tc(const fastb200_model* m, const void* packed, const float* x, int B, int T, int n_win,
                         int step, float* feat, int n_terms, void* stream);

/*
 * Training form of fastb200_head_fwd_tc: additionally writes, straight from the TMEM epilogues, what the tensor-core backward
 * needs, into `saved` (fastb200_head_saved_floats: per item three slots of F * roundup4(W-4) floats):
 *   slot 0 / 1: y2 / y3 as split-bf16 OPERAND IMAGES [term 2][channel group 4][t < W-4][8 x bf16] (128 (W-4) bytes of the slot) --
 *               the A tiles of the weight-gradient MMAs, so the backward has no fp32 -> bf16 conversion pass;
 *   slot 2:     y4 as fp32 [F][roundup4(W-4)] (the input-gradient chain starts with gelu'(y4)).
 * This `saved` is for fastb200_head_bwd_tc ONLY; fastb200_head_bwd wants the all-fp32 `saved` of fastb200_head_fwd.
 */
int fastb200_head_fwd_tc_train(const fastb200_model* m, const void* packed, const float* x, int B, int T, int n_win,
                               int step, float* feat, float* saved, int n_terms, void* stream);

/* ---- zone embedding head backward with the cnn3 / cnn4 weight gradients on the tensor cores ------------------- */
/*
 * Autograd of Conv4Layers / Head / forward_head (fast.py:111-119, 210, 247-251) on the tensor cores, two launches:
 *   1. head_dgrad_tc: dL/dy4 = dfeat / T * gelu'(y4), then the two transposed convolutions as implicit GEMMs; leaves dL/dy4,
 *      dL/dy3, dL/dy2 in `work` as split-bf16 operand images (same format as slots 0 / 1 of `saved`);
 *   2. head_wgrad_img: every weight gradient of the head -- per tap a (32 x T) x (T x 32) contraction over time, operands
 *      bulk-copied from the images straight into the MMA tiles, two taps per MMA, fp32 accumulation in TMEM across all items;
 *      the collapsed-embed layer gathers its A operand (the zone's electrodes) from x and also forms the bias gradient.
 * `saved` must come from fastb200_head_fwd_tc_train.  dx must be NULL: an input gradient (SHAP) is computed by
 * fastb200_head_fwd + fastb200_head_bwd.  Shapes: dim_cnn == 32, 132 <= window_len <= 252, zones of 1..16 electrodes
 * (fastb200_head_bwd_tc_supported() returns 1).
 *   work: fastb200_head_bwd_tc_work_floats(m, B, n_win) floats.
 */
int     fastb200_head_bwd_tc_supported(const fastb200_model* m);
int64_t fastb200_head_bwd_tc_work_floats(const fastb200_model* m, int B, int n_win);
int fastb200_head_bwd_tc(const fastb200_model* m, const float* packed, const float* x, const float* saved,
                         const float* dfeat, int B, int T, int n_win, int step, float* gpacked, float* dx,
                         float* work, void* stream);
/*
 * Allocation-free form of fastb200_head_bwd_tc.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot allocate device memory (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc).  Same arguments and results as fastb200_head_bwd_tc with dx == NULL, but the workspace is a
 * file-scope device array of FASTB200_HEAD_BWD_TC_STATIC_BYTES inside the library (every GPU that loads the library holds
 * it until exit); the call performs no cudaMalloc / cudaFree and calls no allocating library routine.
 *   limit: fastb200_head_bwd_tc_work_floats(m, B, n_win) <= fastb200_head_bwd_tc_static_max_floats()  (production shape:
 *          B <= 8 trials of 5 windows x 8 zones); larger calls return -1 with a message naming both sizes.
 *   The array is shared by every call on a device and keeps the previous call's contents; nothing in it has to start at
 *   zero (each call overwrites what it reads).  Calls that overlap on the GPU (two streams, two threads) RACE on it:
 *   serialise them, or use fastb200_head_bwd_tc with one workspace per stream.
 */
#define FASTB200_HEAD_BWD_TC_STATIC_BYTES (32ll << 20)
int64_t fastb200_head_bwd_tc_static_max_floats(void);
int fastb200_head_bwd_tc_static(const fastb200_model* m, const float* packed, const float* x, const float* saved,
                                const float* dfeat, int B, int T, int n_win, int step, float* gpacked, void* stream);

/* ---- alternative zone heads, eval mode ---------------------------------------------------------------------- */
/*
 * Replaces Head.forward over CVBlock / EEGNet_Encoder / HeadConv_Paper_Version encoders (fast.py:32-100, 122-167, 170-196,
 * 208-210) inside FAST.forward_head (fast.py:242-252) when the model is in eval mode: BatchNorm on its running statistics
 * folds into a per-channel (scale, shift), Dropout is the identity, so each head is a chain of at most
 * FASTB200_ALT_MAX_STAGES stages
 *     out[o][t'] = pool( act( scale[o] * sum_{i,k} w[o][i][k] * in[i][t + k - pad] + shift[o] ) )
 * (dense or depth-wise temporal conv, zero padded; pool stride = size in {1, 2, 4, 8}, floor) and a final temporal mean, mean + Linear or
 * flatten + Linear.  All offsets are fp32 elements relative to zone_base[z] inside `packed`; stage 0 reads the zone's
 * electrodes (in_ch = zone_size[z]), so its weights (out_ch, Cz, k) differ in size per zone.  Training mode (batch
 * statistics, running-stat updates) is not covered: the Python modules run stock torch ops there.
 */
#define FASTB200_ALT_MAX_STAGES 6
#define FASTB200_ALT_DENSE 0
#define FASTB200_ALT_DEPTHWISE 1
#define FASTB200_ALT_ACT_NONE 0
#define FASTB200_ALT_ACT_ELU 1
#define FASTB200_ALT_ACT_GELU 2
#define FASTB200_ALT_POOL_NONE 0
#define FASTB200_ALT_POOL_AVG 1
#define FASTB200_ALT_POOL_MAX 2
#define FASTB200_ALT_FINAL_MEAN 0
#define FASTB200_ALT_FINAL_MEAN_LINEAR 1
#define FASTB200_ALT_FINAL_FLATTEN_LINEAR 2
typedef struct fastb200_alt_stage {
    int32_t kind, out_ch, k, pad, act, pool_kind, pool;
    int32_t w_off;                    /* output channel fastest -- dense: [in_ch][k][out_ch]; depth-wise: [k][out_ch], channel o reads input o / (out_ch / in_ch) */
    int32_t scale_off, shift_off;     /* (out_ch) each; scale_off < 0: no affine */
} fastb200_alt_stage;
typedef struct fastb200_alt_head {
    int32_t window_len, n_zones, n_stages, final_kind;
    int32_t lin_in, lin_out, lin_w_off, lin_b_off;     /* Linear (lin_out, lin_in) + bias; FINAL_MEAN: lin_out = channels */
    fastb200_alt_stage stage[FASTB200_ALT_MAX_STAGES];
    int32_t zone_size[FASTB200_MAX_ZONES], zone_off[FASTB200_MAX_ZONES], zone_chan[FASTB200_MAX_CHANS];
    int64_t zone_base[FASTB200_MAX_ZONES];
} fastb200_alt_head;
int fastb200_sizeof_alt_head(void);
/*   x (B, C, T) -> feat (B, n_win, Z, lin_out); window n of trial b starts at sample n * step. */
int fastb200_alt_head_fwd(const fastb200_alt_head* h, const float* packed, const float* x, int B, int C, int T, int n_win,
                          int step, float* feat, void* stream);

/* ---- transformer (input_layer, CLS, pos, L x AttentionBlock, last_layer) --------------------------------- */
/* Floats saved per forward for the backward: B * (n_layers + 1) * n_tok * D  (block inputs + final tokens). */
int64_t fastb200_transformer_saved_floats(const fastb200_model* m, int B, int n_win);
/*
 * Replaces FAST.forward_transformer (fast.py:260-268) and the 'train_head' branch (fast.py:273-278).
 *   feat (B, n_win, Z, F) -> logits (B, n_classes).   mode: FASTB200_MODE_*.
 *   dropout_p > 0 enables train-mode dropout at the four sites of fast.py:14,19,21,267 with a counter-based
 *   generator keyed by (seed, offset); the same (seed, offset) must be given to transformer_bwd.
 *   dropout_step_dev: NULL, or a device int32 whose value is added to `offset` when the kernel runs (the optimiser
 *   step counter of fastb200_adamw_flat: a replayed CUDA graph then draws fresh masks every step).
 *   saved: NULL for inference.
 */
int fastb200_transformer_fwd(const fastb200_model* m, const float* params, const float* feat, int B, int n_win,
                             int mode, float dropout_p, uint64_t seed, uint64_t offset,
                             const int32_t* dropout_step_dev, float* logits, float* saved, void* stream);
/*
 *   dlogits (B, n_classes)   grads: flat parameter-layout buffer, ACCUMULATED into (caller zeroes)
 *   dfeat (B, n_win, Z, F) written (may be NULL when the head is frozen: 'train_transformer', fast.py:279-282)
 */
int fastb200_transformer_bwd(const fastb200_model* m, const float* params, const float* feat, const float* saved,
                             const float* dlogits, int B, int n_win, int mode, float dropout_p, uint64_t seed,
                             uint64_t offset, const int32_t* dropout_step_dev, float* grads, float* dfeat,
                             void* stream);

/* ---- loss / metrics / optimiser --------------------------------------------------------------------------- */
/*
 * Replaces nn.CrossEntropyLoss()(logits, y) + its backward (trainer.py:37,59) and the argmax of
 * inference_on_loader (trainer.py:89).  labels: uint8 (label_bytes = 1) or int64 (label_bytes = 8).
 *   loss (1): mean CE.   dlogits (B, NC) = (softmax - onehot) * grad_scale / B, may be NULL.
 *   pred (B) int64 argmax, may be NULL.   n_correct (1) int32, may be NULL.
 */
int fastb200_ce_fwd_bwd(const float* logits, const void* labels, int label_bytes, int B, int n_classes,
                        float grad_scale, float* loss, float* dlogits, int64_t* pred, int32_t* n_correct,
                        void* stream);
/*
 * Replaces torch.optim.AdamW(lr=5e-4).step() over all tensors (trainer.py:49): decoupled weight decay on every
 * element, bias-corrected moments.  lr and step are read from DEVICE memory so the whole training step can sit in
 * one CUDA graph: step_dev (1) int32 = number of updates already applied (this call applies update step+1 and,
 * when bump_step != 0, stores step+1 back); lr_dev fp32: the learning rate of this update is lr_dev[0] when
 * lr_by_step == 0, else lr_dev[step] (a whole per-update schedule resident on the device -- the LambdaLR of
 * trainer.py:50-53, base_lr * factor per optimiser step).  lr == 0 is legal.
 */
int fastb200_adamw_flat(float* p, const float* g, float* m, float* v, int64_t n, const float* lr_dev,
                        int lr_by_step, double beta1, double beta2, double eps, double weight_decay,
                        int32_t* step_dev, int bump_step, void* stream);

/*
 * Data-parallel form of the above for the large-batch config (BASELINE.json configs[4]; the reference itself is
 * single-device, scripts/train_fast.py:127-131): the all-reduce of the flat gradient is FUSED into the update.
 * peer_grads: HOST array of n_peers DEVICE pointers, entry r = rank r's flat gradient buffer mapped into this
 * process (NVLink peer / symmetric memory; entry [own rank] is the local buffer).  Element i of the gradient is
 * sum_r peer_grads[r][i] in rank order, so all ranks compute bit-identical updates.  The caller orders the launch
 * after every rank's gradient is complete and keeps the buffers unmodified until every rank's launch finished
 * (two symmetric-memory barriers on the stream); the kernel itself does no inter-GPU synchronisation.
 */
#define FASTB200_MAX_PEERS 16
int fastb200_adamw_flat_peers(float* p, const float* const* peer_grads, int n_peers, float* m, float* v, int64_t n,
                              const float* lr_dev, int lr_by_step, double beta1, double beta2, double eps,
                              double weight_decay, int32_t* step_dev, int bump_step, void* stream);

/* ---- preprocessing that feeds the model ------------------------------------------------------------------ */
/*
 * Replaces the array math of load_subject_train_val / load_training_set / proc_one
 * (src/fast/data/preprocess.py:57-62, 197-203): (T, C, N) MATLAB layout, fp64 (src_bytes = 8) or fp32
 * (src_bytes = 4) -> (N, C, T_pad) fp32 with the last sample repeated (np.pad 'edge').  Bit exact.
 */
int fastb200_trials_from_mat(const void* src, int src_bytes, float* dst, int T, int C, int N, int T_pad,
                             void* stream);
/*
 * Zero-phase IIR: scipy.signal.sosfiltfilt(sos, x, axis=-1, padtype='odd', padlen) (north_star (c); no
 * reference implementation -- SURVEY.md section 8c).  x (rows, T) fp32 or fp64 (x_bytes), y (rows, T) fp32 or
 * fp64 (y_bytes); sos: HOST pointer (n_sections, 6) fp64, n_sections <= 4.  Arithmetic in fp64; one warp per
 * row, the recursion is a chunked warp-level scan.
 */
int fastb200_sosfiltfilt(const void* x, int x_bytes, void* y, int y_bytes, int64_t rows, int T,
                         const double* sos_host, int n_sections, int padlen, void* stream);
/*
 * scipy.signal.resample_poly(x, up, down, axis=-1) with padtype='constant' (zero) edges.  h: DEVICE pointer to
 * the already up-scaled, pre-padded fp64 FIR (len_h taps); y[j] = sum_i x[i] * h[(j+n_pre_remove)*down - i*up].
 * h_host: the same FIR in HOST memory, or NULL.  When it is given and (up, down) = (4, 5) with scipy's default FIR
 * alignment (the 250 -> 200 Hz case), a phase-major kernel runs whose coefficients sit in the kernel parameter block
 * (constant-bank operands); any other shape runs the generic kernel from h_dev.
 */
int fastb200_resample_poly(const void* x, int x_bytes, void* y, int y_bytes, int64_t rows, int T_in, int T_out,
                           const double* h_dev, int len_h, int up, int down, int n_pre_remove, const double* h_host,
                           void* stream);
/*
 * The chain resample_poly(sosfiltfilt(sos, x), up, down) in ONE kernel: the fp64 filtered row stays in shared
 * memory (the two-call chain writes and re-reads it through HBM).  Same arithmetic, same results as the two calls
 * with an fp64 intermediate.  x (rows, T), y (rows, T_out); sos_host / h_host: HOST pointers.  Supported shapes:
 * fastb200_sosfiltfilt_resample_supported(...) == 1 (up/down = 4/5, default FIR alignment, padlen >= the
 * resampler's edge reach); otherwise the call returns -3 and the caller uses the two-call chain.
 */
int fastb200_sosfiltfilt_resample_supported(int T, int n_sections, int padlen, int T_out, const double* h_host,
                                            int len_h, int up, int down, int n_pre_remove);
int fastb200_sosfiltfilt_resample(const void* x, int x_bytes, void* y, int y_bytes, int64_t rows, int T,
                                  const double* sos_host, int n_sections, int padlen, int T_out,
                                  const double* h_host, int len_h, int up, int down, int n_pre_remove, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* FASTB200_H */
