/*
 * goalnet_b200.h -- C-ABI of the B200-native per-frame scoring path of CVML-GoalNet.
 *
 * The reference (Vasilispapg/CVML-GoalNet) is pure Python and has no FFI / plugin registry: its
 * "boundary" is the set of plain functions main.py:11 imports from utils.py.  Each entry point
 * below replaces the arithmetic of one of those functions (cited per declaration, paths relative to
 * the reference root); the Python mirror of the reference interface (cvml_goalnet_b200/utils.py)
 * binds them through ctypes, and INTEGRATION.md shows the stub a maintainer of the reference adds.
 *
 * Conventions
 *   - plain pointers and sizes only; no torch / C++ types cross this boundary;
 *   - every `const T* d_...` / `T* d_...` argument is a DEVICE pointer owned by the caller;
 *     `h_...` arguments are HOST pointers;
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream); all work is
 *     enqueued on it and the call returns without synchronising unless stated;
 *   - return value: 0 = ok, <0 = argument/shape error (GN_E_*), >0 = a cudaError_t / CUresult
 *     value; nothing throws; gn_status_string() describes any code;
 *   - the library allocates nothing persistent except the two opaque objects created by
 *     gn_create() / gn_model_create(); it is re-entrant per handle, starts no threads and has NO
 *     CPU fallback: without a CUDA device every compute entry point returns an error.
 */
#ifndef GOALNET_B200_H_
#define GOALNET_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GN_ABI_VERSION 1

#define GN_OK 0
#define GN_E_ARG (-1)       /* null pointer / negative size / inconsistent arguments */
#define GN_E_SHAPE (-2)     /* a shape this build does not support (see each entry point) */
#define GN_E_WORKSPACE (-3) /* caller workspace too small */
#define GN_E_NODEVICE (-4)  /* no usable CUDA device / driver entry point missing */
#define GN_E_RANGE (-5)     /* integer overflow risk (values do not fit the device integer type) */
#define GN_E_STATIC_LIMIT (-6) /* gn_knapsack_static: the problem needs more than GN_KNAPSACK_STATIC_MAX_BYTES */

typedef struct gn_handle gn_handle; /* per-device context: SM count, driver entry points, tables */
typedef struct gn_model gn_model;   /* packed AVM weights on the device */

int gn_abi_version(void);
const char* gn_status_string(int status);

/* Create / destroy the per-device context.  `device` is a CUDA ordinal. */
int gn_create(gn_handle** out, int device);
int gn_destroy(gn_handle* h);
/* Kernel launches issued through `h` since creation (bench.py's `gpu_launches`). */
int64_t gn_launch_count(const gn_handle* h);

/* ---------------------------------------------------------------------------------------------
 * (a1) frame preprocessing -- replaces the arithmetic of utils.py:284-285,292
 *      (`extract_condensed_frame_tensor`): per-frame global min/max over H*W*3 bytes,
 *      (x-min)/(max-min+1e-7), cv2.resize INTER_LINEAR to SxS, HWC->CHW.  BGR order is kept.
 * d_frames : uint8 [n][H][W][3];  d_out_nchw : float32 [n][3][S][S] or NULL;
 * d_out_nhwc_bf16 : bf16 [n][S][S][3] or NULL (at least one output must be given).
 * ------------------------------------------------------------------------------------------- */
int gn_preprocess_u8_bgr(gn_handle* h, const uint8_t* d_frames, int64_t n, int H, int W, int S,
                         float* d_out_nchw, void* d_out_nhwc_bf16, void* stream);

/* ---------------------------------------------------------------------------------------------
 * (a2) audio features -- replaces utils.py:322-349 (`extract_audio_features` after librosa.load):
 *      per frame-slot MFCC(30) of its sample segment (librosa.feature.mfcc defaults: STFT 2048/512
 *      periodic hann, centre zero padding, power 2, 128 Slaney mels, power_to_db top_db 80 per
 *      segment, DCT-II ortho) followed by the cubic time resampling to B columns, given as the
 *      linear operator W (B x q).
 * d_y        : float32 mono samples [n_samples]
 * d_seg_start: int64 [n]  first sample of each slot;  d_seg_len : int32 [n] samples per slot
 *              (q_i = 1 + len_i/512 STFT columns, 4 <= q_i <= GN_MFCC_MAX_Q)
 * d_w_offset : int32 [n]  offset (in floats) of slot i's row-major W (B x q_i) inside d_w
 * d_mel_*    : the mel filterbank in CSR-like form: for mel m, bins [start[m], start[m]+count[m])
 *              with weights d_mel_w[offset[m] ...]  (128 entries each); n_mel_w = length of d_mel_w (<= 2114: a bin
 *              feeds at most two triangular filters)
 * d_dct_t    : float32 [128][32] DCT-II ortho rows TRANSPOSED (element [m][c] = row c, mel m; columns 30, 31 zero)
 * d_out      : float32 [n][30][B]
 * ------------------------------------------------------------------------------------------- */
#define GN_MFCC_MAX_Q 128
int gn_mfcc_resampled(gn_handle* h, const float* d_y, int64_t n_samples, const int64_t* d_seg_start,
                      const int32_t* d_seg_len, const int32_t* d_w_offset, const float* d_w, int64_t n,
                      int B, int max_q, const int32_t* d_mel_start, const int32_t* d_mel_count,
                      const int32_t* d_mel_offset, const float* d_mel_w, int n_mel_w, const float* d_dct_t,
                      float* d_out, void* stream);

/* ---------------------------------------------------------------------------------------------
 * (a3-a5) AVM forward, eval mode -- replaces utils.py:172-195 (VisBl), 214-227 (AudBl),
 *      260-272 (AVM.forward).  Weights are given in the reference's own state_dict layout
 *      (float32, PyTorch memory order); gn_model_create packs them for the kernels.
 * ------------------------------------------------------------------------------------------- */
typedef struct gn_avm_state {
    /* visbl.conv{1,2,3}.weight [O][I][3][3], .bias [O] */
    const float *conv1_w, *conv1_b, *conv2_w, *conv2_b, *conv3_w, *conv3_b;
    /* visbl.bnorm{1,2,3}: weight, bias, running_mean, running_var [C] */
    const float *bn1_w, *bn1_b, *bn1_m, *bn1_v;
    const float *bn2_w, *bn2_b, *bn2_m, *bn2_v;
    const float *bn3_w, *bn3_b, *bn3_m, *bn3_v;
    /* visbl.linear5.weight [512][512*P*P] (C,H,W flatten order), .bias [512] */
    const float *linear5_w, *linear5_b;
    /* audbl.* (NULL when audio_included == 0): conv1 [64][30][3], conv2 [128][64][3], linear3 [128][128*L2] */
    const float *aconv1_w, *aconv1_b, *aconv2_w, *aconv2_b, *alinear3_w, *alinear3_b;
    /* fusion.{0,3,6,9,12}: [512][Fin], [512][512], [256][512], [128][256], [1][128] + biases */
    const float *f0_w, *f0_b, *f3_w, *f3_b, *f6_w, *f6_b, *f9_w, *f9_b, *f12_w, *f12_b;
} gn_avm_state;

/* All pointers in `sd` are DEVICE pointers.  S = frame side (40 in the reference, utils.py:285),
 * B = audio bins (30).  S = 40 runs frame-aligned conv tiles; any other S in 16..1024 (e.g. 224, BASELINE config 5:
 * the reference's layers are Lazy*, so only the literal at utils.py:285 changes) runs overlapping 16x16 conv tiles.
 * linear5_w is then [512][512*P3*P3] with P3 = (S+3)/3 - 5.  B must give 128*L2 = 1024 (B in 29..32).
 * Synchronises `stream` before returning (the caller may free `sd` tensors afterwards). */
int gn_model_create(gn_handle* h, gn_model** out, const gn_avm_state* sd, int audio_included, int S, int B,
                    void* stream);
int gn_model_destroy(gn_model* m);

/* Classifier variant of the fusion head (CAVM / CVM in the report's Table 1).  In the reference it
 * survives as commented lines: utils.py:257 `# nn.Softmax(dim = 1),`, main.py:69 `# criterion = nn.CrossEntropyLoss()`,
 * main.py:97 `# batch_predictions = torch.argmax(batch_predictions, axis = 1) + 1`.  fusion.12 becomes
 * Linear(128, 5) and Softmax(dim = 1) replaces Sigmoid and 4y+1.
 *   weight : device float32 [5][128] (nn.Linear layout), bias : device float32 [5]; n_classes must be 5.
 * After this call gn_avm_forward writes float32 [n][5] class probabilities to d_scores (the caller sizes it so); the
 * predicted importance of a frame is argmax + 1.  The copy is ordered on `stream`. */
int gn_model_set_classifier_head(gn_model* m, const float* weight, const float* bias, int n_classes, void* stream);

/* Bytes of caller workspace gn_avm_forward needs for n frames. */
int64_t gn_avm_workspace_bytes(const gn_model* m, int64_t n);

/* d_visual : float32 [n][3][S][S] (the tensor `extract_condensed_frame_tensor` returns)
 * d_audio  : float32 [n][30][B]  (ignored / may be NULL when the model has no audio branch)
 * d_scores : float32 [n]  (the reference's (n,1) output of utils.py:270, values in [1,5])
 * d_visual_features : optional float32 [n][512] copy of VisBl's output (the TF32-rounded values the fusion network
 *            consumes), or NULL.
 * A frame's score does not depend on the batch it is scored in (no data-dependent work split anywhere in the
 * forward): chunked, sharded and whole-video calls give bit-identical scores. */
int gn_avm_forward(gn_model* m, const float* d_visual, const float* d_audio, int64_t n, float* d_scores,
                   float* d_visual_features, void* d_workspace, int64_t workspace_bytes, void* stream);

/* Frame-sharded scoring of a batch of matches on several GPUs (BASELINE config 3; SURVEY 8e: "scores written straight
 * into peers from the fusion-MLP epilogue").  The reference scores a whole video in one forward (main.py:95,331) and
 * its selection needs every frame's score (utils.py:627-631); here rank r scores frames [a_r, b_r) of each of M
 * matches, and the LAST GEMM'S EPILOGUE stores each score into the (M, n) score matrix of EVERY rank through
 * peer-mapped pointers (NVLink stores; no separate all-gather launch -- only a cross-rank barrier before the
 * selection reads).  Local row i (rows are [match][local frame], rows_per_match = b_r - a_r) of a forward that
 * starts at local row `row0` goes to element
 *     ((row0 + i) / rows_per_match) * dst_match_stride + dst_offset + (row0 + i) % rows_per_match
 * of each dst[p], p < n_dst.  dst_match_stride = n (frames per match), dst_offset = a_r. */
typedef struct gn_score_sink {
    int32_t n_dst;              /* 1..GN_MAX_PEERS destinations (all ranks incl. this one) */
    int32_t reserved;
    float* dst[16];             /* device pointers, peer-mapped where they belong to another GPU */
    int64_t rows_per_match;
    int64_t dst_match_stride;
    int64_t dst_offset;
    int64_t row0;
} gn_score_sink;
#define GN_MAX_PEERS 16
/* gn_avm_forward with the scores scattered through `sink` instead of written to a local [n] vector (regression head
 * only). */
int gn_avm_forward_scatter(gn_model* m, const float* d_visual, const float* d_audio, int64_t n,
                           const gn_score_sink* sink, void* d_workspace, int64_t workspace_bytes, void* stream);
/* The same exchange after a library all-gather (the NCCL variant of the path above): d_src float32
 * [world][n_matches][per] (rank-major, what all_gather_into_tensor of [n_matches][per] rows produces, per =
 * ceil(n / world) padded rows) -> d_dst float32 [n_matches][n], d_dst[m][r * per + j] = d_src[r][m][j]. */
int gn_unshard_rows(gn_handle* h, const float* d_src, int world, int64_t n_matches, int64_t per, int64_t n,
                    float* d_dst, void* stream);

/* Per-stage device timing of gn_avm_forward (CUDA events recorded on the forward's stream).
 * GN_AVM_STAGES stages: 0 VisBl block 1 (conv1), 1 AudBl, 2 conv2 block, 3 conv3 block, 4 linear5,
 * 5 fusion MLP.  gn_model_get_timing synchronises on the last forward and writes milliseconds. */
#define GN_AVM_STAGES 6
int gn_model_set_timing(gn_model* m, int enable);
int gn_model_get_timing(gn_model* m, float* h_ms, int n_stages);

/* ---------------------------------------------------------------------------------------------
 * (a6-a9) keyshot selection -- replaces utils.py:608-613,627-641 (`postprocess` without its HDF5
 *      reads), 396-410 (`expand_array`), 445-464 (`get_clip_information`), 466-510 (`knapsack`).
 *      Integer arithmetic, bit-exact with the reference.
 * ------------------------------------------------------------------------------------------- */

/* scores -> rounded int8 importances (torch.round half-to-even, utils.py:611) -> per-clip value and
 * length over the expanded raw-rate importance vector (utils.py:613,627).  Clips are inclusive [a,b]
 * pairs whose statistics use the end-EXCLUSIVE slice imp[a:b] (utils.py:460-462).
 * d_scores float32 [n]; d_clips int32 [n_clips][2]; outputs int32 [n_clips] each. */
int gn_clip_information(gn_handle* h, const float* d_scores, int64_t n, int skip_frames, int64_t n_raw,
                        const int32_t* d_clips, int n_clips, int32_t* d_values, int32_t* d_lengths,
                        void* stream);

/* Bytes of workspace gn_knapsack needs (decision bits + row buffers) for one problem. */
int64_t gn_knapsack_workspace_bytes(int n_items, int64_t capacity_columns);

/* 0/1 knapsack of utils.py:466-510 for `batch` independent problems that share n_items and
 * capacity (one thread-block per problem).  `d_weights` are the ALREADY SCALED integer weights
 * (int(w*scale_factor), utils.py:478) divided by `unit`, and `capacity` is
 * int(capacity*scale_factor)/unit, where `unit` is a common divisor of all scaled weights chosen by
 * the caller (1 is always valid); `capacity_rem` = int(capacity*scale_factor) % unit.  Dividing by
 * a common divisor leaves the table cells the reference's backtrack visits unchanged.
 * d_values int32 [batch][n_items]; d_weights int32 [batch][n_items];
 * d_selected uint8 [batch][n_items] (1 = clip chosen, ascending order as utils.py:509);
 * d_best int32 [batch] = K[n][capacity]. */
int gn_knapsack(gn_handle* h, const int32_t* d_values, const int32_t* d_weights, int n_items,
                int64_t capacity, int capacity_rem, int batch, uint8_t* d_selected, int32_t* d_best,
                void* d_workspace, int64_t workspace_bytes, void* stream);

/* gn_knapsack for callers that cannot allocate (CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  NO counterpart in the reference.  Same arguments, results
 * and kernels as gn_knapsack, minus the workspace: the scratch is a static device array of
 * GN_KNAPSACK_STATIC_MAX_BYTES inside the library (every GPU that loads the library holds it until exit), and the
 * call itself allocates nothing (no cudaMalloc / cudaFree / allocating library routine).
 * Limit: batch * gn_knapsack_workspace_bytes(n_items, capacity) <= GN_KNAPSACK_STATIC_MAX_BYTES, else
 * GN_E_STATIC_LIMIT (config 2's problem -- about 380 clips, 24 300 capacity columns -- needs 1.2 MB).
 * All calls on one device share the array, and it keeps the previous call's contents (nothing in it has to start at
 * zero; the status words are cleared per call): two calls that OVERLAP on the GPU (different streams, or concurrent
 * host threads) RACE -- serialise them on one stream, or use gn_knapsack with separate workspaces. */
#define GN_KNAPSACK_STATIC_MAX_BYTES (8u << 20)
int gn_knapsack_static(gn_handle* h, const int32_t* d_values, const int32_t* d_weights, int n_items,
                       int64_t capacity, int capacity_rem, int batch, uint8_t* d_selected, int32_t* d_best,
                       void* stream);

/* utils.py:638-641: uint8 indicator over raw frames, 1 on a..b INCLUSIVE of every selected clip. */
int gn_summary_indicator(gn_handle* h, const int32_t* d_clips, const uint8_t* d_selected, int n_clips,
                         int64_t n_raw, uint8_t* d_indicator, void* stream);

/* ---------------------------------------------------------------------------------------------
 * (a10) evaluation -- replaces the counting in utils.py:571-573 (`get_fscore`): for each
 *      annotator u: |S and G_u|, sum(G_u); plus sum(S).  d_counts int64 [2*n_users + 1] =
 *      {overlap_0.., gsum_0.., ssum}.  The float64 tail (utils.py:572-580) stays on the host.
 * ------------------------------------------------------------------------------------------- */
int gn_fscore_counts(gn_handle* h, const uint8_t* d_pred, const uint8_t* d_gd, int n_users, int64_t n_raw,
                     int64_t* d_counts, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Fusion-network training step (BASELINE config 4) -- replaces, for `AVM.fusion`
 * (utils.py:242-258), the reference's train step main.py:187-193:
 *     optimizer.zero_grad(); pred = model(...); loss = nn.MSELoss()(pred, labels); loss.backward();
 *     optimizer.step()          with optimizer = optim.Adam(params, lr = 0.001)  (main.py:70)
 * on precomputed fusion inputs (audio128 | visual512 feature rows).  As written, MSELoss receives
 * pred (N,1) and labels (N,) and broadcasts to N x N: loss = mean_ij (p_i - l_j)^2 (SURVEY quirk
 * Q5); fixed_loss = 1 selects mean_i (p_i - l_i)^2.  Dropout(0.2) of utils.py:245-254 is applied
 * when dropout_p > 0, with caller-supplied keep-masks (uint8 [n][width], widths 512,512,256,128)
 * or, when d_masks is NULL, a counter-based generator keyed by `seed`.
 * Parameters, gradients and Adam moments are flat float32 vectors in the reference's parameter
 * order: fusion.0.weight [512][Fin], fusion.0.bias, fusion.3.*, fusion.6.*, fusion.9.*,
 * fusion.12.weight [1][128], fusion.12.bias  (gn_fusion_param_count values).
 * Data parallel: call gn_fusion_forward_backward on every rank, all-reduce (sum) d_grads over
 * NCCL, then gn_fusion_adam_step with grad_scale = 1 / world_size.
 * ------------------------------------------------------------------------------------------- */
typedef struct gn_fusion_trainer gn_fusion_trainer;
typedef struct gn_fusion_state {
    const float* weight[5];   /* DEVICE pointers: fusion.{0,3,6,9,12}.weight, PyTorch [out][in] layout */
    const float* bias[5];     /* fusion.{0,3,6,9,12}.bias */
} gn_fusion_state;

/* fin = 640 (audio included) or 512; max_rows = largest batch a step will be called with. */
int gn_fusion_trainer_create(gn_handle* h, gn_fusion_trainer** out, const gn_fusion_state* sd, int fin,
                             int64_t max_rows, void* stream);
/* The same with a choice of head: n_classes = 1 is the regression head above (fusion.12 = Linear(128,1)
 * + Sigmoid + 4y+1); n_classes = 5 is the classifier head the reference keeps as commented lines (utils.py:257
 * `nn.Softmax(dim = 1)`, main.py:69 `nn.CrossEntropyLoss()`, main.py:96 `criterion(batch_predictions,
 * (batch_labels-1).long())`): fusion.12 = Linear(128,5) ([5][128] weight, [5] bias) + Softmax. */
int gn_fusion_trainer_create_head(gn_handle* h, gn_fusion_trainer** out, const gn_fusion_state* sd, int fin,
                                  int64_t max_rows, int n_classes, void* stream);
int gn_fusion_trainer_destroy(gn_fusion_trainer* t);
int64_t gn_fusion_param_count(const gn_fusion_trainer* t);
/* torch.optim.Adam hyper-parameters; defaults lr 1e-3, betas (0.9, 0.999), eps 1e-8. */
int gn_fusion_set_hyper(gn_fusion_trainer* t, float lr, float beta1, float beta2, float eps);
/* Copy the float32 master parameters (flat, order above) to d_out. */
int gn_fusion_get_params(gn_fusion_trainer* t, float* d_out, void* stream);
/* d_features float32 [n][fin]; d_labels float32 [n] (importance 1..5); d_grads float32 [param_count] (out);
 * d_scores float32 [n] predictions in [1,5] -- or [n][5] class probabilities with the classifier head -- (out, may be
 * NULL); d_loss double[1] on the device (out, may be NULL); d_masks: NULL or 4 device pointers.
 * fixed_loss selects the loss: regression head 0 = nn.MSELoss as written (pred (N,1) vs labels (N,) broadcast to N x N,
 * SURVEY quirk Q5), 1 = element-wise MSE; classifier head 2 = nn.CrossEntropyLoss as written (applied to the model's
 * softmax OUTPUT with target labels-1, main.py:96), 3 = cross-entropy on the logits. */
int gn_fusion_forward_backward(gn_fusion_trainer* t, const float* d_features, const float* d_labels, int64_t n,
                               int fixed_loss, float dropout_p, const uint8_t* const* d_masks,
                               unsigned long long seed, float* d_grads, float* d_scores, double* d_loss,
                               void* stream);
int gn_fusion_adam_step(gn_fusion_trainer* t, const float* d_grads, float grad_scale, void* stream);
/* forward_backward followed by adam_step(grad_scale = 1): the single-GPU train step. */
int gn_fusion_train_step(gn_fusion_trainer* t, const float* d_features, const float* d_labels, int64_t n,
                         int fixed_loss, float dropout_p, const uint8_t* const* d_masks, unsigned long long seed,
                         float* d_grads, float* d_scores, double* d_loss, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Full-AVM training sub-step -- replaces main.py:187-193 for the WHOLE network in train
 * mode, as the reference runs it (quirk Q4): BatchNorm2d normalises with the sub-batch statistics
 * and updates its running buffers (momentum 0.1, unbiased running variance), Dropout(0.2) follows
 * linear5 and fusion.0/3/6/9 (utils.py:168,245-254), the loss is MSELoss as written (N x N
 * broadcast, quirk Q5; fixed_loss = 1 for the element-wise form) and the optimizer is
 * optim.Adam(model.parameters(), lr = 0.001) (main.py:70).  Sub-batches are 10 frames (main.py:47).
 * Parameters live in one flat float32 vector in model.parameters() order (each tensor padded to a
 * multiple of 4 elements); its gradient vector is exposed for data-parallel all-reduce.
 * d_masks: NULL or 5 device pointers to uint8 keep-masks [n][512],[n][512],[n][512],[n][256],
 * [n][128] (linear5, fusion.0/3/6/9); NULL entries / NULL array -> counter-based generator keyed by
 * `seed`.  dropout_p = 0 disables dropout.
 * ------------------------------------------------------------------------------------------- */
typedef struct gn_avm_trainer gn_avm_trainer;
int gn_avm_trainer_create(gn_handle* h, gn_avm_trainer** out, const gn_avm_state* sd, int audio_included, int S,
                          int B, int64_t max_n, void* stream);
int gn_avm_trainer_destroy(gn_avm_trainer* t);
int64_t gn_avm_trainer_param_count(const gn_avm_trainer* t);
int gn_avm_trainer_set_hyper(gn_avm_trainer* t, float lr, float beta1, float beta2, float eps);
/* d_visual float32 [n][3][S][S]; d_audio float32 [n][30][B] (NULL without audio branch); d_labels float32 [n];
 * d_scores float32 [n] train-mode predictions (out, may be NULL); d_loss double[1] (out, may be NULL). */
int gn_avm_trainer_forward_backward(gn_avm_trainer* t, const float* d_visual, const float* d_audio,
                                    const float* d_labels, int64_t n, int fixed_loss, float dropout_p,
                                    const uint8_t* const* d_masks, unsigned long long seed, float* d_scores,
                                    double* d_loss, void* stream);
/* Device pointer to the flat gradient of the last forward_backward (all-reduce it for data parallel). */
int gn_avm_trainer_grads(gn_avm_trainer* t, float** d_grads);
int gn_avm_trainer_adam_step(gn_avm_trainer* t, float grad_scale, void* stream);
/* forward_backward + adam_step(1): the single-GPU train sub-step. */
int gn_avm_train_step(gn_avm_trainer* t, const float* d_visual, const float* d_audio, const float* d_labels,
                      int64_t n, int fixed_loss, float dropout_p, const uint8_t* const* d_masks,
                      unsigned long long seed, float* d_scores, double* d_loss, void* stream);
/* Copy the trained parameters and BatchNorm running buffers into the WRITABLE device tensors `dst`
 * points at (same shapes as at creation; NULL entries are skipped). */
int gn_avm_trainer_export(gn_avm_trainer* t, const gn_avm_state* dst, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Validation kernels (plain CUDA-core fp32 accumulate).  NOT a backend: tests use them to localise
 * a tensor-core kernel fault layer by layer on the GPU.  Same argument meaning as gn_avm_forward.
 * ------------------------------------------------------------------------------------------- */
int gn_dbg_avm_forward_simt(gn_model* m, const float* d_visual, const float* d_audio, int64_t n,
                            float* d_scores, float* d_visual_features, void* d_workspace,
                            int64_t workspace_bytes, void* stream);
/* Copy an intermediate activation of the LAST forward on `m` out of the workspace as float32:
 * which = 1,2,3 -> VisBl block outputs in NCHW (utils.py:177,182,187); 4 -> fusion input [n][Fin]. */
int gn_dbg_read_activation(gn_model* m, int which, int64_t n, const void* d_workspace, float* d_out,
                           void* stream);

/* d_c[M][N] = d_a[M][K] . d_b[N][K]^T, fp32 operands multiplied as TF32 by tcgen05 (K % 32 == 0, N % 128 == 0): the
 * GEMM behind the full-AVM trainer's convolutions, exported so that tests can check it in isolation. */
int gn_dbg_gemm_tf32(gn_handle* h, const float* d_a, int64_t M, int K, const float* d_b, int N, float* d_c,
                     void* stream);

#ifdef __cplusplus
}
#endif
#endif /* GOALNET_B200_H_ */
