// Shared declarations for the goalnet_b200 CUDA sources (sm_100a only).
#pragma once

#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/goalnet_b200.h"

struct gn_handle {
    int device;
    int num_sms;
    int64_t launches;
    // cuTensorMapEncodeTiled, resolved through cudaGetDriverEntryPoint so that the library does not
    // link against libcuda (it must load, and export its symbols, on a box without a driver).
    void* encode_tiled;
    void* pre_plans;             // gn_preprocess.cu: cached tap plans per (H, W, S)
    // kernels whose function attributes (dynamic shared memory limit, non-portable cluster size) have been set for
    // THIS handle's device: the attributes are per device, so the record lives in the handle, not in the process
    const void* attr_done[96];
    int n_attr_done;
};

// true exactly once per (handle, kernel): the caller then sets the kernel's function attributes on the handle's device
static inline bool gn_attr_first(gn_handle* h, const void* fn) {
    for (int i = 0; i < h->n_attr_done; ++i)
        if (h->attr_done[i] == fn) return false;
    if (h->n_attr_done < 96) h->attr_done[h->n_attr_done++] = fn;
    return true;      // (a full table only costs a repeated, idempotent cudaFuncSetAttribute)
}
void gn_preprocess_free_plans(gn_handle* h);

#define GN_CUDA(expr)                                   \
    do {                                                \
        cudaError_t _e = (expr);                        \
        if (_e != cudaSuccess) return (int)_e;          \
    } while (0)

#define GN_LAUNCHED(h)                                  \
    do {                                                \
        (h)->launches++;                                \
        cudaError_t _e = cudaGetLastError();            \
        if (_e != cudaSuccess) return (int)_e;          \
    } while (0)

static inline cudaStream_t gn_stream(void* s) { return reinterpret_cast<cudaStream_t>(s); }

// Per-trainer optimizer state that lives ON THE DEVICE so that a whole train step can be captured once in a CUDA graph
// and replayed: the Adam step count with its bias corrections, and the bump added to the dropout seed each step
// (a value baked into a captured kernel's arguments would repeat the same step forever).
struct gn_step_state {
    unsigned long long seed_bump;   // added to every dropout seed of the step; advanced by gn_launch_step_advance
    long long step;                 // Adam step count t
    float bc1, bc2;                 // 1 - beta1^t, 1 - beta2^t
};
int gn_launch_step_advance(gn_handle* h, gn_step_state* d_state, float beta1, float beta2, cudaStream_t st);   // gn_train.cu

// ---- model object ------------------------------------------------------------------------------
// Geometry of VisBl at S=40 (reference utils.py:151-164): conv1 k3 s3 p3 -> 15x15, pool -> 13x13,
// conv2 p1 -> 13x13, pool -> 11x11, conv3 p1 -> 11x11, pool -> 9x9, flatten 512*81 = 41472.
struct gn_model {
    gn_handle* h;
    int audio_included, S, B;
    int C1, P1, P2, P3;          // 15, 13, 11, 9
    int L1, L2;                  // AudBl lengths 15, 8
    int fin;                     // 640 or 512
    // fp32 (CUDA-core layers)
    float *conv1_w, *conv1_b;    // [27][64] (k-major: (ky,kx,ci) x co), [64]
    float *bn_scale[3], *bn_shift[3];   // gamma/sqrt(var+eps), beta - mean*scale
    float *conv2_b, *conv3_b, *linear5_b;
    float *aconv1_w, *aconv1_b, *aconv2_w, *aconv2_b, *alinear3_w, *alinear3_b;
    float *f_b[4];               // fusion.0/3/6/9 biases
    float *f12_w, *f12_b;        // [128], [1]
    int n_cls;                   // 0/1 = regression head; 5 = classifier head (gn_model_set_classifier_head)
    float *cls_w, *cls_b;        // [5][128], [5]  (one separate allocation at cls_w)
    // bf16 K-major GEMM operands of the visual backbone (99 % of the FLOPs): [N_out][K]
    __nv_bfloat16 *conv2_wk;     // [256][9*64]    k = (ky*3+kx)*64 + ci
    __nv_bfloat16 *conv3_wk;     // [512][9*256]
    __nv_bfloat16 *linear5_wk;   // K-blocked (gn_w5_off), k = p*512 + c   (reference flatten is c*P3^2 + p)
    // TF32 K-major GEMM operands of the "head" (AudBl conv2 / linear3 and the fusion MLP: 0.6 % of the FLOPs, but the
    // layers whose bf16 rounding dominated the score error -- raw MFCCs span +-600 and the last layers amplify):
    // fp32 storage, values rounded to TF32 (cvt.rna) when packed so that the tensor core's truncation is a no-op
    float *f_wk[4];              // fusion.0 [512][fin], .3 [512][512], .6 [256][512], .9 [128][256]
    float *aconv2_wk;            // [128][3*64]    k = tap*64 + ci
    float *alinear3_wk;          // [128][L2*128]  k = l*128 + c   (reference flatten is c*L2 + l)
    void* arena;                 // one allocation backing everything above
    int timing;                  // record per-stage events in gn_avm_forward
    cudaEvent_t ev[GN_AVM_STAGES + 1];
};
#define GN_STAGE_MARK(m, k, st) do { if ((m)->timing) cudaEventRecord((m)->ev[k], st); } while (0)

// K-blocked operand layout of linear5 (K = 512 * P3^2 runs into the millions at large frame sizes, where K-major
// rows are megabytes apart and a 128-row TMA box would touch 128 DRAM pages): K is cut into blocks of 64 elements
// (one 128-byte swizzle row) and rows are grouped so that every TMA box is ONE contiguous chunk.
//   a3 : [ceil(n/128)][K/64][128 frames][64]     (16 KB per box)     w5 : [K/64][512][64]   (32 KB per 256-row box)
__host__ __device__ __forceinline__ int64_t gn_a3_off(int64_t frame, int64_t k, int64_t kblocks) {
    return ((((frame >> 7) * kblocks + (k >> 6)) << 7) + (frame & 127)) * 64 + (k & 63);
}
__host__ __device__ __forceinline__ int64_t gn_w5_off(int64_t row, int64_t k) {
    return (((k >> 6) << 9) + row) * 64 + (k & 63);
}

// Workspace layout for n frames; offsets are computed by gn_ws_layout().  The visual backbone's activations are bf16,
// the head's (AudBl, fusion input, fusion hidden layers) are fp32 holding TF32-rounded values.
struct gn_ws {
    __nv_bfloat16* a1p;   // [n][15][15][64]   block-1 output, zero border (conv2 input)
    __nv_bfloat16* a2p;   // [n][13][13][256]  block-2 output, zero border (conv3 input)
    __nv_bfloat16* a3;    // block-3 output, k = p*512 + c, K-blocked (gn_a3_off): linear5 input
    float* ac1;           // [n][L1+2][64]     AudBl conv1 output, (l, c) order, zero rows at l = -1 and l = L1
    float* ac2;           // [n][L2][128]      AudBl conv2 output, (l, c) order (linear3 input)
    float* h0;            // [n][fin]          [audio128 | visual512]
    float* h1;            // [n][512]
    float* h2;            // [n][512]
    float* h3;            // [n][256]
    float* part;          // [splits][n][512] fp32 partial sums of linear5's K ranges (gn_linear5_splits)
    int64_t bytes;
};

#if defined(__CUDACC__)
// value rounded to TF32 (10-bit mantissa, round to nearest, ties away): what the head's GEMM operands hold
__device__ __forceinline__ float gn_round_tf32(float x) {
    uint32_t r;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
    return __uint_as_float(r);
}
#endif
gn_ws gn_ws_layout(const gn_model* m, int64_t n, void* base);
int gn_tc_gemm_tf32(gn_handle* h, const float* A, int64_t M, int K, int64_t lda, const float* B, int N, int64_t ldb,
                    float* C, cudaStream_t st, int k_splits = 1);   // gn_tc_gemm.cu: C[split][M][N] = A[M][K range] . B[N][K range]^T, TF32 tensor cores
int gn_launch_zero_border(gn_handle* h, __nv_bfloat16* buf, int64_t n, int PW, int C, cudaStream_t st);   // gn_tc_gemm.cu
int gn_linear5_splits(const gn_model* m);   // gn_tc_gemm.cu: number of K ranges of linear5 (a function of K only: batch-invariant sums)

// ---- stage launchers (defined in the per-stage .cu files) ----------------------------------------
int gn_launch_visbl_front(gn_model* m, const float* d_visual, int64_t n, const gn_ws& ws, cudaStream_t st, bool simt = false);
int gn_launch_visbl_block1_tc(gn_model* m, const float* d_visual, int64_t n, const gn_ws& ws, cudaStream_t st);   // gn_front_tc.cu
int gn_launch_audbl(gn_model* m, const float* d_audio, int64_t n, const gn_ws& ws, cudaStream_t st, bool simt);
int gn_launch_audbl_conv1_tc(gn_model* m, const float* d_audio, int64_t n, const gn_ws& ws, cudaStream_t st);   // gn_front_tc.cu
int gn_launch_tc_audbl(gn_model* m, int64_t n, const gn_ws& ws, cudaStream_t st);   // conv2 + linear3 on tensor cores
int gn_launch_tc_backbone(gn_model* m, int64_t n, const gn_ws& ws, float* d_scores, float* d_visfeat,
                          cudaStream_t st, const gn_score_sink* sink = nullptr);
int gn_launch_simt_backbone(gn_model* m, int64_t n, const gn_ws& ws, float* d_scores, float* d_visfeat,
                            cudaStream_t st);
