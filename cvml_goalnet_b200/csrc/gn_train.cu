// Fusion-network training step (BASELINE config 4; SURVEY 8b `gn_fusion_train_step`, 8e cfg 4): forward,
// as-written MSE loss, backward and Adam for AVM.fusion (reference utils.py:242-258, main.py:68-70,187-193).
//   forward  : h_{k+1} = dropout(relu(h_k W_k^T + b_k)), k = 0..3, on tcgen05 (gn_tc_train_gemm mode 3);
//              p = 4 sigmoid(h_4 w_4 + b_4) + 1
//   loss     : nn.MSELoss()(pred (N,1), label (N,)) BROADCASTS to N x N:
//              L = mean_ij (p_i - l_j)^2, dL/dp_i = (2/N)(p_i - mean(l)); `fixed_loss` gives mean_i (p_i - l_i)^2
//   backward : dW_k = dz_k^T h_k (mode 5), dz_{k-1} = (dz_k W_k) * dropout scale * [h_k > 0] (mode 4)
//   Adam     : torch.optim.Adam defaults (lr 1e-3, betas 0.9/0.999, eps 1e-8, no weight decay), fp32 master
//              weights; the bf16 K-major operand copies W_k [N][K] and W_k^T [K][N] are rewritten after the step.
// Gradients live in ONE flat fp32 buffer in the reference's parameter order (fusion.0.weight, fusion.0.bias,
// ..., fusion.12.bias; 755 201 values for the audio-visual model), so data-parallel training is one NCCL
// all-reduce of that buffer between gn_fusion_forward_backward and gn_fusion_adam_step.
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include "gn_common.cuh"

int gn_tc_train_gemm(gn_handle* h, int mode, const __nv_bfloat16* A, int64_t M, int K, int64_t lda,
                     const __nv_bfloat16* B, int N, int64_t ldb, const float* bias, __nv_bfloat16* out,
                     __nv_bfloat16* out_t, int64_t ld_t, float* out_f32, const __nv_bfloat16* gate,
                     const uint8_t* mask, float scale, float keep_prob, unsigned long long seed, cudaStream_t st,
                     const gn_step_state* step_state = nullptr);

struct gn_fusion_trainer {
    gn_handle* h;
    int fin;
    int64_t max_n;                 // rows the workspace was sized for (multiple of 128)
    int kin[5], nout[5];           // layer k: [nout[k]][kin[k]]
    int64_t off_w[5], off_b[5], n_params;
    float *params, *m, *v;         // flat fp32 [n_params]
    __nv_bfloat16 *wk[4], *wt[4];  // hidden layers: W [N][K] and W^T [K][N] in bf16
    __nv_bfloat16 *a[5];           // a[k] = input of layer k, [max_n][kin[k]]
    __nv_bfloat16 *aT[4];          // a[k]^T, [kin[k]][max_n]
    __nv_bfloat16 *dz[4], *dzT[4]; // gradient w.r.t. the pre-activation of layer k: [max_n][nout[k]] and transposed
    float *y, *dz4;                // sigmoid output and gradient of the last pre-activation, [max_n] ([max_n][5] classifier)
    double* sums;                  // [8] sum p, p^2, l, l^2, p*l
    float* partial;                // [blocks][kPartStride] per-block partial sums of dW_4 / db_4
    double* lpart;                 // [blocks] per-block partial sums of the classifier loss
    int partial_blocks;
    int n_cls;                     // 1 = regression head (fusion.12 = Linear(128,1) + Sigmoid, 4y+1); 5 = classifier head
    long long step;                // host copy of the count (the arithmetic uses the device copy in `state`)
    gn_step_state* state;          // device: Adam step count, bias corrections, dropout seed bump (CUDA-graph safe)
    float lr, beta1, beta2, eps;
    void* arena;
};

__global__ void step_advance_kernel(gn_step_state* s, float b1, float b2) {
    const long long t = s->step + 1;
    s->step = t;
    s->bc1 = (float)(1.0 - pow((double)b1, (double)t));
    s->bc2 = (float)(1.0 - pow((double)b2, (double)t));
    s->seed_bump += 0x9E3779B97F4A7C15ull;
}

int gn_launch_step_advance(gn_handle* h, gn_step_state* d_state, float beta1, float beta2, cudaStream_t st) {
    step_advance_kernel<<<1, 1, 0, st>>>(d_state, beta1, beta2);
    GN_LAUNCHED(h);
    return GN_OK;
}

namespace {

constexpr int kLastThreads = 256;
constexpr int kCls = 5;                       // importance levels 1..5 (utils.py:257, main.py:69,96-97)
constexpr int kPartStride = 656;              // >= 5 * 128 + 5 partial sums per block (regression uses the first 129)

// fp32 [n][K] -> bf16 [n][K] and bf16 [K][ld_t] (32 x 32 tiles through shared memory)
__global__ void __launch_bounds__(256)
convert_transpose_kernel(const float* __restrict__ x, int64_t n, int K, __nv_bfloat16* __restrict__ out,
                         __nv_bfloat16* __restrict__ out_t, int64_t ld_t) {
    __shared__ float tile[32][33];
    const int64_t r0 = (int64_t)blockIdx.y * 32;
    const int c0 = blockIdx.x * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;     // 32 x 8
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int64_t r = r0 + ty + 8 * j;
        const int c = c0 + tx;
        float vv = 0.f;
        if (r < n && c < K) {
            vv = x[r * K + c];
            out[r * K + c] = __float2bfloat16(vv);
        }
        tile[ty + 8 * j][tx] = vv;
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int c = c0 + ty + 8 * j;
        const int64_t r = r0 + tx;
        if (c < K && r < ld_t) out_t[(int64_t)c * ld_t + r] = __float2bfloat16(r < n ? tile[tx][ty + 8 * j] : 0.f);
    }
}

// last layer forward: z = a4 . w4 + b4, y = sigmoid(z), p = 4y + 1 (utils.py:256,270); loss statistics
__global__ void __launch_bounds__(kLastThreads)
last_forward_kernel(const __nv_bfloat16* __restrict__ a4, const float* __restrict__ w4, const float* __restrict__ b4,
                    const float* __restrict__ labels, int64_t n, float* __restrict__ y, float* __restrict__ scores,
                    double* __restrict__ sums) {
    __shared__ float s_w[128];
    __shared__ double s_red[5][kLastThreads / 32];
    if (threadIdx.x < 128) s_w[threadIdx.x] = w4[threadIdx.x];
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    double acc[5] = {0, 0, 0, 0, 0};
    // one warp per row: 128 bf16 = 4 per lane
    for (int64_t row = (int64_t)blockIdx.x * (kLastThreads / 32) + warp; row < n; row += (int64_t)gridDim.x * (kLastThreads / 32)) {
        const uint2 hv = *reinterpret_cast<const uint2*>(a4 + row * 128 + lane * 4);
        const float2 h01 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&hv.x));
        const float2 h23 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&hv.y));
        float d = h01.x * s_w[lane * 4] + h01.y * s_w[lane * 4 + 1] + h23.x * s_w[lane * 4 + 2] + h23.y * s_w[lane * 4 + 3];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) d += __shfl_xor_sync(0xffffffffu, d, o);
        if (lane == 0) {
            const float yy = 1.f / (1.f + expf(-(d + b4[0])));
            const float p = 4.f * yy + 1.f;
            y[row] = yy;
            if (scores) scores[row] = p;
            const double l = labels ? (double)labels[row] : 0.0;
            acc[0] += p; acc[1] += (double)p * p; acc[2] += l; acc[3] += l * l; acc[4] += (double)p * l;
        }
    }
    if (lane == 0)
#pragma unroll
        for (int k = 0; k < 5; ++k) s_red[k][warp] = acc[k];
    __syncthreads();
    if (threadIdx.x < 5) {
        double t = 0;
        for (int w = 0; w < kLastThreads / 32; ++w) t += s_red[threadIdx.x][w];
        atomicAdd(sums + threadIdx.x, t);
    }
}

// last layer backward: dL/dp (as written: pulls every prediction to the label MEAN, quirk Q5), through
// 4y+1 and the sigmoid, then into layer 3's pre-activation gradient; per-block partial sums of dW_4, db_4
__global__ void __launch_bounds__(kLastThreads)
last_backward_kernel(const __nv_bfloat16* __restrict__ a4, const float* __restrict__ w4, const float* __restrict__ labels,
                     const float* __restrict__ y, const double* __restrict__ sums, int64_t n, int fixed_loss,
                     float scale, __nv_bfloat16* __restrict__ dz3, __nv_bfloat16* __restrict__ dzT3, int64_t ld_t,
                     float* __restrict__ dz4_out, float* __restrict__ partial) {
    __shared__ float s_w[128];
    __shared__ float s_part[kLastThreads / 32][132];
    if (threadIdx.x < 128) s_w[threadIdx.x] = w4[threadIdx.x];
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const float mean_l = (float)(sums[2] / (double)n);
    const float two_over_n = 2.f / (float)n;
    float gw[4] = {0.f, 0.f, 0.f, 0.f}, gb = 0.f;
    for (int64_t row = (int64_t)blockIdx.x * (kLastThreads / 32) + warp; row < n; row += (int64_t)gridDim.x * (kLastThreads / 32)) {
        const float yy = y[row];
        const float p = 4.f * yy + 1.f;
        const float dp = two_over_n * (p - (fixed_loss ? labels[row] : mean_l));
        const float dz = dp * 4.f * yy * (1.f - yy);
        if (lane == 0) dz4_out[row] = dz;
        gb += dz;
        const uint2 hv = *reinterpret_cast<const uint2*>(a4 + row * 128 + lane * 4);
        const float2 h01 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&hv.x));
        const float2 h23 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&hv.y));
        const float hx[4] = {h01.x, h01.y, h23.x, h23.y};
        __nv_bfloat16 o[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            gw[k] = fmaf(dz, hx[k], gw[k]);
            o[k] = __float2bfloat16(hx[k] > 0.f ? dz * s_w[lane * 4 + k] * scale : 0.f);
            dzT3[(int64_t)(lane * 4 + k) * ld_t + row] = o[k];
        }
        *reinterpret_cast<uint2*>(dz3 + row * 128 + lane * 4) = *reinterpret_cast<const uint2*>(o);
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) s_part[warp][lane * 4 + k] = gw[k];
    if (lane == 0) s_part[warp][128] = gb;
    __syncthreads();
    if (threadIdx.x < 129) {
        float t = 0.f;
        for (int w = 0; w < kLastThreads / 32; ++w) t += s_part[w][threadIdx.x];
        partial[(size_t)blockIdx.x * kPartStride + threadIdx.x] = t;
    }
}

// Classifier head (the training half): fusion.12 = Linear(128, 5) + Softmax(dim = 1) (utils.py:257,
// commented) trained with nn.CrossEntropyLoss (main.py:69,96: `criterion(batch_predictions, (batch_labels-1).long())`).
//   loss_kind 2 -- AS WRITTEN: the criterion receives the model's OUTPUT, i.e. the softmax probabilities p, so the loss is
//                  L_i = -log softmax(p_i)[y_i] (a second softmax inside CrossEntropyLoss);
//                  dL/dp = (softmax(p) - onehot) / N, then through the model's softmax: dz_j = p_j (g_j - sum_k p_k g_k)
//   loss_kind 3 -- cross-entropy on the logits: L_i = -log p_i[y_i], dz = (p - onehot) / N
// One warp per row (4 features per lane); forward, loss, last-layer backward and layer 3's pre-activation gradient in
// one pass; per-block partial sums of dW_4 [5][128], db_4 [5] and the loss, reduced in a fixed order afterwards.
__global__ void __launch_bounds__(kLastThreads)
cls_forward_backward_kernel(const __nv_bfloat16* __restrict__ a4, const float* __restrict__ w4, const float* __restrict__ b4,
                            const float* __restrict__ labels, int64_t n, int loss_kind, float scale,
                            float* __restrict__ probs, __nv_bfloat16* __restrict__ dz3, __nv_bfloat16* __restrict__ dzT3,
                            int64_t ld_t, float* __restrict__ partial, double* __restrict__ lpart) {
    __shared__ float s_w[kCls][128];
    __shared__ float s_part[kLastThreads / 32][kCls * 128 + 8];
    __shared__ double s_loss[kLastThreads / 32];
    for (int i = threadIdx.x; i < kCls * 128; i += kLastThreads) s_w[i / 128][i % 128] = w4[i];
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const float inv_n = 1.f / (float)n;
    float gw[kCls][4], gb[kCls];
#pragma unroll
    for (int k = 0; k < kCls; ++k) { gb[k] = 0.f; gw[k][0] = gw[k][1] = gw[k][2] = gw[k][3] = 0.f; }
    double loss = 0.0;
    for (int64_t row = (int64_t)blockIdx.x * (kLastThreads / 32) + warp; row < n; row += (int64_t)gridDim.x * (kLastThreads / 32)) {
        const uint2 hv = *reinterpret_cast<const uint2*>(a4 + row * 128 + lane * 4);
        const float2 h01 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&hv.x));
        const float2 h23 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&hv.y));
        const float hx[4] = {h01.x, h01.y, h23.x, h23.y};
        float z[kCls];
#pragma unroll
        for (int k = 0; k < kCls; ++k) {
            float d = hx[0] * s_w[k][lane * 4] + hx[1] * s_w[k][lane * 4 + 1] + hx[2] * s_w[k][lane * 4 + 2] +
                      hx[3] * s_w[k][lane * 4 + 3];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) d += __shfl_xor_sync(0xffffffffu, d, o);
            z[k] = d + b4[k];
        }
        float zmax = z[0];
#pragma unroll
        for (int k = 1; k < kCls; ++k) zmax = fmaxf(zmax, z[k]);
        float p[kCls], sum = 0.f;
#pragma unroll
        for (int k = 0; k < kCls; ++k) { p[k] = expf(z[k] - zmax); sum += p[k]; }
        const float inv = 1.f / sum;
#pragma unroll
        for (int k = 0; k < kCls; ++k) p[k] *= inv;
        int yk = (int)labels[row] - 1;
        yk = yk < 0 ? 0 : (yk > kCls - 1 ? kCls - 1 : yk);
        float dz[kCls];
        if (loss_kind == 2) {
            // CrossEntropyLoss applied to the probabilities (values in [0,1]: no max subtraction needed)
            float s2[kCls], sum2 = 0.f, py = 0.f;
#pragma unroll
            for (int k = 0; k < kCls; ++k) { s2[k] = expf(p[k]); sum2 += s2[k]; py = (k == yk) ? p[k] : py; }
            const float inv2 = 1.f / sum2;
            float dot = 0.f, g[kCls];
#pragma unroll
            for (int k = 0; k < kCls; ++k) {
                g[k] = (s2[k] * inv2 - (k == yk ? 1.f : 0.f)) * inv_n;
                dot = fmaf(p[k], g[k], dot);
            }
#pragma unroll
            for (int k = 0; k < kCls; ++k) dz[k] = p[k] * (g[k] - dot);
            if (lane == 0) loss += (double)(logf(sum2) - py);
        } else {
            float zy = 0.f;
#pragma unroll
            for (int k = 0; k < kCls; ++k) { dz[k] = (p[k] - (k == yk ? 1.f : 0.f)) * inv_n; zy = (k == yk) ? z[k] : zy; }
            if (lane == 0) loss += (double)(logf(sum) + zmax - zy);
        }
        if (lane == 0 && probs)
#pragma unroll
            for (int k = 0; k < kCls; ++k) probs[row * kCls + k] = p[k];
        __nv_bfloat16 o[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            float back = 0.f;
#pragma unroll
            for (int k = 0; k < kCls; ++k) {
                gw[k][i] = fmaf(dz[k], hx[i], gw[k][i]);
                back = fmaf(dz[k], s_w[k][lane * 4 + i], back);
            }
            o[i] = __float2bfloat16(hx[i] > 0.f ? back * scale : 0.f);
            dzT3[(int64_t)(lane * 4 + i) * ld_t + row] = o[i];
        }
        *reinterpret_cast<uint2*>(dz3 + row * 128 + lane * 4) = *reinterpret_cast<const uint2*>(o);
#pragma unroll
        for (int k = 0; k < kCls; ++k) gb[k] += dz[k];
    }
#pragma unroll
    for (int k = 0; k < kCls; ++k) {
#pragma unroll
        for (int i = 0; i < 4; ++i) s_part[warp][k * 128 + lane * 4 + i] = gw[k][i];
        if (lane == 0) s_part[warp][kCls * 128 + k] = gb[k];
    }
    if (lane == 0) s_loss[warp] = loss;
    __syncthreads();
    for (int j = threadIdx.x; j < kCls * 128 + kCls; j += kLastThreads) {
        float t = 0.f;
        for (int w = 0; w < kLastThreads / 32; ++w) t += s_part[w][j];
        partial[(size_t)blockIdx.x * kPartStride + j] = t;
    }
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int w = 0; w < kLastThreads / 32; ++w) t += s_loss[w];
        lpart[blockIdx.x] = t;
    }
}

// deterministic reduction of the classifier head's per-block partials into grads[W_4 | b_4]; mean loss
__global__ void __launch_bounds__(256)
cls_finalize_kernel(const float* __restrict__ partial, const double* __restrict__ lpart, int blocks, int64_t n,
                    float* __restrict__ g_w4, float* __restrict__ g_b4, double* __restrict__ loss) {
    for (int j = threadIdx.x; j < kCls * 128 + kCls; j += 256) {
        float t = 0.f;
        for (int b = 0; b < blocks; ++b) t += partial[(size_t)b * kPartStride + j];
        if (j < kCls * 128) g_w4[j] = t; else g_b4[j - kCls * 128] = t;
    }
    if (threadIdx.x == 255 && loss) {
        double t = 0.0;
        for (int b = 0; b < blocks; ++b) t += lpart[b];
        *loss = t / (double)n;
    }
}

// deterministic reduction of the per-block partials into grads[W_4 | b_4]; loss value
__global__ void __launch_bounds__(160)
last_finalize_kernel(const float* __restrict__ partial, int blocks, const double* __restrict__ sums, int64_t n,
                     int fixed_loss, float* __restrict__ g_w4, float* __restrict__ g_b4, double* __restrict__ loss) {
    const int j = threadIdx.x;
    if (j < 129) {
        float t = 0.f;
        for (int b = 0; b < blocks; ++b) t += partial[(size_t)b * kPartStride + j];
        if (j < 128) g_w4[j] = t; else g_b4[0] = t;
    }
    if (j == 159 && loss) {
        const double N = (double)n, sp = sums[0], spp = sums[1], sl = sums[2], sll = sums[3], spl = sums[4];
        // mean_ij (p_i - l_j)^2 = mean p^2 - 2 mean p mean l + mean l^2 ;  mean_i (p_i - l_i)^2
        *loss = fixed_loss ? (spp - 2.0 * spl + sll) / N : spp / N - 2.0 * (sp / N) * (sl / N) + sll / N;
    }
}

// bias gradient: db[j] = sum over rows of dz^T[j][:]  (one warp per output feature)
__global__ void __launch_bounds__(256)
rowsum_kernel(const __nv_bfloat16* __restrict__ xt, int rows, int64_t n, int64_t ld, float* __restrict__ out) {
    const int lane = threadIdx.x & 31;
    const int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (row >= rows) return;
    const __nv_bfloat16* p = xt + (int64_t)row * ld;
    float acc = 0.f;
    for (int64_t i = lane * 8; i + 7 < n; i += 256) {
        const uint4 vv = *reinterpret_cast<const uint4*>(p + i);
        const uint32_t w4[4] = {vv.x, vv.y, vv.z, vv.w};
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const float2 f = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&w4[k]));
            acc += f.x + f.y;
        }
    }
    for (int64_t i = (n & ~7LL) + lane; i < n; i += 32) acc += __bfloat162float(p[i]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (lane == 0) out[row] = acc;
}

// torch.optim.Adam (no weight decay, no amsgrad) on the flat parameter vector
__global__ void __launch_bounds__(256)
adam_kernel(float* __restrict__ p, float* __restrict__ m, float* __restrict__ v, const float* __restrict__ g,
            int64_t count, float grad_scale, float lr, float b1, float b2, float eps, const gn_step_state* __restrict__ state) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    const float bc1 = state->bc1, bc2_sqrt = sqrtf(state->bc2);
    const float gi = g[i] * grad_scale;
    const float mi = b1 * m[i] + (1.f - b1) * gi;
    const float vi = b2 * v[i] + (1.f - b2) * gi * gi;
    m[i] = mi;
    v[i] = vi;
    const float denom = sqrtf(vi) / bc2_sqrt + eps;
    p[i] -= (lr / bc1) * (mi / denom);
}

// fp32 master W [N][K] -> bf16 W [N][K] and bf16 W^T [K][N]
__global__ void __launch_bounds__(256)
repack_kernel(const float* __restrict__ w, int N, int K, __nv_bfloat16* __restrict__ wk, __nv_bfloat16* __restrict__ wt) {
    __shared__ float tile[32][33];
    const int n0 = blockIdx.y * 32, k0 = blockIdx.x * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int nn = n0 + ty + 8 * j, kk = k0 + tx;
        float vv = 0.f;
        if (nn < N && kk < K) {
            vv = w[(size_t)nn * K + kk];
            wk[(size_t)nn * K + kk] = __float2bfloat16(vv);
        }
        tile[ty + 8 * j][tx] = vv;
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int kk = k0 + ty + 8 * j, nn = n0 + tx;
        if (kk < K && nn < N) wt[(size_t)kk * N + nn] = __float2bfloat16(tile[tx][ty + 8 * j]);
    }
}

int repack_all(gn_fusion_trainer* t, cudaStream_t st) {
    for (int k = 0; k < 4; ++k) {
        dim3 grid((t->kin[k] + 31) / 32, (t->nout[k] + 31) / 32);
        repack_kernel<<<grid, 256, 0, st>>>(t->params + t->off_w[k], t->nout[k], t->kin[k], t->wk[k], t->wt[k]);
        GN_LAUNCHED(t->h);
    }
    return GN_OK;
}

}  // namespace

// ================================================================================================
// C ABI
// ================================================================================================
extern "C" int gn_fusion_trainer_create(gn_handle* h, gn_fusion_trainer** out, const gn_fusion_state* sd, int fin,
                                        int64_t max_rows, void* stream) {
    return gn_fusion_trainer_create_head(h, out, sd, fin, max_rows, 1, stream);
}

extern "C" int gn_fusion_trainer_create_head(gn_handle* h, gn_fusion_trainer** out, const gn_fusion_state* sd, int fin,
                                             int64_t max_rows, int n_classes, void* stream) {
    if (!h || !out || !sd || max_rows <= 0) return GN_E_ARG;
    if (fin != 640 && fin != 512) return GN_E_SHAPE;
    if (n_classes != 1 && n_classes != kCls) return GN_E_SHAPE;
    *out = nullptr;
    for (int k = 0; k < 5; ++k)
        if (!sd->weight[k] || !sd->bias[k]) return GN_E_ARG;
    gn_fusion_trainer* t = reinterpret_cast<gn_fusion_trainer*>(calloc(1, sizeof(gn_fusion_trainer)));
    if (!t) return GN_E_ARG;
    t->h = h; t->fin = fin; t->n_cls = n_classes;
    t->max_n = (max_rows + 127) / 128 * 128;
    const int kin[5] = {fin, 512, 512, 256, 128}, nout[5] = {512, 512, 256, 128, n_classes};
    int64_t off = 0;
    for (int k = 0; k < 5; ++k) {
        t->kin[k] = kin[k]; t->nout[k] = nout[k];
        t->off_w[k] = off; off += (int64_t)nout[k] * kin[k];
        t->off_b[k] = off; off += nout[k];
    }
    t->n_params = off;
    t->lr = 1e-3f; t->beta1 = 0.9f; t->beta2 = 0.999f; t->eps = 1e-8f; t->step = 0;
    t->partial_blocks = 2 * h->num_sms;
    // one arena: everything 256-byte aligned
    size_t bytes = 0;
    auto take = [&](size_t b) { const size_t o = bytes; bytes += (b + 255) & ~(size_t)255; return o; };
    const size_t o_params = take(off * 4), o_m = take(off * 4), o_v = take(off * 4);
    size_t o_wk[4], o_wt[4], o_a[5], o_aT[4], o_dz[4], o_dzT[4];
    for (int k = 0; k < 4; ++k) { o_wk[k] = take((size_t)nout[k] * kin[k] * 2); o_wt[k] = take((size_t)nout[k] * kin[k] * 2); }
    for (int k = 0; k < 5; ++k) o_a[k] = take((size_t)t->max_n * kin[k] * 2);
    for (int k = 0; k < 4; ++k) {
        o_aT[k] = take((size_t)t->max_n * kin[k] * 2);
        o_dz[k] = take((size_t)t->max_n * nout[k] * 2);
        o_dzT[k] = take((size_t)t->max_n * nout[k] * 2);
    }
    const size_t o_state = take(sizeof(gn_step_state));
    const size_t o_y = take(t->max_n * 4 * kCls), o_dz4 = take(t->max_n * 4), o_sums = take(64),
                 o_part = take((size_t)t->partial_blocks * kPartStride * 4), o_lpart = take((size_t)t->partial_blocks * 8);
    cudaError_t e = cudaMalloc(&t->arena, bytes);
    if (e != cudaSuccess) { free(t); return (int)e; }
    char* base = reinterpret_cast<char*>(t->arena);
    cudaStream_t st = gn_stream(stream);
    e = cudaMemsetAsync(base, 0, bytes, st);
    if (e != cudaSuccess) { cudaFree(t->arena); free(t); return (int)e; }
    t->params = reinterpret_cast<float*>(base + o_params);
    t->m = reinterpret_cast<float*>(base + o_m);
    t->v = reinterpret_cast<float*>(base + o_v);
    for (int k = 0; k < 4; ++k) {
        t->wk[k] = reinterpret_cast<__nv_bfloat16*>(base + o_wk[k]);
        t->wt[k] = reinterpret_cast<__nv_bfloat16*>(base + o_wt[k]);
        t->aT[k] = reinterpret_cast<__nv_bfloat16*>(base + o_aT[k]);
        t->dz[k] = reinterpret_cast<__nv_bfloat16*>(base + o_dz[k]);
        t->dzT[k] = reinterpret_cast<__nv_bfloat16*>(base + o_dzT[k]);
    }
    for (int k = 0; k < 5; ++k) t->a[k] = reinterpret_cast<__nv_bfloat16*>(base + o_a[k]);
    t->y = reinterpret_cast<float*>(base + o_y);
    t->dz4 = reinterpret_cast<float*>(base + o_dz4);
    t->sums = reinterpret_cast<double*>(base + o_sums);
    t->partial = reinterpret_cast<float*>(base + o_part);
    t->lpart = reinterpret_cast<double*>(base + o_lpart);
    t->state = reinterpret_cast<gn_step_state*>(base + o_state);       // zeroed with the arena: step 0, no seed bump
    for (int k = 0; k < 5 && e == cudaSuccess; ++k) {
        e = cudaMemcpyAsync(t->params + t->off_w[k], sd->weight[k], (size_t)nout[k] * kin[k] * 4, cudaMemcpyDeviceToDevice, st);
        if (e == cudaSuccess)
            e = cudaMemcpyAsync(t->params + t->off_b[k], sd->bias[k], (size_t)nout[k] * 4, cudaMemcpyDeviceToDevice, st);
    }
    if (e != cudaSuccess) { cudaFree(t->arena); free(t); return (int)e; }
    const int rc = repack_all(t, st);
    if (rc != GN_OK) { cudaFree(t->arena); free(t); return rc; }
    *out = t;
    return GN_OK;
}

extern "C" int gn_fusion_trainer_destroy(gn_fusion_trainer* t) {
    if (!t) return GN_OK;
    cudaFree(t->arena);
    free(t);
    return GN_OK;
}

extern "C" int64_t gn_fusion_param_count(const gn_fusion_trainer* t) { return t ? t->n_params : GN_E_ARG; }

extern "C" int gn_fusion_set_hyper(gn_fusion_trainer* t, float lr, float beta1, float beta2, float eps) {
    if (!t || lr < 0 || beta1 < 0 || beta1 >= 1 || beta2 < 0 || beta2 >= 1 || eps < 0) return GN_E_ARG;
    t->lr = lr; t->beta1 = beta1; t->beta2 = beta2; t->eps = eps;
    return GN_OK;
}

extern "C" int gn_fusion_get_params(gn_fusion_trainer* t, float* d_out, void* stream) {
    if (!t || !d_out) return GN_E_ARG;
    GN_CUDA(cudaMemcpyAsync(d_out, t->params, t->n_params * 4, cudaMemcpyDeviceToDevice, gn_stream(stream)));
    return GN_OK;
}

extern "C" int gn_fusion_forward_backward(gn_fusion_trainer* t, const float* d_features, const float* d_labels, int64_t n,
                                          int fixed_loss, float dropout_p, const uint8_t* const* d_masks,
                                          unsigned long long seed, float* d_grads, float* d_scores, double* d_loss,
                                          void* stream) {
    if (!t || n <= 0 || !d_features || !d_labels || !d_grads) return GN_E_ARG;
    if (n > t->max_n) return GN_E_WORKSPACE;
    if (dropout_p < 0.f || dropout_p >= 1.f) return GN_E_ARG;
    // fixed_loss doubles as the loss kind: 0 = MSE as written (N x N broadcast), 1 = element-wise MSE (regression head);
    // 2 = CrossEntropyLoss on the softmax output as written, 3 = cross-entropy on the logits (classifier head)
    if (t->n_cls == 1 ? (fixed_loss != 0 && fixed_loss != 1) : (fixed_loss != 2 && fixed_loss != 3)) return GN_E_ARG;
    gn_handle* h = t->h;
    cudaStream_t st = gn_stream(stream);
    const float keep = 1.f - dropout_p;
    const float scale = dropout_p > 0.f ? 1.f / keep : 1.f;
    const int64_t ld = t->max_n;
    int rc;
    // features -> bf16 a0 and a0^T
    {
        dim3 grid((t->fin + 31) / 32, (unsigned)((n + 31) / 32));
        convert_transpose_kernel<<<grid, 256, 0, st>>>(d_features, n, t->fin, t->a[0], t->aT[0], ld);
        GN_LAUNCHED(h);
    }
    // forward through the four hidden layers (Linear + ReLU + Dropout, utils.py:243-254)
    for (int k = 0; k < 4; ++k) {
        const uint8_t* mask = (d_masks && dropout_p > 0.f) ? d_masks[k] : nullptr;
        rc = gn_tc_train_gemm(h, 3, t->a[k], n, t->kin[k], t->kin[k], t->wk[k], t->nout[k], t->kin[k],
                              t->params + t->off_b[k], t->a[k + 1], k < 3 ? t->aT[k + 1] : nullptr, ld, nullptr, nullptr,
                              mask, scale, keep, seed + 0x632BE59BD9B4E019ull * (unsigned long long)(k + 1), st, t->state);
        if (rc) return rc;
    }
    GN_CUDA(cudaMemsetAsync(t->sums, 0, 64, st));
    int blocks = (int)((n + kLastThreads / 32 - 1) / (kLastThreads / 32));
    if (blocks > t->partial_blocks) blocks = t->partial_blocks;
    if (t->n_cls == kCls) {
        cls_forward_backward_kernel<<<blocks, kLastThreads, 0, st>>>(
            t->a[4], t->params + t->off_w[4], t->params + t->off_b[4], d_labels, n, fixed_loss, scale, d_scores,
            t->dz[3], t->dzT[3], ld, t->partial, t->lpart);
        GN_LAUNCHED(h);
        cls_finalize_kernel<<<1, 256, 0, st>>>(t->partial, t->lpart, blocks, n, d_grads + t->off_w[4],
                                               d_grads + t->off_b[4], d_loss);
        GN_LAUNCHED(h);
    } else {
    last_forward_kernel<<<blocks, kLastThreads, 0, st>>>(t->a[4], t->params + t->off_w[4], t->params + t->off_b[4],
                                                        d_labels, n, t->y, d_scores, t->sums);
    GN_LAUNCHED(h);
    last_backward_kernel<<<blocks, kLastThreads, 0, st>>>(t->a[4], t->params + t->off_w[4], d_labels, t->y, t->sums, n,
                                                         fixed_loss, scale, t->dz[3], t->dzT[3], ld, t->dz4, t->partial);
    GN_LAUNCHED(h);
    last_finalize_kernel<<<1, 160, 0, st>>>(t->partial, blocks, t->sums, n, fixed_loss, d_grads + t->off_w[4],
                                            d_grads + t->off_b[4], d_loss);
    GN_LAUNCHED(h);
    }
    // the weight-gradient GEMMs reduce over K = n rounded up to 64: zero the pad columns of the transposed operands
    // (a previous call with a larger n may have left values there)
    const int64_t kred = (n + 63) / 64 * 64;
    if (kred != n)
        for (int k = 0; k < 4; ++k) {
            GN_CUDA(cudaMemset2DAsync(t->aT[k] + n, ld * 2, 0, (kred - n) * 2, t->kin[k], st));
            GN_CUDA(cudaMemset2DAsync(t->dzT[k] + n, ld * 2, 0, (kred - n) * 2, t->nout[k], st));
        }
    for (int k = 3; k >= 0; --k) {
        // dW_k = dz_k^T a_k : A = dz_k^T [nout][n], B = a_k^T [kin][n], fp32 [nout][kin] straight into the flat gradient
        rc = gn_tc_train_gemm(h, 5, t->dzT[k], t->nout[k], (int)kred, ld, t->aT[k], t->kin[k], ld, nullptr, nullptr,
                              nullptr, 0, d_grads + t->off_w[k], nullptr, nullptr, 1.f, 1.f, 0ull, st);
        if (rc) return rc;
        rowsum_kernel<<<(t->nout[k] * 32 + 255) / 256, 256, 0, st>>>(t->dzT[k], t->nout[k], n, ld, d_grads + t->off_b[k]);
        GN_LAUNCHED(h);
        if (k > 0) {
            // dz_{k-1} = (dz_k W_k) * dropout scale * [a_k > 0] : A = dz_k [n][nout], B = W_k^T [kin][nout]
            rc = gn_tc_train_gemm(h, 4, t->dz[k], n, t->nout[k], t->nout[k], t->wt[k], t->kin[k], t->nout[k], nullptr,
                                  t->dz[k - 1], t->dzT[k - 1], ld, nullptr, t->a[k], nullptr, scale, 1.f, 0ull, st);
            if (rc) return rc;
        }
    }
    return GN_OK;
}

extern "C" int gn_fusion_adam_step(gn_fusion_trainer* t, const float* d_grads, float grad_scale, void* stream) {
    if (!t || !d_grads) return GN_E_ARG;
    cudaStream_t st = gn_stream(stream);
    t->step += 1;
    // step count, bias corrections and the next step's dropout seed advance ON THE DEVICE (a captured graph replays them)
    int rc0 = gn_launch_step_advance(t->h, t->state, t->beta1, t->beta2, st);
    if (rc0) return rc0;
    adam_kernel<<<(unsigned)((t->n_params + 255) / 256), 256, 0, st>>>(t->params, t->m, t->v, d_grads, t->n_params, grad_scale,
                                                                       t->lr, t->beta1, t->beta2, t->eps, t->state);
    GN_LAUNCHED(t->h);
    return repack_all(t, st);
}

extern "C" int gn_fusion_train_step(gn_fusion_trainer* t, const float* d_features, const float* d_labels, int64_t n,
                                    int fixed_loss, float dropout_p, const uint8_t* const* d_masks, unsigned long long seed,
                                    float* d_grads, float* d_scores, double* d_loss, void* stream) {
    const int rc = gn_fusion_forward_backward(t, d_features, d_labels, n, fixed_loss, dropout_p, d_masks, seed, d_grads,
                                              d_scores, d_loss, stream);
    if (rc) return rc;
    return gn_fusion_adam_step(t, d_grads, 1.f, stream);
}
