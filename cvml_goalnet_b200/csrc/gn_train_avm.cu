// Full-AVM training step -- replaces the reference's train sub-step main.py:187-193
//     optimizer.zero_grad(); pred = model(audios, frames); loss = MSELoss()(pred, labels); loss.backward(); optimizer.step()
// for the WHOLE network in train mode, as the reference runs it (it never calls .eval(), quirk Q4):
// BatchNorm2d uses the statistics of the sub-batch and updates its running buffers (momentum 0.1, unbiased running
// variance), Dropout(0.2) is active after linear5 and after fusion.0/3/6/9 (utils.py:168,245-254).
// Sub-batches are 10 frames (main.py:47), so every layer is tiny and latency-bound; this first version is plain
// fp32 CUDA-core kernels in the reference's own NCHW / [out][in] layouts (no repacking, tight parity with torch
// autograd).  The tensor-core training GEMMs of gn_train.cu cover the fusion network at batch 4096 (config 4).
#include <math.h>
#include <stdlib.h>

#include "gn_common.cuh"

namespace {

constexpr int kT = 256;
inline int grid_for(int64_t total, int cap) { int64_t b = (total + kT - 1) / kT; return (int)(b < 1 ? 1 : (b > cap ? cap : b)); }

// ---- forward ------------------------------------------------------------------------------------------------
// y[n][co][oy][ox] = b[co] + sum_{ci,kh,kw} x[n][ci][oy*sh-ph+kh][ox*sw-pw+kw] * w[co][ci][kh][kw]   (+ ReLU)
__global__ void __launch_bounds__(kT)
conv_fwd_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ b, int N, int Ci, int H,
                int W, int Co, int KH, int KW, int sh, int sw, int ph, int pw, int OH, int OW, int relu,
                float* __restrict__ y) {
    const int64_t total = (int64_t)N * Co * OH * OW;
    for (int64_t i = blockIdx.x * (int64_t)kT + threadIdx.x; i < total; i += (int64_t)gridDim.x * kT) {
        const int ox = (int)(i % OW), oy = (int)((i / OW) % OH), co = (int)((i / ((int64_t)OW * OH)) % Co);
        const int n = (int)(i / ((int64_t)OW * OH * Co));
        float acc = b[co];
        for (int ci = 0; ci < Ci; ++ci)
            for (int kh = 0; kh < KH; ++kh) {
                const int iy = oy * sh - ph + kh;
                if (iy < 0 || iy >= H) continue;
                for (int kw = 0; kw < KW; ++kw) {
                    const int ix = ox * sw - pw + kw;
                    if (ix < 0 || ix >= W) continue;
                    acc = fmaf(x[(((int64_t)n * Ci + ci) * H + iy) * W + ix], w[((co * Ci + ci) * KH + kh) * KW + kw], acc);
                }
            }
        y[i] = relu ? fmaxf(acc, 0.f) : acc;
    }
}

// p = MaxPool(3, stride 1) of y (already ReLU'd); arg = window offset (dy*3+dx) of the first maximum (torch scan order)
__global__ void __launch_bounds__(kT)
pool_fwd_kernel(const float* __restrict__ y, int NC, int OH, int OW, float* __restrict__ p, uint8_t* __restrict__ arg) {
    const int PH = OH - 2, PW = OW - 2;
    const int64_t total = (int64_t)NC * PH * PW;
    for (int64_t i = blockIdx.x * (int64_t)kT + threadIdx.x; i < total; i += (int64_t)gridDim.x * kT) {
        const int px = (int)(i % PW), py = (int)((i / PW) % PH);
        const int64_t nc = i / ((int64_t)PW * PH);
        const float* base = y + (nc * OH + py) * OW + px;
        float m = base[0];
        int a = 0;
        for (int d = 1; d < 9; ++d) {
            const float v = base[(d / 3) * OW + d % 3];
            if (v > m) { m = v; a = d; }
        }
        p[i] = m;
        arg[i] = (uint8_t)a;
    }
}

// BatchNorm2d in training mode, one CTA per channel: batch mean / biased variance over (N, H, W), normalise, affine;
// running_mean/var <- 0.9 * old + 0.1 * (mean, unbiased variance).  Keeps xhat and invstd for the backward.
__global__ void __launch_bounds__(kT)
bn_fwd_kernel(const float* __restrict__ p, int N, int C, int HW, const float* __restrict__ gamma,
              const float* __restrict__ beta, float* __restrict__ rmean, float* __restrict__ rvar, float* __restrict__ xhat,
              float* __restrict__ invstd_out, float* __restrict__ out) {
    __shared__ double s1[kT], s2[kT];
    const int c = blockIdx.x, m = N * HW;
    double a = 0.0, q = 0.0;
    for (int i = threadIdx.x; i < m; i += kT) {
        const float v = p[((int64_t)(i / HW) * C + c) * HW + i % HW];
        a += v;
        q += (double)v * v;
    }
    s1[threadIdx.x] = a; s2[threadIdx.x] = q;
    __syncthreads();
    for (int o = kT / 2; o > 0; o >>= 1) {
        if (threadIdx.x < o) { s1[threadIdx.x] += s1[threadIdx.x + o]; s2[threadIdx.x] += s2[threadIdx.x + o]; }
        __syncthreads();
    }
    const double mean = s1[0] / m;
    double var = s2[0] / m - mean * mean;
    if (var < 0.0) var = 0.0;
    const float istd = (float)(1.0 / sqrt(var + 1e-5));
    if (threadIdx.x == 0) {
        invstd_out[c] = istd;
        rmean[c] = 0.9f * rmean[c] + 0.1f * (float)mean;
        rvar[c] = 0.9f * rvar[c] + 0.1f * (float)(m > 1 ? var * m / (m - 1) : var);
    }
    const float g = gamma[c], bt = beta[c], mu = (float)mean;
    for (int i = threadIdx.x; i < m; i += kT) {
        const int64_t idx = ((int64_t)(i / HW) * C + c) * HW + i % HW;
        const float xh = (p[idx] - mu) * istd;
        xhat[idx] = xh;
        out[idx] = fmaf(xh, g, bt);
    }
}

// y[n][o] = act(b[o] + x[n] . w[o]) ; one warp per output; optional ReLU and dropout keep-mask * scale
__global__ void __launch_bounds__(kT)
linear_fwd_kernel(const float* __restrict__ x, int ldx, const float* __restrict__ w, const float* __restrict__ b, int N, int K,
                  int O, int relu, const uint8_t* __restrict__ mask, float scale, float* __restrict__ y, int ldy,
                  int col_off) {
    const int lane = threadIdx.x & 31;
    const int64_t gw = (blockIdx.x * (int64_t)kT + threadIdx.x) >> 5, nw = ((int64_t)gridDim.x * kT) >> 5;
    for (int64_t o = gw; o < (int64_t)N * O; o += nw) {
        const int n = (int)(o / O), j = (int)(o - (int64_t)n * O);
        const float* xr = x + (int64_t)n * ldx;
        const float* wr = w + (int64_t)j * K;
        float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;            // four independent chains, loads batched by the unroll
        int k = lane;
#pragma unroll 2
        for (; k + 96 < K; k += 128) {
            a0 = fmaf(xr[k], wr[k], a0);
            a1 = fmaf(xr[k + 32], wr[k + 32], a1);
            a2 = fmaf(xr[k + 64], wr[k + 64], a2);
            a3 = fmaf(xr[k + 96], wr[k + 96], a3);
        }
        for (; k < K; k += 32) a0 = fmaf(xr[k], wr[k], a0);
        float acc = (a0 + a1) + (a2 + a3);
        for (int s = 16; s > 0; s >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, s);
        if (lane == 0) {
            float v = acc + b[j];
            if (relu) v = fmaxf(v, 0.f);
            if (mask) v = mask[(int64_t)n * O + j] ? v * scale : 0.f;
            y[(int64_t)n * ldy + col_off + j] = v;
        }
    }
}

// keep-mask from a counter-based hash (splitmix64 of seed + site offset + element index), P(keep) = keep_prob
__global__ void __launch_bounds__(kT)
make_mask_kernel(uint8_t* __restrict__ mask, int64_t total, unsigned long long seed, float keep_prob,
                 const gn_step_state* __restrict__ state) {
    seed += state->seed_bump;         // advances on the device every step, so a captured graph draws fresh masks on replay
    for (int64_t i = blockIdx.x * (int64_t)kT + threadIdx.x; i < total; i += (int64_t)gridDim.x * kT) {
        unsigned long long z = seed + (unsigned long long)i * 0x9E3779B97F4A7C15ull;
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
        z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
        z ^= z >> 31;
        mask[i] = ((float)(unsigned int)(z >> 40) * (1.f / 16777216.f)) < keep_prob ? 1 : 0;
    }
}

// scores = 4 sigmoid(z) + 1; loss = mean_ij (p_i - l_j)^2 (as written, quirk Q5) or mean_i (p_i - l_i)^2;
// dz_i = dL/dp_i * 4 s (1 - s).  One CTA (n is a sub-batch).
__global__ void __launch_bounds__(kT)
loss_kernel(const float* __restrict__ z, const float* __restrict__ labels, int n, int fixed_loss, float* __restrict__ scores,
            float* __restrict__ dz, double* __restrict__ loss) {
    __shared__ double sl[kT], sl2[kT], sp[kT];
    double a = 0.0, b = 0.0;
    for (int i = threadIdx.x; i < n; i += kT) { a += labels[i]; b += (double)labels[i] * labels[i]; }
    sl[threadIdx.x] = a; sl2[threadIdx.x] = b;
    __syncthreads();
    for (int o = kT / 2; o > 0; o >>= 1) {
        if (threadIdx.x < o) { sl[threadIdx.x] += sl[threadIdx.x + o]; sl2[threadIdx.x] += sl2[threadIdx.x + o]; }
        __syncthreads();
    }
    const double lmean = sl[0] / n, l2mean = sl2[0] / n;
    double part = 0.0;
    for (int i = threadIdx.x; i < n; i += kT) {
        const float s = 1.f / (1.f + expf(-z[i]));
        const float p = 4.f * s + 1.f;
        if (scores) scores[i] = p;
        double dp;
        if (fixed_loss) {
            const double d = (double)p - labels[i];
            part += d * d / n;
            dp = 2.0 * d / n;
        } else {
            // mean_j (p - l_j)^2 = p^2 - 2 p mean(l) + mean(l^2), averaged over i
            part += ((double)p * p - 2.0 * p * lmean + l2mean) / n;
            dp = 2.0 * ((double)p - lmean) / n;
        }
        dz[i] = (float)(dp * 4.0 * s * (1.0 - s));
    }
    sp[threadIdx.x] = part;
    __syncthreads();
    for (int o = kT / 2; o > 0; o >>= 1) {
        if (threadIdx.x < o) sp[threadIdx.x] += sp[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0 && loss) *loss = sp[0];
}

}  // namespace

namespace {

// ---- backward -----------------------------------------------------------------------------------------------
// dx[n][k] = sum_o dy[n][o] * w[o][k]; thread per (n, k)
__global__ void __launch_bounds__(kT)
linear_bwd_x_kernel(const float* __restrict__ dy, int ldy, int col_off, const float* __restrict__ w, int N, int K, int O,
                    float* __restrict__ dx) {
    const int64_t total = (int64_t)N * K;
    for (int64_t i = blockIdx.x * (int64_t)kT + threadIdx.x; i < total; i += (int64_t)gridDim.x * kT) {
        const int n = (int)(i / K), k = (int)(i - (int64_t)n * K);
        const float* d = dy + (int64_t)n * ldy + col_off;
        float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
        int o = 0;
#pragma unroll 2
        for (; o + 3 < O; o += 4) {
            a0 = fmaf(d[o], w[(int64_t)o * K + k], a0);
            a1 = fmaf(d[o + 1], w[(int64_t)(o + 1) * K + k], a1);
            a2 = fmaf(d[o + 2], w[(int64_t)(o + 2) * K + k], a2);
            a3 = fmaf(d[o + 3], w[(int64_t)(o + 3) * K + k], a3);
        }
        for (; o < O; ++o) a0 = fmaf(d[o], w[(int64_t)o * K + k], a0);
        dx[i] = (a0 + a1) + (a2 + a3);
    }
}

// dw[o][k] = sum_n dy[n][o] * x[n][k]; db[o] = sum_n dy[n][o]; thread per (o, k)
__global__ void __launch_bounds__(kT)
linear_bwd_w_kernel(const float* __restrict__ dy, int ldy, int col_off, const float* __restrict__ x, int ldx, int N, int K,
                    int O, float* __restrict__ dw, float* __restrict__ db) {
    const int64_t total = (int64_t)O * K;
    for (int64_t i = blockIdx.x * (int64_t)kT + threadIdx.x; i < total; i += (int64_t)gridDim.x * kT) {
        const int o = (int)(i / K), k = (int)(i - (int64_t)o * K);
        float acc = 0.f, accb = 0.f;
        for (int n = 0; n < N; ++n) {
            const float d = dy[(int64_t)n * ldy + col_off + o];
            acc = fmaf(d, x[(int64_t)n * ldx + k], acc);
            accb += d;
        }
        dw[i] = acc;
        if (k == 0) db[o] = accb;
    }
}

// Sub-batch versions (N <= kMaxRows): every weight is read ONCE and used for all rows, instead of once per row.
constexpr int kMaxRows = 16;
// y[n][o]: one warp per output feature o, all rows at once
__global__ void __launch_bounds__(kT)
linear_fwd_rows_kernel(const float* __restrict__ x, int ldx, const float* __restrict__ w, const float* __restrict__ b, int N,
                       int K, int O, int relu, const uint8_t* __restrict__ mask, float scale, float* __restrict__ y, int ldy,
                       int col_off) {
    const int lane = threadIdx.x & 31;
    const int gw = (int)((blockIdx.x * (int64_t)kT + threadIdx.x) >> 5), nw = (int)(((int64_t)gridDim.x * kT) >> 5);
    for (int j = gw; j < O; j += nw) {
        const float* wr = w + (int64_t)j * K;
        float acc[kMaxRows];
#pragma unroll
        for (int n = 0; n < kMaxRows; ++n) acc[n] = 0.f;
#pragma unroll 4
        for (int k = lane; k < K; k += 32) {
            const float wv = wr[k];
#pragma unroll
            for (int n = 0; n < kMaxRows; ++n)
                if (n < N) acc[n] = fmaf(x[(int64_t)n * ldx + k], wv, acc[n]);
        }
#pragma unroll
        for (int n = 0; n < kMaxRows; ++n) {
            if (n >= N) break;
            float v = acc[n];
            for (int s = 16; s > 0; s >>= 1) v += __shfl_xor_sync(0xffffffffu, v, s);
            if (lane == 0) {
                v += b[j];
                if (relu) v = fmaxf(v, 0.f);
                if (mask) v = mask[(int64_t)n * O + j] ? v * scale : 0.f;
                y[(int64_t)n * ldy + col_off + j] = v;
            }
        }
    }
}
// Long reductions (linear5: K = 41 472): the K axis is cut into chunks of kLinCK, one CTA per (chunk, 64 features); the
// chunk's activations sit in shared memory, every weight is read once and used for all rows; partial sums
// part[chunk][n][o] are reduced (in chunk order: deterministic) by linear_fwd_reduce_kernel with bias / ReLU / dropout.
constexpr int kLinCK = 512;
__global__ void __launch_bounds__(kT)
linear_fwd_chunk_kernel(const float* __restrict__ x, int ldx, const float* __restrict__ w, int N, int K, int O,
                        float* __restrict__ part) {
    extern __shared__ float xs[];                    // [N][kLinCK]
    const int k0 = blockIdx.x * kLinCK, kc = min(kLinCK, K - k0);
    for (int i = threadIdx.x; i < N * kLinCK; i += kT) {
        const int n = i / kLinCK, k = i - n * kLinCK;
        xs[i] = k < kc ? x[(int64_t)n * ldx + k0 + k] : 0.f;
    }
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int jj = warp; jj < 64; jj += kT / 32) {
        const int j = blockIdx.y * 64 + jj;
        if (j >= O) break;
        const float* wr = w + (int64_t)j * K + k0;
        float acc[kMaxRows];
#pragma unroll
        for (int n = 0; n < kMaxRows; ++n) acc[n] = 0.f;
#pragma unroll 4
        for (int k = lane; k < kc; k += 32) {
            const float wv = wr[k];
#pragma unroll
            for (int n = 0; n < kMaxRows; ++n)
                if (n < N) acc[n] = fmaf(xs[n * kLinCK + k], wv, acc[n]);
        }
#pragma unroll
        for (int n = 0; n < kMaxRows; ++n) {
            if (n >= N) break;
            float v = acc[n];
            for (int s = 16; s > 0; s >>= 1) v += __shfl_xor_sync(0xffffffffu, v, s);
            if (lane == 0) part[((int64_t)blockIdx.x * N + n) * O + j] = v;
        }
    }
}
__global__ void __launch_bounds__(kT)
linear_fwd_reduce_kernel(const float* __restrict__ part, int chunks, const float* __restrict__ b, int N, int O, int relu,
                         const uint8_t* __restrict__ mask, float scale, float* __restrict__ y, int ldy, int col_off) {
    for (int i = blockIdx.x * kT + threadIdx.x; i < N * O; i += gridDim.x * kT) {
        const int n = i / O, j = i - n * O;
        float v = b[j];
        for (int c = 0; c < chunks; ++c) v += part[(int64_t)c * N * O + i];
        if (relu) v = fmaxf(v, 0.f);
        if (mask) v = mask[i] ? v * scale : 0.f;
        y[(int64_t)n * ldy + col_off + j] = v;
    }
}

// dx[n][k]: one thread per input feature k, all rows at once
// (gridDim.y > 1: output features are cut into gridDim.y ranges whose partial sums go to dx[range][n][k]; the caller
//  reduces them in range order)
__global__ void __launch_bounds__(kT)
linear_bwd_x_rows_kernel(const float* __restrict__ dy, int ldy, int col_off, const float* __restrict__ w, int N, int K, int O,
                         float* __restrict__ dx) {
    extern __shared__ float sd[];                    // dy[n][o] for the whole sub-batch
    for (int i = threadIdx.x; i < N * O; i += kT) sd[i] = dy[(int64_t)(i / O) * ldy + col_off + i % O];
    __syncthreads();
    const int per = (O + gridDim.y - 1) / gridDim.y, o_lo = blockIdx.y * per, o_hi = min(O, o_lo + per);
    dx += (int64_t)blockIdx.y * N * K;
    for (int k = blockIdx.x * kT + threadIdx.x; k < K; k += gridDim.x * kT) {
        float acc[kMaxRows];
#pragma unroll
        for (int n = 0; n < kMaxRows; ++n) acc[n] = 0.f;
#pragma unroll 8
        for (int o = o_lo; o < o_hi; ++o) {
            const float wv = w[(int64_t)o * K + k];
#pragma unroll
            for (int n = 0; n < kMaxRows; ++n)
                if (n < N) acc[n] = fmaf(sd[n * O + o], wv, acc[n]);
        }
#pragma unroll
        for (int n = 0; n < kMaxRows; ++n)
            if (n < N) dx[(int64_t)n * K + k] = acc[n];
    }
}
// out[i] = sum over r of part[r][i] (range order)
__global__ void __launch_bounds__(kT)
sum_ranges_kernel(const float* __restrict__ part, int ranges, int64_t total, float* __restrict__ out) {
    for (int64_t i = blockIdx.x * (int64_t)kT + threadIdx.x; i < total; i += (int64_t)gridDim.x * kT) {
        float v = 0.f;
        for (int r = 0; r < ranges; ++r) v += part[(int64_t)r * total + i];
        out[i] = v;
    }
}

// dw[o][k..k+3], db[o]: one thread per (o, 4 consecutive k), K % 4 == 0
__global__ void __launch_bounds__(kT)
linear_bwd_w_vec4_kernel(const float* __restrict__ dy, int ldy, int col_off, const float* __restrict__ x, int ldx, int N, int K,
                         int O, float* __restrict__ dw, float* __restrict__ db) {
    const int K4 = K >> 2;
    const int64_t total = (int64_t)O * K4;
    for (int64_t i = blockIdx.x * (int64_t)kT + threadIdx.x; i < total; i += (int64_t)gridDim.x * kT) {
        const int o = (int)(i / K4), k4 = (int)(i - (int64_t)o * K4);
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        float accb = 0.f;
        for (int n = 0; n < N; ++n) {
            const float d = dy[(int64_t)n * ldy + col_off + o];
            const float4 xv = *reinterpret_cast<const float4*>(x + (int64_t)n * ldx + 4 * k4);
            acc.x = fmaf(d, xv.x, acc.x); acc.y = fmaf(d, xv.y, acc.y); acc.z = fmaf(d, xv.z, acc.z); acc.w = fmaf(d, xv.w, acc.w);
            accb += d;
        }
        *reinterpret_cast<float4*>(dw + (int64_t)o * K + 4 * k4) = acc;
        if (k4 == 0) db[o] = accb;
    }
}

// d <- d * [act > 0] * (mask ? mask * scale : 1): backward of Dropout(ReLU(.)) given the layer's OUTPUT act
__global__ void __launch_bounds__(kT)
relu_drop_bwd_kernel(float* __restrict__ d, const float* __restrict__ act, const uint8_t* __restrict__ mask, float scale,
                     int64_t total) {
    for (int64_t i = blockIdx.x * (int64_t)kT + threadIdx.x; i < total; i += (int64_t)gridDim.x * kT) {
        float v = d[i];
        if (mask) v = mask[i] ? v * scale : 0.f;       // act already carries the mask; a dropped unit has act == 0
        d[i] = act[i] > 0.f ? v : 0.f;
    }
}

// BatchNorm2d backward (training statistics), one CTA per channel:
//   dgamma = sum dy*xhat, dbeta = sum dy, dx = gamma*invstd * (dy - dbeta/m - xhat*dgamma/m)
__global__ void __launch_bounds__(kT)
bn_bwd_kernel(const float* __restrict__ dy, const float* __restrict__ xhat, const float* __restrict__ invstd,
              const float* __restrict__ gamma, int N, int C, int HW, float* __restrict__ dx, float* __restrict__ dgamma,
              float* __restrict__ dbeta) {
    __shared__ double s1[kT], s2[kT];
    const int c = blockIdx.x, m = N * HW;
    double a = 0.0, q = 0.0;
    for (int i = threadIdx.x; i < m; i += kT) {
        const int64_t idx = ((int64_t)(i / HW) * C + c) * HW + i % HW;
        a += dy[idx];
        q += (double)dy[idx] * xhat[idx];
    }
    s1[threadIdx.x] = a; s2[threadIdx.x] = q;
    __syncthreads();
    for (int o = kT / 2; o > 0; o >>= 1) {
        if (threadIdx.x < o) { s1[threadIdx.x] += s1[threadIdx.x + o]; s2[threadIdx.x] += s2[threadIdx.x + o]; }
        __syncthreads();
    }
    const float sdy = (float)s1[0], sdx = (float)s2[0];
    if (threadIdx.x == 0) { dbeta[c] = sdy; dgamma[c] = sdx; }
    const float k = gamma[c] * invstd[c], im = 1.f / (float)m;
    for (int i = threadIdx.x; i < m; i += kT) {
        const int64_t idx = ((int64_t)(i / HW) * C + c) * HW + i % HW;
        dx[idx] = k * (dy[idx] - sdy * im - xhat[idx] * sdx * im);
    }
}

// backward of MaxPool(3,1) o ReLU: dy[nc][oy][ox] = [y > 0] * sum of dp over the windows whose argmax is (oy, ox)
__global__ void __launch_bounds__(kT)
pool_relu_bwd_kernel(const float* __restrict__ dp, const uint8_t* __restrict__ arg, const float* __restrict__ y, int NC,
                     int OH, int OW, float* __restrict__ dy) {
    const int PH = OH - 2, PW = OW - 2;
    const int64_t total = (int64_t)NC * OH * OW;
    for (int64_t i = blockIdx.x * (int64_t)kT + threadIdx.x; i < total; i += (int64_t)gridDim.x * kT) {
        const int ox = (int)(i % OW), oy = (int)((i / OW) % OH);
        const int64_t nc = i / ((int64_t)OW * OH);
        float acc = 0.f;
        if (y[i] > 0.f) {
            for (int dy_ = 0; dy_ < 3; ++dy_) {
                const int py = oy - dy_;
                if (py < 0 || py >= PH) continue;
                for (int dx_ = 0; dx_ < 3; ++dx_) {
                    const int px = ox - dx_;
                    if (px < 0 || px >= PW) continue;
                    const int64_t pi = (nc * PH + py) * PW + px;
                    if (arg[pi] == dy_ * 3 + dx_) acc += dp[pi];
                }
            }
        }
        dy[i] = acc;
    }
}

// dx[n][ci][iy][ix] = sum_{co,kh,kw} dy[n][co][oy][ox] * w[co][ci][kh][kw] with oy*sh - ph + kh == iy (same for x)
__global__ void __launch_bounds__(kT)
conv_bwd_x_kernel(const float* __restrict__ dy, const float* __restrict__ w, int N, int Ci, int H, int W, int Co, int KH,
                  int KW, int sh, int sw, int ph, int pw, int OH, int OW, float* __restrict__ dx) {
    const int64_t total = (int64_t)N * Ci * H * W;
    for (int64_t i = blockIdx.x * (int64_t)kT + threadIdx.x; i < total; i += (int64_t)gridDim.x * kT) {
        const int ix = (int)(i % W), iy = (int)((i / W) % H), ci = (int)((i / ((int64_t)W * H)) % Ci);
        const int n = (int)(i / ((int64_t)W * H * Ci));
        float acc = 0.f;
        for (int kh = 0; kh < KH; ++kh) {
            const int ty = iy + ph - kh;
            if (ty < 0 || ty % sh) continue;
            const int oy = ty / sh;
            if (oy >= OH) continue;
            for (int kw = 0; kw < KW; ++kw) {
                const int tx = ix + pw - kw;
                if (tx < 0 || tx % sw) continue;
                const int ox = tx / sw;
                if (ox >= OW) continue;
                const float* d = dy + (((int64_t)n * Co) * OH + oy) * OW + ox;
                const float* wp = w + ((int64_t)ci * KH + kh) * KW + kw;
                for (int co = 0; co < Co; ++co)
                    acc = fmaf(d[(int64_t)co * OH * OW], wp[(int64_t)co * Ci * KH * KW], acc);
            }
        }
        dx[i] = acc;
    }
}

// dw[co][ci][kh][kw] = sum_{n,oy,ox} dy[n][co][oy][ox] * x[n][ci][oy*sh-ph+kh][ox*sw-pw+kw]; one warp per weight
// (lanes stride over the n*OH*OW positions); db[co] = sum dy, from the (ci,kh,kw) = 0 warp
__global__ void __launch_bounds__(kT)
conv_bwd_w_kernel(const float* __restrict__ dy, const float* __restrict__ x, int N, int Ci, int H, int W, int Co, int KH,
                  int KW, int sh, int sw, int ph, int pw, int OH, int OW, float* __restrict__ dw, float* __restrict__ db) {
    const int lane = threadIdx.x & 31;
    const int64_t gw = (blockIdx.x * (int64_t)kT + threadIdx.x) >> 5, nw = ((int64_t)gridDim.x * kT) >> 5;
    const int64_t total = (int64_t)Co * Ci * KH * KW;
    const int npos = N * OH * OW;
    for (int64_t i = gw; i < total; i += nw) {
        const int kw = (int)(i % KW), kh = (int)((i / KW) % KH), ci = (int)((i / (KW * KH)) % Ci);
        const int co = (int)(i / ((int64_t)KW * KH * Ci));
        float acc = 0.f, accb = 0.f;
        for (int p = lane; p < npos; p += 32) {
            const int ox = p % OW, oy = (p / OW) % OH, n = p / (OW * OH);
            const float d = dy[(((int64_t)n * Co + co) * OH + oy) * OW + ox];
            accb += d;
            const int iy = oy * sh - ph + kh, ix = ox * sw - pw + kw;
            if (iy >= 0 && iy < H && ix >= 0 && ix < W) acc = fmaf(d, x[(((int64_t)n * Ci + ci) * H + iy) * W + ix], acc);
        }
        for (int s = 16; s > 0; s >>= 1) {
            acc += __shfl_xor_sync(0xffffffffu, acc, s);
            accb += __shfl_xor_sync(0xffffffffu, accb, s);
        }
        if (lane == 0) {
            dw[i] = acc;
            if (ci == 0 && kh == 0 && kw == 0) db[co] = accb;
        }
    }
}

// torch.optim.Adam (no weight decay, no amsgrad): p -= lr * mhat / (sqrt(vhat) + eps)
__global__ void __launch_bounds__(kT)
adam_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m, float* __restrict__ v, int64_t total,
            float grad_scale, float lr, float b1, float b2, float eps, const gn_step_state* __restrict__ state) {
    const float bc1 = state->bc1, bc2 = state->bc2;
    for (int64_t i = blockIdx.x * (int64_t)kT + threadIdx.x; i < total; i += (int64_t)gridDim.x * kT) {
        const float gi = g[i] * grad_scale;
        const float mi = b1 * m[i] + (1.f - b1) * gi;
        const float vi = b2 * v[i] + (1.f - b2) * gi * gi;
        m[i] = mi;
        v[i] = vi;
        p[i] -= (lr / bc1) * mi / (sqrtf(vi) / sqrtf(bc2) + eps);
    }
}

}  // namespace

namespace {

// ---- convolutions as fp32 GEMMs ---------------------------------------------------------------------------
// One strided SGEMM serves conv forward, weight gradient and input gradient (through an explicit fp32 im2col buffer;
// a 10-frame sub-batch makes it 11 MB at conv3):  C(i, j) = sum_l A(i, l) * B(j, l).
// An operand element (r, l) lives at base[(r / rdiv) * rs_hi + (r % rdiv) * rs_lo + (l / ldiv) * ls_hi + (l % ldiv) * ls_lo],
// which covers row-major matrices, their transposes and NCHW activations read as (pixel, channel) matrices.
struct GOp {
    const float* base;
    int rdiv; int64_t rs_hi, rs_lo;
    int ldiv; int64_t ls_hi, ls_lo;
    int l_fast;            // 1: consecutive l are contiguous in memory (tile loads run along l), 0: consecutive r are
};
__device__ __forceinline__ int64_t gop_off(const GOp& o, int r, int l) {
    return (int64_t)(r / o.rdiv) * o.rs_hi + (int64_t)(r % o.rdiv) * o.rs_lo + (int64_t)(l / o.ldiv) * o.ls_hi +
           (int64_t)(l % o.ldiv) * o.ls_lo;
}
constexpr int kGM = 64, kGN = 64, kGK = 16;

__global__ void __launch_bounds__(256)
sgemm_strided_kernel(GOp A, GOp B, int M, int N, int K, float* __restrict__ C, int idiv, int64_t is_hi, int64_t is_lo,
                     int jdiv, int64_t js_hi, int64_t js_lo, const float* __restrict__ bias_j, int relu) {
    __shared__ __align__(16) float As[kGK][kGM + 4], Bs[kGK][kGN + 4];
    const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
    const int i0 = blockIdx.y * kGM, j0 = blockIdx.x * kGN;
    float acc[4][4] = {};
    for (int l0 = 0; l0 < K; l0 += kGK) {
        for (int e = tid; e < kGM * kGK; e += 256) {
            int r, l;
            if (A.l_fast) { r = e / kGK; l = e % kGK; } else { l = e / kGM; r = e % kGM; }
            As[l][r] = (i0 + r < M && l0 + l < K) ? A.base[gop_off(A, i0 + r, l0 + l)] : 0.f;
            if (B.l_fast) { r = e / kGK; l = e % kGK; } else { l = e / kGN; r = e % kGN; }
            Bs[l][r] = (j0 + r < N && l0 + l < K) ? B.base[gop_off(B, j0 + r, l0 + l)] : 0.f;
        }
        __syncthreads();
#pragma unroll
        for (int l = 0; l < kGK; ++l) {
            const float4 a = *reinterpret_cast<const float4*>(&As[l][ty * 4]);
            const float4 b = *reinterpret_cast<const float4*>(&Bs[l][tx * 4]);
            const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
            for (int u = 0; u < 4; ++u)
#pragma unroll
                for (int v = 0; v < 4; ++v) acc[u][v] = fmaf(av[u], bv[v], acc[u][v]);
        }
        __syncthreads();
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
        const int i = i0 + ty * 4 + u;
        if (i >= M) continue;
#pragma unroll
        for (int v = 0; v < 4; ++v) {
            const int j = j0 + tx * 4 + v;
            if (j >= N) continue;
            float val = acc[u][v] + (bias_j ? bias_j[j] : 0.f);
            if (relu) val = fmaxf(val, 0.f);
            C[(int64_t)(i / idiv) * is_hi + (int64_t)(i % idiv) * is_lo + (int64_t)(j / jdiv) * js_hi + (int64_t)(j % jdiv) * js_lo] = val;
        }
    }
}

// col[p][ci*KH*KW + kh*KW + kw] = x[n][ci][oy*s-pad+kh][ox*s-pad+kw] (0 outside), p = (n, oy, ox)
__global__ void __launch_bounds__(kT)
im2col_kernel(const float* __restrict__ x, int N, int Ci, int H, int W, int KH, int KW, int s, int pad, int OH, int OW,
              float* __restrict__ col) {
    const int KK = Ci * KH * KW;
    const int64_t total = (int64_t)N * OH * OW * KK;
    for (int64_t i = blockIdx.x * (int64_t)kT + threadIdx.x; i < total; i += (int64_t)gridDim.x * kT) {
        const int k = (int)(i % KK);
        const int64_t p = i / KK;
        const int kw = k % KW, kh = (k / KW) % KH, ci = k / (KW * KH);
        const int ox = (int)(p % OW), oy = (int)((p / OW) % OH), n = (int)(p / ((int64_t)OW * OH));
        const int iy = oy * s - pad + kh, ix = ox * s - pad + kw;
        col[i] = (iy >= 0 && iy < H && ix >= 0 && ix < W) ? x[(((int64_t)n * Ci + ci) * H + iy) * W + ix] : 0.f;
    }
}

// dx[n][ci][iy][ix] = sum over (kh, kw) of dcol[p(oy, ox)][ci*KH*KW + kh*KW + kw] with oy*s - pad + kh == iy (same in x)
__global__ void __launch_bounds__(kT)
col2im_kernel(const float* __restrict__ dcol, int N, int Ci, int H, int W, int KH, int KW, int s, int pad, int OH, int OW,
              float* __restrict__ dx) {
    const int KK = Ci * KH * KW;
    const int64_t total = (int64_t)N * Ci * H * W;
    for (int64_t i = blockIdx.x * (int64_t)kT + threadIdx.x; i < total; i += (int64_t)gridDim.x * kT) {
        const int ix = (int)(i % W), iy = (int)((i / W) % H), ci = (int)((i / ((int64_t)W * H)) % Ci);
        const int n = (int)(i / ((int64_t)W * H * Ci));
        float acc = 0.f;
        for (int kh = 0; kh < KH; ++kh) {
            const int ty = iy + pad - kh;
            if (ty < 0 || ty % s || ty / s >= OH) continue;
            for (int kw = 0; kw < KW; ++kw) {
                const int tx = ix + pad - kw;
                if (tx < 0 || tx % s || tx / s >= OW) continue;
                const int64_t p = ((int64_t)n * OH + ty / s) * OW + tx / s;
                acc += dcol[p * KK + (ci * KH + kh) * KW + kw];
            }
        }
        dx[i] = acc;
    }
}

// Error-compensated TF32 ("3xTF32"): v = hi + lo with hi = v rounded to TF32 (13 low mantissa bits cleared) and lo the
// exact remainder; a.b ~= a_hi.b_hi + a_hi.b_lo + a_lo.b_hi.  The three products become ONE GEMM by tripling K:
// an A operand row is [hi | hi | lo], a B operand row is [hi | lo | hi] (segments of `seg` elements).
__device__ __forceinline__ void store3(float* row, int seg, int k, float v, bool is_a) {
    const float hi = __uint_as_float((__float_as_uint(v) + 0x1000u) & 0xFFFFE000u);
    const float lo = v - hi;
    row[k] = hi;
    row[seg + k] = is_a ? hi : lo;
    row[2 * seg + k] = is_a ? lo : hi;
}

// ---- layout shuffles around the TF32 tensor-core GEMMs (gn_tc_gemm_tf32: C[M][N] = A[M][K] . B[N][K]^T, K-major) ----
// colT[k][ld] : transposed im2col, columns p >= P zero (K padding of the weight-gradient GEMM)
__global__ void __launch_bounds__(kT)
im2col_t_kernel(const float* __restrict__ x, int N, int Ci, int H, int W, int s, int pad, int OH, int OW, int ld,
                float* __restrict__ colT) {
    const int KK = Ci * 9, P = N * OH * OW;
    const int64_t total = (int64_t)KK * ld;
    for (int64_t i = blockIdx.x * (int64_t)kT + threadIdx.x; i < total; i += (int64_t)gridDim.x * kT) {
        const int p = (int)(i % ld), k = (int)(i / ld);
        float v = 0.f;
        if (p < P) {
            const int kw = k % 3, kh = (k / 3) % 3, ci = k / 9;
            const int ox = p % OW, oy = (p / OW) % OH, n = p / (OW * OH);
            const int iy = oy * s - pad + kh, ix = ox * s - pad + kw;
            if (iy >= 0 && iy < H && ix >= 0 && ix < W) v = x[(((int64_t)n * Ci + ci) * H + iy) * W + ix];
        }
        store3(colT + (int64_t)k * 3 * ld, ld, p, v, true);
    }
}
// NCHW [n][C][HW] -> [C][ld] (p = n*HW + pos; columns >= P zero)   or   -> [rows][C] (rows >= P zero)
__global__ void __launch_bounds__(kT)
nchw_to_cp_kernel(const float* __restrict__ t, int N, int C, int HW, int ld, float* __restrict__ out) {
    const int P = N * HW;
    const int64_t total = (int64_t)C * ld;
    for (int64_t i = blockIdx.x * (int64_t)kT + threadIdx.x; i < total; i += (int64_t)gridDim.x * kT) {
        const int p = (int)(i % ld), c = (int)(i / ld);
        store3(out + (int64_t)c * 3 * ld, ld, p, p < P ? t[((int64_t)(p / HW) * C + c) * HW + p % HW] : 0.f, false);
    }
}
__global__ void __launch_bounds__(kT)
nchw_to_pc_kernel(const float* __restrict__ t, int N, int C, int HW, int rows, float* __restrict__ out) {
    const int P = N * HW;
    const int64_t total = (int64_t)rows * C;
    for (int64_t i = blockIdx.x * (int64_t)kT + threadIdx.x; i < total; i += (int64_t)gridDim.x * kT) {
        const int c = (int)(i % C), p = (int)(i / C);
        store3(out + (int64_t)p * 3 * C, C, c, p < P ? t[((int64_t)(p / HW) * C + c) * HW + p % HW] : 0.f, false);
    }
}
// The tensor-core GEMMs run K-split (few tiles, long K): their consumers add the `splits` partial slabs, `slab` floats apart.
__device__ __forceinline__ float sum_slabs(const float* p, int splits, int64_t slab) {
    float v = p[0];
    for (int s = 1; s < splits; ++s) v += p[(int64_t)s * slab];
    return v;
}
// y[n][c][pos] = relu(yT[p][c] + b[c])
__global__ void __launch_bounds__(kT)
pc_bias_relu_to_nchw_kernel(const float* __restrict__ yT, int splits, int64_t slab, const float* __restrict__ b, int N, int C,
                            int HW, float* __restrict__ y) {
    const int64_t total = (int64_t)N * C * HW;
    for (int64_t i = blockIdx.x * (int64_t)kT + threadIdx.x; i < total; i += (int64_t)gridDim.x * kT) {
        const int pos = (int)(i % HW), c = (int)((i / HW) % C), n = (int)(i / ((int64_t)HW * C));
        y[i] = fmaxf(sum_slabs(yT + ((int64_t)n * HW + pos) * C + c, splits, slab) + b[c], 0.f);
    }
}
// out[c][r] = in[r][c]  (split: rows of out are 3xTF32 A operands [hi | hi | lo])
__global__ void __launch_bounds__(kT)
transpose_kernel(const float* __restrict__ in, int R, int Cc, float* __restrict__ out, int split, int splits = 1,
                 int64_t slab = 0) {
    const int64_t total = (int64_t)R * Cc;
    for (int64_t i = blockIdx.x * (int64_t)kT + threadIdx.x; i < total; i += (int64_t)gridDim.x * kT) {
        const int r = (int)(i % R), c = (int)(i / R);
        const float v = sum_slabs(in + (int64_t)r * Cc + c, splits, slab);
        if (split) store3(out + (int64_t)c * 3 * R, R, r, v, true);
        else out[i] = v;
    }
}
// 3xTF32 operand copy of a row-major matrix [R][K] -> [R][3K]
__global__ void __launch_bounds__(kT)
split3_kernel(const float* __restrict__ in, int R, int K, float* __restrict__ out, int is_a) {
    const int64_t total = (int64_t)R * K;
    for (int64_t i = blockIdx.x * (int64_t)kT + threadIdx.x; i < total; i += (int64_t)gridDim.x * kT) {
        const int k = (int)(i % K), r = (int)(i / K);
        store3(out + (int64_t)r * 3 * K, K, k, in[i], is_a != 0);
    }
}
// col[p][3*KK]: im2col as a 3xTF32 A operand
__global__ void __launch_bounds__(kT)
im2col_split_kernel(const float* __restrict__ x, int N, int Ci, int H, int W, int s, int pad, int OH, int OW,
                    float* __restrict__ col) {
    const int KK = Ci * 9;
    const int64_t total = (int64_t)N * OH * OW * KK;
    for (int64_t i = blockIdx.x * (int64_t)kT + threadIdx.x; i < total; i += (int64_t)gridDim.x * kT) {
        const int k = (int)(i % KK);
        const int64_t p = i / KK;
        const int kw = k % 3, kh = (k / 3) % 3, ci = k / 9;
        const int ox = (int)(p % OW), oy = (int)((p / OW) % OH), n = (int)(p / ((int64_t)OW * OH));
        const int iy = oy * s - pad + kh, ix = ox * s - pad + kw;
        const float v = (iy >= 0 && iy < H && ix >= 0 && ix < W) ? x[(((int64_t)n * Ci + ci) * H + iy) * W + ix] : 0.f;
        store3(col + p * 3 * KK, KK, k, v, true);
    }
}
// dx from the transposed column gradient dcolT[k][ld]
__global__ void __launch_bounds__(kT)
col2im_t_kernel(const float* __restrict__ dcolT, int splits, int64_t slab, int N, int Ci, int H, int W, int s, int pad, int OH,
                int OW, int ld, float* __restrict__ dx) {
    const int64_t total = (int64_t)N * Ci * H * W;
    for (int64_t i = blockIdx.x * (int64_t)kT + threadIdx.x; i < total; i += (int64_t)gridDim.x * kT) {
        const int ix = (int)(i % W), iy = (int)((i / W) % H), ci = (int)((i / ((int64_t)W * H)) % Ci);
        const int n = (int)(i / ((int64_t)W * H * Ci));
        float acc = 0.f;
        for (int kh = 0; kh < 3; ++kh) {
            const int ty = iy + pad - kh;
            if (ty < 0 || ty % s || ty / s >= OH) continue;
            for (int kw = 0; kw < 3; ++kw) {
                const int tx = ix + pad - kw;
                if (tx < 0 || tx % s || tx / s >= OW) continue;
                acc += sum_slabs(dcolT + (int64_t)((ci * 3 + kh) * 3 + kw) * ld + ((int64_t)n * OH + ty / s) * OW + tx / s, splits, slab);
            }
        }
        dx[i] = acc;
    }
}

// db[c] = sum over (n, pixel) of dy[n][c][pixel]; one CTA per channel
__global__ void __launch_bounds__(kT)
bias_grad_kernel(const float* __restrict__ dy, int N, int C, int HW, float* __restrict__ db) {
    __shared__ float sred[kT];
    const int c = blockIdx.x;
    float a = 0.f;
    for (int i = threadIdx.x; i < N * HW; i += kT) a += dy[((int64_t)(i / HW) * C + c) * HW + i % HW];
    sred[threadIdx.x] = a;
    __syncthreads();
    for (int o = kT / 2; o > 0; o >>= 1) {
        if (threadIdx.x < o) sred[threadIdx.x] += sred[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) db[c] = sred[0];
}

}  // namespace

// ================================================================================================================
// trainer object
// ================================================================================================================
enum { P_C1W = 0, P_C1B, P_BN1W, P_BN1B, P_C2W, P_C2B, P_BN2W, P_BN2B, P_C3W, P_C3B, P_BN3W, P_BN3B, P_L5W, P_L5B,
       P_AC1W, P_AC1B, P_AC2W, P_AC2B, P_AL3W, P_AL3B, P_F0W, P_F0B, P_F3W, P_F3B, P_F6W, P_F6B, P_F9W, P_F9B, P_F12W,
       P_F12B, P_COUNT };

struct gn_avm_trainer {
    gn_handle* h;
    int audio, S, B, C1, P1, P2, P3, L1, L2, fin;
    int64_t max_n, K5;
    int64_t off[P_COUNT + 1];        // offsets of each tensor in the flat parameter vector (model.parameters() order)
    int64_t n_params, step;
    float lr, b1, b2, eps;
    float *params, *grads, *m, *v;   // flat fp32 [n_params]
    float* rmean[3];                 // BatchNorm running buffers
    float* rvar[3];
    // activations of the last forward (fp32, NCHW) and their gradients
    float *y[3], *p[3], *xh[3], *a[3], *istd[3];     // conv+ReLU, pooled, normalised, BN output, 1/std per block
    uint8_t* arg[3];
    float *h5, *ay1, *ay2, *fin_buf, *f[4], *z, *dz;
    float *d_big0, *d_big1;          // ping-pong gradient buffers sized for the largest activation
    int64_t col_elems;               // capacity of col / dcol in floats
    float *col, *dcol, *opb;         // im2col of a conv input / gradient w.r.t. it (largest layer), transposed B operand
    float* lin_part;                 // partial sums of the K-chunked linear5 forward
    int64_t lin_part_elems;
    int use_tc;                      // conv2 / conv3 GEMMs on the tensor cores as TF32 (default) or the fp32 SGEMM
    float *d_fin, *d_f[4];
    uint8_t* masks[5];               // keep-masks: linear5 (512), fusion.0/3/6/9 (512, 512, 256, 128)
    gn_step_state* state;            // device: Adam step count, bias corrections, dropout seed bump (CUDA-graph safe)
    void* arena;
};

namespace {
const float* src_of(const gn_avm_state* sd, int k) {
    switch (k) {
        case P_C1W: return sd->conv1_w; case P_C1B: return sd->conv1_b; case P_BN1W: return sd->bn1_w; case P_BN1B: return sd->bn1_b;
        case P_C2W: return sd->conv2_w; case P_C2B: return sd->conv2_b; case P_BN2W: return sd->bn2_w; case P_BN2B: return sd->bn2_b;
        case P_C3W: return sd->conv3_w; case P_C3B: return sd->conv3_b; case P_BN3W: return sd->bn3_w; case P_BN3B: return sd->bn3_b;
        case P_L5W: return sd->linear5_w; case P_L5B: return sd->linear5_b;
        case P_AC1W: return sd->aconv1_w; case P_AC1B: return sd->aconv1_b; case P_AC2W: return sd->aconv2_w;
        case P_AC2B: return sd->aconv2_b; case P_AL3W: return sd->alinear3_w; case P_AL3B: return sd->alinear3_b;
        case P_F0W: return sd->f0_w; case P_F0B: return sd->f0_b; case P_F3W: return sd->f3_w; case P_F3B: return sd->f3_b;
        case P_F6W: return sd->f6_w; case P_F6B: return sd->f6_b; case P_F9W: return sd->f9_w; case P_F9B: return sd->f9_b;
        case P_F12W: return sd->f12_w; case P_F12B: return sd->f12_b;
    }
    return nullptr;
}
inline int cout_dim(int n, int k, int s, int p) { return (n + 2 * p - k) / s + 1; }
}  // namespace

extern "C" int gn_avm_trainer_create(gn_handle* h, gn_avm_trainer** out, const gn_avm_state* sd, int audio_included, int S,
                                     int B, int64_t max_n, void* stream) {
    if (!h || !out || !sd || max_n < 1 || S < 16 || S > 1024 || B < 4) return GN_E_ARG;
    *out = nullptr;
    gn_avm_trainer* t = reinterpret_cast<gn_avm_trainer*>(calloc(1, sizeof(gn_avm_trainer)));
    if (!t) return GN_E_ARG;
    t->h = h; t->audio = audio_included ? 1 : 0; t->S = S; t->B = B; t->max_n = max_n;
    t->C1 = cout_dim(S, 3, 3, 3); t->P1 = t->C1 - 2; t->P2 = t->P1 - 2; t->P3 = t->P2 - 2;
    t->L1 = cout_dim(B, 3, 2, 1); t->L2 = cout_dim(t->L1, 3, 2, 1);
    t->fin = 512 + (t->audio ? 128 : 0);
    t->K5 = 512LL * t->P3 * t->P3;
    t->lr = 1e-3f; t->b1 = 0.9f; t->b2 = 0.999f; t->eps = 1e-8f;
    t->use_tc = getenv("GN_TRAIN_FP32") ? 0 : 1;     // GN_TRAIN_FP32=1: exact-fp32 SGEMM convolutions (validation)
    if (t->P3 < 1) { free(t); return GN_E_SHAPE; }
    const int64_t sz[P_COUNT] = {64 * 27, 64, 64, 64, 256 * 576, 256, 256, 256, 512 * 2304, 512, 512, 512, 512 * t->K5, 512,
                                 t->audio ? 64 * 90 : 0, t->audio ? 64 : 0, t->audio ? 128 * 192 : 0, t->audio ? 128 : 0,
                                 t->audio ? 128LL * 128 * t->L2 : 0, t->audio ? 128 : 0,
                                 512LL * t->fin, 512, 512 * 512, 512, 256 * 512, 256, 128 * 256, 128, 128, 1};
    t->off[0] = 0;
    for (int k = 0; k < P_COUNT; ++k) {
        if (sz[k] > 0 && !src_of(sd, k)) { free(t); return GN_E_ARG; }
        t->off[k + 1] = t->off[k] + ((sz[k] + 3) & ~3LL);      // 16-byte aligned tensors
    }
    t->n_params = t->off[P_COUNT];
    const float* bnm[3] = {sd->bn1_m, sd->bn2_m, sd->bn3_m};
    const float* bnv[3] = {sd->bn1_v, sd->bn2_v, sd->bn3_v};
    for (int i = 0; i < 3; ++i) if (!bnm[i] || !bnv[i]) { free(t); return GN_E_ARG; }

    const int C[3] = {64, 256, 512}, OH[3] = {t->C1, t->P1, t->P2}, PH[3] = {t->P1, t->P2, t->P3};
    const int64_t n = max_n;
    int64_t big = 0;
    size_t bytes = 0;
    for (int pass = 0; pass < 2; ++pass) {
        char* base = reinterpret_cast<char*>(t->arena);
        size_t off = 0;
        auto take = [&](int64_t elems, size_t es) -> void* {
            off = (off + 255) & ~size_t(255);
            void* p = pass ? base + off : nullptr;
            off += (size_t)elems * es;
            return p;
        };
        t->params = (float*)take(t->n_params, 4); t->grads = (float*)take(t->n_params, 4);
        t->m = (float*)take(t->n_params, 4); t->v = (float*)take(t->n_params, 4);
        for (int i = 0; i < 3; ++i) {
            t->rmean[i] = (float*)take(C[i], 4); t->rvar[i] = (float*)take(C[i], 4); t->istd[i] = (float*)take(C[i], 4);
            const int64_t ye = n * C[i] * OH[i] * OH[i], pe = n * C[i] * PH[i] * PH[i];
            t->y[i] = (float*)take(ye, 4); t->p[i] = (float*)take(pe, 4); t->xh[i] = (float*)take(pe, 4);
            t->a[i] = (float*)take(pe, 4); t->arg[i] = (uint8_t*)take(pe, 1);
            if (ye > big) big = ye;
        }
        t->h5 = (float*)take(n * 512, 4);
        t->ay1 = (float*)take(n * 64 * t->L1, 4); t->ay2 = (float*)take(n * 128 * t->L2, 4);
        t->fin_buf = (float*)take(n * t->fin, 4);
        const int fw[4] = {512, 512, 256, 128};
        for (int i = 0; i < 4; ++i) { t->f[i] = (float*)take(n * fw[i], 4); t->d_f[i] = (float*)take(n * fw[i], 4); }
        t->z = (float*)take(n, 4); t->dz = (float*)take(n, 4);
        t->d_fin = (float*)take(n * t->fin, 4);
        if (n * 128 * t->L2 > big) big = n * 128 * t->L2;
        t->d_big0 = (float*)take(big, 4); t->d_big1 = (float*)take(big, 4);
        int64_t colmax = 0;
        for (int i = 0; i < 3; ++i) {
            const int64_t ce = n * OH[i] * OH[i] * (i == 0 ? 27 : (i == 1 ? 576 : 2304));
            if (ce > colmax) colmax = ce;
        }
        int64_t bmax = 0;
        for (int i = 1; i < 3; ++i) {      // tensor-core layers: K-padded / N-padded transposed operands
            const int64_t pp = ((n * OH[i] * OH[i] + 127) / 128) * 128;
            const int64_t kk = (i == 1 ? 576 : 2304);
            const int64_t ce = 3 * pp * kk, be = 3 * (pp * C[i] > C[i] * kk ? pp * C[i] : C[i] * kk);   // 3xTF32 operands
            if (ce > colmax) colmax = ce;
            if (be > bmax) bmax = be;
        }
        t->col_elems = colmax;
        t->col = (float*)take(colmax, 4); t->dcol = (float*)take(colmax, 4); t->opb = (float*)take(bmax, 4);
        t->lin_part_elems = ((t->K5 + kLinCK - 1) / kLinCK) * (n < kMaxRows ? n : kMaxRows) * 512;
        t->lin_part = (float*)take(t->lin_part_elems, 4);
        const int mw[5] = {512, 512, 512, 256, 128};
        for (int i = 0; i < 5; ++i) t->masks[i] = (uint8_t*)take(n * mw[i], 1);
        t->state = (gn_step_state*)take(sizeof(gn_step_state), 1);      // zeroed with the arena: step 0, no seed bump
        if (pass == 0) {
            bytes = off + 256;
            cudaError_t e = cudaMalloc(&t->arena, bytes);
            if (e != cudaSuccess) { free(t); return (int)e; }
        }
    }
    cudaStream_t st = gn_stream(stream);
    cudaError_t e = cudaMemsetAsync(t->arena, 0, bytes, st);
    for (int k = 0; k < P_COUNT && e == cudaSuccess; ++k)
        if (sz[k] > 0) e = cudaMemcpyAsync(t->params + t->off[k], src_of(sd, k), sz[k] * 4, cudaMemcpyDeviceToDevice, st);
    for (int i = 0; i < 3 && e == cudaSuccess; ++i) {
        e = cudaMemcpyAsync(t->rmean[i], bnm[i], C[i] * 4, cudaMemcpyDeviceToDevice, st);
        if (e == cudaSuccess) e = cudaMemcpyAsync(t->rvar[i], bnv[i], C[i] * 4, cudaMemcpyDeviceToDevice, st);
    }
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) { cudaFree(t->arena); free(t); return (int)e; }
    *out = t;
    return GN_OK;
}

extern "C" int gn_avm_trainer_destroy(gn_avm_trainer* t) {
    if (!t) return GN_OK;
    cudaFree(t->arena);
    free(t);
    return GN_OK;
}

extern "C" int64_t gn_avm_trainer_param_count(const gn_avm_trainer* t) { return t ? t->n_params : GN_E_ARG; }

extern "C" int gn_avm_trainer_set_hyper(gn_avm_trainer* t, float lr, float beta1, float beta2, float eps) {
    if (!t || lr < 0.f) return GN_E_ARG;
    t->lr = lr; t->b1 = beta1; t->b2 = beta2; t->eps = eps;
    return GN_OK;
}

// Copy parameters and BatchNorm running buffers into the (writable, device) tensors `dst` points at.
extern "C" int gn_avm_trainer_export(gn_avm_trainer* t, const gn_avm_state* dst, void* stream) {
    if (!t || !dst) return GN_E_ARG;
    cudaStream_t st = gn_stream(stream);
    for (int k = 0; k < P_COUNT; ++k) {
        const int64_t cnt = t->off[k + 1] - t->off[k];
        float* d = const_cast<float*>(src_of(dst, k));
        if (cnt == 0 || !d) continue;
        int64_t real = cnt;
        if (k == P_C1W) real = 64 * 27; else if (k == P_F12B) real = 1;      // the only sizes that are not multiples of 4
        GN_CUDA(cudaMemcpyAsync(d, t->params + t->off[k], real * 4, cudaMemcpyDeviceToDevice, st));
    }
    const int C[3] = {64, 256, 512};
    float* bm[3] = {const_cast<float*>(dst->bn1_m), const_cast<float*>(dst->bn2_m), const_cast<float*>(dst->bn3_m)};
    float* bv[3] = {const_cast<float*>(dst->bn1_v), const_cast<float*>(dst->bn2_v), const_cast<float*>(dst->bn3_v)};
    for (int i = 0; i < 3; ++i) {
        if (bm[i]) GN_CUDA(cudaMemcpyAsync(bm[i], t->rmean[i], C[i] * 4, cudaMemcpyDeviceToDevice, st));
        if (bv[i]) GN_CUDA(cudaMemcpyAsync(bv[i], t->rvar[i], C[i] * 4, cudaMemcpyDeviceToDevice, st));
    }
    return GN_OK;
}

extern "C" int gn_avm_trainer_grads(gn_avm_trainer* t, float** d_grads) {
    if (!t || !d_grads) return GN_E_ARG;
    *d_grads = t->grads;
    return GN_OK;
}

// ================================================================================================================
// forward + backward (train mode) and the optimizer step
// ================================================================================================================
namespace {
constexpr int kBig = 1 << 30;
int launch_sgemm(gn_handle* h, const GOp& A, const GOp& B, int M, int N, int K, float* C, int idiv, int64_t is_hi,
                 int64_t is_lo, int jdiv, int64_t js_hi, int64_t js_lo, const float* bias_j, int relu, cudaStream_t st) {
    dim3 grid((N + kGN - 1) / kGN, (M + kGM - 1) / kGM);
    sgemm_strided_kernel<<<grid, 256, 0, st>>>(A, B, M, N, K, C, idiv, is_hi, is_lo, jdiv, js_hi, js_lo, bias_j, relu);
    GN_LAUNCHED(h);
    return GN_OK;
}
// y = relu(conv(x, w) + b), square images / kernels 3x3
int conv_fwd_gemm(gn_avm_trainer* t, const float* x, const float* w, const float* b, int n, int Ci, int H, int Co, int s,
                  int pad, int OH, float* y, cudaStream_t st) {
    const int KK = Ci * 9, OHW = OH * OH, cap = t->h->num_sms * 8;
    im2col_kernel<<<grid_for((int64_t)n * OHW * KK, cap), kT, 0, st>>>(x, n, Ci, H, H, 3, 3, s, pad, OH, OH, t->col);
    GN_LAUNCHED(t->h);
    const GOp A{t->col, kBig, 0, KK, kBig, 0, 1, 1}, B{w, kBig, 0, KK, kBig, 0, 1, 1};
    return launch_sgemm(t->h, A, B, n * OHW, Co, KK, y, OHW, (int64_t)Co * OHW, 1, kBig, 0, OHW, b, 1, st);
}
// dw[co][k] = sum_p dy(p, co) col(p, k); db[co] = sum_p dy(p, co)
int conv_bwd_w_gemm(gn_avm_trainer* t, const float* dy, const float* x, int n, int Ci, int H, int Co, int s, int pad, int OH,
                    float* dw, float* db, cudaStream_t st) {
    const int KK = Ci * 9, OHW = OH * OH, cap = t->h->num_sms * 8;
    im2col_kernel<<<grid_for((int64_t)n * OHW * KK, cap), kT, 0, st>>>(x, n, Ci, H, H, 3, 3, s, pad, OH, OH, t->col);
    GN_LAUNCHED(t->h);
    const GOp A{dy, kBig, 0, OHW, OHW, (int64_t)Co * OHW, 1, 1}, B{t->col, kBig, 0, 1, kBig, 0, KK, 0};
    int rc = launch_sgemm(t->h, A, B, Co, KK, n * OHW, dw, kBig, 0, KK, kBig, 0, 1, nullptr, 0, st);
    if (rc) return rc;
    bias_grad_kernel<<<Co, kT, 0, st>>>(dy, n, Co, OHW, db);
    GN_LAUNCHED(t->h);
    return GN_OK;
}
// dx = col2im(dy(p, co) . w(co, k))
int conv_bwd_x_gemm(gn_avm_trainer* t, const float* dy, const float* w, int n, int Ci, int H, int Co, int s, int pad, int OH,
                    float* dx, cudaStream_t st) {
    const int KK = Ci * 9, OHW = OH * OH, cap = t->h->num_sms * 8;
    const GOp A{dy, OHW, (int64_t)Co * OHW, 1, kBig, 0, OHW, 0}, B{w, kBig, 0, 1, kBig, 0, KK, 0};
    int rc = launch_sgemm(t->h, A, B, n * OHW, KK, Co, t->dcol, kBig, 0, KK, kBig, 0, 1, nullptr, 0, st);
    if (rc) return rc;
    col2im_kernel<<<grid_for((int64_t)n * Ci * H * H, cap), kT, 0, st>>>(t->dcol, n, Ci, H, H, 3, 3, s, pad, OH, OH, dx);
    GN_LAUNCHED(t->h);
    return GN_OK;
}
// ---- the same three products on the tensor cores (kind::tf32), for layers with Cin*9 % 32 == 0 and Cout % 128 == 0 ----
inline int pad_to(int v, int m) { return (v + m - 1) / m * m; }
// K ranges for a TF32 GEMM of M x N with K elements: fill the SMs, keep >= 16 k-steps per range, fit the output buffer
int tf32_splits(const gn_avm_trainer* t, int64_t M, int N, int K) {
    const int64_t tiles = ((M + 127) / 128) * (N / 128), ksteps = K / 32;
    int64_t s = t->h->num_sms / tiles;
    if (s > ksteps / 16) s = ksteps / 16;
    while (s > 1 && s * M * N > t->col_elems) --s;
    if (s < 1) s = 1;
    const int64_t per = (ksteps + s - 1) / s;          // the same normalisation gn_tc_gemm_tf32 applies: no empty range
    return (int)((ksteps + per - 1) / per);
}
int conv_fwd_tc(gn_avm_trainer* t, const float* x, const float* w, const float* b, int n, int Ci, int H, int Co, int s, int pad,
                int OH, float* y, cudaStream_t st) {
    const int KK = Ci * 9, OHW = OH * OH, P = n * OHW, cap = t->h->num_sms * 8;
    im2col_split_kernel<<<grid_for((int64_t)P * KK, cap), kT, 0, st>>>(x, n, Ci, H, H, s, pad, OH, OH, t->col);
    GN_LAUNCHED(t->h);
    split3_kernel<<<grid_for((int64_t)Co * KK, cap), kT, 0, st>>>(w, Co, KK, t->opb, 0);
    GN_LAUNCHED(t->h);
    const int sp = tf32_splits(t, P, Co, 3 * KK);
    int rc = gn_tc_gemm_tf32(t->h, t->col, P, 3 * KK, 3 * KK, t->opb, Co, 3 * KK, t->dcol, st, sp);    // yT[split][p][co]
    if (rc) return rc;
    pc_bias_relu_to_nchw_kernel<<<grid_for((int64_t)P * Co, cap), kT, 0, st>>>(t->dcol, sp, (int64_t)P * Co, b, n, Co, OHW, y);
    GN_LAUNCHED(t->h);
    return GN_OK;
}
int conv_bwd_w_tc(gn_avm_trainer* t, const float* dy, const float* x, int n, int Ci, int H, int Co, int s, int pad, int OH,
                  float* dw, float* db, cudaStream_t st) {
    const int KK = Ci * 9, OHW = OH * OH, P = n * OHW, ld = pad_to(P, 32), cap = t->h->num_sms * 8;
    im2col_t_kernel<<<grid_for((int64_t)KK * ld, cap), kT, 0, st>>>(x, n, Ci, H, H, s, pad, OH, OH, ld, t->col);
    GN_LAUNCHED(t->h);
    nchw_to_cp_kernel<<<grid_for((int64_t)Co * ld, cap), kT, 0, st>>>(dy, n, Co, OHW, ld, t->opb);
    GN_LAUNCHED(t->h);
    const int sp = tf32_splits(t, KK, Co, 3 * ld);
    int rc = gn_tc_gemm_tf32(t->h, t->col, KK, 3 * ld, 3 * ld, t->opb, Co, 3 * ld, t->dcol, st, sp);   // dW^T[split][k][co]
    if (rc) return rc;
    transpose_kernel<<<grid_for((int64_t)KK * Co, cap), kT, 0, st>>>(t->dcol, KK, Co, dw, 0, sp, (int64_t)KK * Co);
    GN_LAUNCHED(t->h);
    bias_grad_kernel<<<Co, kT, 0, st>>>(dy, n, Co, OHW, db);
    GN_LAUNCHED(t->h);
    return GN_OK;
}
int conv_bwd_x_tc(gn_avm_trainer* t, const float* dy, const float* w, int n, int Ci, int H, int Co, int s, int pad, int OH,
                  float* dx, cudaStream_t st) {
    const int KK = Ci * 9, OHW = OH * OH, P = n * OHW, rows = pad_to(P, 128), cap = t->h->num_sms * 8;
    transpose_kernel<<<grid_for((int64_t)KK * Co, cap), kT, 0, st>>>(w, Co, KK, t->col, 1);             // W^T[k][3*co]
    GN_LAUNCHED(t->h);
    nchw_to_pc_kernel<<<grid_for((int64_t)rows * Co, cap), kT, 0, st>>>(dy, n, Co, OHW, rows, t->opb);
    GN_LAUNCHED(t->h);
    const int sp = tf32_splits(t, KK, rows, 3 * Co);
    int rc = gn_tc_gemm_tf32(t->h, t->col, KK, 3 * Co, 3 * Co, t->opb, rows, 3 * Co, t->dcol, st, sp); // dcol^T[split][k][rows]
    if (rc) return rc;
    col2im_t_kernel<<<grid_for((int64_t)n * Ci * H * H, cap), kT, 0, st>>>(t->dcol, sp, (int64_t)KK * rows, n, Ci, H, H, s, pad, OH,
                                                                           OH, rows, dx);
    GN_LAUNCHED(t->h);
    return GN_OK;
}
}  // namespace

namespace {
// linear layers: sub-batch kernels (every weight read once for all rows) when the batch fits their register arrays
int lin_fwd(gn_avm_trainer* t, cudaStream_t st, const float* x, int ldx, const float* w, const float* b, int N, int K, int O,
            int relu, const uint8_t* mask, float scale, float* y, int ldy, int col_off) {
    const int cap = t->h->num_sms * 8;
    if (N <= kMaxRows && K >= 4096 && t->lin_part) {
        const int chunks = (K + kLinCK - 1) / kLinCK;
        if ((int64_t)chunks * N * O <= t->lin_part_elems) {
            linear_fwd_chunk_kernel<<<dim3(chunks, (O + 63) / 64), kT, (size_t)N * kLinCK * 4, st>>>(x, ldx, w, N, K, O, t->lin_part);
            GN_LAUNCHED(t->h);
            linear_fwd_reduce_kernel<<<grid_for((int64_t)N * O, cap), kT, 0, st>>>(t->lin_part, chunks, b, N, O, relu, mask, scale, y, ldy, col_off);
            GN_LAUNCHED(t->h);
            return GN_OK;
        }
    }
    // (the all-rows kernel pays 16 strided activation loads per weight: only worth it for short reductions)
    if (N <= kMaxRows && K <= 1024) linear_fwd_rows_kernel<<<grid_for((int64_t)O * 32, cap), kT, 0, st>>>(x, ldx, w, b, N, K, O, relu, mask, scale, y, ldy, col_off);
    else linear_fwd_kernel<<<grid_for((int64_t)N * O * 32, cap), kT, 0, st>>>(x, ldx, w, b, N, K, O, relu, mask, scale, y, ldy, col_off);
    GN_LAUNCHED(t->h);
    return GN_OK;
}
int lin_bwd_x(gn_avm_trainer* t, cudaStream_t st, const float* dy, int ldy, int col_off, const float* w, int N, int K, int O,
              float* dx) {
    const int cap = t->h->num_sms * 8;
    if (N <= kMaxRows && (size_t)N * O * 4 <= 48 * 1024) {
        // long rows (linear5): cut the output features into 8 ranges so that the grid fills the GPU; partials in t->col
        // (a 512-wide fusion layer alone would give 2-3 CTAs of 512 dependent steps)
        const int gx = grid_for(K, cap);
        int ranges = 592 / gx;
        if (ranges > O / 16) ranges = O / 16;
        if (ranges < 1 || (int64_t)ranges * N * K > t->col_elems) ranges = 1;
        if (ranges > 1) {
            linear_bwd_x_rows_kernel<<<dim3(grid_for(K, cap), ranges), kT, (size_t)N * O * 4, st>>>(dy, ldy, col_off, w, N, K, O, t->col);
            GN_LAUNCHED(t->h);
            sum_ranges_kernel<<<grid_for((int64_t)N * K, cap), kT, 0, st>>>(t->col, ranges, (int64_t)N * K, dx);
        } else {
            linear_bwd_x_rows_kernel<<<grid_for(K, cap), kT, (size_t)N * O * 4, st>>>(dy, ldy, col_off, w, N, K, O, dx);
        }
    }
    else linear_bwd_x_kernel<<<grid_for((int64_t)N * K, cap), kT, 0, st>>>(dy, ldy, col_off, w, N, K, O, dx);
    GN_LAUNCHED(t->h);
    return GN_OK;
}
int lin_bwd_w(gn_avm_trainer* t, cudaStream_t st, const float* dy, int ldy, int col_off, const float* x, int ldx, int N, int K,
              int O, float* dw, float* db) {
    const int cap = t->h->num_sms * 16;
    const bool v4 = (K % 4 == 0) && (ldx % 4 == 0) && ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(dw)) & 15) == 0;
    if (v4) linear_bwd_w_vec4_kernel<<<grid_for((int64_t)O * (K / 4), cap), kT, 0, st>>>(dy, ldy, col_off, x, ldx, N, K, O, dw, db);
    else linear_bwd_w_kernel<<<grid_for((int64_t)O * K, cap), kT, 0, st>>>(dy, ldy, col_off, x, ldx, N, K, O, dw, db);
    GN_LAUNCHED(t->h);
    return GN_OK;
}
}  // namespace
#define LF(...) do { int rc_ = lin_fwd(t, st, __VA_ARGS__); if (rc_) return rc_; } while (0)
#define LBX(...) do { int rc_ = lin_bwd_x(t, st, __VA_ARGS__); if (rc_) return rc_; } while (0)
#define LBW(...) do { int rc_ = lin_bwd_w(t, st, __VA_ARGS__); if (rc_) return rc_; } while (0)

#define TL(kern, total, ...)                                                             \
    do {                                                                                 \
        kern<<<grid_for((total), cap), kT, 0, st>>>(__VA_ARGS__);                        \
        GN_LAUNCHED(t->h);                                                               \
    } while (0)

extern "C" int gn_avm_trainer_forward_backward(gn_avm_trainer* t, const float* d_visual, const float* d_audio,
                                               const float* d_labels, int64_t n64, int fixed_loss, float dropout_p,
                                               const uint8_t* const* d_masks, unsigned long long seed, float* d_scores,
                                               double* d_loss, void* stream) {
    if (!t || !d_visual || !d_labels || n64 < 1 || n64 > t->max_n) return GN_E_ARG;
    if (t->audio && !d_audio) return GN_E_ARG;
    if (dropout_p < 0.f || dropout_p >= 1.f) return GN_E_ARG;
    cudaStream_t st = gn_stream(stream);
    const int n = (int)n64, cap = t->h->num_sms * 8;
    float* P = t->params;
    float* G = t->grads;
    const int64_t* o = t->off;
    const int C[3] = {64, 256, 512}, Cin[3] = {3, 64, 256}, IH[3] = {t->S, t->P1, t->P2};
    const int OH[3] = {t->C1, t->P1, t->P2}, PH[3] = {t->P1, t->P2, t->P3};
    const int cw[3] = {P_C1W, P_C2W, P_C3W}, cb[3] = {P_C1B, P_C2B, P_C3B}, bw[3] = {P_BN1W, P_BN2W, P_BN3W},
              bb[3] = {P_BN1B, P_BN2B, P_BN3B};
    const int stride[3] = {3, 1, 1}, pad[3] = {3, 1, 1};
    const float scale = dropout_p > 0.f ? 1.f / (1.f - dropout_p) : 1.f;
    const int mw[5] = {512, 512, 512, 256, 128};
    const uint8_t* mk[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
    if (dropout_p > 0.f) {
        for (int i = 0; i < 5; ++i) {
            if (d_masks && d_masks[i]) mk[i] = d_masks[i];
            else {
                TL(make_mask_kernel, (int64_t)n * mw[i], t->masks[i], (int64_t)n * mw[i],
                   seed + 0x632BE59BD9B4E019ull * (unsigned long long)(i + 1), 1.f - dropout_p, t->state);
                mk[i] = t->masks[i];
            }
        }
    }
    // (no memset of G: every gradient element is written, not accumulated, by exactly one kernel of the backward)

    // ---------------- forward: VisBl (utils.py:172-195) ----------------
    const float* x = d_visual;
    for (int i = 0; i < 3; ++i) {
        const bool tc = t->use_tc && i > 0;
        if (i == 0) {     // K = 27: direct kernel (a GEMM tile would be mostly padding)
            TL(conv_fwd_kernel, (int64_t)n * C[i] * OH[i] * OH[i], x, P + o[cw[i]], P + o[cb[i]], n, Cin[i], IH[i], IH[i], C[i], 3,
               3, stride[i], stride[i], pad[i], pad[i], OH[i], OH[i], 1, t->y[i]);
        } else {
            int rc = (tc ? conv_fwd_tc : conv_fwd_gemm)(t, x, P + o[cw[i]], P + o[cb[i]], n, Cin[i], IH[i], C[i], stride[i],
                                                        pad[i], OH[i], t->y[i], st);
            if (rc) return rc;
        }
        TL(pool_fwd_kernel, (int64_t)n * C[i] * PH[i] * PH[i], t->y[i], n * C[i], OH[i], OH[i], t->p[i], t->arg[i]);
        bn_fwd_kernel<<<C[i], kT, 0, st>>>(t->p[i], n, C[i], PH[i] * PH[i], P + o[bw[i]], P + o[bb[i]], t->rmean[i],
                                           t->rvar[i], t->xh[i], t->istd[i], t->a[i]);
        GN_LAUNCHED(t->h);
        x = t->a[i];
    }
    const int aoff = t->audio ? 128 : 0;
    // linear5 + ReLU + Dropout -> columns [aoff, aoff+512) of the fusion input; NCHW flatten == the reference's order
    LF( t->a[2], (int)t->K5, P + o[P_L5W], P + o[P_L5B], n, (int)t->K5, 512, 1, mk[0],
       scale, t->fin_buf, t->fin, aoff);
    // ---------------- forward: AudBl (utils.py:214-227) ----------------
    if (t->audio) {
        TL(conv_fwd_kernel, (int64_t)n * 64 * t->L1, d_audio, P + o[P_AC1W], P + o[P_AC1B], n, 30, 1, t->B, 64, 1, 3, 1, 2, 0, 1,
           1, t->L1, 1, t->ay1);
        TL(conv_fwd_kernel, (int64_t)n * 128 * t->L2, t->ay1, P + o[P_AC2W], P + o[P_AC2B], n, 64, 1, t->L1, 128, 1, 3, 1, 2, 0,
           1, 1, t->L2, 1, t->ay2);
        LF( t->ay2, 128 * t->L2, P + o[P_AL3W], P + o[P_AL3B], n, 128 * t->L2, 128, 1,
           (const uint8_t*)nullptr, 1.f, t->fin_buf, t->fin, 0);
    }
    // ---------------- forward: fusion (utils.py:242-258) ----------------
    const int fw[4] = {512, 512, 256, 128}, fk[4] = {t->fin, 512, 512, 256};
    const int fwi[4] = {P_F0W, P_F3W, P_F6W, P_F9W}, fbi[4] = {P_F0B, P_F3B, P_F6B, P_F9B};
    const float* hin = t->fin_buf;
    for (int i = 0; i < 4; ++i) {
        LF( hin, fk[i], P + o[fwi[i]], P + o[fbi[i]], n, fk[i], fw[i], 1, mk[i + 1],
           scale, t->f[i], fw[i], 0);
        hin = t->f[i];
    }
    LF( t->f[3], 128, P + o[P_F12W], P + o[P_F12B], n, 128, 1, 0, (const uint8_t*)nullptr,
       1.f, t->z, 1, 0);
    loss_kernel<<<1, kT, 0, st>>>(t->z, d_labels, n, fixed_loss, d_scores, t->dz, d_loss);
    GN_LAUNCHED(t->h);

    // ---------------- backward: fusion ----------------
    LBW( t->dz, 1, 0, t->f[3], 128, n, 128, 1, G + o[P_F12W], G + o[P_F12B]);
    LBX( t->dz, 1, 0, P + o[P_F12W], n, 128, 1, t->d_f[3]);
    for (int i = 3; i >= 0; --i) {
        TL(relu_drop_bwd_kernel, (int64_t)n * fw[i], t->d_f[i], t->f[i], mk[i + 1], scale, (int64_t)n * fw[i]);
        const float* xin = i ? t->f[i - 1] : t->fin_buf;
        LBW( t->d_f[i], fw[i], 0, xin, fk[i], n, fk[i], fw[i], G + o[fwi[i]],
           G + o[fbi[i]]);
        LBX( t->d_f[i], fw[i], 0, P + o[fwi[i]], n, fk[i], fw[i],
           i ? t->d_f[i - 1] : t->d_fin);
    }
    // d_fin = gradient w.r.t. [audio128 | visual512] (both are post-ReLU outputs; the visual half also post-dropout)
    // ---------------- backward: AudBl ----------------
    if (t->audio) {
        // columns [0,128): gate by the ReLU of linear3 (no dropout)
        float* dl3 = t->d_big0;          // [n][128]
        GN_CUDA(cudaMemcpy2DAsync(dl3, 128 * 4, t->d_fin, (size_t)t->fin * 4, 128 * 4, n, cudaMemcpyDeviceToDevice, st));
        float* al3 = t->d_big1;          // [n][128] copy of the activations for the gate
        GN_CUDA(cudaMemcpy2DAsync(al3, 128 * 4, t->fin_buf, (size_t)t->fin * 4, 128 * 4, n, cudaMemcpyDeviceToDevice, st));
        TL(relu_drop_bwd_kernel, (int64_t)n * 128, dl3, al3, (const uint8_t*)nullptr, 1.f, (int64_t)n * 128);
        const int K3 = 128 * t->L2;
        LBW( dl3, 128, 0, t->ay2, K3, n, K3, 128, G + o[P_AL3W], G + o[P_AL3B]);
        float* day2 = t->d_big1;         // [n][128][L2]  (al3 is dead after the gate)
        LBX( dl3, 128, 0, P + o[P_AL3W], n, K3, 128, day2);
        TL(relu_drop_bwd_kernel, (int64_t)n * K3, day2, t->ay2, (const uint8_t*)nullptr, 1.f, (int64_t)n * K3);
        TL(conv_bwd_w_kernel, (int64_t)128 * 192 * 32, day2, t->ay1, n, 64, 1, t->L1, 128, 1, 3, 1, 2, 0, 1, 1, t->L2,
           G + o[P_AC2W], G + o[P_AC2B]);
        float* day1 = t->d_big0;         // [n][64][L1]
        TL(conv_bwd_x_kernel, (int64_t)n * 64 * t->L1, day2, P + o[P_AC2W], n, 64, 1, t->L1, 128, 1, 3, 1, 2, 0, 1, 1, t->L2,
           day1);
        TL(relu_drop_bwd_kernel, (int64_t)n * 64 * t->L1, day1, t->ay1, (const uint8_t*)nullptr, 1.f, (int64_t)n * 64 * t->L1);
        TL(conv_bwd_w_kernel, (int64_t)64 * 90 * 32, day1, d_audio, n, 30, 1, t->B, 64, 1, 3, 1, 2, 0, 1, 1, t->L1,
           G + o[P_AC1W], G + o[P_AC1B]);
    }
    // ---------------- backward: VisBl ----------------
    float* dh5 = t->d_big0;              // [n][512]
    GN_CUDA(cudaMemcpy2DAsync(dh5, 512 * 4, t->d_fin + aoff, (size_t)t->fin * 4, 512 * 4, n, cudaMemcpyDeviceToDevice, st));
    GN_CUDA(cudaMemcpy2DAsync(t->h5, 512 * 4, t->fin_buf + aoff, (size_t)t->fin * 4, 512 * 4, n, cudaMemcpyDeviceToDevice, st));
    TL(relu_drop_bwd_kernel, (int64_t)n * 512, dh5, t->h5, mk[0], scale, (int64_t)n * 512);
    LBW( dh5, 512, 0, t->a[2], (int)t->K5, n, (int)t->K5, 512, G + o[P_L5W], G + o[P_L5B]);
    float* da = t->d_big1;               // gradient w.r.t. the BN output of block i
    LBX( dh5, 512, 0, P + o[P_L5W], n, (int)t->K5, 512, da);
    for (int i = 2; i >= 0; --i) {
        float* dp = (da == t->d_big1) ? t->d_big0 : t->d_big1;
        bn_bwd_kernel<<<C[i], kT, 0, st>>>(da, t->xh[i], t->istd[i], P + o[bw[i]], n, C[i], PH[i] * PH[i], dp, G + o[bw[i]],
                                           G + o[bb[i]]);
        GN_LAUNCHED(t->h);
        float* dy = da;                  // reuse: da is dead now
        TL(pool_relu_bwd_kernel, (int64_t)n * C[i] * OH[i] * OH[i], dp, t->arg[i], t->y[i], n * C[i], OH[i], OH[i], dy);
        const float* xin = i ? t->a[i - 1] : d_visual;
        const bool tc = t->use_tc && i > 0;
        int rc = GN_OK;
        if (i == 0) {
            TL(conv_bwd_w_kernel, (int64_t)C[i] * Cin[i] * 9 * 32, dy, xin, n, Cin[i], IH[i], IH[i], C[i], 3, 3, stride[i],
               stride[i], pad[i], pad[i], OH[i], OH[i], G + o[cw[i]], G + o[cb[i]]);
        } else {
            rc = (tc ? conv_bwd_w_tc : conv_bwd_w_gemm)(t, dy, xin, n, Cin[i], IH[i], C[i], stride[i], pad[i], OH[i],
                                                        G + o[cw[i]], G + o[cb[i]], st);
            if (rc) return rc;
        }
        if (i > 0) {
            rc = (tc ? conv_bwd_x_tc : conv_bwd_x_gemm)(t, dy, P + o[cw[i]], n, Cin[i], IH[i], C[i], stride[i], pad[i], OH[i],
                                                        dp, st);
            if (rc) return rc;
            da = dp;
        }
    }
    return GN_OK;
}

extern "C" int gn_avm_trainer_adam_step(gn_avm_trainer* t, float grad_scale, void* stream) {
    if (!t) return GN_E_ARG;
    cudaStream_t st = gn_stream(stream);
    t->step += 1;
    // step count and bias corrections advance ON THE DEVICE (a captured graph replays them)
    int rc0 = gn_launch_step_advance(t->h, t->state, t->b1, t->b2, st);
    if (rc0) return rc0;
    adam_kernel<<<grid_for(t->n_params, t->h->num_sms * 8), kT, 0, st>>>(t->params, t->grads, t->m, t->v, t->n_params,
                                                                         grad_scale, t->lr, t->b1, t->b2, t->eps, t->state);
    GN_LAUNCHED(t->h);
    return GN_OK;
}

extern "C" int gn_avm_train_step(gn_avm_trainer* t, const float* d_visual, const float* d_audio, const float* d_labels,
                                 int64_t n, int fixed_loss, float dropout_p, const uint8_t* const* d_masks,
                                 unsigned long long seed, float* d_scores, double* d_loss, void* stream) {
    int rc = gn_avm_trainer_forward_backward(t, d_visual, d_audio, d_labels, n, fixed_loss, dropout_p, d_masks, seed,
                                             d_scores, d_loss, stream);
    if (rc) return rc;
    return gn_avm_trainer_adam_step(t, 1.f, stream);
}
