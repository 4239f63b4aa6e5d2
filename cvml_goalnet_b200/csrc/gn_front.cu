// CUDA-core front layers of the AVM forward (fp32 math; they are <0.5 % of the FLOPs and their inputs
// need fp32 range/precision: normalised pixels and raw MFCCs).
//   * VisBl block 1: conv1 3->64 k3 s3 p3 + ReLU + MaxPool(3,1) + BatchNorm(eval)  (utils.py:174-177)
//     -> bf16 NHWC with a zero border (the padding=1 input of conv2);
//   * AudBl: conv1d 30->64 s2, conv1d 64->128 s2, linear -> 128, ReLUs             (utils.py:216-225)
//     -> bf16 columns [0,128) of the fusion input (audio first, utils.py:266).
#include "gn_common.cuh"

namespace {

constexpr int kVThreads = 256;

__global__ void __launch_bounds__(kVThreads)
visbl_block1_kernel(const float* __restrict__ vis, int64_t n, int S, int C1, const float* __restrict__ w /*[27][64]*/,
                    const float* __restrict__ b, const float* __restrict__ sc, const float* __restrict__ sh,
                    __nv_bfloat16* __restrict__ a1p) {
    extern __shared__ __align__(16) float sm[];
    float* img = sm;                      // [3][S][S]
    float* wsm = img + 3 * S * S;         // [27][64]
    float* c1 = wsm + 27 * 64;            // [C1*C1][64]  conv + ReLU
    const int tid = threadIdx.x;
    for (int i = tid; i < 27 * 64; i += kVThreads) wsm[i] = w[i];
    const int P = C1 - 2;                 // pooled size; the padded output is C1 x C1 again
    const int co = tid & 63, pg = tid >> 6;
    const float bias = b[co], scale = sc[co], shift = sh[co];
    for (int64_t f = blockIdx.x; f < n; f += gridDim.x) {
        __syncthreads();
        const float4* src = reinterpret_cast<const float4*>(vis + f * 3 * S * S);
        for (int i = tid; i < 3 * S * S / 4; i += kVThreads) reinterpret_cast<float4*>(img)[i] = src[i];
        __syncthreads();
        for (int p = pg; p < C1 * C1; p += kVThreads / 64) {
            const int oy = p / C1, ox = p - oy * C1;
            float acc = bias;
#pragma unroll
            for (int ci = 0; ci < 3; ++ci)
#pragma unroll
                for (int ky = 0; ky < 3; ++ky) {
                    const int iy = oy * 3 - 3 + ky;
#pragma unroll
                    for (int kx = 0; kx < 3; ++kx) {
                        const int ix = ox * 3 - 3 + kx;
                        const float v = (iy >= 0 && iy < S && ix >= 0 && ix < S) ? img[(ci * S + iy) * S + ix] : 0.f;
                        acc = fmaf(v, wsm[((ky * 3 + kx) * 3 + ci) * 64 + co], acc);
                    }
                }
            c1[p * 64 + co] = fmaxf(acc, 0.f);
        }
        __syncthreads();
        __nv_bfloat16* o = a1p + f * (int64_t)C1 * C1 * 64;
        for (int p = pg; p < C1 * C1; p += kVThreads / 64) {
            const int y = p / C1, x = p - y * C1;
            float r = 0.f;
            if (y >= 1 && y <= P && x >= 1 && x <= P) {
                float m = -INFINITY;
#pragma unroll
                for (int dy = 0; dy < 3; ++dy)
#pragma unroll
                    for (int dx = 0; dx < 3; ++dx) m = fmaxf(m, c1[((y - 1 + dy) * C1 + x - 1 + dx) * 64 + co]);
                r = fmaf(m, scale, shift);
            }
            o[p * 64 + co] = __float2bfloat16(r);
        }
    }
}

constexpr int kAThreads = 256;
constexpr int kAFrames = 4;    // frames per CTA pass: every weight fetched from L1/L2 is used 4 times

__global__ void __launch_bounds__(kAThreads)
audbl_kernel(const float* __restrict__ aud, int64_t n, int B, int L1, int L2, const float* __restrict__ w1,
             const float* __restrict__ b1, const float* __restrict__ w2, const float* __restrict__ b2,
             const float* __restrict__ w3, const float* __restrict__ b3, __nv_bfloat16* __restrict__ h0, int ld) {
    extern __shared__ __align__(16) float sm[];
    float* in = sm;                              // [F][30][B]
    float* c1 = in + kAFrames * 30 * B;          // [F][64][L1]
    float* c2 = c1 + kAFrames * 64 * L1;         // [F][128*L2]
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int K3 = 128 * L2;
    for (int64_t f0 = (int64_t)blockIdx.x * kAFrames; f0 < n; f0 += (int64_t)gridDim.x * kAFrames) {
        const int nf = (int)((n - f0) < kAFrames ? (n - f0) : kAFrames);
        __syncthreads();
        for (int i = tid; i < kAFrames * 30 * B; i += kAThreads)
            in[i] = (i < nf * 30 * B) ? aud[f0 * 30 * B + i] : 0.f;
        __syncthreads();
        // conv1d 30 -> 64, k3 s2 p1
        for (int o = tid; o < 64 * L1; o += kAThreads) {
            const int co = o / L1, l = o - co * L1;
            float acc[kAFrames];
#pragma unroll
            for (int f = 0; f < kAFrames; ++f) acc[f] = b1[co];
            for (int ci = 0; ci < 30; ++ci)
#pragma unroll
                for (int k = 0; k < 3; ++k) {
                    const int t = 2 * l - 1 + k;
                    if (t < 0 || t >= B) continue;
                    const float wv = __ldg(w1 + (co * 30 + ci) * 3 + k);
#pragma unroll
                    for (int f = 0; f < kAFrames; ++f) acc[f] = fmaf(wv, in[(f * 30 + ci) * B + t], acc[f]);
                }
#pragma unroll
            for (int f = 0; f < kAFrames; ++f) c1[(f * 64 + co) * L1 + l] = fmaxf(acc[f], 0.f);
        }
        __syncthreads();
        // conv1d 64 -> 128, k3 s2 p1
        for (int o = tid; o < 128 * L2; o += kAThreads) {
            const int co = o / L2, l = o - co * L2;
            float acc[kAFrames];
#pragma unroll
            for (int f = 0; f < kAFrames; ++f) acc[f] = b2[co];
            for (int ci = 0; ci < 64; ++ci)
#pragma unroll
                for (int k = 0; k < 3; ++k) {
                    const int t = 2 * l - 1 + k;
                    if (t < 0 || t >= L1) continue;
                    const float wv = __ldg(w2 + (co * 64 + ci) * 3 + k);
#pragma unroll
                    for (int f = 0; f < kAFrames; ++f) acc[f] = fmaf(wv, c1[(f * 64 + ci) * L1 + t], acc[f]);
                }
#pragma unroll
            for (int f = 0; f < kAFrames; ++f) c2[f * K3 + co * L2 + l] = fmaxf(acc[f], 0.f);   // flatten: c*L2 + l
        }
        __syncthreads();
        // linear 128*L2 -> 128: one warp per output feature, lanes stride over K (coalesced weight rows)
        for (int j = warp; j < 128; j += kAThreads / 32) {
            float acc[kAFrames];
#pragma unroll
            for (int f = 0; f < kAFrames; ++f) acc[f] = 0.f;
            const float* wr = w3 + (size_t)j * K3;
            for (int k = lane; k < K3; k += 32) {
                const float wv = __ldg(wr + k);
#pragma unroll
                for (int f = 0; f < kAFrames; ++f) acc[f] = fmaf(wv, c2[f * K3 + k], acc[f]);
            }
#pragma unroll
            for (int f = 0; f < kAFrames; ++f) {
                float v = acc[f];
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
                if (lane == 0 && f < nf) h0[(f0 + f) * ld + j] = __float2bfloat16(fmaxf(v + b3[j], 0.f));
            }
        }
    }
}

}  // namespace

int gn_launch_visbl_front(gn_model* m, const float* d_visual, int64_t n, const gn_ws& ws, cudaStream_t st) {
    const size_t smem = sizeof(float) * (3 * m->S * m->S + 27 * 64 + (size_t)m->C1 * m->C1 * 64);
    static bool attr = false;
    if (!attr) {
        GN_CUDA(cudaFuncSetAttribute(visbl_block1_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
        attr = true;
    }
    if (smem > 100 * 1024) return GN_E_SHAPE;
    const int64_t cap = (int64_t)m->h->num_sms * 2;
    const int grid = (int)(n < cap ? n : cap);
    visbl_block1_kernel<<<grid, kVThreads, smem, st>>>(d_visual, n, m->S, m->C1, m->conv1_w, m->conv1_b,
                                                       m->bn_scale[0], m->bn_shift[0], ws.a1p);
    GN_LAUNCHED(m->h);
    return GN_OK;
}

int gn_launch_audbl(gn_model* m, const float* d_audio, int64_t n, const gn_ws& ws, cudaStream_t st) {
    if (!m->audio_included) return GN_OK;
    const size_t smem = sizeof(float) * kAFrames * (30 * m->B + 64 * m->L1 + 128 * m->L2);
    if (smem > 48 * 1024) return GN_E_SHAPE;
    const int64_t groups = (n + kAFrames - 1) / kAFrames;
    const int64_t cap = (int64_t)m->h->num_sms * 4;
    const int grid = (int)(groups < cap ? groups : cap);
    audbl_kernel<<<grid, kAThreads, smem, st>>>(d_audio, n, m->B, m->L1, m->L2, m->aconv1_w, m->aconv1_b,
                                                m->aconv2_w, m->aconv2_b, m->alinear3_w, m->alinear3_b, ws.h0, m->fin);
    GN_LAUNCHED(m->h);
    return GN_OK;
}
