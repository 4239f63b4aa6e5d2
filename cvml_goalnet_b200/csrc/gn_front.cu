// CUDA-core front layers of the AVM forward (fp32 math; they are <0.5 % of the FLOPs and their inputs
// need fp32 range/precision: normalised pixels and raw MFCCs).  For the reference's shapes (40x40 frames, 30 time bins)
// the default path runs both layers on the tensor cores instead (gn_front_tc.cu); these kernels serve other shapes, the
// validation path (gn_dbg_avm_forward_simt) and the GN_BLOCK1_SIMT / GN_AUDBL_SIMT switches.
//   * VisBl block 1: conv1 3->64 k3 s3 p3 + ReLU + MaxPool(3,1) + BatchNorm(eval)  (utils.py:174-177)
//     -> bf16 NHWC with a zero border (the padding=1 input of conv2);
//   * AudBl: conv1d 30->64 s2, conv1d 64->128 s2, linear -> 128, ReLUs             (utils.py:216-225)
//     -> TF32 columns [0,128) of the fusion input (audio first, utils.py:266).
#include "gn_common.cuh"
#include "gn_fft2.cuh"      // packed fp32 helpers (V2, fma2)

namespace {

constexpr int kVThreads = 256;
constexpr int kS = 40;             // frame side (reference utils.py:285)
constexpr int kC1 = 15;            // conv1 output side: (40 + 2*3 - 3)/3 + 1
constexpr int kPatch = 28;         // 27 taps padded to 7 float4

// conv1 has stride == kernel size (k3 s3 p3), so its im2col is a permutation of the zero-padded image:
// pixel (y,x) of channel ci is tap (ci, y%3, x%3) of output pixel (y/3+1, x/3+1); row/column 0 of the
// output see only padding.  The image is scattered into patch-major shared memory once, after which a
// warp computes one output pixel for all 64 channels (2 per lane, weights in registers) from 7
// broadcast 16-byte loads.
__global__ void __launch_bounds__(kVThreads, 2)
visbl_block1_kernel(const float* __restrict__ vis, int64_t n, const float* __restrict__ w /*[27][64], k=(ky*3+kx)*3+ci*/,
                    const float* __restrict__ b, const float* __restrict__ sc, const float* __restrict__ sh,
                    __nv_bfloat16* __restrict__ a1p) {
    extern __shared__ __align__(16) float sm[];
    float* patch = sm;                                         // [225][28]
    float* c1 = patch + kC1 * kC1 * kPatch;                    // [225][64]  conv + ReLU
    unsigned short* dsto = reinterpret_cast<unsigned short*>(c1 + kC1 * kC1 * 64);   // [4800] scatter offsets
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int i = tid; i < kC1 * kC1 * kPatch; i += kVThreads) patch[i] = 0.f;
    for (int e = tid; e < 3 * kS * kS; e += kVThreads) {
        const int ci = e / (kS * kS), r = e - ci * kS * kS, y = r / kS, x = r - y * kS;
        dsto[e] = (unsigned short)(((y / 3 + 1) * kC1 + (x / 3 + 1)) * kPatch + ci * 9 + (y % 3) * 3 + (x % 3));
    }
    float w0[27], w1[27];                                       // weights of channels 2*lane, 2*lane+1, k = ci*9 + tap
#pragma unroll
    for (int ci = 0; ci < 3; ++ci)
#pragma unroll
        for (int tap = 0; tap < 9; ++tap) {
            const float2 wv = *reinterpret_cast<const float2*>(w + (tap * 3 + ci) * 64 + 2 * lane);
            w0[ci * 9 + tap] = wv.x;
            w1[ci * 9 + tap] = wv.y;
        }
    const float2 bias = *reinterpret_cast<const float2*>(b + 2 * lane);
    const float2 scale = *reinterpret_cast<const float2*>(sc + 2 * lane);
    const float2 shift = *reinterpret_cast<const float2*>(sh + 2 * lane);
    constexpr int P = kC1 - 2;                                  // pooled size; the padded output is C1 x C1 again
    // the next frame's pixels travel to registers while this frame is convolved and pooled
    constexpr int kVec = 3 * kS * kS / 4, kPre = (kVec + kVThreads - 1) / kVThreads;
    float4 pre[kPre];
    auto prefetch = [&](int64_t f) {
        const float4* src = reinterpret_cast<const float4*>(vis + f * 3 * kS * kS);
#pragma unroll
        for (int v = 0; v < kPre; ++v) {
            const int i = tid + v * kVThreads;
            if (i < kVec) pre[v] = __ldg(src + i);
        }
    };
    if ((int64_t)blockIdx.x < n) prefetch(blockIdx.x);
    for (int64_t f = blockIdx.x; f < n; f += gridDim.x) {
        // patch was last read by the previous frame's convolution, which every thread left at the barrier below
#pragma unroll
        for (int v = 0; v < kPre; ++v) {
            const int i = tid + v * kVThreads;
            if (i < kVec) {
                const uint2 o = *reinterpret_cast<const uint2*>(dsto + 4 * i);
                patch[o.x & 0xffffu] = pre[v].x;
                patch[o.x >> 16] = pre[v].y;
                patch[o.y & 0xffffu] = pre[v].z;
                patch[o.y >> 16] = pre[v].w;
            }
        }
        __syncthreads();          // patch complete; the previous frame's pooling is done with c1
        if (f + gridDim.x < n) prefetch(f + gridDim.x);
        for (int p = warp; p < kC1 * kC1; p += kVThreads / 32) {
            const float4* pp = reinterpret_cast<const float4*>(patch + p * kPatch);
            float in[kPatch];
#pragma unroll
            for (int v = 0; v < kPatch / 4; ++v) {
                const float4 t = pp[v];
                in[4 * v] = t.x; in[4 * v + 1] = t.y; in[4 * v + 2] = t.z; in[4 * v + 3] = t.w;
            }
            float a0 = bias.x, a1 = bias.y;
#pragma unroll
            for (int k = 0; k < 27; ++k) {
                a0 = fmaf(in[k], w0[k], a0);
                a1 = fmaf(in[k], w1[k], a1);
            }
            *reinterpret_cast<float2*>(c1 + p * 64 + 2 * lane) = make_float2(fmaxf(a0, 0.f), fmaxf(a1, 0.f));
        }
        __syncthreads();
        uint32_t* o = reinterpret_cast<uint32_t*>(a1p + f * (int64_t)kC1 * kC1 * 64);
        // zero border = conv2's padding
        for (int bi = warp; bi < 4 * kC1 - 4; bi += kVThreads / 32) {
            int y, x;
            if (bi < kC1) { y = 0; x = bi; }
            else if (bi < 2 * kC1) { y = kC1 - 1; x = bi - kC1; }
            else { const int t = bi - 2 * kC1; y = 1 + (t >> 1); x = (t & 1) ? kC1 - 1 : 0; }
            o[(y * kC1 + x) * 32 + lane] = 0u;
        }
        // separable 3x3 max pool: a warp walks a third of one pooled column (lane = channel pair), keeping the
        // horizontal maxima of the last three conv rows -- 3 shared loads per pooled pixel instead of 9
        constexpr int kThirds = 3;
        for (int item = warp; item < P * kThirds; item += kVThreads / 32) {
            const int x = item / kThirds, t = item - x * kThirds;          // pooled column, row range
            const int y0 = (t * P) / kThirds, y1 = ((t + 1) * P) / kThirds;
            const float* col = c1 + x * 64 + 2 * lane;                     // conv columns x .. x+2
            float2 h0 = make_float2(0.f, 0.f), h1 = h0;
            for (int r = y0; r < y1 + 2; ++r) {
                const float* rp = col + r * kC1 * 64;
                const float2 u = *reinterpret_cast<const float2*>(rp), v = *reinterpret_cast<const float2*>(rp + 64),
                             w2 = *reinterpret_cast<const float2*>(rp + 128);
                const float2 h2 = make_float2(fmaxf(fmaxf(u.x, v.x), w2.x), fmaxf(fmaxf(u.y, v.y), w2.y));
                if (r >= y0 + 2) {
                    const float mx = fmaxf(fmaxf(h0.x, h1.x), h2.x), my = fmaxf(fmaxf(h0.y, h1.y), h2.y);
                    const __nv_bfloat162 h = __floats2bfloat162_rn(fmaf(mx, scale.x, shift.x), fmaf(my, scale.y, shift.y));
                    o[((r - 1) * kC1 + x + 1) * 32 + lane] = *reinterpret_cast<const uint32_t*>(&h);
                }
                h0 = h1;
                h1 = h2;
            }
        }
    }
}

// VisBl block 1 for frames of any size S (e.g. 224: conv1 -> 76x76, pool -> 74x74).  One CTA visit = one 16x16 tile of
// conv1 outputs (tiles overlap by the pooling halo -> 14x14 pooled pixels); a warp computes one conv pixel for all 64
// channels (2 per lane, weights in registers) from the tile's input window staged patch-major in shared memory.  This
// block is 0.1 % of the model's FLOPs at 224x224; the zero border of a1p is written by gn_launch_zero_border.
constexpr int kT1 = 16, kT1P = kT1 - 2;

__global__ void __launch_bounds__(kVThreads, 2)
visbl_block1_tiled_kernel(const float* __restrict__ vis, int64_t n, int S, int C1, int tiles,
                          const float* __restrict__ w /*[27][64]*/, const float* __restrict__ b,
                          const float* __restrict__ sc, const float* __restrict__ sh, __nv_bfloat16* __restrict__ a1p) {
    extern __shared__ __align__(16) float sm[];
    float* c1 = sm;                                             // [256][64] conv + ReLU
    float* patch = sm + kT1 * kT1 * 64;                         // [256][28] taps of every conv pixel of the tile (k = ci*9+ky*3+kx)
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    float w0[27], w1[27];                                       // k = ci*9 + ky*3 + kx
#pragma unroll
    for (int ci = 0; ci < 3; ++ci)
#pragma unroll
        for (int tap = 0; tap < 9; ++tap) {
            const float2 wv = *reinterpret_cast<const float2*>(w + (tap * 3 + ci) * 64 + 2 * lane);
            w0[ci * 9 + tap] = wv.x;
            w1[ci * 9 + tap] = wv.y;
        }
    const float2 bias = *reinterpret_cast<const float2*>(b + 2 * lane);
    const float2 scale = *reinterpret_cast<const float2*>(sc + 2 * lane);
    const float2 shift = *reinterpret_cast<const float2*>(sh + 2 * lane);
    const int P1 = C1 - 2;
    constexpr int kIn = 3 * kT1;                                // 48 input rows / columns feed a 16x16 conv tile (k3 s3)
    const int64_t units = n * tiles * tiles;
    for (int64_t u = blockIdx.x; u < units; u += gridDim.x) {
        const int64_t f = u / (tiles * tiles);
        const int t = (int)(u - f * tiles * tiles);
        const int ty = t / tiles, tx = t - ty * tiles;
        const float* src = vis + f * 3 * (int64_t)S * S;
        const int iy0 = 3 * ty * kT1P - 3, ix0 = 3 * tx * kT1P - 3;   // input coordinate of the tile's first tap (padding 3)
        __syncthreads();
        // conv1 has stride == kernel size: its im2col is a permutation of the input window -> scatter it patch-major
        // (thread = one window column and every 5th row: the column part of the index math is loop-invariant)
        if (tid < 5 * kIn) {
            const int x = tid % kIn, yy = tid / kIn, ix = ix0 + x;
            const bool okx = ix >= 0 && ix < S;
            float* pcol = patch + (x / 3) * kPatch + (x % 3);
            for (int y = yy; y < kIn; y += 5) {
                const int iy = iy0 + y;
                const bool ok = okx && iy >= 0 && iy < S;
                float* pd = pcol + (y / 3) * (kT1 * kPatch) + (y % 3) * 3;
                const float* ps = src + (int64_t)iy * S + ix;
#pragma unroll
                for (int ci = 0; ci < 3; ++ci) pd[ci * 9] = ok ? __ldg(ps + (int64_t)ci * S * S) : 0.f;
            }
        }
        __syncthreads();
        for (int p = warp; p < kT1 * kT1; p += kVThreads / 32) {
            const float4* pp = reinterpret_cast<const float4*>(patch + p * kPatch);
            float in[kPatch];
#pragma unroll
            for (int v = 0; v < kPatch / 4; ++v) {
                const float4 q = pp[v];
                in[4 * v] = q.x; in[4 * v + 1] = q.y; in[4 * v + 2] = q.z; in[4 * v + 3] = q.w;
            }
            float a0 = bias.x, a1 = bias.y;
#pragma unroll
            for (int k = 0; k < 27; ++k) {
                a0 = fmaf(in[k], w0[k], a0);
                a1 = fmaf(in[k], w1[k], a1);
            }
            *reinterpret_cast<float2*>(c1 + p * 64 + 2 * lane) = make_float2(fmaxf(a0, 0.f), fmaxf(a1, 0.f));
        }
        __syncthreads();
        uint32_t* o = reinterpret_cast<uint32_t*>(a1p + f * (int64_t)C1 * C1 * 64);
        for (int p = warp; p < kT1P * kT1P; p += kVThreads / 32) {
            const int oy = p / kT1P, ox = p - oy * kT1P;
            const int gy = ty * kT1P + oy, gx = tx * kT1P + ox;
            if (gy >= P1 || gx >= P1) continue;
            const float* base = c1 + (oy * kT1 + ox) * 64 + 2 * lane;
            float2 m = *reinterpret_cast<const float2*>(base);
#pragma unroll
            for (int dy = 0; dy < 3; ++dy)
#pragma unroll
                for (int dx = 0; dx < 3; ++dx) {
                    if (dy == 0 && dx == 0) continue;
                    const float2 q = *reinterpret_cast<const float2*>(base + (dy * kT1 + dx) * 64);
                    m.x = fmaxf(m.x, q.x);
                    m.y = fmaxf(m.y, q.y);
                }
            const __nv_bfloat162 h = __floats2bfloat162_rn(fmaf(m.x, scale.x, shift.x), fmaf(m.y, scale.y, shift.y));
            o[((int64_t)(gy + 1) * C1 + gx + 1) * 32 + lane] = *reinterpret_cast<const uint32_t*>(&h);
        }
    }
}

constexpr int kAThreads = 256;
constexpr int kAFrames = 4;    // frames per CTA pass: every weight fetched from L1/L2 is used 4 times

__global__ void __launch_bounds__(kAThreads)
audbl_kernel(const float* __restrict__ aud, int64_t n, int B, int L1, int L2, const float* __restrict__ w1,
             const float* __restrict__ b1, const float* __restrict__ w2, const float* __restrict__ b2,
             const float* __restrict__ w3, const float* __restrict__ b3, float* __restrict__ h0, int ld) {
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
                if (lane == 0 && f < nf) h0[(f0 + f) * ld + j] = gn_round_tf32(fmaxf(v + b3[j], 0.f));
            }
        }
    }
}

// AudBl conv1 (30 -> 64, k3 s2 p1) + ReLU on CUDA cores (fp32: the raw MFCCs span +-600), written as
// TF32 (position, channel)-major rows with a zero row on each side for the tensor-core conv2.
// thread = (output channel, position group); its 90 weights live in registers.  TWO frames ride in the two lanes of
// the packed fp32 instructions (a scalar FFMA issues every other cycle): the inputs of kCF frames are staged as
// [frame pair][t][ci][2], so that one 16-byte load is two input channels of both frames and a tap is 15 broadcast
// loads; the 16-byte groups of a row are XOR-swizzled with the row so that the transposing stores do not all hit one
// bank (the plain [t][ci] layout was a 32-way conflict).
constexpr int kCThreads = 256;
constexpr int kCF = 4;

__global__ void __launch_bounds__(kCThreads, 2)
audbl_conv1_kernel(const float* __restrict__ aud, int64_t n, int B, int L1, const float* __restrict__ w1,
                   const float* __restrict__ b1, float* __restrict__ ac1) {
    using gnfft2::V2;
    using gnfft2::bc;
    using gnfft2::mk;
    using gnfft2::fma2;
    extern __shared__ __align__(16) float sm_all[];        // 2 x [kCF / 2][B + 2][32 ci][2 frames]
    const int tid = threadIdx.x, co = tid & 63, g = tid >> 6;
    const int rows = B + 2;
    const int buf_floats = (kCF / 2) * rows * 64;
    const unsigned b_magic = 0xFFFFFFFFu / (unsigned)B + 1u;      // __umulhi(r, b_magic) == r / B for r < 2^16
    float wr[90];
#pragma unroll
    for (int i = 0; i < 90; ++i) wr[i] = __ldg(w1 + co * 90 + i);     // [ci][k]
    const float bias = __ldg(b1 + co);
    for (int i = tid; i < 2 * buf_floats; i += kCThreads) sm_all[i] = 0.f;     // pad rows stay zero
    __syncthreads();
    // the inputs of pass p+1 travel (cp.async, 4 bytes each: the copy transposes) while pass p is computed
    auto stage = [&](int64_t f0, int buf) {
        const int nf = (int)((n - f0) < kCF ? (n - f0) : kCF);
        const uint32_t base = (uint32_t)__cvta_generic_to_shared(sm_all + buf * buf_floats);
        for (int f = 0; f < kCF; ++f) {
            const bool live = f < nf;                          // frames beyond the batch: zero fill (src-size 0)
            const float* src = aud + (live ? (f0 + f) * 30 * B : 0);
            const uint32_t fbase = base + 4u * (uint32_t)((f >> 1) * rows * 64 + (f & 1));
            for (int r = tid; r < 30 * B; r += kCThreads) {
                const int ci = (int)__umulhi((unsigned)r, b_magic), t = r - ci * B;      // r / B, r % B (r < 2^16)
                const int row = t + 1;
                const uint32_t dst = fbase + 4u * (uint32_t)(row * 64 + (((ci >> 1) ^ (row & 15)) << 2) + ((ci & 1) << 1));
                asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(dst), "l"(src + (live ? r : 0)),
                             "r"(live ? 4 : 0) : "memory");
            }
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    const int64_t fstep = (int64_t)gridDim.x * kCF;
    int buf = 0;
    if ((int64_t)blockIdx.x * kCF < n) stage((int64_t)blockIdx.x * kCF, 0);
    for (int64_t f0 = (int64_t)blockIdx.x * kCF; f0 < n; f0 += fstep, buf ^= 1) {
        const int nf = (int)((n - f0) < kCF ? (n - f0) : kCF);
        const bool more = f0 + fstep < n;
        if (more) stage(f0 + fstep, buf ^ 1);                  // (its previous readers passed the barrier below)
        if (more) asm volatile("cp.async.wait_group 1;" ::: "memory");
        else asm volatile("cp.async.wait_group 0;" ::: "memory");
        __syncthreads();
        const float* sm = sm_all + buf * buf_floats;
        for (int fp = 0; 2 * fp < nf; ++fp) {
            float* o0 = ac1 + (f0 + 2 * fp) * (int64_t)(L1 + 2) * 64;
            float* o1 = o0 + (int64_t)(L1 + 2) * 64;
            const bool two = 2 * fp + 1 < nf;
            if (g == 0) {
                o0[co] = 0.f; o0[(L1 + 1) * 64 + co] = 0.f;
                if (two) { o1[co] = 0.f; o1[(L1 + 1) * 64 + co] = 0.f; }
            }
            for (int l = g; l < L1; l += kCThreads / 64) {
                V2 acc0 = mk(bias, bias), acc1 = bc(0.f);
#pragma unroll
                for (int k = 0; k < 3; ++k) {
                    const int row = 2 * l + k;                                                   // t = 2l-1+k
                    const float4* xr = reinterpret_cast<const float4*>(sm + (fp * rows + row) * 64);
                    const int sw = row & 15;
#pragma unroll
                    for (int v = 0; v < 15; ++v) {
                        const float4 x = xr[v ^ sw];
                        acc0 = fma2(mk(x.x, x.y), bc(wr[(2 * v) * 3 + k]), acc0);
                        acc1 = fma2(mk(x.z, x.w), bc(wr[(2 * v + 1) * 3 + k]), acc1);
                    }
                }
                o0[(l + 1) * 64 + co] = gn_round_tf32(fmaxf(acc0.x + acc1.x, 0.f));
                if (two) o1[(l + 1) * 64 + co] = gn_round_tf32(fmaxf(acc0.y + acc1.y, 0.f));
            }
        }
        __syncthreads();                                       // this buffer is refilled by the next pass's stage()
    }
}

}  // namespace

int gn_launch_visbl_front(gn_model* m, const float* d_visual, int64_t n, const gn_ws& ws, cudaStream_t st, bool simt) {
    // the 40x40 model's block 1 runs on the tensor cores (gn_front_tc.cu); the CUDA-core kernel below stays as the
    // validation path (gn_dbg_avm_forward_simt) and for A/B timing (GN_BLOCK1_SIMT)
    static const bool force_simt = getenv("GN_BLOCK1_SIMT") != nullptr;
    if (m->S == kS && m->C1 == kC1 && !simt && !force_simt) return gn_launch_visbl_block1_tc(m, d_visual, n, ws, st);
    if (m->S != kS || m->C1 != kC1) {
        const size_t smem_t = sizeof(float) * kT1 * kT1 * (64 + kPatch);
        if (gn_attr_first(m->h, reinterpret_cast<const void*>(visbl_block1_tiled_kernel)))
            GN_CUDA(cudaFuncSetAttribute(visbl_block1_tiled_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_t));
        const int P1 = m->C1 - 2, tiles = (P1 + kT1P - 1) / kT1P;
        int rc = gn_launch_zero_border(m->h, ws.a1p, n, m->C1, 64, st);
        if (rc) return rc;
        const int64_t units = n * tiles * tiles, cap_t = (int64_t)m->h->num_sms * 2;
        visbl_block1_tiled_kernel<<<(int)(units < cap_t ? units : cap_t), kVThreads, smem_t, st>>>(
            d_visual, n, m->S, m->C1, tiles, m->conv1_w, m->conv1_b, m->bn_scale[0], m->bn_shift[0], ws.a1p);
        GN_LAUNCHED(m->h);
        return GN_OK;
    }
    const size_t smem = sizeof(float) * (size_t)(kC1 * kC1 * (kPatch + 64)) + sizeof(unsigned short) * 3 * kS * kS;
    if (gn_attr_first(m->h, reinterpret_cast<const void*>(visbl_block1_kernel)))
        GN_CUDA(cudaFuncSetAttribute(visbl_block1_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
    const int64_t cap = (int64_t)m->h->num_sms * 2;
    const int grid = (int)(n < cap ? n : cap);
    visbl_block1_kernel<<<grid, kVThreads, smem, st>>>(d_visual, n, m->conv1_w, m->conv1_b, m->bn_scale[0],
                                                       m->bn_shift[0], ws.a1p);
    GN_LAUNCHED(m->h);
    return GN_OK;
}

int gn_launch_audbl(gn_model* m, const float* d_audio, int64_t n, const gn_ws& ws, cudaStream_t st, bool simt) {
    if (!m->audio_included) return GN_OK;
    if (simt) {
        // validation path: the whole AudBl on CUDA cores in fp32
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
    // the reference's 30 time bins: conv1 on the tensor cores too (gn_front_tc.cu); GN_AUDBL_SIMT keeps the CUDA-core kernel
    static const bool conv1_simt = getenv("GN_AUDBL_SIMT") != nullptr;
    if (m->B == 30 && m->L1 == 15 && !conv1_simt) {
        const int rc = gn_launch_audbl_conv1_tc(m, d_audio, n, ws, st);
        if (rc) return rc;
        return gn_launch_tc_audbl(m, n, ws, st);
    }
    const size_t smem = sizeof(float) * 2 * (kCF / 2) * (m->B + 2) * 64;      // two staging buffers
    if (smem > 48 * 1024) return GN_E_SHAPE;
    const int64_t groups = (n + kCF - 1) / kCF;
    const int64_t cap = (int64_t)m->h->num_sms * 2;
    const int grid = (int)(groups < cap ? groups : cap);
    audbl_conv1_kernel<<<grid, kCThreads, smem, st>>>(d_audio, n, m->B, m->L1, m->aconv1_w, m->aconv1_b, ws.ac1);
    GN_LAUNCHED(m->h);
    return gn_launch_tc_audbl(m, n, ws, st);
}
