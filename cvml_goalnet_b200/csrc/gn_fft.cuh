// 2048-point complex FFT as four Stockham (autosort) passes of radix 4, 8, 8, 8 over split
// real/imaginary shared-memory arrays.  256 threads cooperate on one transform; every thread keeps
// its 8 (or 2 x 4) points in registers for the butterfly, so a transform costs four shared-memory
// round trips instead of eleven radix-2 ones.  The intermediate buffers use per-pass index padding
// that makes every strided write bank-conflict free (32-bit accesses, one element per bank):
//   pass 1 (R=4, Ns=1)   writes idx -> idx + idx/32          (stride-4 writes)
//   pass 2 (R=8, Ns=4)   writes idx -> idx + 4*(idx/32)      (4 consecutive, then stride 32)
//   pass 3 (R=8, Ns=32)  and pass 4 (R=8, Ns=256) write unpadded (warp-contiguous).
// The functions are __host__ __device__ so that tests can run the same index/twiddle logic on the CPU.
#pragma once
#include <cuda_runtime.h>

namespace gnfft {

constexpr int kN = 2048;
constexpr int kThreads = 256;
constexpr int kBufA = kN + kN / 32;        // pass-1 padding
constexpr int kBufB = kN + 4 * (kN / 32);  // pass-2 padding
// twiddle table layout (float2): [0,256) exp(-2 pi i k/2048), [256,288) exp(-2 pi i k/256), [288,292) exp(-2 pi i k/32)
constexpr int kTw2048 = 0, kTw256 = 256, kTw32 = 288, kTwCount = 292;

__host__ __device__ __forceinline__ float2 cmul(float2 a, float2 b) {
    return make_float2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x);
}
__host__ __device__ __forceinline__ float2 cadd(float2 a, float2 b) { return make_float2(a.x + b.x, a.y + b.y); }
__host__ __device__ __forceinline__ float2 csub(float2 a, float2 b) { return make_float2(a.x - b.x, a.y - b.y); }
__host__ __device__ __forceinline__ float2 mul_mi(float2 a) { return make_float2(a.y, -a.x); }   // a * (-i)

// forward DFT-4, natural order in and out
__host__ __device__ __forceinline__ void dft4(float2& x0, float2& x1, float2& x2, float2& x3) {
    const float2 a0 = cadd(x0, x2), a1 = csub(x0, x2);
    const float2 b0 = cadd(x1, x3), b1 = mul_mi(csub(x1, x3));
    x0 = cadd(a0, b0);
    x2 = csub(a0, b0);
    x1 = cadd(a1, b1);
    x3 = csub(a1, b1);
}
// forward DFT-8, natural order in and out
__host__ __device__ __forceinline__ void dft8(float2 (&v)[8]) {
    float2 e0 = v[0], e1 = v[2], e2 = v[4], e3 = v[6];
    float2 o0 = v[1], o1 = v[3], o2 = v[5], o3 = v[7];
    dft4(e0, e1, e2, e3);
    dft4(o0, o1, o2, o3);
    const float s = 0.70710678118654752440f;
    o1 = make_float2(s * (o1.x + o1.y), s * (o1.y - o1.x));     // * exp(-i pi/4)
    o2 = mul_mi(o2);                                            // * exp(-i pi/2)
    o3 = make_float2(s * (o3.y - o3.x), -s * (o3.x + o3.y));    // * exp(-3 i pi/4)
    v[0] = cadd(e0, o0); v[4] = csub(e0, o0);
    v[1] = cadd(e1, o1); v[5] = csub(e1, o1);
    v[2] = cadd(e2, o2); v[6] = csub(e2, o2);
    v[3] = cadd(e3, o3); v[7] = csub(e3, o3);
}

__host__ __device__ __forceinline__ int pad1(int a) { return a + (a >> 5); }
__host__ __device__ __forceinline__ int pad2(int a) { return a + ((a >> 5) << 2); }

// pass 1: radix 4, Ns = 1 (no twiddles).  Butterfly j in [0,512) takes x[j + 512 r]; this thread does
// j = tid and j = tid + 256, i.e. input points z[tid + 256 s], s = 0..7 (z[s] below).
__host__ __device__ __forceinline__ void pass1(int tid, const float2 (&z)[8], float* are, float* aim) {
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        float2 x0 = z[h], x1 = z[h + 2], x2 = z[h + 4], x3 = z[h + 6];
        dft4(x0, x1, x2, x3);
        const int j0 = (tid + 256 * h) * 4;
        are[pad1(j0)] = x0.x;     aim[pad1(j0)] = x0.y;
        are[pad1(j0 + 1)] = x1.x; aim[pad1(j0 + 1)] = x1.y;
        are[pad1(j0 + 2)] = x2.x; aim[pad1(j0 + 2)] = x2.y;
        are[pad1(j0 + 3)] = x3.x; aim[pad1(j0 + 3)] = x3.y;
    }
}

// generic radix-8 Stockham pass: v[r] = in[j + 256 r] * w^r, DFT-8, out[j0 + r Ns]
template <int NS, int PAD_IN, int PAD_OUT>
__host__ __device__ __forceinline__ void pass8(int j, const float* ire, const float* iim, float* ore, float* oim,
                                               const float2* tw /* table for this pass, index k = j % NS */) {
    float2 v[8];
#pragma unroll
    for (int r = 0; r < 8; ++r) {
        const int a = j + 256 * r;
        const int ia = PAD_IN == 1 ? pad1(a) : (PAD_IN == 2 ? pad2(a) : a);
        v[r] = make_float2(ire[ia], iim[ia]);
    }
    const int k = j & (NS - 1);
    const float2 w1 = tw[k];
    float2 w = w1;
    v[1] = cmul(v[1], w);
#pragma unroll
    for (int r = 2; r < 8; ++r) {
        w = cmul(w, w1);
        v[r] = cmul(v[r], w);
    }
    dft8(v);
    const int j0 = (j / NS) * (NS * 8) + k;
#pragma unroll
    for (int r = 0; r < 8; ++r) {
        const int a = j0 + r * NS;
        const int oa = PAD_OUT == 1 ? pad1(a) : (PAD_OUT == 2 ? pad2(a) : a);
        ore[oa] = v[r].x;
        oim[oa] = v[r].y;
    }
}

}  // namespace gnfft
