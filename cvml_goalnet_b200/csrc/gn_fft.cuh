// 2048-point complex FFT computed by ONE WARP with the points in registers: 2048 = 64 x 32.
//   stage 1: lane n2 holds x[32 n1 + n2], n1 = 0..63, runs a 64-point DFT over n1 in registers (8 x 8,
//            all indices compile-time) and multiplies Y[k1] by W_2048^(n2 k1);
//   transpose through a per-warp shared-memory tile T[k1][n2] (rows padded to 33 float2, so both the
//            row-wise writes and the column-wise reads are bank-conflict free);
//   stage 2: lane l reads the two rows k1 = l and k1 = 64 - l (lane 0: rows 0 and 32) and runs two
//            32-point DFTs over n2 in registers (8 x 4), giving X[k1 + 64 k2], k2 = 0..31.
// The row pairing puts X[k] and X[2048 - k] in the SAME lane, so the two real spectra packed into one
// complex transform (column j in the real part, column j+1 in the imaginary part) are separated and
// squared without any further exchange.  One shared-memory round trip per transform, no block barrier.
// Everything is __host__ __device__ so that tests/native/fft_host_check.cu runs the same code on the CPU.
#pragma once
#include <cuda_runtime.h>

#if defined(__CUDACC__)
#define GNFFT_UNROLL _Pragma("unroll")
#else
#define GNFFT_UNROLL
#endif

namespace gnfft {

constexpr int kN = 2048;
constexpr int kRow = 33;                    // padded row length of the transpose tile (float2)
constexpr int kTile = 64 * kRow;            // float2 per warp
constexpr int kTwLane = 64 * 32;            // twiddle table tw[k1 * 32 + n2] = W_2048^(n2 k1)

__host__ __device__ __forceinline__ float2 cmul(float2 a, float2 b) {
    return make_float2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x);
}
__host__ __device__ __forceinline__ float2 cadd(float2 a, float2 b) { return make_float2(a.x + b.x, a.y + b.y); }
__host__ __device__ __forceinline__ float2 csub(float2 a, float2 b) { return make_float2(a.x - b.x, a.y - b.y); }
__host__ __device__ __forceinline__ float2 mul_mi(float2 a) { return make_float2(a.y, -a.x); }   // a * (-i)

// W_64^m = exp(-2 pi i m / 64); m is a compile-time constant after unrolling, so this folds to immediates
__host__ __device__ __forceinline__ float2 w64(int m) {
    m &= 63;
    const int q = m >> 4, r = m & 15;
    float c = 1.f, s = 0.f;                 // cos, sin of 2 pi r / 64
    switch (r) {
        case 1: c = 0.9951847266721969f; s = 0.0980171403295606f; break;
        case 2: c = 0.9807852804032304f; s = 0.19509032201612825f; break;
        case 3: c = 0.9569403357322088f; s = 0.29028467725446233f; break;
        case 4: c = 0.9238795325112867f; s = 0.3826834323650898f; break;
        case 5: c = 0.881921264348355f; s = 0.47139673682599764f; break;
        case 6: c = 0.8314696123025452f; s = 0.5555702330196022f; break;
        case 7: c = 0.773010453362737f; s = 0.6343932841636455f; break;
        case 8: c = 0.7071067811865476f; s = 0.7071067811865476f; break;
        case 9: c = 0.6343932841636455f; s = 0.773010453362737f; break;
        case 10: c = 0.5555702330196023f; s = 0.8314696123025452f; break;
        case 11: c = 0.4713967368259978f; s = 0.8819212643483549f; break;
        case 12: c = 0.38268343236508984f; s = 0.9238795325112867f; break;
        case 13: c = 0.29028467725446233f; s = 0.9569403357322089f; break;
        case 14: c = 0.19509032201612833f; s = 0.9807852804032304f; break;
        case 15: c = 0.09801714032956077f; s = 0.9951847266721968f; break;
        default: break;
    }
    // (c, -s) times (-i)^q
    if (q == 0) return make_float2(c, -s);
    if (q == 1) return make_float2(-s, -c);
    if (q == 2) return make_float2(-c, s);
    return make_float2(s, c);
}

// v * W_64^m with the trivial factors special-cased (m compile-time)
__host__ __device__ __forceinline__ float2 mul_w64(float2 v, int m) {
    m &= 63;
    if (m == 0) return v;
    if (m == 16) return mul_mi(v);
    if (m == 32) return make_float2(-v.x, -v.y);
    if (m == 48) return make_float2(-v.y, v.x);
    if (m == 8) { const float s = 0.70710678118654752440f; return make_float2(s * (v.x + v.y), s * (v.y - v.x)); }
    if (m == 24) { const float s = 0.70710678118654752440f; return make_float2(s * (v.y - v.x), -s * (v.x + v.y)); }
    return cmul(v, w64(m));
}

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

// In-register 64-point DFT (n = 8a + b, k = c + 8d).  Output is permuted: z[8c + d] = Y[c + 8d].
__host__ __device__ __forceinline__ constexpr int perm64(int p) { return (p >> 3) + 8 * (p & 7); }
__host__ __device__ __forceinline__ void dft64(float2 (&z)[64]) {
    GNFFT_UNROLL
    for (int b = 0; b < 8; ++b) {
        float2 v[8];
        GNFFT_UNROLL
        for (int a = 0; a < 8; ++a) v[a] = z[8 * a + b];
        dft8(v);
        GNFFT_UNROLL
        for (int c = 0; c < 8; ++c) z[8 * c + b] = mul_w64(v[c], b * c);      // t[b][c] * W_64^(bc), kept at z[8c + b]
    }
    GNFFT_UNROLL
    for (int c = 0; c < 8; ++c) {
        float2 v[8];
        GNFFT_UNROLL
        for (int b = 0; b < 8; ++b) v[b] = z[8 * c + b];
        dft8(v);
        GNFFT_UNROLL
        for (int d = 0; d < 8; ++d) z[8 * c + d] = v[d];
    }
}

// In-register 32-point DFT (n = 4a + b, k = c + 8d).  Output is permuted: v[4c + d] = X[c + 8d].
__host__ __device__ __forceinline__ constexpr int perm32(int p) { return (p >> 2) + 8 * (p & 3); }
__host__ __device__ __forceinline__ constexpr int inv32(int k) { return 4 * (k & 7) + (k >> 3); }
__host__ __device__ __forceinline__ void dft32(float2 (&x)[32]) {
    GNFFT_UNROLL
    for (int b = 0; b < 4; ++b) {
        float2 v[8];
        GNFFT_UNROLL
        for (int a = 0; a < 8; ++a) v[a] = x[4 * a + b];
        dft8(v);
        GNFFT_UNROLL
        for (int c = 0; c < 8; ++c) x[4 * c + b] = mul_w64(v[c], 2 * b * c);  // W_32^(bc) = W_64^(2bc)
    }
    GNFFT_UNROLL
    for (int c = 0; c < 8; ++c) dft4(x[4 * c], x[4 * c + 1], x[4 * c + 2], x[4 * c + 3]);
}

// Stage 1 for lane n2: z[n1] = x[32 n1 + n2] on entry.  tw = the lane-major twiddle table (kTwLane entries).
__host__ __device__ __forceinline__ void warp_fft_stage1(int lane, float2 (&z)[64], const float2* tw, float2* tile) {
    dft64(z);
    GNFFT_UNROLL
    for (int p = 0; p < 64; ++p) {
        const int k1 = perm64(p);
        tile[k1 * kRow + lane] = k1 == 0 ? z[p] : cmul(z[p], tw[k1 * 32 + lane]);
    }
}

// rows of the transpose tile that lane l transforms in stage 2
__host__ __device__ __forceinline__ int stage2_row_a(int lane) { return lane; }
__host__ __device__ __forceinline__ int stage2_row_b(int lane) { return lane == 0 ? 32 : 64 - lane; }

// Stage 2 for lane l: a[inv32(k2)] = X[rowA + 64 k2], b[inv32(k2)] = X[rowB + 64 k2].
__host__ __device__ __forceinline__ void warp_fft_stage2(int lane, const float2* tile, float2 (&a)[32], float2 (&b)[32]) {
    const float2* ra = tile + stage2_row_a(lane) * kRow;
    const float2* rb = tile + stage2_row_b(lane) * kRow;
    GNFFT_UNROLL
    for (int n = 0; n < 32; ++n) { a[n] = ra[n]; b[n] = rb[n]; }
    dft32(a);
    dft32(b);
}

// Power spectra of the two real sequences packed as z = x1 + i x2, from Z[k] and Z[N - k]:
//   X1 = (Z[k] + conj Z[N-k]) / 2,  X2 = (Z[k] - conj Z[N-k]) / (2i)
__host__ __device__ __forceinline__ void split_power(float2 zk, float2 zn, float& p1, float& p2) {
    const float ar = zk.x + zn.x, ai = zk.y - zn.y;
    const float br = zk.y + zn.y, bi = zk.x - zn.x;
    p1 = 0.25f * (ar * ar + ai * ai);
    p2 = 0.25f * (br * br + bi * bi);
}

// After stage 2: write |X1[k]|^2 -> pw1[k], |X2[k]|^2 -> pw2[k] for this lane's bins k <= 1024.
__host__ __device__ __forceinline__ void warp_fft_power(int lane, const float2 (&a)[32], const float2 (&b)[32],
                                                        float* pw1, float* pw2) {
    if (lane != 0) {
        // a = row l, b = row 64 - l; X[N - (l + 64 k2)] = b[31 - k2] and vice versa
        GNFFT_UNROLL
        for (int k2 = 0; k2 < 16; ++k2) {
            const int ka = lane + 64 * k2, kb = 64 - lane + 64 * k2;
            split_power(a[inv32(k2)], b[inv32(31 - k2)], pw1[ka], pw2[ka]);
            split_power(b[inv32(k2)], a[inv32(31 - k2)], pw1[kb], pw2[kb]);
        }
    } else {
        // a = row 0 (self-paired with k2 -> 32 - k2), b = row 32 (self-paired with k2 -> 31 - k2)
        GNFFT_UNROLL
        for (int k2 = 0; k2 <= 16; ++k2) split_power(a[inv32(k2)], a[inv32((32 - k2) & 31)], pw1[64 * k2], pw2[64 * k2]);
        GNFFT_UNROLL
        for (int k2 = 0; k2 < 16; ++k2)
            split_power(b[inv32(k2)], b[inv32(31 - k2)], pw1[32 + 64 * k2], pw2[32 + 64 * k2]);
    }
}

}  // namespace gnfft
