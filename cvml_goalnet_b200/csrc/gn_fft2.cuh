// 2048-point complex FFT for the MFCC kernel, computed by a 128-thread CTA, TWO transforms at a time in the two
// lanes of Blackwell's packed fp32 instructions (add/sub/mul/fma .f32x2 -> FADD2 / FMUL2 / FFMA2: the only way to
// reach the SM's full fp32 rate; a scalar FFMA issues every other cycle).
//
//   N = 16 x 16 x 8,  n = 128 n1 + 8 n2 + n3,  k = k1 + 16 k2 + 256 k3:
//     W_N^(nk) = W_16^(n1 k1) . W_2048^((8 n2 + n3) k1) . W_16^(n2 k2) . W_128^(n3 k2) . W_8^(n3 k3)
//   pass 1: thread t = 8 n2 + n3 holds x[t + 128 n1] (coalesced loads), DFT-16 over n1, twiddle W_2048^(t k1)
//   exchange through shared memory, layout [k1][n2][n3]
//   pass 2: thread t = 8 k1 + n3, DFT-16 over n2, twiddle W_128^(n3 k2)
//   exchange, layout [k1][k2][n3 ^ (k1 & 7)]
//   pass 3: thread u handles two units p = 16 k1 + k2 (DFT-8 over n3 -> X[k1 + 16 k2 + 256 k3]): a unit and the unit of
//           its mirror frequencies N - k, so that the power spectra of the packed real columns come straight out of
//           the registers (no spectrum round trip), stored in natural order
// Every shared-memory element is a float4 {re(A), re(B), im(A), im(B)} of the two transforms A and B, and every access
// pattern above is conflict-free for 16-byte accesses (8 lanes of a quarter-warp hit 8 different 16-byte bank groups).
//
// Everything is __host__ __device__: tests/native/fft2_host_check.cu runs the same code for the 128 "threads"
// sequentially on the CPU (packed operations emulated lane by lane) against a direct DFT in double.
#pragma once
#include <cuda_runtime.h>

#if defined(__CUDACC__)
#define GNF2_UNROLL _Pragma("unroll")
#else
#define GNF2_UNROLL
#endif
#define GNF2_HD __host__ __device__ __forceinline__

namespace gnfft2 {

constexpr int kN = 2048;
constexpr int kThreads = 128;

// two fp32 lanes: x = transform A, y = transform B
struct __align__(8) V2 { float x, y; };

GNF2_HD V2 mk(float a, float b) { V2 r; r.x = a; r.y = b; return r; }
GNF2_HD V2 bc(float a) { return mk(a, a); }

#if defined(__CUDA_ARCH__)
__device__ __forceinline__ unsigned long long as_u64(V2 a) { return *reinterpret_cast<unsigned long long*>(&a); }
__device__ __forceinline__ V2 as_v2(unsigned long long a) { return *reinterpret_cast<V2*>(&a); }
__device__ __forceinline__ V2 add2(V2 a, V2 b) {
    unsigned long long d;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(as_u64(a)), "l"(as_u64(b)));
    return as_v2(d);
}
__device__ __forceinline__ V2 sub2(V2 a, V2 b) {
    unsigned long long d;
    asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(as_u64(a)), "l"(as_u64(b)));
    return as_v2(d);
}
__device__ __forceinline__ V2 mul2(V2 a, V2 b) {
    unsigned long long d;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(as_u64(a)), "l"(as_u64(b)));
    return as_v2(d);
}
__device__ __forceinline__ V2 fma2(V2 a, V2 b, V2 c) {      // a * b + c
    unsigned long long d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(as_u64(a)), "l"(as_u64(b)), "l"(as_u64(c)));
    return as_v2(d);
}
#else
inline V2 add2(V2 a, V2 b) { return mk(a.x + b.x, a.y + b.y); }
inline V2 sub2(V2 a, V2 b) { return mk(a.x - b.x, a.y - b.y); }
inline V2 mul2(V2 a, V2 b) { return mk(a.x * b.x, a.y * b.y); }
inline V2 fma2(V2 a, V2 b, V2 c) { return mk(fmaf(a.x, b.x, c.x), fmaf(a.y, b.y, c.y)); }
#endif
GNF2_HD V2 neg2(V2 a) { return mk(-a.x, -a.y); }      // sign flips: integer pipe, not the fp32 pipe

// a complex value of both transforms
struct C2 { V2 re, im; };
GNF2_HD C2 cadd(C2 a, C2 b) { C2 r; r.re = add2(a.re, b.re); r.im = add2(a.im, b.im); return r; }
GNF2_HD C2 csub(C2 a, C2 b) { C2 r; r.re = sub2(a.re, b.re); r.im = sub2(a.im, b.im); return r; }
// (re + i im) * (wr + i wi), the factor given as broadcast pairs
GNF2_HD C2 cmul(C2 a, V2 wr, V2 wi) {
    C2 r;
    r.re = fma2(a.re, wr, mul2(a.im, neg2(wi)));
    r.im = fma2(a.re, wi, mul2(a.im, wr));
    return r;
}

// forward DFT-4, natural order in and out: 16 packed adds
GNF2_HD void dft4(C2& x0, C2& x1, C2& x2, C2& x3) {
    const C2 a0 = cadd(x0, x2), a1 = csub(x0, x2), b0 = cadd(x1, x3), d = csub(x1, x3);
    // b1 = d * (-i) = (d.im, -d.re)
    x0 = cadd(a0, b0);
    x2 = csub(a0, b0);
    x1.re = add2(a1.re, d.im); x1.im = sub2(a1.im, d.re);
    x3.re = sub2(a1.re, d.im); x3.im = add2(a1.im, d.re);
}

// v * W_16^m (m compile-time after unrolling); W_16 = exp(-2 pi i / 16)
GNF2_HD C2 mul_w16(C2 v, int m) {
    m &= 15;
    const float c1 = 0.92387953251128675613f, s1 = 0.38268343236508977173f, h = 0.70710678118654752440f;
    switch (m) {
        case 0: return v;
        case 4: { C2 r; r.re = v.im; r.im = neg2(v.re); return r; }                  // -i
        case 8: { C2 r; r.re = neg2(v.re); r.im = neg2(v.im); return r; }            // -1
        case 12: { C2 r; r.re = neg2(v.im); r.im = v.re; return r; }                 // +i
        case 2: { C2 r; r.re = mul2(add2(v.re, v.im), bc(h)); r.im = mul2(sub2(v.im, v.re), bc(h)); return r; }      // (1 - i)/sqrt2
        case 6: { C2 r; r.re = mul2(sub2(v.im, v.re), bc(h)); r.im = mul2(add2(v.re, v.im), bc(-h)); return r; }     // (-1 - i)/sqrt2
        case 10: { C2 r; r.re = mul2(add2(v.re, v.im), bc(-h)); r.im = mul2(sub2(v.re, v.im), bc(h)); return r; }    // (-1 + i)/sqrt2
        case 14: { C2 r; r.re = mul2(sub2(v.re, v.im), bc(h)); r.im = mul2(add2(v.re, v.im), bc(h)); return r; }     // (1 + i)/sqrt2
        case 1: return cmul(v, bc(c1), bc(-s1));
        case 3: return cmul(v, bc(s1), bc(-c1));
        case 5: return cmul(v, bc(-s1), bc(-c1));
        case 7: return cmul(v, bc(-c1), bc(-s1));
        case 9: return cmul(v, bc(-c1), bc(s1));
        case 11: return cmul(v, bc(-s1), bc(c1));
        case 13: return cmul(v, bc(s1), bc(c1));
        default: return cmul(v, bc(c1), bc(s1));   // 15
    }
}

// In-register DFT-16 (n = 4a + b, k = c + 4d).  Output is permuted: z[4c + d] = X[c + 4d].
GNF2_HD constexpr int perm16(int p) { return (p >> 2) + 4 * (p & 3); }      // index in z -> frequency
GNF2_HD void dft16(C2 (&z)[16]) {
    GNF2_UNROLL
    for (int b = 0; b < 4; ++b) {
        dft4(z[b], z[4 + b], z[8 + b], z[12 + b]);                           // over a: z[4c + b] = t[b][c]
        GNF2_UNROLL
        for (int c = 1; c < 4; ++c) z[4 * c + b] = mul_w16(z[4 * c + b], b * c);
    }
    GNF2_UNROLL
    for (int c = 0; c < 4; ++c) dft4(z[4 * c], z[4 * c + 1], z[4 * c + 2], z[4 * c + 3]);   // over b: z[4c + d]
}

// In-register DFT-8, natural order in and out
GNF2_HD void dft8(C2 (&v)[8]) {
    C2 e0 = v[0], e1 = v[2], e2 = v[4], e3 = v[6];
    C2 o0 = v[1], o1 = v[3], o2 = v[5], o3 = v[7];
    dft4(e0, e1, e2, e3);
    dft4(o0, o1, o2, o3);
    o1 = mul_w16(o1, 2);
    o2 = mul_w16(o2, 4);
    o3 = mul_w16(o3, 6);
    v[0] = cadd(e0, o0); v[4] = csub(e0, o0);
    v[1] = cadd(e1, o1); v[5] = csub(e1, o1);
    v[2] = cadd(e2, o2); v[6] = csub(e2, o2);
    v[3] = cadd(e3, o3); v[7] = csub(e3, o3);
}

GNF2_HD float4 pack4(C2 v) { return make_float4(v.re.x, v.re.y, v.im.x, v.im.y); }
GNF2_HD C2 unpack4(float4 f) { C2 v; v.re = mk(f.x, f.y); v.im = mk(f.z, f.w); return v; }

// ---- tables (built on the host in double) ------------------------------------------------------------------------
//   tw1[(k1 - 1) * 128 + t] = {wr, wi} of W_2048^(t k1),  k1 = 1..15
//   tw2[(k2 - 1) * 8 + n3]  = {wr, wi} of W_128^(n3 k2),  k2 = 1..15
// (scalars: the packed instructions broadcast a 32-bit operand to both lanes for free)
// A thread reads only the rows k = 1, 2, 4, 8 (four loads, issued before the butterflies that hide their latency) and
// forms the other eleven powers as products of two of them, at most three multiplications deep: fifteen dependent table
// loads per pass were the kernel's largest single stall.
constexpr int kTw1 = 15 * 128, kTw2 = 15 * 8;

struct TwBase { V2 w1, w2, w4, w8; };      // (re, im) pairs of w^1, w^2, w^4, w^8
GNF2_HD TwBase tw_load(const float2* tw, int stride, int i) {
    const float2 a = tw[i], b = tw[stride + i], c = tw[3 * stride + i], d = tw[7 * stride + i];
    TwBase r;
    r.w1 = mk(a.x, a.y); r.w2 = mk(b.x, b.y); r.w4 = mk(c.x, c.y); r.w8 = mk(d.x, d.y);
    return r;
}
// a * b for (re, im) pairs; b_rot = (-b.im, b.re)
GNF2_HD V2 cpair_mul(V2 a, V2 b, V2 b_rot) { return fma2(bc(a.x), b, mul2(bc(a.y), b_rot)); }
// w^k, k = 1..15 (k is a compile-time constant after unrolling; common sub-products are merged by the compiler)
GNF2_HD V2 tw_power(const TwBase& b, int k) {
    const V2 r2 = mk(-b.w2.y, b.w2.x), r4 = mk(-b.w4.y, b.w4.x), r8 = mk(-b.w8.y, b.w8.x);
    V2 lo;                                  // w^(k & 7)
    switch (k & 7) {
        case 1: lo = b.w1; break;
        case 2: lo = b.w2; break;
        case 3: lo = cpair_mul(b.w1, b.w2, r2); break;
        case 4: lo = b.w4; break;
        case 5: lo = cpair_mul(b.w1, b.w4, r4); break;
        case 6: lo = cpair_mul(b.w2, b.w4, r4); break;
        case 7: lo = cpair_mul(cpair_mul(b.w1, b.w2, r2), b.w4, r4); break;
        default: return b.w8;               // k == 8
    }
    return (k & 8) ? cpair_mul(lo, b.w8, r8) : lo;
}

// pass 1 for thread t: z[n1] = x[t + 128 n1] on entry; results to the exchange buffer
GNF2_HD void pass1(int t, C2 (&z)[16], const float2* tw1, float4* xbuf) {
    const TwBase tb = tw_load(tw1, 128, t);
    dft16(z);
    const int n2 = t >> 3, n3 = t & 7;
    GNF2_UNROLL
    for (int p = 0; p < 16; ++p) {
        const int k1 = perm16(p);
        C2 v = z[p];
        if (k1 != 0) {
            const V2 w = tw_power(tb, k1);
            v = cmul(v, bc(w.x), bc(w.y));
        }
        xbuf[(k1 * 16 + n2) * 8 + n3] = pack4(v);
    }
}

// pass 2 for thread t = 8 k1 + n3
GNF2_HD void pass2(int t, const float2* tw2, float4* xbuf_in, float4* xbuf_out) {
    const int k1 = t >> 3, n3 = t & 7;
    const TwBase tb = tw_load(tw2, 8, n3);
    C2 z[16];
    GNF2_UNROLL
    for (int n2 = 0; n2 < 16; ++n2) z[n2] = unpack4(xbuf_in[(k1 * 16 + n2) * 8 + n3]);
    dft16(z);
    GNF2_UNROLL
    for (int p = 0; p < 16; ++p) {
        const int k2 = perm16(p);
        C2 v = z[p];
        if (k2 != 0) {
            const V2 w = tw_power(tb, k2);
            v = cmul(v, bc(w.x), bc(w.y));
        }
        xbuf_out[(k1 * 16 + k2) * 8 + (n3 ^ (k1 & 7))] = pack4(v);
    }
}

// ---- pass 3, fused with the split of the packed real columns ---------------------------------------------------
// Thread u transforms TWO units p = 16 k1 + k2 (DFT-8 over n3 -> X[k1 + 16 k2 + 256 k3]): unit pa and the unit pb that
// holds the mirror frequencies, N - (ka + 256 k3) = kb + 256 (7 - k3) with ka = k1 + 16 k2 of pa and kb of pb.  Both
// halves of every pair (Z[k], Z[N-k]) are then in the thread's registers, and the power spectra of the real columns
//   X1 = (Z[k] + conj Z[N-k]) / 2,  X2 = (Z[k] - conj Z[N-k]) / (2i)
// are formed without writing the spectrum to shared memory and reading it back.
// With u = 8 k2 + j the eight lanes of a quarter-warp hold eight different k1 & 7 in either half (so the reads, swizzled
// by k1 & 7, are conflict-free) and store eight consecutive bins (so the power spectrum needs no swizzle):
//   j = 1..7          : pa = (k1 = j, k2),       pb = (16 - j, 15 - k2)
//   j = 0, k2 = 0..7  : pa = (8, k2),            pb = (8, 15 - k2)
//   j = 0, k2 = 9..15 : pa = (0, k2 - 8),        pb = (0, 24 - k2)
//   j = 0, k2 = 8     : the two self-mirrored units (0, 0) and (0, 8): k = 256 k3 and k = 128 + 256 k3
constexpr int kSelfMirrorThread = 64;
GNF2_HD void pass3_units(int u, int& pa, int& pb) {
    const int k2 = u >> 3, j = u & 7;
    if (j != 0) { pa = 16 * j + k2; pb = 16 * (16 - j) + (15 - k2); }
    else if (k2 < 8) { pa = 128 + k2; pb = 143 - k2; }
    else if (k2 == 8) { pa = 0; pb = 8; }
    else { pa = k2 - 8; pb = 24 - k2; }
}

GNF2_HD void pass3_load(int u, const float4* xbuf_in, C2 (&va)[8], C2 (&vb)[8]) {
    int pa, pb;
    pass3_units(u, pa, pb);
    const int sa = (pa >> 4) & 7, sb = (pb >> 4) & 7;
    GNF2_UNROLL
    for (int n3 = 0; n3 < 8; ++n3) {
        va[n3] = unpack4(xbuf_in[pa * 8 + (n3 ^ sa)]);
        vb[n3] = unpack4(xbuf_in[pb * 8 + (n3 ^ sb)]);
    }
}

// {|X1|^2 (A), |X1|^2 (B), |X2|^2 (A), |X2|^2 (B)} from zk = Z[k], zn = Z[N-k] (symmetric in the two: bin k == bin N-k)
GNF2_HD float4 power_of(C2 zk, C2 zn) {
    const V2 ar = add2(zk.re, zn.re), ai = sub2(zk.im, zn.im), br = add2(zk.im, zn.im), bi = sub2(zk.re, zn.re);
    const V2 q = bc(0.25f);
    const V2 p1 = mul2(fma2(ar, ar, mul2(ai, ai)), q), p2 = mul2(fma2(br, br, mul2(bi, bi)), q);
    return make_float4(p1.x, p1.y, p2.x, p2.y);
}

GNF2_HD void pass3_power(int u, C2 (&va)[8], C2 (&vb)[8], float4* pw) {
    int pa, pb;
    pass3_units(u, pa, pb);
    dft8(va);
    dft8(vb);
    const int ka = (pa >> 4) + 16 * (pa & 15), kb = (pb >> 4) + 16 * (pb & 15);
    if (u != kSelfMirrorThread) {
        GNF2_UNROLL
        for (int i = 0; i < 8; ++i) {
            const int bin = (i < 4) ? ka + 256 * i : kb + 256 * (7 - i);
            pw[bin] = power_of(va[i], vb[7 - i]);
        }
    } else {
        pw[0] = power_of(va[0], va[0]);
        pw[256] = power_of(va[1], va[7]);
        pw[512] = power_of(va[2], va[6]);
        pw[768] = power_of(va[3], va[5]);
        pw[1024] = power_of(va[4], va[4]);
        pw[128] = power_of(vb[0], vb[7]);
        pw[384] = power_of(vb[1], vb[6]);
        pw[640] = power_of(vb[2], vb[5]);
        pw[896] = power_of(vb[3], vb[4]);
    }
}

}  // namespace gnfft2
