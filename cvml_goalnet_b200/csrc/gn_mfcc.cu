// (a2) audio features: per frame-slot STFT(2048/512, hann, centred) -> |.|^2 -> 128 Slaney mels ->
// 10 log10 with an 80 dB floor under the segment max -> DCT-II ortho (30 rows) -> cubic time
// resampling W (B x q).  Replaces reference utils.py:322-349 (librosa.feature.mfcc + interp1d loop).
//
// One 128-thread CTA per slot, several CTAs per SM (so that one CTA's barriers and latencies hide under
// another's arithmetic).  The transform is the stage's cost (8 complex FFT-2048 per slot: ~1 MFLOP of
// 1.1): FOUR STFT columns go through the butterflies at once -- two real columns share one complex
// transform (column j in the real part, j+1 in the imaginary part), and two such transforms ride in
// the two lanes of the packed fp32 instructions (gn_fft2.cuh: FADD2 / FMUL2 / FFMA2).  The samples of
// four consecutive columns overlap (hop 512 = 4 rows of 128), so a thread loads 28 values instead of
// 64.  One 32 KB shared-memory buffer serves the two exchanges, the spectrum and the power spectra;
// the tables (twiddles, window, mel weights, DCT rows, W) are read through L1.  HBM traffic is the
// slot's samples in and 30*B floats out.  Every sum is formed in a fixed order: a slot's features do
// not depend on the launch it is part of.
#include <type_traits>

#include "gn_common.cuh"
#include "gn_fft2.cuh"

namespace {

using namespace gnfft2;

constexpr int kFFT = 2048;
constexpr int kHop = 512;
constexpr int kMels = 128;
constexpr int kMfcc = 30;
constexpr int kT = 128;                     // threads per CTA == gnfft2::kThreads
constexpr int kXBytes = kFFT * 16;          // exchange / power buffer (8192 floats); in a slot's tail:
constexpr int kDctCols = 32;                //   DCT partial sums [4 warps][kDctCols][32] = the first 4096 floats,
constexpr int kDctPart = 4 * kDctCols * 32; //   then the output rows [30][B] over them,
constexpr int kWStage = kFFT * 4 - kDctPart;    //   and the staged resampling operator W [B][q] in the other 4096

// a sample is read once: it bypasses L1, which then keeps the tables (twiddles, window, DCT rows, W)
__device__ __forceinline__ float ld_stream(const float* p) {
    float v;
    asm volatile("ld.global.nc.L1::no_allocate.f32 %0, [%1];" : "=f"(v) : "l"(p));
    return v;
}

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

__global__ void __launch_bounds__(kT, 4)
mfcc_kernel(const float* __restrict__ y, const int64_t* __restrict__ seg_start, const int32_t* __restrict__ seg_len,
            const int32_t* __restrict__ w_offset, const float* __restrict__ w_all, int64_t n, int B, int mq,
            const int32_t* __restrict__ mel_start, const int32_t* __restrict__ mel_count,
            const int32_t* __restrict__ mel_offset, const float* __restrict__ mel_w, int n_mel_w,
            const float* __restrict__ dctT /*[128][32]*/, const float2* __restrict__ tw1, const float2* __restrict__ tw2,
            const float* __restrict__ hann, float* __restrict__ out) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float4* xb = reinterpret_cast<float4*>(smem_raw);                      // [2048] exchange / spectrum / power
    float* mel = reinterpret_cast<float*>(smem_raw + kXBytes);             // [128][mq]  (mq = max_q rounded up to 4)
    float* cepT = mel + kMels * mq;                                        // [mq][32]   cepstra, transposed (row = column j)
    float* mws = cepT + 32 * mq;                                           // [n_mel_w]  mel filter weights (CSR order)
    __shared__ float red[kT / 32];

    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    // this thread's mel band: contiguous bins [mst, mst + mcn), weights mw[0..mcn); the weights are staged once per CTA
    const int mst = __ldg(mel_start + t), mcn = __ldg(mel_count + t);
    const float* mw = mws + __ldg(mel_offset + t);
    const int mrot = (mcn > 0) ? (((t & 7) - mst) & 7) % mcn : 0;
    for (int i = t; i < n_mel_w; i += kT) mws[i] = __ldg(mel_w + i);
    __syncthreads();

    for (int64_t slot = blockIdx.x; slot < n; slot += gridDim.x) {
        const int64_t s0 = seg_start[slot];
        const int L = seg_len[slot];
        const int q = 1 + L / kHop;
        const float* seg = y + s0;
        float emax = 0.f;                                      // largest mel energy seen by this thread
        // sample rows of 128: row m of a group holds samples j0*512 - 1024 + 128 m + t; column j0+c starts at row 4c.
        // Samples outside [0, L) are the centre zero padding; interior groups skip the bounds checks.  Rows 16..27 of a
        // group are rows 0..11 of the next one (hop 512 = 4 rows), so a group only loads 16 new rows, and those are
        // requested while the current group is between its passes (they are only consumed one group later).
        float raw[28];
        auto load_rows = [&](int j0, auto first) {
            constexpr int kFirst = decltype(first)::value;
            const int base = j0 * kHop - kFFT / 2 + t;
            if (base - t + kFirst * 128 >= 0 && base - t + 28 * 128 <= L) {
                const float* p = seg + base;
#pragma unroll
                for (int m = kFirst; m < 28; ++m) raw[m] = ld_stream(p + 128 * m);
            } else {
#pragma unroll
                for (int m = kFirst; m < 28; ++m) {
                    const int i = base + 128 * m;
                    raw[m] = ((unsigned)i < (unsigned)L) ? ld_stream(seg + i) : 0.f;
                }
            }
        };
        load_rows(0, std::integral_constant<int, 0>{});
        for (int j0 = 0; j0 < q; j0 += 4) {
            // ---- windowed columns j0..j0+3: transform A = (j0, j0+1), transform B = (j0+2, j0+3) ----
            {
                C2 z[16];
#pragma unroll
                for (int n1 = 0; n1 < 16; ++n1) {
                    const V2 h = bc(__ldg(hann + t + 128 * n1));
                    z[n1].re = mul2(mk(raw[n1], raw[n1 + 8]), h);
                    z[n1].im = mul2(mk(raw[n1 + 4], raw[n1 + 12]), h);
                }
                if (j0 + 3 >= q) {                             // the last group of a slot: columns >= q count as silence
                    const V2 mre = mk(1.f, (j0 + 2 < q) ? 1.f : 0.f);
                    const V2 mim = mk((j0 + 1 < q) ? 1.f : 0.f, 0.f);
#pragma unroll
                    for (int n1 = 0; n1 < 16; ++n1) {
                        z[n1].re = mul2(z[n1].re, mre);
                        z[n1].im = mul2(z[n1].im, mim);
                    }
                }
                if (j0 + 4 < q) {
#pragma unroll
                    for (int m = 0; m < 12; ++m) raw[m] = raw[m + 16];
                    load_rows(j0 + 4, std::integral_constant<int, 12>{});
                }
                pass1(t, z, tw1, xb);                          // (the previous group's readers left at the barrier below)
            }
            __syncthreads();
            {
                // pass 2 in place: a quarter-warp (same k1) reads and writes the same 128 elements
                const int k1 = t >> 3, n3 = t & 7;
                const TwBase tb = tw_load(tw2, 8, n3);
                C2 z[16];
#pragma unroll
                for (int n2 = 0; n2 < 16; ++n2) z[n2] = unpack4(xb[(k1 * 16 + n2) * 8 + n3]);
                __syncwarp();
                dft16(z);
#pragma unroll
                for (int p = 0; p < 16; ++p) {
                    const int k2 = perm16(p);
                    C2 v = z[p];
                    if (k2 != 0) {
                        const V2 w = tw_power(tb, k2);
                        v = cmul(v, bc(w.x), bc(w.y));
                    }
                    xb[(k1 * 16 + k2) * 8 + (n3 ^ (k1 & 7))] = pack4(v);
                }
            }
            __syncthreads();
            {
                // pass 3 on a unit and its mirror unit, fused with the split into the four real columns' power spectra:
                // all reads, barrier, then pw[bin] = {col j0, col j0+2, col j0+1, col j0+3} over the same buffer
                C2 va[8], vb[8];
                pass3_load(t, xb, va, vb);
                __syncthreads();
                pass3_power(t, va, vb, xb);
            }
            __syncthreads();
            {
                // mel filterbank: thread = band (its bins are contiguous); four independent accumulator pairs.  A band is
                // walked from its bin mrot round to mrot - 1, mrot chosen so that the lanes of a quarter-warp start on
                // eight different 16-byte slots and move on in step: the 128-bit reads do not collide.
                V2 a02 = bc(0.f), a13 = bc(0.f), b02 = bc(0.f), b13 = bc(0.f);
                int i = 0, ix = mrot;
                for (; i + 1 < mcn; i += 2) {
                    int ix1 = ix + 1;
                    ix1 = (ix1 >= mcn) ? 0 : ix1;
                    const float w0 = mw[ix], w1 = mw[ix1];
                    const float4 p0 = xb[mst + ix], p1 = xb[mst + ix1];
                    a02 = fma2(bc(w0), mk(p0.x, p0.y), a02);
                    a13 = fma2(bc(w0), mk(p0.z, p0.w), a13);
                    b02 = fma2(bc(w1), mk(p1.x, p1.y), b02);
                    b13 = fma2(bc(w1), mk(p1.z, p1.w), b13);
                    ix = ix1 + 1;
                    ix = (ix >= mcn) ? 0 : ix;
                }
                if (i < mcn) {
                    const float w0 = mw[ix];
                    const float4 p0 = xb[mst + ix];
                    a02 = fma2(bc(w0), mk(p0.x, p0.y), a02);
                    a13 = fma2(bc(w0), mk(p0.z, p0.w), a13);
                }
                a02 = add2(a02, b02);
                a13 = add2(a13, b13);
                // columns j0, j0+1, j0+2, j0+3 (columns >= q were transformed as zeros: 0 energy, within the padded row)
                *reinterpret_cast<float4*>(mel + t * mq + j0) = make_float4(a02.x, a13.x, a02.y, a13.y);
                emax = fmaxf(emax, fmaxf(fmaxf(a02.x, a13.x), fmaxf(a02.y, a13.y)));
            }
            __syncthreads();                                   // pw reads done before the next group's pass 1 writes
        }
        emax = warp_max(emax);
        if (lane == 0) red[warp] = emax;
        // Requested now, so that they arrive while the log stage runs: this warp's 32 rows of the (transposed) DCT table
        // (the warp takes the mel rows 32 warp .. 32 warp + 31 of every column) and the slot's resampling operator W,
        // staged behind the DCT partial sums when it fits
        float dct_r[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) dct_r[i] = __ldg(dctT + (32 * warp + i) * 32 + lane);
        float* fx = reinterpret_cast<float*>(xb);
        const float* W = w_all + w_offset[slot];
        const bool w_staged = B * q <= kWStage && kMfcc * B <= kDctPart;
        if (w_staged)
            for (int i = t; i < B * q; i += kT) fx[kDctPart + i] = __ldg(W + i);
        __syncthreads();
        // power_to_db: 10 log10(max(1e-10, S)), floored at (segment max - 80 dB); log10 is monotone, so the
        // segment max of the log is the log of the largest energy
        const float smax = fmaxf(fmaxf(red[0], red[1]), fmaxf(red[2], red[3]));
        const float floor_db = 10.f * log10f(fmaxf(1e-10f, smax)) - 80.f;
        const int qp = (q + 3) & ~3;
        for (int idx = t; idx < kMels * (qp >> 2); idx += kT) {
            const int band = idx / (qp >> 2), jq = idx - band * (qp >> 2);
            float4* mp = reinterpret_cast<float4*>(mel + band * mq + 4 * jq);
            float4 v = *mp;
            v.x = fmaxf(3.0102999566398120f * __log2f(fmaxf(1e-10f, v.x)), floor_db);
            v.y = fmaxf(3.0102999566398120f * __log2f(fmaxf(1e-10f, v.y)), floor_db);
            v.z = fmaxf(3.0102999566398120f * __log2f(fmaxf(1e-10f, v.z)), floor_db);
            v.w = fmaxf(3.0102999566398120f * __log2f(fmaxf(1e-10f, v.w)), floor_db);
            *mp = v;
        }
        __syncthreads();
        // DCT-II (ortho) over the mel axis: cep[c][j] = sum_m dct[c][m] * logmel[m][j]; lane = row c.  Each warp sums its
        // 32 mel rows for every column (table rows in registers, log-mels as broadcast reads); the four partial sums are
        // added in warp order.  Columns go in chunks of kDctCols.
        for (int c0 = 0; c0 < qp; c0 += kDctCols) {
            const int nq = ((qp - c0 < kDctCols) ? qp - c0 : kDctCols) >> 2;
            for (int jq = 0; jq < nq; ++jq) {
                V2 a01 = bc(0.f), a23 = bc(0.f);
                const float* mp = mel + (32 * warp) * mq + c0 + 4 * jq;
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    const float4 v = *reinterpret_cast<const float4*>(mp + i * mq);      // broadcast
                    a01 = fma2(bc(dct_r[i]), mk(v.x, v.y), a01);
                    a23 = fma2(bc(dct_r[i]), mk(v.z, v.w), a23);
                }
                float* pp = fx + (warp * kDctCols + 4 * jq) * 32 + lane;
                pp[0] = a01.x; pp[32] = a01.y; pp[64] = a23.x; pp[96] = a23.y;
            }
            __syncthreads();
            for (int i = t; i < nq * 128; i += kT)
                cepT[c0 * 32 + i] = ((fx[i] + fx[kDctCols * 32 + i]) + fx[2 * kDctCols * 32 + i]) + fx[3 * kDctCols * 32 + i];
            __syncthreads();
        }
        // cubic time resampling (utils.py:336-343) as the linear operator W (B x q): warp = output column b (its W row is
        // a broadcast read), lane = cepstral row c; results are staged row-major over the (dead) partial sums and stored
        // as whole lines
        {
            if (w_staged) {
                for (int bb = warp; bb < B; bb += kT / 32) {
                    const float* wr = fx + kDctPart + bb * q;
                    float acc = 0.f;
#pragma unroll 8
                    for (int j = 0; j < q; ++j) acc = fmaf(wr[j], cepT[j * 32 + lane], acc);
                    if (lane < kMfcc) fx[lane * B + bb] = acc;
                }
            } else {
                for (int bb = warp; bb < B; bb += kT / 32) {
                    const float* wr = W + bb * q;
                    float acc = 0.f;
                    for (int j = 0; j < q; ++j) acc = fmaf(__ldg(wr + j), cepT[j * 32 + lane], acc);
                    if (lane < kMfcc) fx[lane * B + bb] = acc;
                }
            }
            __syncthreads();
            float* o = out + slot * (int64_t)(kMfcc * B);
            for (int i = t; i < kMfcc * B; i += kT) o[i] = fx[i];
        }
        // (the next slot's first pass-1 store is behind that slot's loads; cep / mel are rewritten only after barriers)
        __syncthreads();
    }
}

}  // namespace

struct gn_mfcc_tables { const float2* tw1; const float2* tw2; const float* hann; };
gn_mfcc_tables gn_get_mfcc_tables(gn_handle* h);   // gn_model.cu

extern "C" int gn_mfcc_resampled(gn_handle* h, const float* d_y, int64_t n_samples, const int64_t* d_seg_start,
                                 const int32_t* d_seg_len, const int32_t* d_w_offset, const float* d_w, int64_t n,
                                 int B, int max_q, const int32_t* d_mel_start, const int32_t* d_mel_count,
                                 const int32_t* d_mel_offset, const float* d_mel_w, int n_mel_w_given,
                                 const float* d_dct, float* d_out, void* stream) {
    if (!h || n < 0 || n_samples < 0 || B <= 0) return GN_E_ARG;
    if (max_q < 1 || max_q > GN_MFCC_MAX_Q || B > 4096) return GN_E_SHAPE;
    if (n_mel_w_given < 0 || n_mel_w_given > 2 * (kFFT / 2 + 1) + 64) return GN_E_SHAPE;
    if (n == 0) return GN_OK;
    if (!d_y || !d_seg_start || !d_seg_len || !d_w_offset || !d_w || !d_mel_start || !d_mel_count ||
        !d_mel_offset || !d_mel_w || !d_dct || !d_out) return GN_E_ARG;
    const gn_mfcc_tables t = gn_get_mfcc_tables(h);
    const int mq = (max_q + 3) & ~3;
    if (B * kMfcc * 4 > kXBytes) return GN_E_SHAPE;       // the output rows are staged in the exchange buffer
    const int n_mel_w = 2 * (kFFT / 2 + 1) + 64;           // upper bound the caller's CSR must respect (two filters per bin)
    const size_t smem = (size_t)kXBytes + sizeof(float) * ((size_t)(kMels + 32) * mq + n_mel_w);
    if (gn_attr_first(h, reinterpret_cast<const void*>(mfcc_kernel)))
        GN_CUDA(cudaFuncSetAttribute(mfcc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
    // persistent: as many CTAs as are resident at once (4 per SM at 16 columns per slot), slots strided over them
    int per_sm = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, mfcc_kernel, kT, smem) != cudaSuccess || per_sm < 1) {
        cudaGetLastError();
        per_sm = 1;
    }
    const int64_t cap = (int64_t)h->num_sms * per_sm;
    const int grid = (int)(n < cap ? n : cap);
    mfcc_kernel<<<grid, kT, smem, gn_stream(stream)>>>(
        d_y, d_seg_start, d_seg_len, d_w_offset, d_w, n, B, mq, d_mel_start, d_mel_count, d_mel_offset,
        d_mel_w, n_mel_w_given, d_dct, t.tw1, t.tw2, t.hann, d_out);
    GN_LAUNCHED(h);
    return GN_OK;
}
