// (a2) audio features: per frame-slot STFT(2048/512, hann, centred) -> |.|^2 -> 128 Slaney mels ->
// 10 log10 with an 80 dB floor under the segment max -> DCT-II ortho (30 rows) -> cubic time
// resampling W (B x q).  Replaces reference utils.py:322-349 (librosa.feature.mfcc + interp1d loop).
//
// One CTA per slot.  Two STFT columns share one complex 2048-point FFT (column j in the real part,
// column j+1 in the imaginary part), four Stockham passes of radix 4/8/8/8 with register butterflies
// (gn_fft.cuh); mel sums and the segment max are warp-shuffle reductions.  HBM traffic is the slot's samples in and 30*B floats out.
#include "gn_common.cuh"
#include "gn_fft.cuh"

namespace {

constexpr int kFFT = 2048;
constexpr int kHop = 512;
constexpr int kMels = 128;
constexpr int kMfcc = 30;
constexpr int kThreads = 256;
constexpr int kMelWMax = 2304;              // >= 2 * 1025: every FFT bin feeds at most two triangular filters
static_assert(kThreads == 2 * kMels, "one thread per (mel band, STFT column of the pair)");

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

__global__ void __launch_bounds__(kThreads)
mfcc_kernel(const float* __restrict__ y, const int64_t* __restrict__ seg_start, const int32_t* __restrict__ seg_len,
            const int32_t* __restrict__ w_offset, const float* __restrict__ w_all, int64_t n, int B, int max_q,
            const int32_t* __restrict__ mel_start, const int32_t* __restrict__ mel_count,
            const int32_t* __restrict__ mel_offset, const float* __restrict__ mel_w,
            const float* __restrict__ dct, const float2* __restrict__ twiddle, const float* __restrict__ hann,
            float* __restrict__ out) {
    using gnfft::kBufA; using gnfft::kBufB; using gnfft::kTwCount; using gnfft::kTw2048; using gnfft::kTw256;
    using gnfft::kTw32; using gnfft::pass1; using gnfft::pass8;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float* are = reinterpret_cast<float*>(smem_raw);             // FFT buffer A (pass-1 padding), re / im
    float* aim = are + kBufA;
    float* bre = aim + kBufA;                                    // FFT buffer B (pass-2 padding), re / im
    float* bim = bre + kBufB;
    float2* tw = reinterpret_cast<float2*>(bim + kBufB);         // [kTwCount] per-pass twiddle tables
    float* pw0 = reinterpret_cast<float*>(tw + kTwCount);        // [1025] power of column j
    float* pw1 = pw0 + (kFFT / 2 + 1);                           // [1025] power of column j+1
    float* mel = pw1 + (kFFT / 2 + 1);                           // [128][max_q] log-mel
    float* cep = mel + kMels * max_q;                            // [30][max_q]
    float* melw = cep + kMfcc * max_q;                           // [kMelWMax] mel filter weights
    __shared__ float red[kThreads / 32];
    __shared__ float s_max;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // twiddle[k] = exp(-2 pi i k / 2048), k < 1024
    tw[kTw2048 + tid] = twiddle[tid];
    if (tid < 32) tw[kTw256 + tid] = twiddle[8 * tid];
    if (tid < 4) tw[kTw32 + tid] = twiddle[64 * tid];
    // mel filterbank weights (CSR) in shared memory when they fit; this thread owns band tid & 127 of column tid >> 7
    const int n_mel_w = mel_offset[kMels - 1] + mel_count[kMels - 1];
    const bool melw_fits = n_mel_w <= kMelWMax;
    if (melw_fits)
        for (int i = tid; i < n_mel_w; i += kThreads) melw[i] = mel_w[i];
    const int my_band = tid & (kMels - 1), my_col = tid >> 7;
    const int my_st = mel_start[my_band], my_cnt = mel_count[my_band];
    const float* my_w = (melw_fits ? melw : mel_w) + mel_offset[my_band];
    float hw[8];                                                 // this thread's window taps t = tid + 256 s
#pragma unroll
    for (int s = 0; s < 8; ++s) hw[s] = hann[tid + 256 * s];

    for (int64_t slot = blockIdx.x; slot < n; slot += gridDim.x) {
        const int64_t s0 = seg_start[slot];
        const int L = seg_len[slot];
        const int q = 1 + L / kHop;
        const float* seg = y + s0;
        __syncthreads();
        for (int j = 0; j < q; j += 2) {
            const bool two = (j + 1 < q);
            // windowed columns j (re) and j+1 (im); samples outside [0,L) are the centre zero padding.
            // Pass 1 consumes the points straight from registers (t = tid + 256 s).
            float2 z[8];
#pragma unroll
            for (int s = 0; s < 8; ++s) {
                const int t = tid + 256 * s;
                const int i0 = j * kHop - kFFT / 2 + t;
                const int i1 = i0 + kHop;
                const float a = (i0 >= 0 && i0 < L) ? __ldg(seg + i0) * hw[s] : 0.f;
                const float b = (two && i1 >= 0 && i1 < L) ? __ldg(seg + i1) * hw[s] : 0.f;
                z[s] = make_float2(a, b);
            }
            pass1(tid, z, are, aim);
            __syncthreads();
            pass8<4, 1, 2>(tid, are, aim, bre, bim, tw + kTw32);
            __syncthreads();
            pass8<32, 2, 0>(tid, bre, bim, are, aim, tw + kTw256);
            __syncthreads();
            pass8<256, 0, 0>(tid, are, aim, bre, bim, tw + kTw2048);
            __syncthreads();
            // split the two real spectra: X1[k] = (Z[k] + conj Z[N-k])/2, X2[k] = (Z[k] - conj Z[N-k])/(2i)
            for (int k = tid; k <= kFFT / 2; k += kThreads) {
                const int kn = (kFFT - k) & (kFFT - 1);
                const float zkx = bre[k], zky = bim[k], znx = bre[kn], zny = bim[kn];
                const float x1r = 0.5f * (zkx + znx), x1i = 0.5f * (zky - zny);
                const float x2r = 0.5f * (zky + zny), x2i = -0.5f * (zkx - znx);
                pw0[k] = x1r * x1r + x1i * x1i;
                pw1[k] = x2r * x2r + x2i * x2i;
            }
            __syncthreads();
            // mel filterbank: thread = (band, column); the band's bins are contiguous (triangular filter)
            {
                const float* pwc = my_col ? pw1 : pw0;
                float acc = 0.f;
                for (int i = 0; i < my_cnt; ++i) acc = fmaf(my_w[i], pwc[my_st + i], acc);
                if (my_col == 0 || two) mel[my_band * max_q + j + my_col] = acc;
            }
            // no barrier needed here: the next column's passes 1-4 do not touch pw0/pw1/mel, and their
            // barriers order this iteration's pw reads before the next split overwrites pw0/pw1
        }
        __syncthreads();
        // power_to_db: 10 log10(max(1e-10, S)), floored at (segment max - 80 dB)
        float lmax = -INFINITY;
        for (int i = tid; i < kMels * q; i += kThreads) {
            const int mband = i / q, j = i - mband * q;
            const float v = 10.f * log10f(fmaxf(1e-10f, mel[mband * max_q + j]));
            mel[mband * max_q + j] = v;
            lmax = fmaxf(lmax, v);
        }
        lmax = warp_max(lmax);
        if (lane == 0) red[warp] = lmax;
        __syncthreads();
        if (warp == 0) {
            float v = lane < kThreads / 32 ? red[lane] : -INFINITY;
            v = warp_max(v);
            if (lane == 0) s_max = v;
        }
        __syncthreads();
        const float floor_db = s_max - 80.f;
        // DCT-II ortho over the mel axis: cep[c][j] = sum_m dct[c][m] * max(logmel[m][j], floor)
        for (int i = tid; i < kMfcc * q; i += kThreads) {
            const int c = i / q, j = i - c * q;
            const float* d = dct + c * kMels;
            float acc = 0.f;
#pragma unroll 8
            for (int mband = 0; mband < kMels; ++mband)
                acc = fmaf(__ldg(d + mband), fmaxf(mel[mband * max_q + j], floor_db), acc);
            cep[c * max_q + j] = acc;
        }
        __syncthreads();
        // cubic time resampling (utils.py:336-343) as the linear operator W (B x q)
        const float* W = w_all + w_offset[slot];
        float* o = out + slot * (int64_t)(kMfcc * B);
        for (int i = tid; i < kMfcc * B; i += kThreads) {
            const int c = i / B, b = i - c * B;
            float acc = 0.f;
            for (int j = 0; j < q; ++j) acc = fmaf(__ldg(W + b * q + j), cep[c * max_q + j], acc);
            o[i] = acc;
        }
    }
}

}  // namespace

struct gn_mfcc_tables { const float2* twiddle; const float* hann; };
gn_mfcc_tables gn_get_mfcc_tables(gn_handle* h);   // gn_api.cu

extern "C" int gn_mfcc_resampled(gn_handle* h, const float* d_y, int64_t n_samples, const int64_t* d_seg_start,
                                 const int32_t* d_seg_len, const int32_t* d_w_offset, const float* d_w, int64_t n,
                                 int B, int max_q, const int32_t* d_mel_start, const int32_t* d_mel_count,
                                 const int32_t* d_mel_offset, const float* d_mel_w, const float* d_dct,
                                 float* d_out, void* stream) {
    if (!h || n < 0 || n_samples < 0 || B <= 0) return GN_E_ARG;
    if (max_q < 1 || max_q > GN_MFCC_MAX_Q || B > 4096) return GN_E_SHAPE;
    if (n == 0) return GN_OK;
    if (!d_y || !d_seg_start || !d_seg_len || !d_w_offset || !d_w || !d_mel_start || !d_mel_count ||
        !d_mel_offset || !d_mel_w || !d_dct || !d_out) return GN_E_ARG;
    const gn_mfcc_tables t = gn_get_mfcc_tables(h);
    const size_t smem = sizeof(float) * 2 * (gnfft::kBufA + gnfft::kBufB) + sizeof(float2) * gnfft::kTwCount +
                        sizeof(float) * (2 * (kFFT / 2 + 1)) + sizeof(float) * (size_t)(kMels + kMfcc) * max_q + sizeof(float) * kMelWMax + 16;
    static bool attr_set = false;
    if (!attr_set) {
        GN_CUDA(cudaFuncSetAttribute(mfcc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
        attr_set = true;
    }
    const int64_t max_grid = (int64_t)h->num_sms * 4;
    const int grid = (int)(n < max_grid ? n : max_grid);
    mfcc_kernel<<<grid, kThreads, smem, gn_stream(stream)>>>(
        d_y, d_seg_start, d_seg_len, d_w_offset, d_w, n, B, max_q, d_mel_start, d_mel_count, d_mel_offset,
        d_mel_w, d_dct, t.twiddle, t.hann, d_out);
    GN_LAUNCHED(h);
    return GN_OK;
}
