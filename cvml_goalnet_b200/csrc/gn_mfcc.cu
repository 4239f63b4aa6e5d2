// (a2) audio features: per frame-slot STFT(2048/512, hann, centred) -> |.|^2 -> 128 Slaney mels ->
// 10 log10 with an 80 dB floor under the segment max -> DCT-II ortho (30 rows) -> cubic time
// resampling W (B x q).  Replaces reference utils.py:322-349 (librosa.feature.mfcc + interp1d loop).
//
// One CTA (8 warps) per slot, one warp per pair of STFT columns.  Two columns share one complex
// 2048-point FFT (column j in the real part, column j+1 in the imaginary part) that the warp computes
// with the points in registers (gn_fft.cuh: 64-point DFTs, one transpose through a private
// shared-memory tile, 32-point DFTs); X[k] and X[N-k] land in the same lane, so the two power spectra
// are formed in registers and only then written to the warp's tile for the mel filterbank.  No block
// barrier inside the transform; HBM traffic is the slot's samples in and 30*B floats out.
#include "gn_common.cuh"
#include "gn_fft.cuh"

namespace {

constexpr int kFFT = 2048;
constexpr int kHop = 512;
constexpr int kMels = 128;
constexpr int kMfcc = 30;
constexpr int kWarps = 8;
constexpr int kThreads = kWarps * 32;
constexpr int kMelWMax = 2304;              // >= 2 * 1025: every FFT bin feeds at most two triangular filters
constexpr int kPw2 = 1056;                  // offset (floats) of the second power spectrum inside a warp tile
static_assert(kPw2 >= kFFT / 2 + 1 && kPw2 + kFFT / 2 + 1 <= 2 * gnfft::kTile, "power spectra alias the tile");

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

__global__ void __launch_bounds__(kThreads, 1)
mfcc_kernel(const float* __restrict__ y, const int64_t* __restrict__ seg_start, const int32_t* __restrict__ seg_len,
            const int32_t* __restrict__ w_offset, const float* __restrict__ w_all, int64_t n, int B, int max_q,
            const int32_t* __restrict__ mel_start, const int32_t* __restrict__ mel_count,
            const int32_t* __restrict__ mel_offset, const float* __restrict__ mel_w,
            const float* __restrict__ dct, const float2* __restrict__ twiddle, const float* __restrict__ hann,
            float* __restrict__ out, int dct_smem, int w_smem) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float2* tiles = reinterpret_cast<float2*>(smem_raw);                  // [kWarps][kTile] transpose tiles
    float2* tw = tiles + kWarps * gnfft::kTile;                           // [64][32] lane-major twiddles
    float* hw = reinterpret_cast<float*>(tw + gnfft::kTwLane);            // [2048] hann window
    float* melw = hw + kFFT;                                              // [kMelWMax] mel filter weights
    int* mst = reinterpret_cast<int*>(melw + kMelWMax);                   // [128] first bin of each filter
    int* mcn = mst + kMels;                                               // [128] bins per filter
    int* mof = mcn + kMels;                                               // [128] offset into melw
    float* mel = reinterpret_cast<float*>(mof + kMels);                   // [128][max_q] log-mel
    float* cep = mel + kMels * max_q;                                     // [30][max_q]
    float* dctT = cep + kMfcc * max_q;                                    // [128][32] DCT rows, transposed (if dct_smem)
    float* wsm = dctT + (dct_smem ? kMels * 32 : 0);                      // [B][max_q + 4] this slot's W (if w_smem)
    __shared__ float red[kWarps];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int i = tid; i < gnfft::kTwLane; i += kThreads) tw[i] = twiddle[i];
    for (int i = tid; i < kFFT; i += kThreads) hw[i] = hann[i];
    const int n_mel_w = mel_offset[kMels - 1] + mel_count[kMels - 1];
    const bool melw_fits = n_mel_w <= kMelWMax;
    if (melw_fits)
        for (int i = tid; i < n_mel_w; i += kThreads) melw[i] = mel_w[i];
    if (tid < kMels) { mst[tid] = mel_start[tid]; mcn[tid] = mel_count[tid]; mof[tid] = mel_offset[tid]; }
    if (dct_smem)
        for (int i = tid; i < kMels * 32; i += kThreads) {
            const int m = i >> 5, c = i & 31;
            dctT[i] = c < kMfcc ? dct[c * kMels + m] : 0.f;
        }
    const float* mw = melw_fits ? melw : mel_w;
    float2* tile = tiles + warp * gnfft::kTile;
    float* pw1 = reinterpret_cast<float*>(tile);
    float* pw2 = pw1 + kPw2;

    for (int64_t slot = blockIdx.x; slot < n; slot += gridDim.x) {
        const int64_t s0 = seg_start[slot];
        const int L = seg_len[slot];
        const int q = 1 + L / kHop;
        const float* seg = y + s0;
        const float* W = w_all + w_offset[slot];
        const int wq = q + 4;                                  // padded row of the staged W: conflict-free 16-byte reads
        __syncthreads();          // tables loaded; previous slot's resampling is done with cep / wsm
        if (w_smem)
            for (int i = tid; i < B * q; i += kThreads) {
                const int bb = i / q;
                wsm[bb * wq + (i - bb * q)] = __ldg(W + i);
            }
        float emax = 0.f;                                      // largest mel energy seen by this lane
        for (int j = 2 * warp; j < q; j += 2 * kWarps) {
            const bool two = (j + 1 < q);
            {
                // windowed columns j (re) and j+1 (im); samples outside [0,L) are the centre zero padding.
                // Column j+1 is column j shifted by one hop = 16 rows of 32 samples, so 80 rows feed both.
                float2 z[64];
                const int base = j * kHop - kFFT / 2 + lane;
                float raw[16];
#pragma unroll
                for (int r = 0; r < 16; ++r) {
                    const int i = base + 32 * r;
                    raw[r] = ((unsigned)i < (unsigned)L) ? __ldg(seg + i) : 0.f;
                }
#pragma unroll
                for (int n1 = 0; n1 < 64; ++n1) {
                    const int i = base + 32 * (n1 + 16);
                    const float nxt = ((unsigned)i < (unsigned)L) ? __ldg(seg + i) : 0.f;
                    const float h = hw[32 * n1 + lane];
                    z[n1] = make_float2(raw[n1 & 15] * h, two ? nxt * h : 0.f);
                    raw[n1 & 15] = nxt;
                }
                gnfft::warp_fft_stage1(lane, z, tw, tile);
            }
            __syncwarp();
            {
                float2 a[32], b[32];
                gnfft::warp_fft_stage2(lane, tile, a, b);
                __syncwarp();                                  // every lane has read its rows: the tile becomes pw1 / pw2
                gnfft::warp_fft_power(lane, a, b, pw1, pw2);
            }
            __syncwarp();
            // mel filterbank: lane owns bands lane, lane + 32, ... (a band's bins are contiguous)
#pragma unroll 1
            for (int band = lane; band < kMels; band += 32) {
                const int st = mst[band], cnt = mcn[band];
                const float* wv = mw + mof[band];
                float a0 = 0.f, a1 = 0.f;
                for (int i = 0; i < cnt; ++i) {
                    const float wgt = wv[i];
                    a0 = fmaf(wgt, pw1[st + i], a0);
                    a1 = fmaf(wgt, pw2[st + i], a1);
                }
                mel[band * max_q + j] = a0;
                emax = fmaxf(emax, a0);
                if (two) { mel[band * max_q + j + 1] = a1; emax = fmaxf(emax, a1); }
            }
            __syncwarp();                                      // pw reads done before the next pair reuses the tile
        }
        emax = warp_max(emax);
        if (lane == 0) red[warp] = emax;
        __syncthreads();
        // power_to_db: 10 log10(max(1e-10, S)), floored at (segment max - 80 dB); log10 is monotone, so the
        // segment max of the log is the log of the largest energy
        float smax = red[0];
#pragma unroll
        for (int w = 1; w < kWarps; ++w) smax = fmaxf(smax, red[w]);
        const float floor_db = 10.f * log10f(fmaxf(1e-10f, smax)) - 80.f;
        // every warp finishes its own columns: log + floor in place, then the DCT-II (ortho) over the mel axis,
        // cep[c][j] = sum_m dct[c][m] * logmel[m][j], lane = cepstral row c
        for (int j = 2 * warp; j < q; j += 2 * kWarps) {
            const bool two = (j + 1 < q);
            for (int band = lane; band < kMels; band += 32) {
                float* mp = mel + band * max_q + j;
                mp[0] = fmaxf(10.f * log10f(fmaxf(1e-10f, mp[0])), floor_db);
                if (two) mp[1] = fmaxf(10.f * log10f(fmaxf(1e-10f, mp[1])), floor_db);
            }
            __syncwarp();
            if (lane < kMfcc) {
                float c0 = 0.f, c1 = 0.f;
                const float* mp = mel + j;
                const int j1 = two ? 1 : 0;
                if (dct_smem) {
#pragma unroll 8
                    for (int m = 0; m < kMels; ++m) {
                        const float d = dctT[m * 32 + lane];
                        c0 = fmaf(d, mp[m * max_q], c0);
                        c1 = fmaf(d, mp[m * max_q + j1], c1);
                    }
                } else {
                    const float* dg = dct + lane * kMels;
#pragma unroll 8
                    for (int m = 0; m < kMels; ++m) {
                        const float d = __ldg(dg + m);
                        c0 = fmaf(d, mp[m * max_q], c0);
                        c1 = fmaf(d, mp[m * max_q + j1], c1);
                    }
                }
                cep[lane * max_q + j] = c0;
                if (two) cep[lane * max_q + j + 1] = c1;
            }
        }
        __syncthreads();
        // cubic time resampling (utils.py:336-343) as the linear operator W (B x q)
        float* o = out + slot * (int64_t)(kMfcc * B);
        if (w_smem && (q & 3) == 0 && (max_q & 3) == 0) {
            for (int i = tid; i < kMfcc * B; i += kThreads) {
                const int c = i / B, bb = i - c * B;
                const float4* wr = reinterpret_cast<const float4*>(wsm + bb * wq);
                const float4* cr = reinterpret_cast<const float4*>(cep + c * max_q);
                float acc = 0.f;
                for (int j4 = 0; j4 < (q >> 2); ++j4) {
                    const float4 wv = wr[j4], cv = cr[j4];
                    acc = fmaf(wv.x, cv.x, acc); acc = fmaf(wv.y, cv.y, acc);
                    acc = fmaf(wv.z, cv.z, acc); acc = fmaf(wv.w, cv.w, acc);
                }
                o[i] = acc;
            }
        } else {
            for (int i = tid; i < kMfcc * B; i += kThreads) {
                const int c = i / B, bb = i - c * B;
                float acc = 0.f;
                if (w_smem) for (int j = 0; j < q; ++j) acc = fmaf(wsm[bb * wq + j], cep[c * max_q + j], acc);
                else        for (int j = 0; j < q; ++j) acc = fmaf(__ldg(W + bb * q + j), cep[c * max_q + j], acc);
                o[i] = acc;
            }
        }
    }
}

}  // namespace

struct gn_mfcc_tables { const float2* twiddle; const float* hann; };
gn_mfcc_tables gn_get_mfcc_tables(gn_handle* h);   // gn_model.cu

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
    size_t smem = sizeof(float2) * (size_t)(kWarps * gnfft::kTile + gnfft::kTwLane) + sizeof(float) * (kFFT + kMelWMax) +
                  sizeof(int) * 3 * kMels + sizeof(float) * (size_t)(kMels + kMfcc) * max_q;
    constexpr size_t kSmemMax = 220 * 1024;
    // optional staging (dropped for very long segments): the DCT rows and the slot's resampling operator
    const size_t dct_bytes = sizeof(float) * kMels * 32, w_bytes = sizeof(float) * (size_t)B * (max_q + 4);
    const int dct_smem = smem + dct_bytes <= kSmemMax ? 1 : 0;
    if (dct_smem) smem += dct_bytes;
    const int w_smem = (w_bytes <= 16 * 1024 && smem + w_bytes <= kSmemMax) ? 1 : 0;
    if (w_smem) smem += w_bytes;
    if (gn_attr_first(h, reinterpret_cast<const void*>(mfcc_kernel)))
        GN_CUDA(cudaFuncSetAttribute(mfcc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemMax));
    // persistent: one 8-warp CTA per SM (the register-resident transforms need ~all of the register file)
    const int grid = (int)(n < (int64_t)h->num_sms ? n : (int64_t)h->num_sms);
    mfcc_kernel<<<grid, kThreads, smem, gn_stream(stream)>>>(
        d_y, d_seg_start, d_seg_len, d_w_offset, d_w, n, B, max_q, d_mel_start, d_mel_count, d_mel_offset,
        d_mel_w, d_dct, t.twiddle, t.hann, d_out, dct_smem, w_smem);
    GN_LAUNCHED(h);
    return GN_OK;
}
