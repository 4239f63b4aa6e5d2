// (a1) frame preprocessing: uint8 BGR HWC -> min/max normalise -> bilinear SxS -> fp32 NCHW / bf16 NHWC.
// Replaces the arithmetic of reference utils.py:284-285,292.  HBM-bound: every source byte is read
// exactly once from DRAM (the min/max pass); the 4*S*S*3 bilinear taps re-hit L1/L2.
#include "gn_common.cuh"

namespace {

constexpr int kThreads = 512;

__device__ __forceinline__ uint4 ld_stream(const uint4* p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}

// cv2.resize INTER_LINEAR taps for one axis, formed in double like OpenCV does (oracle: _bilinear_taps).
__device__ __forceinline__ void axis_tap(int d, int n_src, int n_dst, int& i0, int& i1, float& frac) {
    const double scale = 1.0 / ((double)n_dst / (double)n_src);
    const double src = ((double)d + 0.5) * scale - 0.5;
    int i = (int)floor(src);
    float f = (float)(src - (double)i);
    if (i < 0) { i = 0; f = 0.f; }
    if (i >= n_src - 1) { i = n_src - 1; f = 0.f; }
    i0 = i;
    i1 = min(i + 1, n_src - 1);
    frac = f;
}

__global__ void __launch_bounds__(kThreads)
preprocess_kernel(const uint8_t* __restrict__ frames, int64_t n, int H, int W, int S,
                  float* __restrict__ out_nchw, __nv_bfloat16* __restrict__ out_nhwc) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    int* y0 = reinterpret_cast<int*>(smem_raw);
    int* y1 = y0 + S;
    int* x0 = y1 + S;
    int* x1 = x0 + S;
    float* fy = reinterpret_cast<float*>(x1 + S);
    float* fx = fy + S;
    float* lut = fx + S;                       // 256 entries: normalised value of every byte level
    __shared__ unsigned int red_min[kThreads / 32], red_max[kThreads / 32];
    __shared__ unsigned int s_mn, s_mx;

    const int tid = threadIdx.x;
    for (int d = tid; d < 2 * S; d += kThreads) {
        if (d < S) axis_tap(d, H, S, y0[d], y1[d], fy[d]);
        else       axis_tap(d - S, W, S, x0[d - S], x1[d - S], fx[d - S]);
    }

    const int64_t frame_bytes = (int64_t)H * W * 3;
    for (int64_t f = blockIdx.x; f < n; f += gridDim.x) {
        const uint8_t* p = frames + f * frame_bytes;
        // ---- pass 1: global min / max over all H*W*3 bytes (packed byte-wise SIMD min/max) ----
        unsigned int vmin = 0xFFFFFFFFu, vmax = 0u;
        const uintptr_t addr = reinterpret_cast<uintptr_t>(p);
        int64_t head = (int64_t)((16 - (addr & 15)) & 15);
        if (head > frame_bytes) head = frame_bytes;
        const int64_t nvec = (frame_bytes - head) / 16;
        const int64_t tail0 = head + nvec * 16;
        for (int64_t i = tid; i < head; i += kThreads) {
            unsigned int b = p[i] * 0x01010101u;
            vmin = __vminu4(vmin, b); vmax = __vmaxu4(vmax, b);
        }
        for (int64_t i = tail0 + tid; i < frame_bytes; i += kThreads) {
            unsigned int b = p[i] * 0x01010101u;
            vmin = __vminu4(vmin, b); vmax = __vmaxu4(vmax, b);
        }
        const uint4* pv = reinterpret_cast<const uint4*>(p + head);
        int64_t i = tid;
        for (; i + 3 * kThreads < nvec; i += 4 * kThreads) {          // 4 x 16 B in flight per thread
            uint4 a = ld_stream(pv + i), b = ld_stream(pv + i + kThreads);
            uint4 c = ld_stream(pv + i + 2 * kThreads), d = ld_stream(pv + i + 3 * kThreads);
            unsigned int lo = __vminu4(__vminu4(__vminu4(a.x, a.y), __vminu4(a.z, a.w)),
                                       __vminu4(__vminu4(b.x, b.y), __vminu4(b.z, b.w)));
            unsigned int lo2 = __vminu4(__vminu4(__vminu4(c.x, c.y), __vminu4(c.z, c.w)),
                                        __vminu4(__vminu4(d.x, d.y), __vminu4(d.z, d.w)));
            unsigned int hi = __vmaxu4(__vmaxu4(__vmaxu4(a.x, a.y), __vmaxu4(a.z, a.w)),
                                       __vmaxu4(__vmaxu4(b.x, b.y), __vmaxu4(b.z, b.w)));
            unsigned int hi2 = __vmaxu4(__vmaxu4(__vmaxu4(c.x, c.y), __vmaxu4(c.z, c.w)),
                                        __vmaxu4(__vmaxu4(d.x, d.y), __vmaxu4(d.z, d.w)));
            vmin = __vminu4(vmin, __vminu4(lo, lo2));
            vmax = __vmaxu4(vmax, __vmaxu4(hi, hi2));
        }
        for (; i < nvec; i += kThreads) {
            uint4 a = ld_stream(pv + i);
            vmin = __vminu4(vmin, __vminu4(__vminu4(a.x, a.y), __vminu4(a.z, a.w)));
            vmax = __vmaxu4(vmax, __vmaxu4(__vmaxu4(a.x, a.y), __vmaxu4(a.z, a.w)));
        }
        // fold the 4 byte lanes, then warp-shuffle and block reduce
        unsigned int mn = min(min(vmin & 0xFF, (vmin >> 8) & 0xFF), min((vmin >> 16) & 0xFF, vmin >> 24));
        unsigned int mx = max(max(vmax & 0xFF, (vmax >> 8) & 0xFF), max((vmax >> 16) & 0xFF, vmax >> 24));
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            mn = min(mn, __shfl_xor_sync(0xffffffffu, mn, o));
            mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        }
        if ((tid & 31) == 0) { red_min[tid >> 5] = mn; red_max[tid >> 5] = mx; }
        __syncthreads();
        if (tid < 32) {
            mn = tid < kThreads / 32 ? red_min[tid] : 255u;
            mx = tid < kThreads / 32 ? red_max[tid] : 0u;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                mn = min(mn, __shfl_xor_sync(0xffffffffu, mn, o));
                mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
            }
            if (tid == 0) { s_mn = mn; s_mx = mx; }
        }
        __syncthreads();
        // ---- normalisation LUT: float32((v - min) / (max - min + 1e-7)) formed in double (utils.py:284) ----
        if (tid < 256) {
            const double rng = (double)((int)s_mx - (int)s_mn) + 1e-7;
            lut[tid] = (float)(((double)tid - (double)s_mn) / rng);
        }
        __syncthreads();
        // ---- pass 2: bilinear taps (horizontal then vertical, like cv2's hresize/vresize) ----
        const int SS = S * S;
        if (out_nchw != nullptr) {
            float* o = out_nchw + f * 3 * SS;
            for (int idx = tid; idx < 3 * SS; idx += kThreads) {        // idx = c*S*S + y*S + x
                const int c = idx / SS, r = idx - c * SS, y = r / S, x = r - y * S;
                const uint8_t* r0 = p + ((int64_t)y0[y] * W) * 3 + c;
                const uint8_t* r1 = p + ((int64_t)y1[y] * W) * 3 + c;
                const float ax = fx[x], ay = fy[y];
                const float top = lut[r0[x0[x] * 3]] * (1.f - ax) + lut[r0[x1[x] * 3]] * ax;
                const float bot = lut[r1[x0[x] * 3]] * (1.f - ax) + lut[r1[x1[x] * 3]] * ax;
                o[idx] = top * (1.f - ay) + bot * ay;
            }
        }
        if (out_nhwc != nullptr) {
            __nv_bfloat16* o = out_nhwc + f * 3 * SS;
            for (int idx = tid; idx < 3 * SS; idx += kThreads) {        // idx = (y*S + x)*3 + c
                const int r = idx / 3, c = idx - r * 3, y = r / S, x = r - y * S;
                const uint8_t* r0 = p + ((int64_t)y0[y] * W) * 3 + c;
                const uint8_t* r1 = p + ((int64_t)y1[y] * W) * 3 + c;
                const float ax = fx[x], ay = fy[y];
                const float top = lut[r0[x0[x] * 3]] * (1.f - ax) + lut[r0[x1[x] * 3]] * ax;
                const float bot = lut[r1[x0[x] * 3]] * (1.f - ax) + lut[r1[x1[x] * 3]] * ax;
                o[idx] = __float2bfloat16(top * (1.f - ay) + bot * ay);
            }
        }
        __syncthreads();   // lut / s_mn reused by the next frame
    }
}

}  // namespace

extern "C" int gn_preprocess_u8_bgr(gn_handle* h, const uint8_t* d_frames, int64_t n, int H, int W, int S,
                                    float* d_out_nchw, void* d_out_nhwc_bf16, void* stream) {
    if (!h || n < 0 || H <= 0 || W <= 0 || S <= 0) return GN_E_ARG;
    if (!d_out_nchw && !d_out_nhwc_bf16) return GN_E_ARG;
    if (S > 1024) return GN_E_SHAPE;
    if (n == 0) return GN_OK;
    if (!d_frames) return GN_E_ARG;
    const size_t smem = (size_t)S * 6 * sizeof(int) + 256 * sizeof(float);
    // One CTA per frame, 4 resident CTAs (2048 threads) per SM; frames beyond that are grid-strided.
    const int64_t max_grid = (int64_t)h->num_sms * 4;
    const int grid = (int)(n < max_grid ? n : max_grid);
    preprocess_kernel<<<grid, kThreads, smem, gn_stream(stream)>>>(
        d_frames, n, H, W, S, d_out_nchw, reinterpret_cast<__nv_bfloat16*>(d_out_nhwc_bf16));
    GN_LAUNCHED(h);
    return GN_OK;
}
