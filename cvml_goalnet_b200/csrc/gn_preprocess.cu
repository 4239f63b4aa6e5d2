// (a1) frame preprocessing: uint8 BGR HWC -> min/max normalise -> bilinear SxS -> fp32 NCHW / bf16 NHWC.
// Replaces the arithmetic of reference utils.py:284-285,292.  HBM-bound: every source byte is read
// exactly once.  preprocess_tap_kernel streams a frame through the min/max reduction and, on the way,
// keeps the 32-bit words that hold bilinear taps in shared memory (only 2 of every H/S rows and 2 of
// every W/S columns are taps), so the resize never goes back to L2/DRAM.  Big frames are split over a
// thread-block cluster (row bands; min/max exchanged through distributed shared memory).
// preprocess_l2_kernel is the layout-agnostic version (unaligned rows / views): same arithmetic, taps
// re-read through L2.
#include <stdlib.h>
#include <string.h>

#include <vector>

#include "gn_common.cuh"

namespace {

constexpr int kThreads = 512;

__device__ __forceinline__ uint4 ld_stream(const uint4* p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}


// Byte-wise min/max.  The packed-byte video instructions (__vminu4/__vmaxu4) are emulated on sm_100 (~7 SASS
// instructions each); the packed 16-bit three-input min/max (VIMNMX3.U16x2) is native.  A word is split into
// its even and odd bytes as two u16x2 values (one PRMT each) and folded into u16x2 accumulators.
struct ByteMinMax {
    unsigned int mn, mx;                      // two u16 lanes each
    __device__ __forceinline__ ByteMinMax() : mn(0x00FF00FFu), mx(0u) {}
    __device__ __forceinline__ void word(unsigned int w) {
        const unsigned int e = __byte_perm(w, 0u, 0x4240), o = __byte_perm(w, 0u, 0x4341);
        mn = __vimin3_u16x2(mn, e, o);
        mx = __vimax3_u16x2(mx, e, o);
    }
    __device__ __forceinline__ void vec(uint4 v) { word(v.x); word(v.y); word(v.z); word(v.w); }
    __device__ __forceinline__ void byte(unsigned int b) { word(b * 0x01010101u); }
    __device__ __forceinline__ unsigned int lo() const { return min(mn & 0xFFFFu, mn >> 16); }
    __device__ __forceinline__ unsigned int hi() const { return max(mx & 0xFFFFu, mx >> 16); }
};

// cv2.resize INTER_LINEAR taps for one axis, formed in double like OpenCV does (oracle: _bilinear_taps).
__host__ __device__ __forceinline__ void axis_tap(int d, int n_src, int n_dst, int& i0, int& i1, float& frac) {
    const double scale = 1.0 / ((double)n_dst / (double)n_src);
    const double src = ((double)d + 0.5) * scale - 0.5;
    int i = (int)floor(src);
    float f = (float)(src - (double)i);
    if (i < 0) { i = 0; f = 0.f; }
    if (i >= n_src - 1) { i = n_src - 1; f = 0.f; }
    i0 = i;
    i1 = (i + 1 < n_src - 1) ? i + 1 : n_src - 1;
    frac = f;
}

__global__ void __launch_bounds__(kThreads)
preprocess_l2_kernel(const uint8_t* __restrict__ frames, int64_t n, int H, int W, int S,
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
        ByteMinMax mm;
        const uintptr_t addr = reinterpret_cast<uintptr_t>(p);
        int64_t head = (int64_t)((16 - (addr & 15)) & 15);
        if (head > frame_bytes) head = frame_bytes;
        const int64_t nvec = (frame_bytes - head) / 16;
        const int64_t tail0 = head + nvec * 16;
        for (int64_t i = tid; i < head; i += kThreads) mm.byte(p[i]);
        for (int64_t i = tail0 + tid; i < frame_bytes; i += kThreads) mm.byte(p[i]);
        const uint4* pv = reinterpret_cast<const uint4*>(p + head);
        int64_t i = tid;
        for (; i + 3 * kThreads < nvec; i += 4 * kThreads) {          // 4 x 16 B in flight per thread
            const uint4 a = ld_stream(pv + i), b = ld_stream(pv + i + kThreads);
            const uint4 c = ld_stream(pv + i + 2 * kThreads), d = ld_stream(pv + i + 3 * kThreads);
            mm.vec(a); mm.vec(b); mm.vec(c); mm.vec(d);
        }
        for (; i < nvec; i += kThreads) mm.vec(ld_stream(pv + i));
        unsigned int mn = mm.lo(), mx = mm.hi();
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


// ---------------------------------------------------------------------------------------------
// tap-capturing version.  A host-built plan (one per (H, W, S), cached in the handle) lists, as 32-bit
// words:  header[20] | ly0[S] ly1[S] fy[S] | cb0 cb1 fx on oh [3S each] | rowslot[H] | wordmap[3W/4] | rank[CL][5]
//   ly0/ly1  compact index of the two source rows of output row y       rowslot  source row -> compact index or -1
//   cb0/cb1  byte offset inside a compact row of tap xc = c*S + x       wordmap  source word -> compact word or -1
//   fx, on, oh (per xc): horizontal weight, NCHW offset c*S*S + x, NHWC offset x*3 + c
//   rank r   {ya, yb, sa, se, first}: output rows [ya,yb), streamed source rows [sa,se), first compact row
// ---------------------------------------------------------------------------------------------
constexpr int kTapThreads = 512;
constexpr int kTapWarps = kTapThreads / 32;
enum { PP_S = 0, PP_H, PP_W, PP_CL, PP_NWC, PP_ROWSMAX, PP_LY0, PP_LY1, PP_FY, PP_CB0, PP_CB1, PP_FX, PP_ON, PP_OH,
       PP_ROWSLOT, PP_WORDMAP, PP_RANK, PP_TOTAL, PP_HEADER = 20 };

__device__ __forceinline__ uint32_t pp_cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ uint32_t pp_cluster_id() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%clusterid.x;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void pp_st_cluster(uint32_t local_addr, uint32_t rank, uint32_t v) {
    uint32_t ra;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(ra) : "r"(local_addr), "r"(rank));
    asm volatile("st.shared::cluster.u32 [%0], %1;" ::"r"(ra), "r"(v) : "memory");
}
__device__ __forceinline__ void pp_cluster_barrier() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void tap_scatter(uint4 d, const int32_t* wm, uint32_t* dst) {
    const int4 m = *reinterpret_cast<const int4*>(wm);
    if (m.x >= 0) dst[m.x] = d.x;
    if (m.y >= 0) dst[m.y] = d.y;
    if (m.z >= 0) dst[m.z] = d.z;
    if (m.w >= 0) dst[m.w] = d.w;
}

__global__ void __launch_bounds__(kTapThreads, 3)
preprocess_tap_kernel(const uint8_t* __restrict__ frames, int64_t n, const int32_t* __restrict__ plan,
                      float* __restrict__ out_nchw, __nv_bfloat16* __restrict__ out_nhwc) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    int32_t* sp = reinterpret_cast<int32_t*>(smem_raw);
    __shared__ unsigned int red_min[kTapWarps], red_max[kTapWarps];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int total = plan[PP_TOTAL];
    for (int i = tid; i < total; i += kTapThreads) sp[i] = plan[i];
    __syncthreads();
    const int S = sp[PP_S], H = sp[PP_H], W = sp[PP_W], CL = sp[PP_CL], nWc = sp[PP_NWC], rowsMax = sp[PP_ROWSMAX];
    const int* ly0 = sp + sp[PP_LY0];
    const int* ly1 = sp + sp[PP_LY1];
    const float* fy = reinterpret_cast<const float*>(sp + sp[PP_FY]);
    const int* cb0 = sp + sp[PP_CB0];
    const int* cb1 = sp + sp[PP_CB1];
    const float* fx = reinterpret_cast<const float*>(sp + sp[PP_FX]);
    const int* on = sp + sp[PP_ON];
    const int* oh = sp + sp[PP_OH];
    const int* rowslot = sp + sp[PP_ROWSLOT];
    const int* wordmap = sp + sp[PP_WORDMAP];
    float* lut = reinterpret_cast<float*>(sp + total);                 // 256 normalised byte levels
    unsigned int* xch = reinterpret_cast<unsigned int*>(lut + 256);    // [2 parities][8 ranks][min, max]
    uint32_t* tap = xch + 32;                                          // [rowsMax][nWc] captured words
    const uint8_t* tapb = reinterpret_cast<const uint8_t*>(tap);

    const int rank = CL > 1 ? (int)pp_cluster_ctarank() : 0;
    const int64_t cid = CL > 1 ? (int64_t)pp_cluster_id() : (int64_t)blockIdx.x;
    const int64_t ncl = gridDim.x / CL;
    const int* rk = sp + sp[PP_RANK] + 5 * rank;
    const int ya = rk[0], yb = rk[1], sa = rk[2], se = rk[3], first = rk[4];
    const int RB = 3 * W, vpr = RB >> 4;
    const int64_t frame_bytes = (int64_t)H * RB;
    const uint32_t xch_addr = (uint32_t)__cvta_generic_to_shared(xch);

    // thread-constant starting points and steps of the two flat loops (no division inside them)
    const int row0 = tid / vpr, v0 = tid - row0 * vpr;                    // pass 1: vector tid of the band
    const int drow = kTapThreads / vpr, dv = kTapThreads - drow * vpr;
    const int XC = 3 * S, ny = yb - ya;
    const int yy0 = tid / XC, xc0 = tid - yy0 * XC;                       // pass 2: item tid = (y - ya, xc)
    const int dyy = kTapThreads / XC, dxc = kTapThreads - dyy * XC;

    int parity = 0;
    for (int64_t f = cid; f < n; f += ncl, parity ^= 1) {
        const uint8_t* p = frames + f * frame_bytes;
        // ---- pass 1: stream rows [sa, se): packed-byte min/max, tap words -> shared memory ----
        ByteMinMax mm;
        {
            const uint4* pv = reinterpret_cast<const uint4*>(p + (int64_t)sa * RB);
            const int nv = (se - sa) * vpr;
            int row = sa + row0, v = v0;                                  // (row, vector in row) of vector i
            for (int i = tid; i < nv; i += 4 * kTapThreads) {             // 4 x 16 B in flight per thread
                uint4 d[4];
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    d[k] = (i + k * kTapThreads < nv) ? ld_stream(pv + i + k * kTapThreads) : d[0];
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    mm.vec(d[k]);
                    if (i + k * kTapThreads < nv) {
                        const int slot = rowslot[row];
                        const int l = slot - first;
                        if (slot >= 0 && l >= 0 && l < rowsMax) tap_scatter(d[k], wordmap + 4 * v, tap + l * nWc);
                    }
                    row += drow; v += dv;
                    if (v >= vpr) { v -= vpr; ++row; }
                }
            }
        }
        unsigned int mn = mm.lo(), mx = mm.hi();
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            mn = min(mn, __shfl_xor_sync(0xffffffffu, mn, o));
            mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        }
        if (lane == 0) { red_min[warp] = mn; red_max[warp] = mx; }
        __syncthreads();
        if (warp == 0) {
            mn = lane < kTapWarps ? red_min[lane] : 255u;
            mx = lane < kTapWarps ? red_max[lane] : 0u;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                mn = min(mn, __shfl_xor_sync(0xffffffffu, mn, o));
                mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
            }
            if (CL > 1) {
                // every CTA of the cluster publishes its (min, max) into every CTA's exchange slot
                if (lane < CL) {
                    const uint32_t slot = xch_addr + (uint32_t)((parity * 8 + rank) * 2) * 4u;
                    pp_st_cluster(slot, (uint32_t)lane, mn);
                    pp_st_cluster(slot + 4u, (uint32_t)lane, mx);
                }
            } else if (lane == 0) {
                xch[0] = mn;
                xch[1] = mx;
            }
        }
        if (CL > 1) pp_cluster_barrier(); else __syncthreads();
        // ---- normalisation LUT: float32((v - min) / (max - min + 1e-7)) formed in double (utils.py:284) ----
        if (tid < 256) {
            unsigned int gmn = 255u, gmx = 0u;
            if (CL > 1) {
                for (int r = 0; r < CL; ++r) {
                    gmn = min(gmn, xch[(parity * 8 + r) * 2]);
                    gmx = max(gmx, xch[(parity * 8 + r) * 2 + 1]);
                }
            } else {
                gmn = xch[0];
                gmx = xch[1];
            }
            const double rng = (double)((int)gmx - (int)gmn) + 1e-7;
            lut[tid] = (float)(((double)tid - (double)gmn) / rng);
        }
        __syncthreads();
        // ---- pass 2: bilinear taps from the captured words (horizontal then vertical, like cv2) ----
        const int tapRB = nWc * 4;
        float* onc = out_nchw ? out_nchw + f * 3 * S * S : nullptr;
        __nv_bfloat16* onh = out_nhwc ? out_nhwc + f * 3 * S * S : nullptr;
        int yy = yy0, xc = xc0;
        for (int idx = tid; idx < ny * XC; idx += kTapThreads) {
            const int y = ya + yy;
            const uint8_t* r0 = tapb + (ly0[y] - first) * tapRB;
            const uint8_t* r1 = tapb + (ly1[y] - first) * tapRB;
            const int o0 = cb0[xc], o1 = cb1[xc];
            const float ax = fx[xc], ay = fy[y];
            const float top = lut[r0[o0]] * (1.f - ax) + lut[r0[o1]] * ax;
            const float bot = lut[r1[o0]] * (1.f - ax) + lut[r1[o1]] * ax;
            const float val = top * (1.f - ay) + bot * ay;
            if (onc) onc[y * S + on[xc]] = val;
            if (onh) onh[y * XC + oh[xc]] = __float2bfloat16(val);
            yy += dyy; xc += dxc;
            if (xc >= XC) { xc -= XC; ++yy; }
        }
        __syncthreads();   // lut / tap / xch[0..1] are reused by the next frame
    }
}

// ---------------------------------------------------------------------------------------------
// host-side plan
// ---------------------------------------------------------------------------------------------
struct PrePlan {
    int H, W, S, CL;
    int32_t* d_plan;
    size_t smem;
    int ctas_per_sm;
};
struct PrePlanCache { std::vector<PrePlan> plans; };

constexpr size_t kTapSmemTwo = 110 * 1024;     // two CTAs per SM
constexpr size_t kTapSmemOne = 220 * 1024;     // one CTA per SM

// Returns false when no cluster size up to 8 keeps the captured taps in shared memory.
bool build_plan(int H, int W, int S, std::vector<int32_t>& out, int& CL_out, size_t& smem_out) {
    std::vector<int> y0(S), y1(S), x0(S), x1(S);
    std::vector<float> fyv(S), fxv(S);
    for (int d = 0; d < S; ++d) {
        axis_tap(d, H, S, y0[d], y1[d], fyv[d]);
        axis_tap(d, W, S, x0[d], x1[d], fxv[d]);
    }
    std::vector<int32_t> rowslot(H, -1), wordmap((3 * W) / 4, -1);
    for (int d = 0; d < S; ++d) { rowslot[y0[d]] = 0; rowslot[y1[d]] = 0; }
    int nR = 0;
    for (int r = 0; r < H; ++r) if (rowslot[r] == 0) rowslot[r] = nR++;
    for (int d = 0; d < S; ++d)
        for (int c = 0; c < 3; ++c) { wordmap[(x0[d] * 3 + c) >> 2] = 0; wordmap[(x1[d] * 3 + c) >> 2] = 0; }
    int nWc = 0;
    for (size_t w = 0; w < wordmap.size(); ++w) if (wordmap[w] == 0) wordmap[w] = nWc++;
    for (int pass = 0; pass < 2; ++pass) {
        const size_t limit = pass == 0 ? kTapSmemTwo : kTapSmemOne;
        for (int CL = 1; CL <= 8; CL *= 2) {
            if (CL > S) break;
            std::vector<int32_t> rk(5 * CL);
            int rowsMax = 0;
            for (int r = 0; r < CL; ++r) {
                const int ya = (int)((int64_t)r * S / CL), yb = (int)((int64_t)(r + 1) * S / CL);
                const int ya_next = yb < S ? yb : S - 1;
                const int sa = r == 0 ? 0 : y0[ya];
                int se = r == CL - 1 ? H : y0[ya_next];
                if (se < y1[yb - 1] + 1) se = y1[yb - 1] + 1;
                const int first = rowslot[y0[ya]];
                const int cnt = rowslot[y1[yb - 1]] - first + 1;
                if (cnt > rowsMax) rowsMax = cnt;
                rk[5 * r] = ya; rk[5 * r + 1] = yb; rk[5 * r + 2] = sa; rk[5 * r + 3] = se; rk[5 * r + 4] = first;
            }
            // layout (offsets in words; wordmap 16-byte aligned for the int4 reads)
            int off = PP_HEADER;
            const int o_ly0 = off; off += S;
            const int o_ly1 = off; off += S;
            const int o_fy = off; off += S;
            const int o_cb0 = off; off += 3 * S;
            const int o_cb1 = off; off += 3 * S;
            const int o_fx = off; off += 3 * S;
            const int o_on = off; off += 3 * S;
            const int o_oh = off; off += 3 * S;
            const int o_rowslot = off; off += H;
            off = (off + 3) & ~3;
            const int o_wordmap = off; off += (int)wordmap.size();
            const int o_rank = off; off += 5 * CL;
            const int total = (off + 3) & ~3;
            const size_t smem = (size_t)(total + 256 + 32) * 4 + (size_t)rowsMax * nWc * 4;
            if (smem > limit) continue;
            out.assign(total, 0);
            out[PP_S] = S; out[PP_H] = H; out[PP_W] = W; out[PP_CL] = CL; out[PP_NWC] = nWc; out[PP_ROWSMAX] = rowsMax;
            out[PP_LY0] = o_ly0; out[PP_LY1] = o_ly1; out[PP_FY] = o_fy; out[PP_CB0] = o_cb0; out[PP_CB1] = o_cb1;
            out[PP_FX] = o_fx; out[PP_ON] = o_on; out[PP_OH] = o_oh; out[PP_ROWSLOT] = o_rowslot; out[PP_WORDMAP] = o_wordmap; out[PP_RANK] = o_rank;
            out[PP_TOTAL] = total;
            for (int d = 0; d < S; ++d) {
                out[o_ly0 + d] = rowslot[y0[d]];
                out[o_ly1 + d] = rowslot[y1[d]];
                memcpy(&out[o_fy + d], &fyv[d], 4);
                for (int c = 0; c < 3; ++c) {
                    const int b0 = x0[d] * 3 + c, b1 = x1[d] * 3 + c, xc = c * S + d;
                    out[o_cb0 + xc] = wordmap[b0 >> 2] * 4 + (b0 & 3);
                    out[o_cb1 + xc] = wordmap[b1 >> 2] * 4 + (b1 & 3);
                    memcpy(&out[o_fx + xc], &fxv[d], 4);
                    out[o_on + xc] = c * S * S + d;
                    out[o_oh + xc] = d * 3 + c;
                }
            }
            for (int r = 0; r < H; ++r) out[o_rowslot + r] = rowslot[r];
            for (size_t w = 0; w < wordmap.size(); ++w) out[o_wordmap + w] = wordmap[w];
            for (int i = 0; i < 5 * CL; ++i) out[o_rank + i] = rk[i];
            CL_out = CL;
            smem_out = smem;
            return true;
        }
    }
    return false;
}

const PrePlan* get_plan(gn_handle* h, int H, int W, int S, int* status) {
    *status = GN_OK;
    if (!h->pre_plans) h->pre_plans = new PrePlanCache();
    PrePlanCache* cache = reinterpret_cast<PrePlanCache*>(h->pre_plans);
    for (const PrePlan& pl : cache->plans)
        if (pl.H == H && pl.W == W && pl.S == S) return pl.d_plan ? &pl : nullptr;
    PrePlan pl{H, W, S, 0, nullptr, 0, 0};
    std::vector<int32_t> words;
    if (build_plan(H, W, S, words, pl.CL, pl.smem)) {
        cudaError_t e = cudaMalloc(&pl.d_plan, words.size() * 4);
        if (e == cudaSuccess) e = cudaMemcpy(pl.d_plan, words.data(), words.size() * 4, cudaMemcpyHostToDevice);
        if (e == cudaSuccess)
            e = cudaFuncSetAttribute(preprocess_tap_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kTapSmemOne);
        if (e != cudaSuccess) { *status = (int)e; return nullptr; }
        int per = (int)((227 * 1024) / (pl.smem + 1024));
        pl.ctas_per_sm = per > 3 ? 3 : (per < 1 ? 1 : per);   // 39 registers x 512 threads: three CTAs per SM
    }
    cache->plans.push_back(pl);                    // a plan without tables records "use the L2 kernel"
    const PrePlan& back = cache->plans.back();
    return back.d_plan ? &back : nullptr;
}

}  // namespace

void gn_preprocess_free_plans(gn_handle* h) {
    if (!h || !h->pre_plans) return;
    PrePlanCache* cache = reinterpret_cast<PrePlanCache*>(h->pre_plans);
    for (PrePlan& pl : cache->plans) if (pl.d_plan) cudaFree(pl.d_plan);
    delete cache;
    h->pre_plans = nullptr;
}

extern "C" int gn_preprocess_u8_bgr(gn_handle* h, const uint8_t* d_frames, int64_t n, int H, int W, int S,
                                    float* d_out_nchw, void* d_out_nhwc_bf16, void* stream) {
    if (!h || n < 0 || H <= 0 || W <= 0 || S <= 0) return GN_E_ARG;
    if (!d_out_nchw && !d_out_nhwc_bf16) return GN_E_ARG;
    if (S > 1024) return GN_E_SHAPE;
    if (n == 0) return GN_OK;
    if (!d_frames) return GN_E_ARG;
    // tap-capturing kernel when rows are whole 16-byte vectors; otherwise (odd widths, unaligned views) the L2 version
    const bool aligned = ((3 * (int64_t)W) % 16 == 0) && ((reinterpret_cast<uintptr_t>(d_frames) & 15) == 0) && S >= 2;
    if (aligned && !getenv("GN_PREPROCESS_L2")) {
        int st = GN_OK;
        const PrePlan* pl = get_plan(h, H, W, S, &st);
        if (st != GN_OK) return st;
        if (pl) {
            int64_t clusters = (int64_t)h->num_sms * pl->ctas_per_sm / pl->CL;
            if (clusters < 1) clusters = 1;
            if (clusters > n) clusters = n;
            cudaLaunchConfig_t cfg{};
            cfg.gridDim = dim3((unsigned)(clusters * pl->CL));
            cfg.blockDim = dim3(kTapThreads);
            cfg.dynamicSmemBytes = pl->smem;
            cfg.stream = gn_stream(stream);
            cudaLaunchAttribute at[1];
            at[0].id = cudaLaunchAttributeClusterDimension;
            at[0].val.clusterDim.x = (unsigned)pl->CL;
            at[0].val.clusterDim.y = 1;
            at[0].val.clusterDim.z = 1;
            cfg.attrs = at;
            cfg.numAttrs = pl->CL > 1 ? 1 : 0;
            GN_CUDA(cudaLaunchKernelEx(&cfg, preprocess_tap_kernel, d_frames, n, (const int32_t*)pl->d_plan, d_out_nchw,
                                       reinterpret_cast<__nv_bfloat16*>(d_out_nhwc_bf16)));
            GN_LAUNCHED(h);
            return GN_OK;
        }
    }
    const size_t smem = (size_t)S * 6 * sizeof(int) + 256 * sizeof(float);
    // One CTA per frame, 4 resident CTAs (2048 threads) per SM; frames beyond that are grid-strided.
    const int64_t max_grid = (int64_t)h->num_sms * 4;
    const int grid = (int)(n < max_grid ? n : max_grid);
    preprocess_l2_kernel<<<grid, kThreads, smem, gn_stream(stream)>>>(
        d_frames, n, H, W, S, d_out_nchw, reinterpret_cast<__nv_bfloat16*>(d_out_nhwc_bf16));
    GN_LAUNCHED(h);
    return GN_OK;
}
