// (a1) frame preprocessing: uint8 BGR HWC -> min/max normalise -> bilinear SxS -> fp32 NCHW / bf16 NHWC.
// Replaces the arithmetic of reference utils.py:284-285,292.  HBM-bound: every source byte is read
// exactly once.  preprocess_tap_kernel streams a frame through the min/max reduction and, on the way,
// keeps the 32-bit words that hold bilinear taps in shared memory (only 2 of every H/S rows and 2 of
// every W/S columns are taps), so the resize never goes back to L2/DRAM.  Big frames are split over a
// thread-block cluster (row bands; min/max exchanged through distributed shared memory).
// preprocess_l2_kernel is the layout-agnostic version (unaligned rows / views): same arithmetic, taps
// re-read through L2.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <vector>

#include "gn_common.cuh"
#include "gn_tc_ptx.cuh"

namespace {

constexpr int kThreads = 512;

__device__ __forceinline__ uint4 ld_stream(const uint4* p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}


// Byte-wise min/max.  The packed-byte video instructions (__vminu4/__vmaxu4) are emulated on sm_100 (~7 SASS
// instructions each); the packed 16-bit three-input min/max (VIMNMX3.U16x2) is native.  The accumulators are
// u16x2 values whose HIGH bytes carry the result (a u16 compare is decided by its high byte first, so the low
// bytes may hold anything): a word w contributes its odd bytes as it is and its even bytes as w << 8 --
// three instructions per four bytes.
struct ByteMinMax {
    unsigned int mn, mx;
    __device__ __forceinline__ ByteMinMax() : mn(0xFFFFFFFFu), mx(0u) {}
    __device__ __forceinline__ void word(unsigned int w) {
        const unsigned int e = w << 8;
        mn = __vimin3_u16x2(mn, w, e);
        mx = __vimax3_u16x2(mx, w, e);
    }
    __device__ __forceinline__ void vec(uint4 v) { word(v.x); word(v.y); word(v.z); word(v.w); }
    __device__ __forceinline__ void byte(unsigned int b) { word(b * 0x01010101u); }
    __device__ __forceinline__ unsigned int lo() const { return min((mn >> 8) & 0xFFu, mn >> 24); }
    __device__ __forceinline__ unsigned int hi() const { return max((mx >> 8) & 0xFFu, mx >> 24); }
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
// words:  header[24] | ly0[S] ly1[S] fy[S] | cb0[S] cb1[S] fx[S] | rowslot[H] | rank[CL][5]
//   ly0/ly1  compact index of the two source rows of output row y       rowslot  source row -> compact index or -1
//   cb0/cb1  byte offset inside a source row of the B byte of the two tap pixels of output column x
//   rank r   {ya, yb, sa, se, first}: output rows [ya,yb), streamed source rows [sa,se), first compact row
//   narrow outputs (S <= 64) also carry  tapbefore[H+1] (tap rows above source row r)  and  taprow[nR] (compact
//   index -> source row): a chunk's tap rows are the compact indices [tapbefore[r0], tapbefore[r0 + nr))
// A compact tap row holds, per output column, the 6 bytes the resize needs (8-byte slots), so the resize pass reads
// two conflict-free 8-byte words per pixel and looks the bytes up in a lane-private copy of the normalisation LUT.
// ---------------------------------------------------------------------------------------------
constexpr int kTapThreads = 512;
constexpr int kTapWarps = kTapThreads / 32;
enum { PP_S = 0, PP_H, PP_W, PP_CL, PP_NWC, PP_ROWSMAX, PP_LY0, PP_LY1, PP_FY, PP_CB0, PP_CB1, PP_FX, PP_K,
       PP_ROWSLOT, PP_SRCWORD, PP_RANK, PP_TOTAL, PP_NSLOT, PP_SLOTB, PP_TAPBEFORE, PP_TAPROW, PP_HEADER = 24 };

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
// Streaming structure: warp 16 is a producer whose lane 0 walks the CTA's (frame, chunk of K whole rows <= 16 KB)
// sequence and issues cp.async.bulk copies into a ring of kRing shared-memory slots (full/empty mbarriers), so
// that loads for the next chunks -- and for the next frame, during the resize pass -- are always in flight.
// The 16 consumer warps take each chunk from shared memory: byte min/max over all of it, then the words of
// its tap rows are copied to the tap buffer (one warp per row).
constexpr int kChunkBytes = 16 * 1024;
constexpr int kRingMax = 8;   // ring slots: as many as shared memory allows (bytes in flight per SM set the HBM rate)
constexpr int kTapLaunchThreads = kTapThreads + 32;

__device__ __forceinline__ void bar_consumers() { asm volatile("bar.sync 1, %0;" ::"n"(kTapThreads) : "memory"); }
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_arrive_remote(uint32_t local_bar, uint32_t rank) {
    uint32_t ra;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(ra) : "r"(local_bar), "r"(rank));
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(ra) : "memory");
}
__device__ __forceinline__ void mbar_wait_cluster(uint32_t bar, uint32_t parity) {
    asm volatile(
        "{\n\t"
        ".reg .pred P1;\n\t"
        "WAIT_LOOP:\n\t"
        "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 P1, [%0], %1;\n\t"
        "@P1 bra WAIT_DONE;\n\t"
        "bra WAIT_LOOP;\n\t"
        "WAIT_DONE:\n\t"
        "}\n" ::"r"(bar), "r"(parity) : "memory");
}

template <bool wide>   // wide: many output columns (S > 64): the resize pass is heavy and a chunk holds only a few rows
__global__ void __launch_bounds__(kTapLaunchThreads, 2)
preprocess_tap_kernel(const uint8_t* __restrict__ frames, int64_t n, const int32_t* __restrict__ plan,
                      float* __restrict__ out_nchw, __nv_bfloat16* __restrict__ out_nhwc) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int kRing = plan[PP_NSLOT], slotB = plan[PP_SLOTB];
    unsigned char* ring = smem_raw;                                    // [kRing][slotB]
    int32_t* sp = reinterpret_cast<int32_t*>(smem_raw + kRing * slotB);
    __shared__ unsigned int red_min[kTapWarps], red_max[kTapWarps];
    __shared__ __align__(8) uint64_t bars[2 * kRingMax + 1];           // full[kRing], empty[kRing], exchange
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int total = plan[PP_TOTAL];
    for (int i = tid; i < total; i += kTapLaunchThreads) sp[i] = plan[i];
    const uint32_t bar0 = tcx::smem_u32(bars);
    const int CLk = plan[PP_CL];
    if (tid == 0) {
        for (int i = 0; i < kRing; ++i) {
            tcx::mbar_init(bar0 + 8 * i, 1);                           // full: the producer's expect_tx arrival
            tcx::mbar_init(bar0 + 8 * (kRing + i), kTapWarps);         // empty: one arrival per consumer warp
        }
        tcx::mbar_init(bar0 + 8 * 2 * kRing, CLk);                     // exchange: one arrival per CTA of the cluster
        tcx::fence_barrier_init();
    }
    __syncthreads();
    if (CLk > 1) pp_cluster_barrier();                                 // peers' exchange barriers exist before anyone arrives
    const int S = sp[PP_S], H = sp[PP_H], W = sp[PP_W], CL = sp[PP_CL], rowsMax = sp[PP_ROWSMAX];
    const int* ly0 = sp + sp[PP_LY0];
    const int* ly1 = sp + sp[PP_LY1];
    const float* fy = reinterpret_cast<const float*>(sp + sp[PP_FY]);
    const int* cb0 = sp + sp[PP_CB0];
    const int* cb1 = sp + sp[PP_CB1];
    const float* fx = reinterpret_cast<const float*>(sp + sp[PP_FX]);
    const int* rowslot = sp + sp[PP_ROWSLOT];
    const int K = sp[PP_K];                                            // rows per chunk
    const int* tapbefore = sp + sp[PP_TAPBEFORE];                      // (narrow outputs only)
    const int* taprow = sp + sp[PP_TAPROW];
    float* lut = reinterpret_cast<float*>(sp + total);                 // 256 normalised byte levels
    unsigned int* xch = reinterpret_cast<unsigned int*>(lut + 256);    // [2 parities][8 ranks][min, max]
    float* lutr = reinterpret_cast<float*>(xch + 32);                  // [256][32]: one copy of the LUT per lane = per bank
    uint2* tap = reinterpret_cast<uint2*>(lutr + (wide ? 256 * 32 : 0));            // [rowsMax][S] captured taps: 6 bytes per output column
                                                                       //   {B,G,R of column x0, B,G,R of column x1} of one tap row

    const int rank = CL > 1 ? (int)pp_cluster_ctarank() : 0;
    const int64_t cid = CL > 1 ? (int64_t)pp_cluster_id() : (int64_t)blockIdx.x;
    const int64_t ncl = gridDim.x / CL;
    const int* rk = sp + sp[PP_RANK] + 5 * rank;
    const int ya = rk[0], yb = rk[1], sa = rk[2], se = rk[3], first = rk[4];
    const int RB = 3 * W, vpr = RB >> 4;
    const int64_t frame_bytes = (int64_t)H * RB;
    const int nrows = se - sa;                                         // rows this CTA streams per frame
    const int nchunk = (nrows + K - 1) / K;
    const uint32_t ring_addr = tcx::smem_u32(ring);

    if (warp == kTapWarps) {
        // ---------------- producer ----------------
        if (lane == 0) {
            int slot = 0;
            uint32_t phase = 0;
            for (int64_t f = cid; f < n; f += ncl) {
                const uint8_t* src = frames + f * frame_bytes + (int64_t)sa * RB;
                for (int c = 0; c < nchunk; ++c) {
                    tcx::mbar_wait(bar0 + 8 * (kRing + slot), phase ^ 1);          // slot drained by all consumer warps
                    const int r0 = c * K;
                    const uint32_t bytes = (uint32_t)((nrows - r0 < K ? nrows - r0 : K) * RB);
                    tcx::mbar_expect_tx(bar0 + 8 * slot, bytes);
                    bulk_g2s(ring_addr + slot * slotB, src + (int64_t)r0 * RB, bytes, bar0 + 8 * slot);
                    if (++slot == kRing) { slot = 0; phase ^= 1; }
                }
            }
        }
        return;
    }

    // ---------------- consumers ----------------
    // thread-constant starting point and step of the resize loop (no division inside it)
    const int ny = yb - ya;
    const int yy0 = tid / S, x0s = tid - yy0 * S;                         // item tid = pixel (y - ya, x)
    const int dyy = kTapThreads / S, dxs = kTapThreads - dyy * S;
    const uint32_t xch_addr = tcx::smem_u32(xch);
    const uint32_t xbar = bar0 + 8 * 2 * kRing;

    int slot = 0;
    uint32_t phase = 0, parity = 0;
    for (int64_t f = cid; f < n; f += ncl, parity ^= 1) {
        // ---- pass 1: chunks of rows [sa, se) from the ring: byte min/max, tap words -> tap buffer ----
        ByteMinMax mm;
        for (int c = 0; c < nchunk; ++c) {
            const int r0 = c * K;
            const int nr = nrows - r0 < K ? nrows - r0 : K;
            const int nvc = nr * vpr;
            tcx::mbar_wait(bar0 + 8 * slot, phase);
            const uint4* cv = reinterpret_cast<const uint4*>(ring + slot * slotB);
            for (int j = tid; j < nvc; j += kTapThreads) mm.vec(cv[j]);
            // tap rows of this chunk: gather the 6 source bytes of every output column into the compact tap buffer
            const uint8_t* cbytes = reinterpret_cast<const uint8_t*>(cv);
            if (wide) {
                // wide outputs, where a chunk is only a few rows: all threads per row
                for (int r = 0; r < nr; ++r) {
                    const int sl = rowslot[sa + r0 + r];
                    const int l = sl - first;
                    if (sl < 0 || l < 0 || l >= rowsMax) continue;
                    for (int x = tid; x < S; x += kTapThreads) {
                        const uint8_t* p0 = cbytes + r * RB + cb0[x];
                        const uint8_t* p1 = cbytes + r * RB + cb1[x];
                        uint2 v;
                        v.x = (uint32_t)p0[0] | ((uint32_t)p0[1] << 8) | ((uint32_t)p0[2] << 16) | ((uint32_t)p1[0] << 24);
                        v.y = (uint32_t)p1[1] | ((uint32_t)p1[2] << 8);
                        tap[l * S + x] = v;
                    }
                }
            } else {
                // narrow outputs: the chunk's (tap row, output column) pairs are one flat list over all consumer threads
                // (at 224 -> 40 a 24-row chunk holds 8-9 tap rows: 340 pairs, less than one pass of the 512 threads)
                int lo = tapbefore[sa + r0], hi = tapbefore[sa + r0 + nr];
                lo = lo < first ? first : lo;
                hi = hi > first + rowsMax ? first + rowsMax : hi;
                const int items = (hi - lo) * S;
                int k = yy0, x = x0s;                                   // item tid = (tid / S, tid % S)
                for (int idx = tid; idx < items; idx += kTapThreads) {
                    const int l = lo + k;
                    const uint8_t* rowp = cbytes + (taprow[l] - sa - r0) * RB;
                    const uint8_t* p0 = rowp + cb0[x];
                    const uint8_t* p1 = rowp + cb1[x];
                    uint2 v;
                    v.x = (uint32_t)p0[0] | ((uint32_t)p0[1] << 8) | ((uint32_t)p0[2] << 16) | ((uint32_t)p1[0] << 24);
                    v.y = (uint32_t)p1[1] | ((uint32_t)p1[2] << 8);
                    tap[(l - first) * S + x] = v;
                    k += dyy; x += dxs;
                    if (x >= S) { x -= S; ++k; }
                }
            }
            __syncwarp();
            if (lane == 0) tcx::mbar_arrive(bar0 + 8 * (kRing + slot));
            if (++slot == kRing) { slot = 0; phase ^= 1; }
        }
        unsigned int mn = mm.lo(), mx = mm.hi();
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            mn = min(mn, __shfl_xor_sync(0xffffffffu, mn, o));
            mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        }
        if (lane == 0) { red_min[warp] = mn; red_max[warp] = mx; }
        bar_consumers();
        if (warp == 0) {
            mn = lane < kTapWarps ? red_min[lane] : 255u;
            mx = lane < kTapWarps ? red_max[lane] : 0u;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                mn = min(mn, __shfl_xor_sync(0xffffffffu, mn, o));
                mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
            }
            if (CL > 1) {
                // publish (min, max) into every CTA's exchange slot, then arrive on that CTA's exchange barrier
                if (lane < CL) {
                    const uint32_t dst = xch_addr + (uint32_t)((parity * 8 + rank) * 2) * 4u;
                    pp_st_cluster(dst, (uint32_t)lane, mn);
                    pp_st_cluster(dst + 4u, (uint32_t)lane, mx);
                    mbar_arrive_remote(xbar, (uint32_t)lane);
                }
            } else if (lane == 0) {
                xch[0] = mn;
                xch[1] = mx;
            }
        }
        if (CL > 1) mbar_wait_cluster(xbar, parity); else bar_consumers();
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
        bar_consumers();
        if (wide) {
            for (int i = tid; i < 256 * 32; i += kTapThreads) lutr[i] = lut[i >> 5];   // lane-private copies: conflict-free lookups
            bar_consumers();
        }
        // ---- pass 2: bilinear taps from the captured bytes (horizontal then vertical, like cv2) ----
        const int SS = S * S;
        float* onc = out_nchw ? out_nchw + f * 3 * SS : nullptr;
        __nv_bfloat16* onh = out_nhwc ? out_nhwc + f * 3 * SS : nullptr;
        const float* lt = wide ? lutr + lane : lut;
        const int ls = wide ? 5 : 0;
        int yy = yy0, x = x0s;
        for (int idx = tid; idx < ny * S; idx += kTapThreads) {           // one pixel = its three channels
            const int y = ya + yy;
            const uint2 a = tap[(ly0[y] - first) * S + x];
            const uint2 b = tap[(ly1[y] - first) * S + x];
            const float ax = fx[x], ay = fy[y], bx = 1.f - ax, by = 1.f - ay;
            const unsigned int a0[3] = {a.x & 0xFFu, (a.x >> 8) & 0xFFu, (a.x >> 16) & 0xFFu};
            const unsigned int a1[3] = {a.x >> 24, a.y & 0xFFu, (a.y >> 8) & 0xFFu};
            const unsigned int b0[3] = {b.x & 0xFFu, (b.x >> 8) & 0xFFu, (b.x >> 16) & 0xFFu};
            const unsigned int b1[3] = {b.x >> 24, b.y & 0xFFu, (b.y >> 8) & 0xFFu};
            float val[3];
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                const float top = lt[a0[c] << ls] * bx + lt[a1[c] << ls] * ax;
                const float bot = lt[b0[c] << ls] * bx + lt[b1[c] << ls] * ax;
                val[c] = top * by + bot * ay;
            }
            if (onc) {
                float* o = onc + y * S + x;
                o[0] = val[0]; o[SS] = val[1]; o[2 * SS] = val[2];
            }
            if (onh) {
                __nv_bfloat16* o = onh + (y * S + x) * 3;
                o[0] = __float2bfloat16(val[0]); o[1] = __float2bfloat16(val[1]); o[2] = __float2bfloat16(val[2]);
            }
            yy += dyy; x += dxs;
            if (x >= S) { x -= S; ++yy; }
        }
        bar_consumers();   // lut / tap / xch[0..1] are reused by the next frame
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
    int max_clusters;      // co-resident clusters (cudaOccupancyMaxActiveClusters), 0 = not queried yet
};
struct PrePlanCache { std::vector<PrePlan> plans; };

constexpr size_t kTapSmemTwo = 112 * 1024;     // two CTAs per SM
constexpr size_t kTapSmemOne = 220 * 1024;     // one CTA per SM

// Returns false when no cluster size up to 8 keeps the captured taps in shared memory.
bool build_plan(int H, int W, int S, std::vector<int32_t>& out, int& CL_out, size_t& smem_out) {
    std::vector<int> y0(S), y1(S), x0(S), x1(S);
    std::vector<float> fyv(S), fxv(S);
    for (int d = 0; d < S; ++d) {
        axis_tap(d, H, S, y0[d], y1[d], fyv[d]);
        axis_tap(d, W, S, x0[d], x1[d], fxv[d]);
    }
    std::vector<int32_t> rowslot(H, -1);
    for (int d = 0; d < S; ++d) { rowslot[y0[d]] = 0; rowslot[y1[d]] = 0; }
    int nR = 0;
    for (int r = 0; r < H; ++r) if (rowslot[r] == 0) rowslot[r] = nR++;
    const int nWc = 2 * S;                               // words per compact tap row (8 bytes per output column)
    const int K = kChunkBytes / (3 * W);                 // whole rows per ring slot
    if (K < 1) return false;
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
            const int o_cb0 = off; off += S;
            const int o_cb1 = off; off += S;
            const int o_fx = off; off += S;
            const int o_rowslot = off; off += H;
            const int o_srcword = off;
            const int o_rank = off; off += 5 * CL;
            const bool narrow = S <= 64;
            const int o_tapbefore = off; off += narrow ? H + 1 : 0;
            const int o_taprow = off; off += narrow ? nR : 0;
            const int total = (off + 3) & ~3;
            const size_t rest = (size_t)(total + 256 + 32 + (S > 64 ? 256 * 32 : 0)) * 4 + (size_t)rowsMax * nWc * 4;
            const size_t slotB = ((size_t)K * 3 * W + 127) & ~(size_t)127;
            if (rest + 2 * slotB > limit) continue;
            size_t nslot = (limit - rest) / slotB;
            if (nslot > (size_t)kRingMax) nslot = kRingMax;
            if (const char* e = getenv("GN_PRE_SLOTS")) { const size_t v = (size_t)atoi(e); if (v >= 2 && v < nslot) nslot = v; }
            const size_t smem = nslot * slotB + rest;
            out.assign(total, 0);
            out[PP_S] = S; out[PP_H] = H; out[PP_W] = W; out[PP_CL] = CL; out[PP_NWC] = nWc; out[PP_ROWSMAX] = rowsMax;
            out[PP_LY0] = o_ly0; out[PP_LY1] = o_ly1; out[PP_FY] = o_fy; out[PP_CB0] = o_cb0; out[PP_CB1] = o_cb1;
            out[PP_FX] = o_fx; out[PP_K] = K; out[PP_ROWSLOT] = o_rowslot; out[PP_SRCWORD] = o_srcword; out[PP_RANK] = o_rank;
            out[PP_TOTAL] = total; out[PP_NSLOT] = (int32_t)nslot; out[PP_SLOTB] = (int32_t)slotB;
            for (int d = 0; d < S; ++d) {
                out[o_ly0 + d] = rowslot[y0[d]];
                out[o_ly1 + d] = rowslot[y1[d]];
                memcpy(&out[o_fy + d], &fyv[d], 4);
                memcpy(&out[o_fx + d], &fxv[d], 4);
                out[o_cb0 + d] = x0[d] * 3;          // byte offsets of the two tap pixels inside a source row
                out[o_cb1 + d] = x1[d] * 3;
            }
            for (int r = 0; r < H; ++r) out[o_rowslot + r] = rowslot[r];
            for (int i = 0; i < 5 * CL; ++i) out[o_rank + i] = rk[i];
            out[PP_TAPBEFORE] = o_tapbefore; out[PP_TAPROW] = o_taprow;
            if (narrow) {
                int seen = 0;
                for (int r = 0; r < H; ++r) {
                    out[o_tapbefore + r] = seen;
                    if (rowslot[r] >= 0) out[o_taprow + seen++] = r;
                }
                out[o_tapbefore + H] = seen;
            }
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
    PrePlan pl{H, W, S, 0, nullptr, 0, 0, 0};
    std::vector<int32_t> words;
    if (build_plan(H, W, S, words, pl.CL, pl.smem)) {
        cudaError_t e = cudaMalloc(&pl.d_plan, words.size() * 4);
        if (e == cudaSuccess) e = cudaMemcpy(pl.d_plan, words.data(), words.size() * 4, cudaMemcpyHostToDevice);
        if (e == cudaSuccess)
            e = cudaFuncSetAttribute(S > 64 ? preprocess_tap_kernel<true> : preprocess_tap_kernel<false>,
                                     cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kTapSmemOne);
        if (e != cudaSuccess) { *status = (int)e; return nullptr; }
        int per = (int)((227 * 1024) / (pl.smem + 1024));
        pl.ctas_per_sm = per > 2 ? 2 : (per < 1 ? 1 : per);
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
            auto kern = S > 64 ? preprocess_tap_kernel<true> : preprocess_tap_kernel<false>;
            cudaLaunchConfig_t cfg{};
            cfg.gridDim = dim3((unsigned)(clusters * pl->CL));
            cfg.blockDim = dim3(kTapLaunchThreads);
            cfg.dynamicSmemBytes = pl->smem;
            cfg.stream = gn_stream(stream);
            cudaLaunchAttribute at[1];
            at[0].id = cudaLaunchAttributeClusterDimension;
            at[0].val.clusterDim.x = (unsigned)pl->CL;
            at[0].val.clusterDim.y = 1;
            at[0].val.clusterDim.z = 1;
            cfg.attrs = at;
            cfg.numAttrs = pl->CL > 1 ? 1 : 0;
            if (pl->CL > 1) {
                // a cluster needs CL free SMs inside one GPC: launch only as many clusters as can be resident at once
                // (frames are strided over clusters, so a cluster that starts late would run its share alone)
                PrePlan* mut = const_cast<PrePlan*>(pl);
                if (mut->max_clusters == 0) {
                    int nc = 0;
                    cfg.gridDim = dim3((unsigned)(h->num_sms / pl->CL * pl->CL));
                    if (cudaOccupancyMaxActiveClusters(&nc, kern, &cfg) != cudaSuccess || nc < 1) {
                        cudaGetLastError();
                        nc = (int)clusters;
                    }
                    mut->max_clusters = nc;
                    if (getenv("GN_DEBUG")) fprintf(stderr, "[gn] preprocess %dx%d->%d: cluster %d, %zu B smem, %d resident clusters\n",
                                                    pl->H, pl->W, pl->S, pl->CL, pl->smem, nc);
                }
                if (clusters > pl->max_clusters) clusters = pl->max_clusters;
                cfg.gridDim = dim3((unsigned)(clusters * pl->CL));
            }
            GN_CUDA(cudaLaunchKernelEx(&cfg, kern, d_frames, n, (const int32_t*)pl->d_plan, d_out_nchw,
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
