// VisBl block 1 of the 40x40 model on the tensor cores: conv1 3->64 k3 s3 p3 + ReLU + MaxPool(3,1) + BatchNorm(eval)
// (utils.py:174-177) -> bf16 NHWC with a zero border (the padding = 1 input of conv2).
//
// conv1 has stride == kernel size, so its im2col is a PERMUTATION of the zero-padded image: pixel (y, x) of channel ci
// is tap k = ci*9 + (y%3)*3 + x%3 of conv pixel (y/3+1, x/3+1).  The CUDA-core kernel (gn_front.cu) spent its time
// feeding 27 taps x 64 channels through the LSU (72 % of the shared-memory pipe) and half-rate scalar FFMAs; here the
// frame is scattered once into a K-major, 128-byte-swizzled patch matrix (one 128-byte row = the 27 taps of a conv
// pixel, padded to 32) and the convolution is one tcgen05 GEMM per frame:
//     D[channel][pixel] = sum_k W[channel][k] * patch[pixel][k]        M = 128, N = 240, K = 32, kind::tf32
//   * rows (TMEM lanes) = channels -- the 64 channels twice, so that every lane quarter has rows to read;
//     columns = conv pixels at 16 y + x (a conv row starts on a 16-column boundary): a thread owns one channel and
//     walks down the conv rows of its quarter of the image, so the separable 3x3/s1 max pool, bias + ReLU (monotone:
//     they commute with the max) and the BatchNorm affine all happen in registers -- no staging tile in shared memory;
//   * precision: operands are rounded to TF32 (cvt.rna).  On the oracle, rounding conv1's pixels and weights to TF32
//     moves the scores by at most 1.3e-5 relative (512 frames, spread weights; the layer's output is rounded to bf16 for
//     conv2 anyway) against the 1e-3 parity tolerance, so one MMA group is the default.  kSplit (GN_BLOCK1_SPLIT) keeps
//     the fp32-accurate variant: x = hi + lo (hi = tf32(x), lo = tf32(x - hi)), three MMA groups hi*hi + lo*hi + hi*lo.
// One CTA = 256 threads, 256 TMEM columns, 56 KB of shared memory (105 KB with kSplit) -> two CTAs per SM (TMEM-limited) overlap each other's phases
// (scatter / MMA / epilogue); the next frame's pixels travel to registers during the MMA and the epilogue.
#include "gn_common.cuh"
#include "gn_tc_ptx.cuh"

namespace {

using namespace tcx;

constexpr int kS = 40;               // frame side (reference utils.py:285)
constexpr int kC1 = 15;              // conv1 output side: (40 + 2*3 - 3)/3 + 1
constexpr int kP = kC1 - 2;          // pooled side
constexpr int kNPix = 16 * kC1;      // GEMM N: conv pixel (y, x) is column 16 y + x
constexpr int kBT = 256;             // threads: 8 warps (warp w reads TMEM lanes 32 (w & 3) ..)
constexpr int kBBytes = kNPix * 128; // one patch matrix (hi or lo)
constexpr int kABytes = 128 * 128;   // one weight matrix (hi or lo): 128 rows = the 64 channels twice
constexpr int kTmemColsB1 = 256;
constexpr int kRowVec = kS / 4;                     // float4 per image row
constexpr int kGroups = 3 * kS / 3;                 // scatter groups of 3 image rows (30 float4: one warp pass)
constexpr int kPre = kGroups / (kBT / 32);          // passes per warp and frame
static_assert(kGroups % (kBT / 32) == 0 && 3 * kRowVec <= 32, "a warp pass = three image rows");
constexpr size_t b1_smem(bool split) {
    return 1024 + (split ? 2 : 1) * (size_t)(kBBytes + kABytes) + sizeof(unsigned short) * 3 * kS * kS + 64;
}

// float index of element (row r, k) in a K-major tile with 128-byte rows and the 128-byte swizzle
__device__ __forceinline__ int sw_index(int r, int k) { return r * 32 + ((((k >> 2) ^ (r & 7))) << 2) + (k & 3); }
// column of conv pixel px inside its row's 16-column block.  The hardware swizzle only sees row & 7, so pixels 8 apart
// would share a bank pattern -- and they are exactly the ones two lanes of a scatter store hit together (lanes 6 apart
// hold pixels 24 = 8 conv columns apart with the same tap): the upper eight columns are rotated by one.
__host__ __device__ constexpr int px_col(int px) { return px < 8 ? px : 8 + ((px + 1) & 7); }

__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr) : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

template <bool kSplit>
__global__ void __launch_bounds__(kBT, 2)
visbl_block1_tc_kernel(const float* __restrict__ vis, int64_t n, const float* __restrict__ w /*[27][64], k=(ky*3+kx)*3+ci*/,
                       const float* __restrict__ b, const float* __restrict__ sc, const float* __restrict__ sh,
                       __nv_bfloat16* __restrict__ a1p) {
    extern __shared__ unsigned char smem_raw[];
    unsigned char* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    constexpr int kParts = kSplit ? 2 : 1;
    float* Bhi = reinterpret_cast<float*>(smem);                               // [240][32] patch matrix (hi parts)
    float* Blo = reinterpret_cast<float*>(smem + kBBytes);                     //           lo parts (kSplit)
    float* Ahi = reinterpret_cast<float*>(smem + kParts * kBBytes);            // [128][32] weights (64 channels twice)
    float* Alo = reinterpret_cast<float*>(smem + kParts * kBBytes + kABytes);  //           lo parts (kSplit)
    unsigned short* dsto = reinterpret_cast<unsigned short*>(smem + kParts * (kBBytes + kABytes));   // [4800] scatter offsets
    uint64_t* bar = reinterpret_cast<uint64_t*>(dsto + 3 * kS * kS);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar + 1);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    // ---- once per CTA: zero patch matrices (padding pixels, pad column 15, k = 27..31 stay zero), weights, offsets ----
    for (int i = tid; i < kParts * kNPix * 32; i += kBT) Bhi[i] = 0.f;         // Bhi | Blo are contiguous
    for (int i = tid; i < 128 * 32; i += kBT) {
        const int m = i >> 5, k = i & 31, ch = m & 63;
        const float v = (k < 27) ? __ldg(w + ((k % 9) * 3 + k / 9) * 64 + ch) : 0.f;      // k = ci*9 + tap
        const float hi = gn_round_tf32(v);
        Ahi[sw_index(m, k)] = hi;
        if (kSplit) Alo[sw_index(m, k)] = gn_round_tf32(v - hi);
    }
    for (int e = tid; e < 3 * kS * kS; e += kBT) {
        const int ci = e / (kS * kS), r = e - ci * kS * kS, y = r / kS, x = r - y * kS;
        dsto[e] = (unsigned short)sw_index(16 * (y / 3 + 1) + px_col(x / 3 + 1), ci * 9 + (y % 3) * 3 + (x % 3));
    }
    if (tid == 0) {
        mbar_init(smem_u32(bar), 1);
        fence_barrier_init();
    }
    if (warp == 0) tmem_alloc(smem_u32(tmem_slot), kTmemColsB1);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    // TMEM lane of this thread = 32 (warp & 3) + lane -> channel (lanes 64..127 repeat the channels); the four warps that
    // share a channel set take a quarter of the pooled rows each
    const int tl = 32 * (warp & 3) + lane, ch = tl & 63, rq = (tl >> 6) + 2 * (warp >> 2);
    const float bias = __ldg(b + ch), scale = __ldg(sc + ch), shift = __ldg(sh + ch);
    const uint32_t taddr = tmem_base + ((uint32_t)(32 * (warp & 3)) << 16);
    const int y_begin = rq == 0 ? 0 : 1 + 3 * rq;               // pooled rows [y_begin, y_end): 0-3, 4-6, 7-9, 10-12
    const int y_end = 4 + 3 * rq;

    // the next frame's pixels travel to registers while this frame is multiplied and pooled
    // a warp pass covers three image rows (30 float4, lanes 30 and 31 idle): their taps differ in ky, so the stores of
    // different rows do not systematically share banks
    float4 pre[kPre];
    const bool lane_on = lane < 3 * kRowVec;
    auto prefetch = [&](int64_t f) {
        const float4* src = reinterpret_cast<const float4*>(vis + f * 3 * kS * kS);
#pragma unroll
        for (int v = 0; v < kPre; ++v)
            if (lane_on) pre[v] = __ldg(src + (warp + v * (kBT / 32)) * (3 * kRowVec) + lane);
    };
    if ((int64_t)blockIdx.x < n) prefetch(blockIdx.x);
    uint32_t phase = 0;
    for (int64_t f = blockIdx.x; f < n; f += gridDim.x) {
        // ---- scatter the frame into the patch matrices (the previous frame's MMAs have completed: every thread waited) ----
#pragma unroll
        for (int v = 0; v < kPre; ++v) {
            const int i = (warp + v * (kBT / 32)) * (3 * kRowVec) + lane;
            if (lane_on) {
                const uint2 o = *reinterpret_cast<const uint2*>(dsto + 4 * i);
                const float x4[4] = {pre[v].x, pre[v].y, pre[v].z, pre[v].w};
                const uint32_t o4[4] = {o.x & 0xffffu, o.x >> 16, o.y & 0xffffu, o.y >> 16};
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const float hi = gn_round_tf32(x4[j]);
                    Bhi[o4[j]] = hi;
                    if (kSplit) Blo[o4[j]] = gn_round_tf32(x4[j] - hi);
                }
            }
        }
        fence_proxy_async();          // generic-proxy stores -> visible to the tensor core's async-proxy reads
        tc_fence_before();            // (and this thread's TMEM reads of the previous frame are ordered before the barrier)
        __syncthreads();
        if (tid == 0) {
            tc_fence_after();
            constexpr uint32_t idesc = make_idesc_tf32(128, kNPix);
            const uint32_t ahi = smem_u32(Ahi), alo = smem_u32(Alo), bhi = smem_u32(Bhi), blo = smem_u32(Blo);
            // hi*hi (+ lo*hi + hi*lo)
#pragma unroll
            for (int p = 0; p < (kSplit ? 3 : 1); ++p) {
                const uint32_t aa = (p == 1) ? alo : ahi, bb = (p == 2) ? blo : bhi;
#pragma unroll
                for (int k4 = 0; k4 < 4; ++k4)
                    umma_tf32(tmem_base, make_desc_sw128(aa + k4 * 32), make_desc_sw128(bb + k4 * 32), idesc,
                              (p > 0 || k4 > 0) ? 1u : 0u);
            }
            umma_commit(smem_u32(bar));
        }
        if (f + gridDim.x < n) prefetch(f + gridDim.x);
        uint32_t* o32 = reinterpret_cast<uint32_t*>(a1p + f * (int64_t)kC1 * kC1 * 64);
        // zero border = conv2's padding
        for (int bi = warp; bi < 4 * kC1 - 4; bi += kBT / 32) {
            int y, x;
            if (bi < kC1) { y = 0; x = bi; }
            else if (bi < 2 * kC1) { y = kC1 - 1; x = bi - kC1; }
            else { const int t = bi - 2 * kC1; y = 1 + (t >> 1); x = (t & 1) ? kC1 - 1 : 0; }
            o32[(y * kC1 + x) * 32 + lane] = 0u;
        }
        mbar_wait(smem_u32(bar), phase);
        phase ^= 1;
        tc_fence_after();
        // ---- epilogue: this thread's channel, conv rows y_begin .. y_end + 1 of D: separable max pool in registers ----
        __nv_bfloat16* ob = a1p + f * (int64_t)kC1 * kC1 * 64 + ch;
        float h0[kP], h1[kP];
#pragma unroll
        for (int i = 0; i < 6; ++i) {
            const int y = y_begin + i;                           // conv row (warp-uniform)
            if (y < y_end + 2) {
                uint32_t r[16];
                tmem_ld16(taddr + (uint32_t)(16 * y), r);
                tmem_ld_wait();
                float h2[kP];
#pragma unroll
                for (int x = 0; x < kP; ++x)
                    h2[x] = fmaxf(fmaxf(__uint_as_float(r[px_col(x)]), __uint_as_float(r[px_col(x + 1)])),
                                  __uint_as_float(r[px_col(x + 2)]));
                if (i >= 2) {
                    __nv_bfloat16* orow = ob + ((y - 1) * kC1 + 1) * 64;       // pooled row y - 2 -> padded row y - 1
#pragma unroll
                    for (int x = 0; x < kP; ++x) {
                        const float m = fmaxf(fmaxf(h0[x], h1[x]), h2[x]);
                        orow[x * 64] = __float2bfloat16(fmaf(fmaxf(m + bias, 0.f), scale, shift));
                    }
                }
#pragma unroll
                for (int x = 0; x < kP; ++x) { h0[x] = h1[x]; h1[x] = h2[x]; }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) {
        tc_fence_after();
        tmem_dealloc(tmem_base, kTmemColsB1);
    }
}


// =====================================================================================================================
// AudBl conv1 (conv1d 30 -> 64, k3 s2 p1) + ReLU on the tensor cores (utils.py:216-217), for the reference's 30 time
// bins (15 output positions).  The CUDA-core kernel (gn_front.cu) is bound by the LSU: every (frame pair, position)
// task of a warp reads its 90 inputs as 45 broadcast 16-byte loads, two wavefronts each (ncu: LSU data pipe 67 %).
// Here 8 frames are scattered into an im2col matrix (row = 16 frame + position, K = ci*3 + k padded 90 -> 96: three
// 128-byte-swizzled k-blocks; an input lands in one or two rows because the windows overlap) and multiplied by the
// weights in one tcgen05 kind::tf32 GEMM group, M = 128, N = 64, K = 96.
// Precision: the raw MFCCs span +-600 and plain TF32 operands would cost 1.5e-4 of the 1e-3 score tolerance (oracle
// experiment), so BOTH operands are split x = hi + lo (hi = tf32(x), lo = tf32(x - hi)) and three groups accumulate
// hi*hi + lo*hi + hi*lo (the dropped lo*lo term is 2^-22 relative): fp32-equivalent, as the CUDA-core kernel.
// Epilogue: lane = (frame, position) row of D, 32 channels per warp: + bias, ReLU, TF32 rounding (the next GEMM's operand
// format), written as (position, channel) rows with the zero row on each side that the tensor-core conv2 reads.
// =====================================================================================================================
constexpr int kAB = 30;                    // time bins (utils.py:313: bin_length)
constexpr int kAL1 = 15;                   // conv1 output positions
constexpr int kAFr = 8;                    // frames per M tile: 8 x 16 rows
constexpr int kAIn = 30 * kAB;             // inputs per frame
constexpr int kAT = 256;                   // threads
constexpr int kAPre = (kAFr * kAIn + kAT - 1) / kAT;
constexpr int kAMat = 3 * 128 * 32;        // floats of one im2col matrix (3 k-blocks x 128 rows x 32)
constexpr int kAWMat = 3 * 64 * 32;        // floats of one weight matrix
constexpr size_t kAudSmem = 1024 + sizeof(float) * (2 * kAMat + 2 * kAWMat) + sizeof(uint32_t) * kAIn + sizeof(float) * 64 + 64;

__global__ void __launch_bounds__(kAT, 1)
audbl_conv1_tc_kernel(const float* __restrict__ aud, int64_t n, const float* __restrict__ w1 /*[64][30][3]*/,
                      const float* __restrict__ b1, float* __restrict__ ac1 /*[n][17][64]*/) {
    extern __shared__ unsigned char smem_raw[];
    unsigned char* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    float* Ahi = reinterpret_cast<float*>(smem);               // [3][128][32] im2col, hi parts
    float* Alo = Ahi + kAMat;
    float* Whi = Alo + kAMat;                                  // [3][64][32] weights
    float* Wlo = Whi + kAWMat;
    uint32_t* dst2 = reinterpret_cast<uint32_t*>(Wlo + kAWMat);    // [900] two 16-bit im2col offsets per input (0xffff = none)
    float* sbias = reinterpret_cast<float*>(dst2 + kAIn);      // [64]
    uint64_t* bar = reinterpret_cast<uint64_t*>(sbias + 64);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar + 1);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    // ---- once per CTA ----
    for (int i = tid; i < 2 * kAMat; i += kAT) Ahi[i] = 0.f;   // padding (t = -1), pad rows (position 15) and k = 90..95 stay zero
    for (int i = tid; i < 64 * 96; i += kAT) {
        const int co = i / 96, kk = i - co * 96;
        const float v = kk < 90 ? __ldg(w1 + co * 90 + kk) : 0.f;
        const float hi = gn_round_tf32(v);
        const int o = (kk >> 5) * (64 * 32) + sw_index(co, kk & 31);
        Whi[o] = hi;
        Wlo[o] = gn_round_tf32(v - hi);
    }
    for (int e = tid; e < kAIn; e += kAT) {
        // input (ci, t) is tap k of position l where t = 2 l - 1 + k: t even -> (t/2, 1); t odd -> ((t+1)/2, 0) and ((t-1)/2, 2)
        const int ci = e / kAB, t = e - ci * kAB;
        auto off = [&](int l, int k) -> uint32_t {
            if (l < 0 || l >= kAL1) return 0xffffu;
            const int kk = ci * 3 + k;
            return (uint32_t)((kk >> 5) * (128 * 32) + sw_index(l, kk & 31));
        };
        const uint32_t d0 = (t & 1) ? off((t + 1) >> 1, 0) : off(t >> 1, 1);
        const uint32_t d1 = (t & 1) ? off((t - 1) >> 1, 2) : 0xffffu;
        dst2[e] = d0 | (d1 << 16);
    }
    if (tid < 64) sbias[tid] = __ldg(b1 + tid);
    if (tid == 0) {
        mbar_init(smem_u32(bar), 1);
        fence_barrier_init();
    }
    if (warp == 0) tmem_alloc(smem_u32(tmem_slot), 64);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    // the next tile's inputs travel to registers while this tile is multiplied and written out
    float pre[kAPre];
    auto prefetch = [&](int64_t f0) {
        const int64_t lim = (n - f0) * kAIn;                   // inputs of frames beyond the batch read as silence
        const float* src = aud + f0 * kAIn;
#pragma unroll
        for (int v = 0; v < kAPre; ++v) {
            const int i = tid + v * kAT;
            pre[v] = (i < kAFr * kAIn && i < lim) ? __ldg(src + i) : 0.f;
        }
    };
    const int64_t tiles = (n + kAFr - 1) / kAFr;
    if ((int64_t)blockIdx.x < tiles) prefetch((int64_t)blockIdx.x * kAFr);
    uint32_t phase = 0;
    for (int64_t tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
        const int64_t f0 = tile * kAFr;
        // ---- scatter (the previous tile's MMAs have completed: every thread waited on the barrier) ----
#pragma unroll
        for (int v = 0; v < kAPre; ++v) {
            const int i = tid + v * kAT;
            if (i < kAFr * kAIn) {
                const int j = i / kAIn, e = i - j * kAIn;
                const uint32_t d = dst2[e];
                const float x = pre[v], hi = gn_round_tf32(x), lo = gn_round_tf32(x - hi);
                const int fo = j * (16 * 32);                  // 16 rows per frame (16 j = 0 mod 8: the swizzle phase is unchanged)
                const uint32_t d0 = d & 0xffffu, d1 = d >> 16;
                if (d0 != 0xffffu) { Ahi[fo + d0] = hi; Alo[fo + d0] = lo; }
                if (d1 != 0xffffu) { Ahi[fo + d1] = hi; Alo[fo + d1] = lo; }
            }
        }
        fence_proxy_async();
        tc_fence_before();
        __syncthreads();
        if (tid == 0) {
            tc_fence_after();
            constexpr uint32_t idesc = make_idesc_tf32(128, 64);
            const uint32_t a_hi = smem_u32(Ahi), a_lo = smem_u32(Alo), w_hi = smem_u32(Whi), w_lo = smem_u32(Wlo);
#pragma unroll
            for (int p = 0; p < 3; ++p) {                       // hi*hi + lo*hi + hi*lo
                const uint32_t aa = (p == 1) ? a_lo : a_hi, ww = (p == 2) ? w_lo : w_hi;
#pragma unroll
                for (int kb = 0; kb < 3; ++kb)
#pragma unroll
                    for (int k8 = 0; k8 < 4; ++k8)
                        umma_tf32(tmem_base, make_desc_sw128(aa + kb * (128 * 128) + k8 * 32),
                                  make_desc_sw128(ww + kb * (64 * 128) + k8 * 32), idesc, (p > 0 || kb > 0 || k8 > 0) ? 1u : 0u);
            }
            umma_commit(smem_u32(bar));
        }
        if (tile + gridDim.x < tiles) prefetch((tile + gridDim.x) * kAFr);
        mbar_wait(smem_u32(bar), phase);
        phase ^= 1;
        tc_fence_after();
        // ---- epilogue: lane = row (frame j, position l) of D; the two warps of a lane quarter take 32 channels each ----
        {
            const int q = warp & 3, ch0 = 32 * (warp >> 2);
            const int r = 32 * q + lane, j = r >> 4, l = r & 15;
            uint32_t acc[32];
            tmem_ld32(tmem_base + ((uint32_t)(32 * q) << 16) + (uint32_t)ch0, acc);
            tmem_ld_wait();
            const int64_t f = f0 + j;
            if (f < n) {
                float* o = ac1 + f * (int64_t)(kAL1 + 2) * 64 + ch0;
                if (l < kAL1) {
                    float4* orow = reinterpret_cast<float4*>(o + (l + 1) * 64);
                    const float4* bb = reinterpret_cast<const float4*>(sbias + ch0);
#pragma unroll
                    for (int v = 0; v < 8; ++v) {
                        const float4 bv = bb[v];
                        orow[v] = make_float4(gn_round_tf32(fmaxf(__uint_as_float(acc[4 * v]) + bv.x, 0.f)),
                                              gn_round_tf32(fmaxf(__uint_as_float(acc[4 * v + 1]) + bv.y, 0.f)),
                                              gn_round_tf32(fmaxf(__uint_as_float(acc[4 * v + 2]) + bv.z, 0.f)),
                                              gn_round_tf32(fmaxf(__uint_as_float(acc[4 * v + 3]) + bv.w, 0.f)));
                    }
                } else {
                    // the pad row's lane writes the zero rows on either side of the frame's positions
                    float4* z0 = reinterpret_cast<float4*>(o);
                    float4* z1 = reinterpret_cast<float4*>(o + (kAL1 + 1) * 64);
#pragma unroll
                    for (int v = 0; v < 8; ++v) { z0[v] = make_float4(0.f, 0.f, 0.f, 0.f); z1[v] = make_float4(0.f, 0.f, 0.f, 0.f); }
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 64);
    }
}

}  // namespace

int gn_launch_visbl_block1_tc(gn_model* m, const float* d_visual, int64_t n, const gn_ws& ws, cudaStream_t st) {
    if (m->S != kS || m->C1 != kC1) return GN_E_SHAPE;
    static const bool split = getenv("GN_BLOCK1_SPLIT") != nullptr;
    auto kern = split ? visbl_block1_tc_kernel<true> : visbl_block1_tc_kernel<false>;
    const size_t smem = b1_smem(split);
    if (gn_attr_first(m->h, reinterpret_cast<const void*>(kern)))
        GN_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int64_t cap = (int64_t)m->h->num_sms * 2;
    const int grid = (int)(n < cap ? n : cap);
    if (grid <= 0) return GN_OK;
    kern<<<grid, kBT, smem, st>>>(d_visual, n, m->conv1_w, m->conv1_b, m->bn_scale[0], m->bn_shift[0], ws.a1p);
    GN_LAUNCHED(m->h);
    return GN_OK;
}

int gn_launch_audbl_conv1_tc(gn_model* m, const float* d_audio, int64_t n, const gn_ws& ws, cudaStream_t st) {
    if (m->B != kAB || m->L1 != kAL1) return GN_E_SHAPE;
    if (gn_attr_first(m->h, reinterpret_cast<const void*>(audbl_conv1_tc_kernel)))
        GN_CUDA(cudaFuncSetAttribute(audbl_conv1_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kAudSmem));
    const int64_t tiles = (n + kAFr - 1) / kAFr;
    const int grid = (int)(tiles < m->h->num_sms ? tiles : m->h->num_sms);
    if (grid <= 0) return GN_OK;
    audbl_conv1_tc_kernel<<<grid, kAT, kAudSmem, st>>>(d_audio, n, m->aconv1_w, m->aconv1_b, ws.ac1);
    GN_LAUNCHED(m->h);
    return GN_OK;
}
