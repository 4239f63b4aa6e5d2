// Tensor-core layers of the AVM forward (reference utils.py:179-191 and 242-258) as persistent,
// warp-specialised tcgen05 kernels:
//   * implicit-GEMM 3x3/s1/p1 convolution (conv2, conv3) with the bias + ReLU + MaxPool(3,1) +
//     BatchNorm(eval) block fused into the epilogue;
//   * K-major GEMM  C = relu(A . W^T + b)  (linear5, fusion.0/3/6), and the same with the last
//     layer (fusion.12), the sigmoid and 4y+1 folded in (fusion.9 -> scores).
// Operands are staged by TMA into 128B-swizzled shared memory, multiplied by tcgen05.mma
// (cta_group::1, M=128) into fp32 accumulators in TMEM (two 256-column buffers so that the epilogue of
// tile i overlaps the MMAs of tile i+1), and read back with tcgen05.ld.  The visual backbone (conv2,
// conv3, linear5: 99 % of the FLOPs) multiplies bf16 operands (kind::f16); the head (AudBl conv2 /
// linear3, fusion MLP) multiplies fp32 operands holding TF32-rounded values (kind::tf32), because its
// bf16 rounding dominated the score error (raw MFCCs span +-600; the last layers amplify).
//
// Warp roles (320 threads): warp 0 = TMA producer, warp 1 = TMEM owner + MMA issuer,
// warps 2..9 = epilogue (TMEM lane quarter = warp_id % 4; the two warps of a quarter split the
// 32-column chunks).  The conv epilogue is the larger half of a conv tile's work, so it gets two
// warps per scheduler and 16-byte shared/global accesses throughout.
#include <stdlib.h>

#include "gn_common.cuh"
#include "gn_tc_ptx.cuh"

namespace {

using namespace tcx;

constexpr int kEpiWarps = 8;             // two epilogue warps per TMEM lane quarter (they split the column chunks)
constexpr int kEpiThreads = kEpiWarps * 32;
constexpr int kThreads = 64 + kEpiThreads;
constexpr int kBK = 64;                  // bf16 elements per k-step = one 128-byte swizzle row
constexpr int kTmemCols = 512;
constexpr int kAccCols = 256;            // columns per accumulator buffer (MT * BN)
constexpr int kCls = 5;                  // classifier head: importance levels 1..5 (utils.py:257, main.py:69,97)
constexpr int kMaxN = 512;               // widest layer: per-channel parameter vectors staged in shared memory

enum Mode { kConv = 0, kLinear = 1, kScore = 2,    // kLinear / kScore: TF32 operands, fp32 (TF32-rounded) out
            // training GEMMs of the fusion MLP (gn_train.cu): all are C = A . B^T with K-major bf16 operands
            kTrainFwd = 3,   // relu(acc + bias), dropout, bf16 out [M][N] and its transpose [N][ld_t]
            kDgrad = 4,      // acc * scale * [gate > 0], bf16 out [M][N] and its transpose
            kWgrad = 5,      // fp32 out [M][N] = acc
            kSplitK = 6,     // fp32 partial sums out_f32[split][M][N] of one K range (bias/ReLU in splitk_reduce_kernel)
            kGemmTF32 = 7 }; // fp32 (TF32) operands, fp32 out [M][N] = acc: the conv GEMMs of the full-AVM trainer

struct ConvGeom {
    int OH, OW;        // conv output == input interior (k3 s1 p1)
    int P;             // pooled size OH-2
    int out_pad;       // 1: write into a (P+2)x(P+2) zero-bordered buffer; 0: dense PxP
    int cin_blocks;    // Cin / 64
    int cout;          // total output channels
    int tiles;         // spatial tiles per frame side (1: the tile is the frame)
    int pfull;         // pooled size of the whole frame (== P when tiles == 1)
    int blocked;       // 1: dense output in the K-blocked linear5 operand layout (gn_a3_off), k = pixel * cout + channel
};

struct GemmArgs {
    int64_t M;               // rows (frames) for linear modes; frames for conv
    int N, K;                // N = total output features, K = reduction length
    const float* bias;       // [N]
    const float* bn_scale;   // conv: [N]
    const float* bn_shift;   // conv: [N]
    __nv_bfloat16* out;      // conv: activation buffer; training modes: [M][ld_out]
    float* out_f;            // kLinear: [M][ld_out] fp32, values rounded to TF32 (the next GEMM's operand)
    int ld_out, col_off;     // linear: leading dimension and first column of the output
    float* out_f32;          // linear: optional fp32 copy [M][N]
    const float* w_last;     // score: fusion.12 weight [128]
    const float* b_last;     // score: fusion.12 bias [1]
    float* scores;           // score: [M] regression scores, or [M][kCls] class probabilities
    int n_cls;               // score: 0/1 = sigmoid regression head (utils.py:255-256,270); kCls = softmax classifier head
    gn_score_sink sink;      // score: n_dst > 0 -> store each score into every destination (peer-mapped) instead of `scores`
    int a3d_rows;            // linear: 0 = A is a 2-D [M][K] map; else A is a 3-D (k, position, frame) map with this many
                             //         positions per frame (AudBl conv2: overlapping 192-element windows, 128/rows frames per tile)
    // training modes
    __nv_bfloat16* out_t;    // transposed bf16 output [N][ld_t] (or null)
    int64_t ld_t;
    const __nv_bfloat16* gate;   // kDgrad: activations [M][N] of the layer below (gradient flows where > 0)
    const uint8_t* mask;     // kTrainFwd: dropout keep-mask [M][N] (1 = keep); null -> counter-based hash of `seed`
    float scale;             // dropout scale 1/(1-p) (1 = no dropout)
    float keep_prob;         // kTrainFwd without mask: P(keep); >= 1 disables dropout
    unsigned long long seed;
    const gn_step_state* step_state;   // kTrainFwd without mask: device-side seed bump (CUDA-graph replay), or null
    int kblocked;            // linear modes: A and W are K-blocked (gn_a3_off / gn_w5_off): every TMA box is contiguous
    int k_splits;            // kSplitK: number of K ranges (1 otherwise)
    int64_t split_stride;    // kSplitK: elements between the partial-sum slabs (M * N)
    ConvGeom g;
};

template <int MT, int BN, int STAGES, int MODE, int SUBC = BN, int STG_ROWS = MT * 128>
struct Cfg {
    static constexpr int kABytes = MT * 128 * 128;          // MT x 128 rows x 128 B
    static constexpr int kBBytes = BN * 128;
    static constexpr int kStageBytes = kABytes + kBBytes;
    // conv: the pooling stage holds SUBC of the tile's BN channels at a time (the epilogue makes BN / SUBC passes);
    // a narrow stage frees shared memory for one more operand stage, which is what hides the TMA round trip
    static constexpr int kStgStride = SUBC * 2 + 16;        // padded bf16 row of the pooling stage
    static constexpr int kStgBytes = (MODE == kConv) ? (((STG_ROWS * kStgStride) + 127) & ~127) : 0;   // one row per conv pixel of the tile
    static constexpr int kParamBytes = 3 * kMaxN * 4;       // bias | bn_scale | bn_shift (or bias | w_last)
    static constexpr int kSmemBytes = 1024 /*align slack*/ + STAGES * kStageBytes + kStgBytes + kParamBytes + 256 /*barriers*/;
    static_assert(MT * BN <= kAccCols, "an accumulator buffer is at most 256 TMEM columns");
    static_assert(MODE == kConv || MT == 1, "linear modes use one 128-row M tile");
    static_assert(kSmemBytes <= 232448, "exceeds the 227 KB shared memory of one CTA");
};

__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
    __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ uint32_t hmax2_u32(uint32_t a, uint32_t b) {
    __nv_bfloat162 r = __hmax2(*reinterpret_cast<__nv_bfloat162*>(&a), *reinterpret_cast<__nv_bfloat162*>(&b));
    return *reinterpret_cast<uint32_t*>(&r);
}
__device__ __forceinline__ uint4 hmax2_u4(uint4 a, uint4 b) {
    return make_uint4(hmax2_u32(a.x, b.x), hmax2_u32(a.y, b.y), hmax2_u32(a.z, b.z), hmax2_u32(a.w, b.w));
}
__device__ __forceinline__ float2 bf2_to_f2(uint32_t v) {
    return __bfloat1622float2(*reinterpret_cast<__nv_bfloat162*>(&v));
}

template <int MT, int BN, int STAGES, int MODE, int OHT, int SUBC = BN>
__global__ void __launch_bounds__(kThreads, 1)
tc_gemm_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
               const GemmArgs args) {
    using C = Cfg<MT, BN, STAGES, MODE, SUBC, (MODE == kConv ? OHT * OHT : MT * 128)>;
    static_assert(BN % SUBC == 0 && SUBC % 64 == 0, "channel passes are whole pairs of 32-column TMEM chunks");
    extern __shared__ unsigned char smem_raw[];
    // 1024-byte alignment: SWIZZLE_128B atoms are 8 rows x 128 B and the swizzle is address based.
    // (offset arithmetic on the shared pointer keeps the address space known to the compiler)
    unsigned char* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    unsigned char* stage_base = smem;
    unsigned char* stg = smem + STAGES * C::kStageBytes;
    float* s_par = reinterpret_cast<float*>(stg + C::kStgBytes);      // [3][kMaxN]
    uint64_t* bars = reinterpret_cast<uint64_t*>(stg + C::kStgBytes + C::kParamBytes);
    uint64_t* full_bar = bars;                    // [STAGES]
    uint64_t* empty_bar = bars + STAGES;          // [STAGES]
    uint64_t* tfull_bar = bars + 2 * STAGES;      // [2]
    uint64_t* tempty_bar = bars + 2 * STAGES + 2; // [2]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * STAGES + 4);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    const int n_blocks = args.N / BN;
    const int tiles_pf = (MODE == kConv) ? args.g.tiles * args.g.tiles : 1;      // spatial tiles per frame
    const int64_t m_blocks = (MODE == kConv) ? args.M * tiles_pf : (args.M + 127) / 128;
    const int k_splits = (MODE == kSplitK || MODE == kGemmTF32) ? (args.k_splits > 1 ? args.k_splits : 1) : 1;
    const int64_t num_tiles = m_blocks * n_blocks * k_splits;
    constexpr bool T32 = (MODE == kGemmTF32 || MODE == kLinear || MODE == kScore);   // fp32 (TF32) operands
    constexpr int kbk = T32 ? 32 : kBK;                      // elements per 128-byte operand row
    const int k_total = args.K / kbk;
    const int k_per = (k_total + k_splits - 1) / k_splits;
    // tile -> (m_blk, n_blk, split): split fastest, then n_blk, so that CTAs running side by side share A and W in L2
#define GN_TILE_DECODE()                                                                  \
    const int64_t mn_ = tile / k_splits;                                                  \
    const int ks = (int)(tile - mn_ * k_splits);                                          \
    const int64_t m_blk = mn_ / n_blocks;                                                 \
    const int n_blk = (int)(mn_ - m_blk * n_blocks);                                      \
    const int k_begin = ks * k_per;                                                       \
    const int k_iters = (k_total - k_begin) < k_per ? (k_total - k_begin) : k_per;

    // per-channel epilogue parameters: global -> shared once per CTA
    for (int i = threadIdx.x; i < args.N && MODE != kDgrad && MODE != kWgrad && MODE != kSplitK && MODE != kGemmTF32; i += kThreads) {
        s_par[i] = __ldg(args.bias + i);
        if (MODE == kConv) {
            s_par[kMaxN + i] = __ldg(args.bn_scale + i);
            s_par[2 * kMaxN + i] = __ldg(args.bn_shift + i);
        } else if (MODE == kScore) {
            s_par[kMaxN + i] = __ldg(args.w_last + i);
        }
    }
    if (MODE == kScore && args.n_cls == kCls)     // rows 1..4 of the [kCls][N] classifier weight (N = 128)
        for (int i = args.N + threadIdx.x; i < kCls * args.N; i += kThreads) s_par[kMaxN + i] = __ldg(args.w_last + i);
    if (threadIdx.x == 0) {
        prefetch_tmap(&tmap_a);
        prefetch_tmap(&tmap_b);
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(smem_u32(&full_bar[s]), 1);
            mbar_init(smem_u32(&empty_bar[s]), 1);
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(smem_u32(&tfull_bar[b]), 1);
            mbar_init(smem_u32(&tempty_bar[b]), kEpiWarps);   // one arrive per epilogue warp
        }
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(smem_u32(tmem_slot), kTmemCols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===================== TMA producer =====================
        if (elect_one()) {
            int stage = 0;
            uint32_t phase = 0;
            for (int64_t tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
                GN_TILE_DECODE();
                int frame = (int)m_blk, tx0 = 0, ty0 = 0;
                if (MODE == kConv && tiles_pf > 1) {
                    frame = (int)(m_blk / tiles_pf);
                    const int t = (int)(m_blk - (int64_t)frame * tiles_pf);
                    const int ty = t / args.g.tiles;
                    ty0 = ty * (OHT - 2);
                    tx0 = (t - ty * args.g.tiles) * (OHT - 2);
                }
                for (int kt = k_begin; kt < k_begin + k_iters; ++kt) {
                    mbar_wait(smem_u32(&empty_bar[stage]), phase ^ 1);
                    const uint32_t fb = smem_u32(&full_bar[stage]);
                    const uint32_t sa = smem_u32(stage_base + stage * C::kStageBytes);
                    const uint32_t sb = sa + C::kABytes;
                    if (MODE == kConv) {
                        const int tap = kt / args.g.cin_blocks;
                        const int cb = kt - tap * args.g.cin_blocks;
                        const int ky = tap / 3, kx = tap - ky * 3;
                        mbar_expect_tx(fb, (uint32_t)(OHT * OHT * 128 + C::kBBytes));
                        // box (64 ch, OW, OH, 1 frame) of the zero-bordered NHWC input, shifted by the tap
                        // (boxes of edge tiles that reach past the frame are zero-filled; those pixels are never pooled)
                        tma_load_4d(sa, &tmap_a, fb, cb * kBK, tx0 + kx, ty0 + ky, frame);
                    } else {
                        mbar_expect_tx(fb, (uint32_t)(C::kABytes + C::kBBytes));
                        if (args.kblocked)      tma_load_2d(sa, &tmap_a, fb, 0, (int)((m_blk * k_total + kt) * 128));
                        else if (args.a3d_rows) tma_load_3d(sa, &tmap_a, fb, kt * kbk, 0, (int)(m_blk * (128 / args.a3d_rows)));
                        else                    tma_load_2d(sa, &tmap_a, fb, kt * kbk, (int)(m_blk * 128));
                    }
                    if (MODE != kConv && args.kblocked) tma_load_2d(sb, &tmap_b, fb, 0, kt * args.N + n_blk * BN);
                    else                                tma_load_2d(sb, &tmap_b, fb, kt * kbk, n_blk * BN);
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        constexpr uint32_t idesc = T32 ? make_idesc_tf32(128, BN) : make_idesc_bf16(128, BN);
        int stage = 0;
        uint32_t phase = 0;
        int ab = 0;
        uint32_t aphase = 0;
        for (int64_t tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
            GN_TILE_DECODE();
            (void)m_blk; (void)n_blk;
            mbar_wait(smem_u32(&tempty_bar[ab]), aphase ^ 1);     // epilogue drained this buffer
            tc_fence_after();
            for (int kt = 0; kt < k_iters; ++kt) {
                mbar_wait(smem_u32(&full_bar[stage]), phase);
                tc_fence_after();
                if (elect_one()) {
                    const uint32_t sa = smem_u32(stage_base + stage * C::kStageBytes);
                    const uint32_t sb = sa + C::kABytes;
#pragma unroll
                    for (int mt = 0; mt < MT; ++mt) {
                        const uint32_t d = tmem_base + (uint32_t)(ab * kAccCols + mt * BN);
#pragma unroll
                        for (int k4 = 0; k4 < kBK / 16; ++k4) {
                            const uint64_t ad = make_desc_sw128(sa + mt * (128 * 128) + k4 * 32);
                            const uint64_t bd = make_desc_sw128(sb + k4 * 32);
                            if (T32) umma_tf32(d, ad, bd, idesc, (kt > 0 || k4 > 0) ? 1u : 0u);
                            else                   umma_bf16(d, ad, bd, idesc, (kt > 0 || k4 > 0) ? 1u : 0u);
                        }
                    }
                    umma_commit(smem_u32(&empty_bar[stage]));              // smem slot free when MMAs retire
                    if (kt == k_iters - 1) umma_commit(smem_u32(&tfull_bar[ab]));  // accumulator ready
                }
                __syncwarp();
                if (++stage == STAGES) { stage = 0; phase ^= 1; }
            }
            if (++ab == 2) { ab = 0; aphase ^= 1; }
        }
    } else {
        // ===================== epilogue (warps 2..9) =====================
        const int q = warp & 3;                     // TMEM lane quarter this warp may read
        const int half = (warp - 2) >> 2;           // which half of the 32-column chunks this warp takes
        const int row = q * 32 + lane;              // accumulator row owned by this thread
        const int et = (warp - 2) * 32 + lane;      // 0..255 among epilogue threads
        const float* s_bias = s_par;
        int ab = 0;
        uint32_t aphase = 0;
        for (int64_t tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
            GN_TILE_DECODE();
            (void)k_iters;
            const int n0 = n_blk * BN;
            mbar_wait(smem_u32(&tfull_bar[ab]), aphase);
            tc_fence_after();
            const uint32_t tbase = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(ab * kAccCols);

            if (MODE == kConv) {
                constexpr int npix = OHT * OHT;
                constexpr int kPasses = BN / SUBC;
#pragma unroll 1
                for (int pass = 0; pass < kPasses; ++pass) {
                const int nsub = n0 + pass * SUBC;       // first channel of this pass
                // ---- phase 1: TMEM -> bf16 staging tile [pixel][SUBC] ----
                // (bias + ReLU are applied AFTER the pool: x -> relu(x + b) is monotone per channel, so
                //  max_window relu(x + b) = relu(max_window x + b); 121 pooled pixels instead of 169 conv pixels, and the
                //  per-thread channel group of phase 2 keeps its bias in registers)
#pragma unroll 1
                for (int mt = 0; mt < MT; ++mt) {
                    const int pix = mt * 128 + row;
                    if (mt * 128 + q * 32 >= npix) continue;      // (warp-uniform) this warp's 32 rows are all padding
#pragma unroll 1
                    for (int c = half; c < SUBC / 32; c += 2) {
                        uint32_t r[32];
                        tmem_ld32(tbase + (uint32_t)(mt * BN + pass * SUBC + c * 32), r);
                        tmem_ld_wait();
                        if (pix < npix) {
                            uint4* dst = reinterpret_cast<uint4*>(stg + (size_t)pix * C::kStgStride + c * 64);
#pragma unroll
                            for (int v = 0; v < 4; ++v) {
                                const int i = v * 8;
                                dst[v] = make_uint4(pack_bf16(__uint_as_float(r[i]), __uint_as_float(r[i + 1])),
                                                    pack_bf16(__uint_as_float(r[i + 2]), __uint_as_float(r[i + 3])),
                                                    pack_bf16(__uint_as_float(r[i + 4]), __uint_as_float(r[i + 5])),
                                                    pack_bf16(__uint_as_float(r[i + 6]), __uint_as_float(r[i + 7])));
                            }
                        }
                    }
                }
                if (pass == kPasses - 1) {
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(smem_u32(&tempty_bar[ab]));  // TMEM buffer may be overwritten
                }
                asm volatile("bar.sync 1, %0;" ::"n"(kEpiThreads) : "memory");   // staging tile complete
                // ---- phase 2: MaxPool(3, stride 1) over the ReLU'd tile, BatchNorm affine, bf16 NHWC store ----
                constexpr int P = OHT - 2;
                constexpr int CG = SUBC / 8;
                const int PW = args.g.pfull + 2 * args.g.out_pad;
                int64_t frame = m_blk;
                int gy0 = 0, gx0 = 0;                 // pooled coordinates of this tile's first output
                if (tiles_pf > 1) {
                    frame = m_blk / tiles_pf;
                    const int t = (int)(m_blk - frame * tiles_pf);
                    const int ty = t / args.g.tiles;
                    gy0 = ty * P;
                    gx0 = (t - ty * args.g.tiles) * P;
                }
                // valid outputs of this tile: the whole P x P when the tile is the frame (compile-time: the walk below is then
                // fully unrolled and its loads scheduled ahead), fewer for an edge tile of a tiled frame
                constexpr bool kWhole = (OHT != 16);
                const int vy = kWhole ? P : min(P, args.g.pfull - gy0), vx = kWhole ? P : min(P, args.g.pfull - gx0);
                __nv_bfloat16* obase = args.out + (size_t)frame * PW * PW * args.g.cout + nsub;
                // bias + ReLU + BatchNorm affine + bf16 store of one pooled (pixel, 8-channel group)
                auto emit = [&](int oy, int ox, int cg, uint4 m, const float4& b0, const float4& b1, const float4& sc0,
                                const float4& sc1, const float4& sh0, const float4& sh1) {
                    const float2 f0 = bf2_to_f2(m.x), f1 = bf2_to_f2(m.y), f2 = bf2_to_f2(m.z), f3 = bf2_to_f2(m.w);
                    const uint4 o = make_uint4(
                        pack_bf16(fmaf(fmaxf(f0.x + b0.x, 0.f), sc0.x, sh0.x), fmaf(fmaxf(f0.y + b0.y, 0.f), sc0.y, sh0.y)),
                        pack_bf16(fmaf(fmaxf(f1.x + b0.z, 0.f), sc0.z, sh0.z), fmaf(fmaxf(f1.y + b0.w, 0.f), sc0.w, sh0.w)),
                        pack_bf16(fmaf(fmaxf(f2.x + b1.x, 0.f), sc1.x, sh1.x), fmaf(fmaxf(f2.y + b1.y, 0.f), sc1.y, sh1.y)),
                        pack_bf16(fmaf(fmaxf(f3.x + b1.z, 0.f), sc1.z, sh1.z), fmaf(fmaxf(f3.y + b1.w, 0.f), sc1.w, sh1.w)));
                    const int opix = (gy0 + oy + args.g.out_pad) * PW + gx0 + ox + args.g.out_pad;
                    if (args.g.blocked)
                        *reinterpret_cast<uint4*>(args.out + gn_a3_off(frame, (int64_t)opix * args.g.cout + nsub + cg * 8,
                                                                       (int64_t)PW * PW * args.g.cout / 64)) = o;
                    else
                        *reinterpret_cast<uint4*>(obase + (size_t)opix * args.g.cout + cg * 8) = o;
                };
                if constexpr (P * CG <= kEpiThreads) {
                    // separable pooling down a column strip: thread = (pooled column, channel group); per conv row one
                    // 3-wide horizontal max (3 loads), per pooled row the max of the last three of those
                    const int ox = et / CG, cg = et - ox * CG;
                    if (ox < vx) {
                        // the thread's channel group is fixed: its BatchNorm parameters are loaded once per tile
                        const float4* sc = reinterpret_cast<const float4*>(s_par + kMaxN + nsub + cg * 8);
                        const float4* sh = reinterpret_cast<const float4*>(s_par + 2 * kMaxN + nsub + cg * 8);
                        const float4 sc0 = sc[0], sc1 = sc[1], sh0 = sh[0], sh1 = sh[1];
                        const float4* bb = reinterpret_cast<const float4*>(s_bias + nsub + cg * 8);
                        const float4 b0 = bb[0], b1 = bb[1];
                        const unsigned char* s0 = stg + (size_t)ox * C::kStgStride + cg * 16;
                        uint4 h0 = make_uint4(0u, 0u, 0u, 0u), h1 = h0;
                        // the three taps of the NEXT conv row are loaded before this row's maxima are combined and stored
                        uint4 t0 = *reinterpret_cast<const uint4*>(s0), t1 = *reinterpret_cast<const uint4*>(s0 + C::kStgStride),
                              t2 = *reinterpret_cast<const uint4*>(s0 + 2 * C::kStgStride);
#pragma unroll
                        for (int r = 0; r < OHT; ++r) {
                            if (r < vy + 2) {
                                uint4 n0 = t0, n1 = t1, n2 = t2;
                                if (r + 1 < OHT && r + 1 < vy + 2) {
                                    const unsigned char* sr = s0 + (size_t)((r + 1) * OHT) * C::kStgStride;
                                    n0 = *reinterpret_cast<const uint4*>(sr);
                                    n1 = *reinterpret_cast<const uint4*>(sr + C::kStgStride);
                                    n2 = *reinterpret_cast<const uint4*>(sr + 2 * C::kStgStride);
                                }
                                const uint4 h2 = hmax2_u4(hmax2_u4(t0, t1), t2);
                                if (r >= 2) emit(r - 2, ox, cg, hmax2_u4(hmax2_u4(h0, h1), h2), b0, b1, sc0, sc1, sh0, sh1);
                                h0 = h1;
                                h1 = h2;
                                t0 = n0; t1 = n1; t2 = n2;
                            }
                        }
                    }
                } else {
                    // item = (pooled pixel, group of 8 channels): 9 x 16-byte shared loads, one 16-byte global store
                    for (int item = et; item < P * P * CG; item += kEpiThreads) {
                        const int op = item / CG, cg = item - op * CG;
                        const int oy = op / P, ox = op - oy * P;
                        if (oy >= vy || ox >= vx) continue;
                        const unsigned char* s0 = stg + (size_t)(oy * OHT + ox) * C::kStgStride + cg * 16;
                        uint4 m = *reinterpret_cast<const uint4*>(s0);
#pragma unroll
                        for (int dy = 0; dy < 3; ++dy)
#pragma unroll
                            for (int dx = 0; dx < 3; ++dx) {
                                if (dy == 0 && dx == 0) continue;
                                m = hmax2_u4(m, *reinterpret_cast<const uint4*>(s0 + (size_t)(dy * OHT + dx) * C::kStgStride));
                            }
                        const float4* sc = reinterpret_cast<const float4*>(s_par + kMaxN + nsub + cg * 8);
                        const float4* sh = reinterpret_cast<const float4*>(s_par + 2 * kMaxN + nsub + cg * 8);
                        const float4* bb = reinterpret_cast<const float4*>(s_bias + nsub + cg * 8);
                        emit(oy, ox, cg, m, bb[0], bb[1], sc[0], sc[1], sh[0], sh[1]);
                    }
                }
                if (args.g.out_pad && tiles_pf == 1) {   // zero border of the next conv's input (padding = 1);
                                                         // tiled frames: zero_border_kernel does it
                    const int nborder = 4 * PW - 4;
                    for (int item = et; item < nborder * CG; item += kEpiThreads) {
                        const int bi = item / CG, cg = item - bi * CG;
                        int by, bx;
                        if (bi < PW) { by = 0; bx = bi; }
                        else if (bi < 2 * PW) { by = PW - 1; bx = bi - PW; }
                        else { const int t = bi - 2 * PW; by = 1 + (t >> 1); bx = (t & 1) ? PW - 1 : 0; }
                        *reinterpret_cast<uint4*>(obase + (size_t)(by * PW + bx) * args.g.cout + cg * 8) =
                            make_uint4(0u, 0u, 0u, 0u);
                    }
                }
                asm volatile("bar.sync 1, %0;" ::"n"(kEpiThreads) : "memory");   // staging tile may be reused
                }   // channel passes
            } else {
                const int64_t grow = m_blk * 128 + row;
                float dot = 0.f;
                float dotc[kCls - 1];
#pragma unroll
                for (int k = 0; k < kCls - 1; ++k) dotc[k] = 0.f;
                // kScore keeps the whole row in one warp (the last layer is a dot product over the row)
                const int c_begin = (MODE == kScore) ? (half == 0 ? 0 : BN / 32) : half;
                const int c_step = (MODE == kScore) ? 1 : 2;
#pragma unroll 1
                for (int c = c_begin; c < BN / 32; c += c_step) {
                    uint32_t r[32];
                    tmem_ld32(tbase + (uint32_t)(c * 32), r);
                    tmem_ld_wait();
                    const float* bp = s_bias + n0 + c * 32;
                    float x[32];
                    if (MODE == kDgrad || MODE == kWgrad || MODE == kSplitK || MODE == kGemmTF32) {
#pragma unroll
                        for (int i = 0; i < 32; ++i) x[i] = __uint_as_float(r[i]);
                    } else {
#pragma unroll
                        for (int i = 0; i < 32; ++i) x[i] = fmaxf(__uint_as_float(r[i]) + bp[i], 0.f);
                    }
                    if (MODE == kTrainFwd && grow < args.M && args.scale != 1.f) {
                        // Dropout(0.2) in training (utils.py:244-254): keep with probability keep_prob, scale by 1/keep_prob
                        const size_t e0 = (size_t)grow * args.N + n0 + c * 32;
                        if (args.mask) {
                            const uint4* mp = reinterpret_cast<const uint4*>(args.mask + e0);
#pragma unroll
                            for (int v = 0; v < 2; ++v) {
                                const uint4 mv = mp[v];
                                const uint32_t w4[4] = {mv.x, mv.y, mv.z, mv.w};
#pragma unroll
                                for (int i = 0; i < 16; ++i)
                                    x[v * 16 + i] = ((w4[i >> 2] >> ((i & 3) * 8)) & 0xFFu) ? x[v * 16 + i] * args.scale : 0.f;
                            }
                        } else {
                            const unsigned long long seed_now = args.seed + (args.step_state ? args.step_state->seed_bump : 0ull);
#pragma unroll
                            for (int i = 0; i < 32; ++i) {
                                // splitmix64 of (seed, element index) -> 24 uniform bits
                                unsigned long long zz = seed_now + (unsigned long long)(e0 + i) * 0x9E3779B97F4A7C15ull;
                                zz = (zz ^ (zz >> 30)) * 0xBF58476D1CE4E5B9ull;
                                zz = (zz ^ (zz >> 27)) * 0x94D049BB133111EBull;
                                zz ^= zz >> 31;
                                const float u = (float)(unsigned int)(zz >> 40) * (1.f / 16777216.f);
                                x[i] = u < args.keep_prob ? x[i] * args.scale : 0.f;
                            }
                        }
                    }
                    if (MODE == kDgrad && grow < args.M) {
                        const uint4* gp = reinterpret_cast<const uint4*>(args.gate + (size_t)grow * args.N + n0 + c * 32);
#pragma unroll
                        for (int v = 0; v < 4; ++v) {
                            const uint4 gv = gp[v];
                            const uint32_t w4[4] = {gv.x, gv.y, gv.z, gv.w};
#pragma unroll
                            for (int i = 0; i < 8; ++i) {
                                const uint32_t hb = (w4[i >> 1] >> ((i & 1) * 16)) & 0xFFFFu;     // bf16 bits of the activation
                                const bool pos = hb != 0u && (hb & 0x8000u) == 0u;               // > 0 (activations are >= 0)
                                x[v * 8 + i] = pos ? x[v * 8 + i] * args.scale : 0.f;
                            }
                        }
                    }
                    if (MODE == kScore) {
                        const float* wl = s_par + kMaxN + n0 + c * 32;
#pragma unroll
                        for (int i = 0; i < 32; ++i) dot = fmaf(x[i], wl[i], dot);
                        if (args.n_cls == kCls) {
#pragma unroll
                            for (int k = 1; k < kCls; ++k)
#pragma unroll
                                for (int i = 0; i < 32; ++i) dotc[k - 1] = fmaf(x[i], wl[k * args.N + i], dotc[k - 1]);
                        }
                    } else if (grow < args.M) {
                        if (MODE == kLinear) {
                            float4* dst = reinterpret_cast<float4*>(args.out_f + (size_t)grow * args.ld_out + args.col_off +
                                                                    n0 + c * 32);
#pragma unroll
                            for (int v = 0; v < 8; ++v)
                                dst[v] = make_float4(gn_round_tf32(x[v * 4]), gn_round_tf32(x[v * 4 + 1]),
                                                     gn_round_tf32(x[v * 4 + 2]), gn_round_tf32(x[v * 4 + 3]));
                        } else if (MODE != kWgrad && MODE != kSplitK && MODE != kGemmTF32 && args.out) {
                            uint4* dst = reinterpret_cast<uint4*>(args.out + (size_t)grow * args.ld_out + args.col_off +
                                                                  n0 + c * 32);
#pragma unroll
                            for (int v = 0; v < 4; ++v)
                                dst[v] = make_uint4(pack_bf16(x[v * 8], x[v * 8 + 1]), pack_bf16(x[v * 8 + 2], x[v * 8 + 3]),
                                                    pack_bf16(x[v * 8 + 4], x[v * 8 + 5]), pack_bf16(x[v * 8 + 6], x[v * 8 + 7]));
                        }
                        if ((MODE == kTrainFwd || MODE == kDgrad) && args.out_t) {
                            // transposed copy for the weight-gradient GEMM: lanes = consecutive rows -> coalesced columns
                            __nv_bfloat16* tp = args.out_t + (size_t)(n0 + c * 32) * args.ld_t + grow;
#pragma unroll
                            for (int i = 0; i < 32; ++i) tp[(size_t)i * args.ld_t] = __float2bfloat16(x[i]);
                        }
                        if (MODE == kWgrad || MODE == kSplitK || MODE == kGemmTF32) {
                            float4* df = reinterpret_cast<float4*>(args.out_f32 + (size_t)ks * args.split_stride +
                                                                   (size_t)grow * args.N + n0 + c * 32);
#pragma unroll
                            for (int v = 0; v < 8; ++v) df[v] = make_float4(x[v * 4], x[v * 4 + 1], x[v * 4 + 2], x[v * 4 + 3]);
                        } else if (args.out_f32) {
                            float4* df = reinterpret_cast<float4*>(args.out_f32 + (size_t)grow * args.N + n0 + c * 32);
#pragma unroll
                            for (int v = 0; v < 8; ++v) {
                                // the fp32 copy carries the bf16-rounded values the next layer consumes
                                df[v] = make_float4(__bfloat162float(__float2bfloat16(x[v * 4])),
                                                    __bfloat162float(__float2bfloat16(x[v * 4 + 1])),
                                                    __bfloat162float(__float2bfloat16(x[v * 4 + 2])),
                                                    __bfloat162float(__float2bfloat16(x[v * 4 + 3])));
                            }
                        }
                    }
                }
                if (MODE == kScore && half == 0 && grow < args.M) {
                    if (args.n_cls == kCls) {
                        // classifier head: Linear(128, 5) + Softmax(dim = 1) (the reference's commented variant, utils.py:257)
                        float z[kCls];
                        z[0] = dot + __ldg(args.b_last);
#pragma unroll
                        for (int k = 1; k < kCls; ++k) z[k] = dotc[k - 1] + __ldg(args.b_last + k);
                        float zmax = z[0];
#pragma unroll
                        for (int k = 1; k < kCls; ++k) zmax = fmaxf(zmax, z[k]);
                        float sum = 0.f;
#pragma unroll
                        for (int k = 0; k < kCls; ++k) { z[k] = expf(z[k] - zmax); sum += z[k]; }
                        const float inv = 1.f / sum;
#pragma unroll
                        for (int k = 0; k < kCls; ++k) args.scores[grow * kCls + k] = z[k] * inv;
                    } else {
                        const float zlin = dot + __ldg(args.b_last);
                        const float sc = 4.f * (1.f / (1.f + expf(-zlin))) + 1.f;      // utils.py:256,270
                        if (args.sink.n_dst > 0) {
                            // all-gather fused into the epilogue: the score goes straight into every rank's (M, n) score
                            // matrix (lanes hold consecutive rows -> 128-byte runs per destination over NVLink)
                            const int64_t i = args.sink.row0 + grow;
                            const int64_t mt = i / args.sink.rows_per_match;
                            const int64_t idx = mt * args.sink.dst_match_stride + args.sink.dst_offset +
                                                (i - mt * args.sink.rows_per_match);
                            for (int p = 0; p < args.sink.n_dst; ++p) args.sink.dst[p][idx] = sc;
                        } else {
                            args.scores[grow] = sc;
                        }
                    }
                }
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(smem_u32(&tempty_bar[ab]));
            }
            if (++ab == 2) { ab = 0; aphase ^= 1; }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, kTmemCols);
    }
}

// =====================================================================================================================
// conv block on CTA PAIRS (tcgen05 cta_group::2): the frame-aligned conv tiles of the 40x40 model, two frames per MMA.
//
// The cta_group::1 kernel above keeps the tensor pipe "active" 95 % of the time on conv3 yet delivers only ~65 % of the
// nominal MMA rate: every MMA reads its whole B tile (256 weight rows) from this SM's shared memory, TMA writes the same
// bytes in, and together with the A tile that saturates the shared-memory port (measured: skipping the pooling pass --
// 6 % of the shared-memory traffic -- speeds the kernel up 18 %).  A CTA pair halves the B side: the two CTAs of a
// cluster sit on one TPC, each stages ITS frame's A tile (128 rows) and HALF of the weight rows, and the leader issues
// one M = 256 MMA whose result rows 0..127 / 128..255 land in the two CTAs' own TMEM.  Per CTA and k-step that is
// 16 KB + 16 KB instead of 16 KB + 32 KB fetched, and 8 KB instead of 12 KB read per MMA; the shorter stage also buys
// two more pipeline stages.  Roles per CTA are as above (TMA producer, MMA issuer -- leader only --, 8 epilogue warps);
// what crosses the pair: the producers count their bytes on the LEADER's full barriers, the MMA commits are multicast to
// both CTAs' empty / accumulator-ready barriers, and both epilogues release the accumulator on the leader's barrier.
// =====================================================================================================================
template <int MT, int BN, int STAGES, int OHT, int SUBC>
struct PairCfg {
    static constexpr int kABytes = MT * 128 * 128;       // this CTA's frame: MT x 128 rows x 64 k
    static constexpr int kBBytes = (BN / 2) * 128;       // this CTA's half of the tile's BN weight rows
    static constexpr int kStageBytes = kABytes + kBBytes;
    static constexpr int kStgStride = SUBC * 2 + 16;
    static constexpr int kStgBytes = ((OHT * OHT * kStgStride) + 127) & ~127;
    static constexpr int kParamBytes = 3 * kMaxN * 4;
    static constexpr int kSmemBytes = 1024 + STAGES * kStageBytes + kStgBytes + kParamBytes + 256;
    static_assert(MT * BN <= kAccCols && BN % 32 == 0 && OHT * OHT <= MT * 128, "MT 128-row M tiles per CTA, <= 256 columns");
    static_assert(kSmemBytes <= 232448, "exceeds the 227 KB shared memory of one CTA");
};

template <int MT, int BN, int STAGES, int OHT, int SUBC>
__global__ void __launch_bounds__(kThreads, 1)
tc_conv_pair_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                    const GemmArgs args) {
    using C = PairCfg<MT, BN, STAGES, OHT, SUBC>;
    static_assert(BN % SUBC == 0 && SUBC % 64 == 0, "channel passes are whole pairs of 32-column TMEM chunks");
    extern __shared__ unsigned char smem_raw[];
    unsigned char* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    unsigned char* stage_base = smem;
    unsigned char* stg = smem + STAGES * C::kStageBytes;
    float* s_par = reinterpret_cast<float*>(stg + C::kStgBytes);      // [3][kMaxN]
    uint64_t* bars = reinterpret_cast<uint64_t*>(stg + C::kStgBytes + C::kParamBytes);
    uint64_t* full_bar = bars;                    // [STAGES]  (the leader's are the ones in use)
    uint64_t* empty_bar = bars + STAGES;          // [STAGES]
    uint64_t* tfull_bar = bars + 2 * STAGES;      // [2]
    uint64_t* tempty_bar = bars + 2 * STAGES + 2; // [2]       (the leader's are the ones in use)
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * STAGES + 4);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();                  // 0 = leader
    const int n_blocks = args.N / BN;
    const int64_t pairs = (args.M + 1) / 2;
    const int64_t num_units = pairs * n_blocks;
    const int64_t cl = blockIdx.x >> 1, ncl = gridDim.x >> 1;
    const int k_total = args.K / kBK;

    for (int i = threadIdx.x; i < args.N; i += kThreads) {
        s_par[i] = __ldg(args.bias + i);
        s_par[kMaxN + i] = __ldg(args.bn_scale + i);
        s_par[2 * kMaxN + i] = __ldg(args.bn_shift + i);
    }
    if (threadIdx.x == 0) {
        prefetch_tmap(&tmap_a);
        prefetch_tmap(&tmap_b);
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(smem_u32(&full_bar[s]), 1);             // the leader's arrive.expect_tx (it expects BOTH CTAs' bytes)
            mbar_init(smem_u32(&empty_bar[s]), 1);            // the multicast MMA commit
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(smem_u32(&tfull_bar[b]), 1);            // the multicast MMA commit
            mbar_init(smem_u32(&tempty_bar[b]), 2 * kEpiWarps);   // every epilogue warp of both CTAs
        }
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc_pair(smem_u32(tmem_slot), kTmemCols);    // all 512 columns, the same base in both CTAs
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();                                       // the peer's barriers are initialised before any remote arrive
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===================== TMA producer (both CTAs) =====================
        if (elect_one()) {
            int stage = 0;
            uint32_t phase = 0;
            for (int64_t unit = cl; unit < num_units; unit += ncl) {
                const int64_t pair = unit / n_blocks;
                const int n_blk = (int)(unit - pair * n_blocks);
                const int frame = (int)(2 * pair + rank);     // beyond the last frame: the box is zero-filled
                for (int kt = 0; kt < k_total; ++kt) {
                    mbar_wait(smem_u32(&empty_bar[stage]), phase ^ 1);
                    const uint32_t fb = mapa(smem_u32(&full_bar[stage]), 0);       // the leader's barrier
                    const uint32_t sa = smem_u32(stage_base + stage * C::kStageBytes);
                    const uint32_t sb = sa + C::kABytes;
                    const int tap = kt / args.g.cin_blocks;
                    const int cb = kt - tap * args.g.cin_blocks;
                    const int ky = tap / 3, kx = tap - ky * 3;
                    // the leader alone arms its barrier, with the bytes of both CTAs' loads: no remote arrive (a
                    // cluster-scope release costs the producer a MEMBAR + ERRBAR per k-step: it starved the MMAs).  The
                    // peer's bytes may land before the leader has armed the phase; the transaction count is signed and
                    // the phase cannot complete before the leader's arrival.
                    if (rank == 0) mbar_expect_tx(smem_u32(&full_bar[stage]), 2u * (uint32_t)(OHT * OHT * 128 + C::kBBytes));
                    tma_load_4d_pair(sa, &tmap_a, fb, cb * kBK, kx, ky, frame);
                    tma_load_2d_pair(sb, &tmap_b, fb, kt * kBK, n_blk * BN + (int)rank * (BN / 2));
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer (leader only) =====================
        if (rank == 0) {
            constexpr uint32_t idesc = make_idesc_bf16(256, BN);
            int stage = 0;
            uint32_t phase = 0;
            int ab = 0;
            uint32_t aphase = 0;
            for (int64_t unit = cl; unit < num_units; unit += ncl) {
                mbar_wait(smem_u32(&tempty_bar[ab]), aphase ^ 1);                  // both epilogues drained this buffer
                tc_fence_after();
                for (int kt = 0; kt < k_total; ++kt) {
                    mbar_wait(smem_u32(&full_bar[stage]), phase);                  // both CTAs' operands have landed
                    tc_fence_after();
                    if (elect_one()) {
                        const uint32_t sa = smem_u32(stage_base + stage * C::kStageBytes);
                        const uint32_t sb = sa + C::kABytes;
#pragma unroll
                        for (int mt = 0; mt < MT; ++mt) {
                            const uint32_t d = tmem_base + (uint32_t)(ab * kAccCols + mt * BN);
#pragma unroll
                            for (int k4 = 0; k4 < kBK / 16; ++k4)
                                umma_bf16_pair(d, make_desc_sw128(sa + mt * (128 * 128) + k4 * 32), make_desc_sw128(sb + k4 * 32),
                                               idesc, (kt > 0 || k4 > 0) ? 1u : 0u);
                        }
                        umma_commit_pair(smem_u32(&empty_bar[stage]));             // both CTAs' stage is free
                        if (kt == k_total - 1) umma_commit_pair(smem_u32(&tfull_bar[ab]));   // both accumulators ready
                    }
                    __syncwarp();
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
                if (++ab == 2) { ab = 0; aphase ^= 1; }
            }
        }
    } else {
        // ===================== epilogue (warps 2..9, both CTAs: each pools and stores ITS frame) =====================
        const int q = warp & 3;
        const int half = (warp - 2) >> 2;
        const int row = q * 32 + lane;
        const int et = (warp - 2) * 32 + lane;
        const float* s_bias = s_par;
        int ab = 0;
        uint32_t aphase = 0;
        constexpr int npix = OHT * OHT;
        constexpr int kPasses = BN / SUBC;
        constexpr int P = OHT - 2;
        constexpr int CG = SUBC / 8;
        static_assert(P * CG <= kEpiThreads, "one thread per (pooled column, channel group)");
        for (int64_t unit = cl; unit < num_units; unit += ncl) {
            const int64_t pair = unit / n_blocks;
            const int n_blk = (int)(unit - pair * n_blocks);
            const int64_t frame = 2 * pair + rank;
            const bool live = frame < args.M;
            const int n0 = n_blk * BN;
            mbar_wait(smem_u32(&tfull_bar[ab]), aphase);
            tc_fence_after();
            const uint32_t tbase = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(ab * kAccCols);
#pragma unroll 1
            for (int pass = 0; pass < kPasses; ++pass) {
                const int nsub = n0 + pass * SUBC;
                // ---- phase 1: TMEM -> bf16 staging tile [pixel][SUBC] (bias + ReLU follow the pool, see above) ----
#pragma unroll 1
                for (int mt = 0; mt < MT; ++mt) {
                    if (mt * 128 + q * 32 >= npix) continue;      // (warp-uniform) this warp's 32 rows are all padding
                    const int pix = mt * 128 + row;
#pragma unroll 1
                    for (int c = half; c < SUBC / 32; c += 2) {
                        uint32_t r[32];
                        tmem_ld32(tbase + (uint32_t)(mt * BN + pass * SUBC + c * 32), r);
                        tmem_ld_wait();
                        if (pix < npix) {
                            uint4* dst = reinterpret_cast<uint4*>(stg + (size_t)pix * C::kStgStride + c * 64);
#pragma unroll
                            for (int v = 0; v < 4; ++v) {
                                const int i = v * 8;
                                dst[v] = make_uint4(pack_bf16(__uint_as_float(r[i]), __uint_as_float(r[i + 1])),
                                                    pack_bf16(__uint_as_float(r[i + 2]), __uint_as_float(r[i + 3])),
                                                    pack_bf16(__uint_as_float(r[i + 4]), __uint_as_float(r[i + 5])),
                                                    pack_bf16(__uint_as_float(r[i + 6]), __uint_as_float(r[i + 7])));
                            }
                        }
                    }
                }
                if (pass == kPasses - 1) {
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive_cluster(mapa(smem_u32(&tempty_bar[ab]), 0));   // on the leader's barrier
                }
                asm volatile("bar.sync 1, %0;" ::"n"(kEpiThreads) : "memory");   // staging tile complete
                // ---- phase 2: MaxPool(3, stride 1), bias + ReLU, BatchNorm affine, bf16 store ----
                const int PW = P + 2 * args.g.out_pad;
                __nv_bfloat16* obase = args.out + (size_t)frame * PW * PW * args.g.cout + nsub;
                const int ox = et / CG, cg = et - ox * CG;
                if (ox < P && live) {
                    const float4* sc = reinterpret_cast<const float4*>(s_par + kMaxN + nsub + cg * 8);
                    const float4* sh = reinterpret_cast<const float4*>(s_par + 2 * kMaxN + nsub + cg * 8);
                    const float4* bb = reinterpret_cast<const float4*>(s_bias + nsub + cg * 8);
                    const float4 sc0 = sc[0], sc1 = sc[1], sh0 = sh[0], sh1 = sh[1], b0 = bb[0], b1 = bb[1];
                    const unsigned char* s0 = stg + (size_t)ox * C::kStgStride + cg * 16;
                    uint4 h0 = make_uint4(0u, 0u, 0u, 0u), h1 = h0;
                    uint4 t0 = *reinterpret_cast<const uint4*>(s0), t1 = *reinterpret_cast<const uint4*>(s0 + C::kStgStride),
                          t2 = *reinterpret_cast<const uint4*>(s0 + 2 * C::kStgStride);
#pragma unroll
                    for (int r = 0; r < OHT; ++r) {
                        uint4 n0v = t0, n1v = t1, n2v = t2;
                        if (r + 1 < OHT) {
                            const unsigned char* sr = s0 + (size_t)((r + 1) * OHT) * C::kStgStride;
                            n0v = *reinterpret_cast<const uint4*>(sr);
                            n1v = *reinterpret_cast<const uint4*>(sr + C::kStgStride);
                            n2v = *reinterpret_cast<const uint4*>(sr + 2 * C::kStgStride);
                        }
                        const uint4 h2 = hmax2_u4(hmax2_u4(t0, t1), t2);
                        if (r >= 2) {
                            const uint4 m = hmax2_u4(hmax2_u4(h0, h1), h2);
                            const float2 f0 = bf2_to_f2(m.x), f1 = bf2_to_f2(m.y), f2 = bf2_to_f2(m.z), f3 = bf2_to_f2(m.w);
                            const uint4 o = make_uint4(
                                pack_bf16(fmaf(fmaxf(f0.x + b0.x, 0.f), sc0.x, sh0.x), fmaf(fmaxf(f0.y + b0.y, 0.f), sc0.y, sh0.y)),
                                pack_bf16(fmaf(fmaxf(f1.x + b0.z, 0.f), sc0.z, sh0.z), fmaf(fmaxf(f1.y + b0.w, 0.f), sc0.w, sh0.w)),
                                pack_bf16(fmaf(fmaxf(f2.x + b1.x, 0.f), sc1.x, sh1.x), fmaf(fmaxf(f2.y + b1.y, 0.f), sc1.y, sh1.y)),
                                pack_bf16(fmaf(fmaxf(f3.x + b1.z, 0.f), sc1.z, sh1.z), fmaf(fmaxf(f3.y + b1.w, 0.f), sc1.w, sh1.w)));
                            const int opix = (r - 2 + args.g.out_pad) * PW + ox + args.g.out_pad;
                            if (args.g.blocked)
                                *reinterpret_cast<uint4*>(args.out + gn_a3_off(frame, (int64_t)opix * args.g.cout + nsub + cg * 8,
                                                                               (int64_t)PW * PW * args.g.cout / 64)) = o;
                            else
                                *reinterpret_cast<uint4*>(obase + (size_t)opix * args.g.cout + cg * 8) = o;
                        }
                        h0 = h1;
                        h1 = h2;
                        t0 = n0v; t1 = n1v; t2 = n2v;
                    }
                }
                if (args.g.out_pad && live) {              // zero border of the next conv's input (padding = 1)
                    const int nborder = 4 * PW - 4;
                    for (int item = et; item < nborder * CG; item += kEpiThreads) {
                        const int bi = item / CG, cgb = item - bi * CG;
                        int by, bx;
                        if (bi < PW) { by = 0; bx = bi; }
                        else if (bi < 2 * PW) { by = PW - 1; bx = bi - PW; }
                        else { const int tt = bi - 2 * PW; by = 1 + (tt >> 1); bx = (tt & 1) ? PW - 1 : 0; }
                        *reinterpret_cast<uint4*>(obase + (size_t)(by * PW + bx) * args.g.cout + cgb * 8) =
                            make_uint4(0u, 0u, 0u, 0u);
                    }
                }
                asm volatile("bar.sync 1, %0;" ::"n"(kEpiThreads) : "memory");   // staging tile may be reused
            }
            if (++ab == 2) { ab = 0; aphase ^= 1; }
        }
    }

    tc_fence_before();
    __syncthreads();
    cluster_sync_all();        // no CTA leaves while its peer may still arrive on its barriers or the leader's MMAs write its TMEM
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc_pair(tmem_base, kTmemCols);
    }
}

// =====================================================================================================================
// linear5 on CTA pairs: the split-K GEMM of run_linear_splitk with one M = 256 MMA per k-step over two 128-row blocks.
// Each CTA of the pair stages ITS row block's A box and HALF of the tile's BN weight rows (32 KB per stage instead of
// 48: the single-CTA kernel read 48 KB of operands from its own shared memory per 128-cycle MMA, more than the port
// delivers).  Same K ranges and the same partial-sum slabs as the single-CTA kernel: the sums a row receives do not
// change (an accumulator row still adds the same products in the same order), so batch invariance is untouched.
// =====================================================================================================================
template <int BN, int STAGES>
struct PairLinCfg {
    static constexpr int kABytes = 128 * 128;            // this CTA's 128 rows x 64 k
    static constexpr int kBBytes = (BN / 2) * 128;       // this CTA's half of the weight rows
    static constexpr int kStageBytes = kABytes + kBBytes;
    static constexpr int kSmemBytes = 1024 + STAGES * kStageBytes + 256;
    static_assert(BN <= kAccCols && BN % 64 == 0, "one 128-row tile per CTA, <= 256 columns");
    static_assert(kSmemBytes <= 232448, "exceeds the 227 KB shared memory of one CTA");
};

template <int BN, int STAGES>
__global__ void __launch_bounds__(kThreads, 1)
tc_splitk_pair_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                      const GemmArgs args) {
    using C = PairLinCfg<BN, STAGES>;
    extern __shared__ unsigned char smem_raw[];
    unsigned char* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    unsigned char* stage_base = smem;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + STAGES * C::kStageBytes);
    uint64_t* full_bar = bars;                    // [STAGES]  (the leader's are the ones in use)
    uint64_t* empty_bar = bars + STAGES;          // [STAGES]
    uint64_t* tfull_bar = bars + 2 * STAGES;      // [2]
    uint64_t* tempty_bar = bars + 2 * STAGES + 2; // [2]       (the leader's are the ones in use)
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * STAGES + 4);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();                  // 0 = leader
    const int n_blocks = args.N / BN;
    const int64_t m_blocks = (args.M + 127) / 128;
    const int64_t m_pairs = (m_blocks + 1) / 2;
    const int k_splits = args.k_splits > 1 ? args.k_splits : 1;
    const int64_t num_units = m_pairs * n_blocks * k_splits;
    const int64_t cl = blockIdx.x >> 1, ncl = gridDim.x >> 1;
    const int k_total = args.K / kBK;
    const int k_per = (k_total + k_splits - 1) / k_splits;
    // unit -> (row-block pair, n_blk, split): split fastest, then n_blk (pairs running side by side share A and W in L2)
#define GN_PAIR_UNIT_DECODE()                                                             \
    const int64_t mn_ = unit / k_splits;                                                  \
    const int ks = (int)(unit - mn_ * k_splits);                                          \
    const int64_t mp = mn_ / n_blocks;                                                    \
    const int n_blk = (int)(mn_ - mp * n_blocks);                                         \
    const int64_t m_blk = 2 * mp + rank;                                                  \
    const int k_begin = ks * k_per;                                                       \
    const int k_iters = (k_total - k_begin) < k_per ? (k_total - k_begin) : k_per;

    if (threadIdx.x == 0) {
        prefetch_tmap(&tmap_a);
        prefetch_tmap(&tmap_b);
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(smem_u32(&full_bar[s]), 1);             // the leader's arrive.expect_tx (it expects BOTH CTAs' bytes)
            mbar_init(smem_u32(&empty_bar[s]), 1);            // the multicast MMA commit
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(smem_u32(&tfull_bar[b]), 1);            // the multicast MMA commit
            mbar_init(smem_u32(&tempty_bar[b]), 2 * kEpiWarps);   // every epilogue warp of both CTAs
        }
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc_pair(smem_u32(tmem_slot), kTmemCols);
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();                                       // the peer's barriers are initialised before any remote arrive
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===================== TMA producer (both CTAs) =====================
        if (elect_one()) {
            int stage = 0;
            uint32_t phase = 0;
            for (int64_t unit = cl; unit < num_units; unit += ncl) {
                GN_PAIR_UNIT_DECODE();
                for (int it = 0; it < k_iters; ++it) {
                    const int kt = k_begin + it;
                    mbar_wait(smem_u32(&empty_bar[stage]), phase ^ 1);
                    const uint32_t fb = mapa(smem_u32(&full_bar[stage]), 0);       // the leader's barrier
                    const uint32_t sa = smem_u32(stage_base + stage * C::kStageBytes);
                    const uint32_t sb = sa + C::kABytes;
                    // the leader alone arms its barrier, with the bytes of both CTAs' loads (see tc_conv_pair_kernel)
                    if (rank == 0) mbar_expect_tx(smem_u32(&full_bar[stage]), 2u * (uint32_t)C::kStageBytes);
                    // K-blocked operands: every box is one contiguous chunk.  A row block beyond the last one (odd number
                    // of blocks) lies outside the map and is zero-filled.
                    tma_load_2d_pair(sa, &tmap_a, fb, 0, (int)((m_blk * k_total + kt) * 128));
                    tma_load_2d_pair(sb, &tmap_b, fb, 0, kt * args.N + n_blk * BN + (int)rank * (BN / 2));
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer (leader only) =====================
        if (rank == 0) {
            constexpr uint32_t idesc = make_idesc_bf16(256, BN);
            int stage = 0;
            uint32_t phase = 0;
            int ab = 0;
            uint32_t aphase = 0;
            for (int64_t unit = cl; unit < num_units; unit += ncl) {
                GN_PAIR_UNIT_DECODE();
                (void)m_blk; (void)n_blk;
                mbar_wait(smem_u32(&tempty_bar[ab]), aphase ^ 1);                  // both epilogues drained this buffer
                tc_fence_after();
                for (int it = 0; it < k_iters; ++it) {
                    mbar_wait(smem_u32(&full_bar[stage]), phase);                  // both CTAs' operands have landed
                    tc_fence_after();
                    if (elect_one()) {
                        const uint32_t sa = smem_u32(stage_base + stage * C::kStageBytes);
                        const uint32_t sb = sa + C::kABytes;
                        const uint32_t d = tmem_base + (uint32_t)(ab * kAccCols);
#pragma unroll
                        for (int k4 = 0; k4 < kBK / 16; ++k4)
                            umma_bf16_pair(d, make_desc_sw128(sa + k4 * 32), make_desc_sw128(sb + k4 * 32), idesc,
                                           (it > 0 || k4 > 0) ? 1u : 0u);
                        umma_commit_pair(smem_u32(&empty_bar[stage]));             // both CTAs' stage is free
                        if (it == k_iters - 1) umma_commit_pair(smem_u32(&tfull_bar[ab]));   // both accumulators ready
                    }
                    __syncwarp();
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
                if (++ab == 2) { ab = 0; aphase ^= 1; }
            }
        }
    } else {
        // ===================== epilogue (warps 2..9, both CTAs: each stores ITS row block's partial sums) ============
        const int q = warp & 3;
        const int half = (warp - 2) >> 2;
        const int row = q * 32 + lane;
        int ab = 0;
        uint32_t aphase = 0;
        for (int64_t unit = cl; unit < num_units; unit += ncl) {
            GN_PAIR_UNIT_DECODE();
            (void)k_iters;
            const int n0 = n_blk * BN;
            const int64_t grow = m_blk * 128 + row;
            mbar_wait(smem_u32(&tfull_bar[ab]), aphase);
            tc_fence_after();
            const uint32_t tbase = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(ab * kAccCols);
#pragma unroll 1
            for (int c = half; c < BN / 32; c += 2) {
                uint32_t r[32];
                tmem_ld32(tbase + (uint32_t)(c * 32), r);
                tmem_ld_wait();
                if (grow < args.M) {
                    float4* df = reinterpret_cast<float4*>(args.out_f32 + (size_t)ks * args.split_stride +
                                                           (size_t)grow * args.N + n0 + c * 32);
#pragma unroll
                    for (int v = 0; v < 8; ++v)
                        df[v] = make_float4(__uint_as_float(r[v * 4]), __uint_as_float(r[v * 4 + 1]),
                                            __uint_as_float(r[v * 4 + 2]), __uint_as_float(r[v * 4 + 3]));
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster(mapa(smem_u32(&tempty_bar[ab]), 0));   // on the leader's barrier
            if (++ab == 2) { ab = 0; aphase ^= 1; }
        }
    }
#undef GN_PAIR_UNIT_DECODE

    tc_fence_before();
    __syncthreads();
    cluster_sync_all();        // no CTA leaves while its peer may still arrive on its barriers or the leader's MMAs write its TMEM
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc_pair(tmem_base, kTmemCols);
    }
}

// ---- host side --------------------------------------------------------------------------------------

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int make_tmap(gn_handle* h, CUtensorMap* m, const void* base, int rank, const uint64_t* dims,
              const uint64_t* strides_bytes, const uint32_t* box, bool f32 = false) {
    if (!h->encode_tiled) return GN_E_NODEVICE;
    cuuint64_t gdim[5], gstr[4];
    cuuint32_t bx[5], es[5];
    for (int i = 0; i < rank; ++i) { gdim[i] = dims[i]; bx[i] = box[i]; es[i] = 1; }
    for (int i = 0; i < rank - 1; ++i) gstr[i] = strides_bytes[i];
    CUresult r = reinterpret_cast<EncodeTiledFn>(h->encode_tiled)(
        m, f32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank,
        const_cast<void*>(base), gdim, gstr, bx, es,
        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS ? GN_OK : (int)r;
}

// [rows][K] row-major -> 2-D map whose box is one 128-byte swizzle row wide: (64 bf16 | 32 fp32, box_rows)
int make_tmap_2d(gn_handle* h, CUtensorMap* m, const void* base, int64_t rows, int64_t K, int64_t ld, int box_rows,
                 bool f32 = false) {
    const uint64_t dims[2] = {(uint64_t)K, (uint64_t)rows};
    const uint64_t str[1] = {(uint64_t)ld * (f32 ? 4 : 2)};
    const uint32_t box[2] = {(uint32_t)(f32 ? 32 : kBK), (uint32_t)box_rows};
    return make_tmap(h, m, base, 2, dims, str, box, f32);
}

template <int MT, int BN, int STAGES, int MODE, int OHT, int SUBC = BN>
int launch(gn_handle* h, const CUtensorMap& ta, const CUtensorMap& tb, const GemmArgs& a, int64_t tiles,
           cudaStream_t st) {
    using C = Cfg<MT, BN, STAGES, MODE, SUBC, (MODE == kConv ? OHT * OHT : MT * 128)>;
    if (a.N > kMaxN && MODE != kDgrad && MODE != kWgrad && MODE != kSplitK && MODE != kGemmTF32) return GN_E_SHAPE;
    auto kern = tc_gemm_kernel<MT, BN, STAGES, MODE, OHT, SUBC>;
    if (gn_attr_first(h, reinterpret_cast<const void*>(kern)))
        GN_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::kSmemBytes));
    const int grid = (int)(tiles < h->num_sms ? tiles : h->num_sms);
    if (grid <= 0) return GN_OK;
    kern<<<grid, kThreads, C::kSmemBytes, st>>>(ta, tb, a);
    GN_LAUNCHED(h);
    return GN_OK;
}

// conv block: in = zero-bordered NHWC [n][OH+2][OW+2][Cin]; weights K-major [Cout][9*Cin]
template <int MT, int BN, int STAGES, int OH, int SUBC = BN>
int run_conv(gn_model* m, int64_t n, const __nv_bfloat16* in, int Cin, const __nv_bfloat16* wk, int Cout,
             const float* bias, const float* sc, const float* sh, __nv_bfloat16* out, int out_pad, cudaStream_t st) {
    CUtensorMap ta, tb;
    const int IW = OH + 2;
    const uint64_t dims[4] = {(uint64_t)Cin, (uint64_t)IW, (uint64_t)IW, (uint64_t)n};
    const uint64_t str[3] = {(uint64_t)Cin * 2, (uint64_t)IW * Cin * 2, (uint64_t)IW * IW * Cin * 2};
    const uint32_t box[4] = {(uint32_t)kBK, (uint32_t)OH, (uint32_t)OH, 1};
    int rc = make_tmap(m->h, &ta, in, 4, dims, str, box);
    if (rc) return rc;
    rc = make_tmap_2d(m->h, &tb, wk, Cout, 9 * Cin, 9 * Cin, BN);
    if (rc) return rc;
    GemmArgs a{};
    a.M = n; a.N = Cout; a.K = 9 * Cin;
    a.bias = bias; a.bn_scale = sc; a.bn_shift = sh; a.out = out;
    a.g.OH = OH; a.g.OW = OH; a.g.P = OH - 2; a.g.out_pad = out_pad; a.g.cin_blocks = Cin / kBK; a.g.cout = Cout;
    a.g.tiles = 1; a.g.pfull = OH - 2; a.g.blocked = out_pad ? 0 : 1;
    if (OH * OH > MT * 128 || Cin % kBK || Cout % BN) return GN_E_SHAPE;
    return launch<MT, BN, STAGES, kConv, OH, SUBC>(m->h, ta, tb, a, n * (Cout / BN), st);
}

// conv block on CTA pairs (see tc_conv_pair_kernel): same operands and result as run_conv<1, BN, ...>
template <int MT, int BN, int STAGES, int OH, int SUBC>
int run_conv_pair(gn_model* m, int64_t n, const __nv_bfloat16* in, int Cin, const __nv_bfloat16* wk, int Cout,
                  const float* bias, const float* sc, const float* sh, __nv_bfloat16* out, int out_pad, cudaStream_t st) {
    using C = PairCfg<MT, BN, STAGES, OH, SUBC>;
    if (Cin % kBK || Cout % BN || Cout > kMaxN) return GN_E_SHAPE;
    CUtensorMap ta, tb;
    const int IW = OH + 2;
    const uint64_t dims[4] = {(uint64_t)Cin, (uint64_t)IW, (uint64_t)IW, (uint64_t)n};
    const uint64_t str[3] = {(uint64_t)Cin * 2, (uint64_t)IW * Cin * 2, (uint64_t)IW * IW * Cin * 2};
    const uint32_t box[4] = {(uint32_t)kBK, (uint32_t)OH, (uint32_t)OH, 1};
    int rc = make_tmap(m->h, &ta, in, 4, dims, str, box);
    if (rc) return rc;
    rc = make_tmap_2d(m->h, &tb, wk, Cout, 9 * Cin, 9 * Cin, BN / 2);      // each CTA of a pair fetches half of the rows
    if (rc) return rc;
    GemmArgs a{};
    a.M = n; a.N = Cout; a.K = 9 * Cin;
    a.bias = bias; a.bn_scale = sc; a.bn_shift = sh; a.out = out;
    a.g.OH = OH; a.g.OW = OH; a.g.P = OH - 2; a.g.out_pad = out_pad; a.g.cin_blocks = Cin / kBK; a.g.cout = Cout;
    a.g.tiles = 1; a.g.pfull = OH - 2; a.g.blocked = out_pad ? 0 : 1;
    auto kern = tc_conv_pair_kernel<MT, BN, STAGES, OH, SUBC>;
    if (gn_attr_first(m->h, reinterpret_cast<const void*>(kern)))
        GN_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::kSmemBytes));
    const int64_t units = ((n + 1) / 2) * (Cout / BN);
    int64_t clusters = m->h->num_sms / 2;
    if (clusters > units) clusters = units;
    if (clusters < 1) return GN_OK;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned)(2 * clusters));
    cfg.blockDim = dim3(kThreads);
    cfg.dynamicSmemBytes = C::kSmemBytes;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = 2;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    GN_CUDA(cudaLaunchKernelEx(&cfg, kern, ta, tb, a));
    GN_LAUNCHED(m->h);
    return GN_OK;
}

// zero the one-pixel border of [n][PW][PW][C] bf16 (the padding the next conv reads); 16 bytes per thread
__global__ void __launch_bounds__(256)
zero_border_kernel(__nv_bfloat16* __restrict__ buf, int64_t n, int PW, int C) {
    const int cg = C / 8, nb = 4 * PW - 4;
    const int64_t total = n * nb * cg;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t f = i / ((int64_t)nb * cg);
        const int r = (int)(i - f * nb * cg);
        const int bi = r / cg, g = r - bi * cg;
        int by, bx;
        if (bi < PW) { by = 0; bx = bi; }
        else if (bi < 2 * PW) { by = PW - 1; bx = bi - PW; }
        else { const int t = bi - 2 * PW; by = 1 + (t >> 1); bx = (t & 1) ? PW - 1 : 0; }
        *reinterpret_cast<uint4*>(buf + ((f * PW + by) * PW + bx) * C + g * 8) = make_uint4(0u, 0u, 0u, 0u);
    }
}

// conv block on frames of any size: 16x16 conv-output tiles (two 128-row MMA tiles x 128 channels) that overlap by
// the pooling halo, so each tile yields 14x14 pooled pixels.  in = zero-bordered NHWC [n][OHF+2][OHF+2][Cin].
constexpr int kTileOH = 16;
int run_conv_tiled(gn_model* m, int64_t n, const __nv_bfloat16* in, int OHF, int Cin, const __nv_bfloat16* wk, int Cout,
                   const float* bias, const float* sc, const float* sh, __nv_bfloat16* out, int out_pad,
                   cudaStream_t st) {
    constexpr int BN = 128;
    CUtensorMap ta, tb;
    const int IW = OHF + 2, PF = OHF - 2, TP = kTileOH - 2;
    if (PF < 1 || Cin % kBK || Cout % BN) return GN_E_SHAPE;
    const uint64_t dims[4] = {(uint64_t)Cin, (uint64_t)IW, (uint64_t)IW, (uint64_t)n};
    const uint64_t str[3] = {(uint64_t)Cin * 2, (uint64_t)IW * Cin * 2, (uint64_t)IW * IW * Cin * 2};
    const uint32_t box[4] = {(uint32_t)kBK, (uint32_t)kTileOH, (uint32_t)kTileOH, 1};
    int rc = make_tmap(m->h, &ta, in, 4, dims, str, box);
    if (rc) return rc;
    rc = make_tmap_2d(m->h, &tb, wk, Cout, 9 * Cin, 9 * Cin, BN);
    if (rc) return rc;
    GemmArgs a{};
    a.M = n; a.N = Cout; a.K = 9 * Cin;
    a.bias = bias; a.bn_scale = sc; a.bn_shift = sh; a.out = out;
    a.g.OH = kTileOH; a.g.OW = kTileOH; a.g.P = TP; a.g.out_pad = out_pad; a.g.cin_blocks = Cin / kBK; a.g.cout = Cout;
    a.g.tiles = (PF + TP - 1) / TP; a.g.pfull = PF; a.g.blocked = out_pad ? 0 : 1;
    if (n * a.g.tiles * a.g.tiles > (1LL << 40)) return GN_E_RANGE;
    if (out_pad) {
        rc = gn_launch_zero_border(m->h, out, n, PF + 2, Cout, st);
        if (rc) return rc;
    }
    return launch<2, BN, 3, kConv, kTileOH>(m->h, ta, tb, a, n * a.g.tiles * a.g.tiles * (Cout / BN), st);
}

// out[r][col_off + j] = tf32(relu(sum_s part[s][r][j] + bias[j])) (+ a dense copy of the same values); 4 columns / thread.
// The ranges are added in a fixed order (s = 0, 1, ...), so a row's result does not depend on the batch it is in.
__global__ void __launch_bounds__(256)
splitk_reduce_kernel(const float* __restrict__ part, int splits, int64_t M, int N, const float* __restrict__ bias,
                     float* __restrict__ out, int ld_out, int col_off, float* __restrict__ out_f32) {
    const int64_t total = M * N / 4;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = (i * 4) / N;
        const int j = (int)(i * 4 - r * N);
        float4 acc = *reinterpret_cast<const float4*>(bias + j);
        for (int s = 0; s < splits; ++s) {
            const float4 v = __ldcs(reinterpret_cast<const float4*>(part + (size_t)s * M * N) + i);
            acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
        }
        const float4 o = make_float4(gn_round_tf32(fmaxf(acc.x, 0.f)), gn_round_tf32(fmaxf(acc.y, 0.f)),
                                     gn_round_tf32(fmaxf(acc.z, 0.f)), gn_round_tf32(fmaxf(acc.w, 0.f)));
        *reinterpret_cast<float4*>(out + r * ld_out + col_off + j) = o;
        if (out_f32) *reinterpret_cast<float4*>(out_f32 + r * N + j) = o;
    }
}

// TMA maps of K-blocked operands (gn_a3_off / gn_w5_off): both are [rows'][64] matrices whose boxes are contiguous
int make_tmaps_kblocked(gn_handle* h, CUtensorMap* ta, CUtensorMap* tb, const __nv_bfloat16* A, int64_t M, int64_t K,
                        const __nv_bfloat16* wk, int N, int BN) {
    const int64_t kb = K / kBK, rows_a = (M + 127) / 128 * 128 * kb, rows_b = kb * N;
    if (rows_a >= (1LL << 31) || rows_b >= (1LL << 31)) return GN_E_RANGE;    // TMA coordinates are int32
    int rc = make_tmap_2d(h, ta, A, rows_a, kBK, kBK, 128);
    if (rc) return rc;
    return make_tmap_2d(h, tb, wk, rows_b, kBK, kBK, BN);
}

// linear5: relu(A . W^T + b) with the reduction split into `splits` K ranges, one (tile, range) unit per CTA visit.
// K = 512 * P3^2 is 41 472 at the reference's 40x40 frames and in the millions at large frame sizes, so the
// (M/128) x (N/256) tiles alone rarely fill the GPU in whole waves.  EVERY row takes this path with the SAME ranges
// (gn_linear5_splits depends on K only): partial sums go to fp32 slabs and one reduction pass adds them in a fixed
// order, so a frame's features -- hence its score -- are bit-identical whatever batch, chunk or GPU shard it is
// scored in (what lets the frame-sharded multi-GPU path reproduce the single-GPU bits).
int run_linear_splitk(gn_model* m, int64_t M, const __nv_bfloat16* A, int64_t K, const __nv_bfloat16* wk, int N,
                      const float* bias, int splits, float* part, float* out, int ld_out, int col_off,
                      float* out_f32, cudaStream_t st) {
    constexpr int BN = 256;
    if (K % kBK || N % BN || K > (1LL << 31) - 64 || splits < 1 || !part) return GN_E_SHAPE;
    CUtensorMap ta, tb;
    int rc = make_tmaps_kblocked(m->h, &ta, &tb, A, M, K, wk, N, BN);
    if (rc) return rc;
    GemmArgs a{};
    a.M = M; a.N = N; a.K = (int)K; a.out_f32 = part; a.k_splits = splits; a.split_stride = M * N; a.kblocked = 1;
    static const bool single = getenv("GN_LINEAR5_SINGLE") != nullptr;      // the single-CTA kernel, kept for A/B timing
    if (single) {
        rc = launch<1, BN, 4, kSplitK, 0>(m->h, ta, tb, a, ((M + 127) / 128) * (N / BN) * splits, st);
        if (rc) return rc;
    } else {
        // CTA pairs: two 128-row blocks per MMA, each CTA fetches half of the weight rows
        constexpr int kStages = 6;
        using C = PairLinCfg<BN, kStages>;
        rc = make_tmaps_kblocked(m->h, &ta, &tb, A, M, K, wk, N, BN / 2);
        if (rc) return rc;
        auto kern = tc_splitk_pair_kernel<BN, kStages>;
        if (gn_attr_first(m->h, reinterpret_cast<const void*>(kern)))
            GN_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::kSmemBytes));
        const int64_t units = (((M + 127) / 128 + 1) / 2) * (N / BN) * splits;
        int64_t clusters = m->h->num_sms / 2;
        if (clusters > units) clusters = units;
        if (clusters >= 1) {
            cudaLaunchConfig_t cfg{};
            cfg.gridDim = dim3((unsigned)(2 * clusters));
            cfg.blockDim = dim3(kThreads);
            cfg.dynamicSmemBytes = C::kSmemBytes;
            cfg.stream = st;
            cudaLaunchAttribute at[1];
            at[0].id = cudaLaunchAttributeClusterDimension;
            at[0].val.clusterDim.x = 2;
            at[0].val.clusterDim.y = 1;
            at[0].val.clusterDim.z = 1;
            cfg.attrs = at;
            cfg.numAttrs = 1;
            GN_CUDA(cudaLaunchKernelEx(&cfg, kern, ta, tb, a));
            GN_LAUNCHED(m->h);
        }
    }
    const int64_t blocks = (M * N / 4 + 255) / 256, cap = (int64_t)m->h->num_sms * 8;
    splitk_reduce_kernel<<<(int)(blocks < cap ? blocks : cap), 256, 0, st>>>(part, splits, M, N, bias, out, ld_out,
                                                                             col_off, out_f32);
    GN_LAUNCHED(m->h);
    return GN_OK;
}

// head layers: out = tf32(relu(A . W^T + b)) (kLinear) or the score epilogue (kScore); A [M][K] and W [N][K] are fp32
// arrays of TF32-rounded values
template <int BN, int STAGES, int MODE>
int run_linear(gn_model* m, int64_t M, const float* A, int K, int lda, const float* wk, int N, const float* bias,
               float* out, int ld_out, int col_off, const float* w_last, const float* b_last, float* scores,
               cudaStream_t st, int n_cls = 1, const gn_score_sink* sink = nullptr) {
    static_assert(MODE == kLinear || MODE == kScore, "head layers only");
    if (K % 32 || N % BN || (lda & 3)) return GN_E_SHAPE;
    CUtensorMap ta, tb;
    int rc = make_tmap_2d(m->h, &ta, A, M, K, lda, 128, true);
    if (rc) return rc;
    rc = make_tmap_2d(m->h, &tb, wk, N, K, K, BN, true);
    if (rc) return rc;
    GemmArgs a{};
    a.M = M; a.N = N; a.K = K; a.bias = bias; a.out_f = out; a.ld_out = ld_out; a.col_off = col_off;
    a.w_last = w_last; a.b_last = b_last; a.scores = scores; a.n_cls = n_cls;
    if (sink) a.sink = *sink;
    return launch<1, BN, STAGES, MODE, 0>(m->h, ta, tb, a, ((M + 127) / 128) * (N / BN), st);
}

}  // namespace

int gn_launch_zero_border(gn_handle* h, __nv_bfloat16* buf, int64_t n, int PW, int C, cudaStream_t st) {
    const int64_t total = n * (4 * PW - 4) * (C / 8);
    if (total <= 0) return GN_OK;
    const int64_t cap = (int64_t)h->num_sms * 8, blocks = (total + 255) / 256;
    zero_border_kernel<<<(int)(blocks < cap ? blocks : cap), 256, 0, st>>>(buf, n, PW, C);
    GN_LAUNCHED(h);
    return GN_OK;
}

// Number of K ranges linear5's reduction is cut into (see run_linear_splitk).  A function of the model only -- never of
// the batch -- so that every frame is summed the same way: about 108 k-steps (of 64) per range, at most 32 ranges.
// 40x40 frames: K = 41 472 = 648 k-steps -> 6 ranges (10 800 frames: 85 x 2 x 6 = 1 020 units over 148 CTAs).
int gn_linear5_splits(const gn_model* m) {
    const int64_t k_total = (int64_t)m->P3 * m->P3 * 512 / kBK;
    int64_t s = (k_total + 107) / 108;
    if (s > 32) s = 32;
    if (s < 1) s = 1;
    const int64_t k_per = (k_total + s - 1) / s;
    return (int)((k_total + k_per - 1) / k_per);   // every range non-empty
}

// AudBl conv2 (64 -> 128, k3 s2 p1) and linear3 on the tensor cores (reference utils.py:219-225).
// conv1's output ws.ac1 is (position, channel)-major with a zero row on each side, so the im2col row of
// output position l is the 192 contiguous elements starting at row 2*l: a 3-D TMA map with overlapping
// rows (row pitch 128 elements) feeds the GEMM directly.  16 frames x 8 positions make one 128-row tile.
int gn_launch_tc_audbl(gn_model* m, int64_t n, const gn_ws& ws, cudaStream_t st) {
    const int L2 = m->L2, rows = m->L1 + 2;
    if (128 % L2) return GN_E_SHAPE;
    CUtensorMap ta, tb;
    const uint64_t dims[3] = {192, (uint64_t)L2, (uint64_t)n};
    const uint64_t str[2] = {128 * 4, (uint64_t)rows * 64 * 4};
    const uint32_t box[3] = {32u, (uint32_t)L2, (uint32_t)(128 / L2)};
    int rc = make_tmap(m->h, &ta, ws.ac1, 3, dims, str, box, true);
    if (rc) return rc;
    rc = make_tmap_2d(m->h, &tb, m->aconv2_wk, 128, 192, 192, 128, true);
    if (rc) return rc;
    GemmArgs a{};
    a.M = n * L2; a.N = 128; a.K = 192; a.bias = m->aconv2_b; a.out_f = ws.ac2; a.ld_out = 128; a.col_off = 0;
    a.a3d_rows = L2;
    rc = launch<1, 128, 4, kLinear, 0>(m->h, ta, tb, a, (a.M + 127) / 128, st);
    if (rc) return rc;
    // linear3 + ReLU -> columns [0,128) of the fusion input (audio first, utils.py:266)
    return run_linear<128, 4, kLinear>(m, n, ws.ac2, 128 * L2, 128 * L2, m->alinear3_wk, 128, m->alinear3_b, ws.h0,
                                       m->fin, 0, nullptr, nullptr, nullptr, st);
}

// conv2 -> conv3 -> linear5 -> fusion MLP -> scores, all on the tensor cores.
int gn_launch_tc_backbone(gn_model* m, int64_t n, const gn_ws& ws, float* d_scores, float* d_visfeat,
                          cudaStream_t st, const gn_score_sink* sink) {
    int rc;
    if (m->P1 == 13 && m->P2 == 11) {
        // the reference's 40x40 frames: one frame per tile
        // block 2: conv2 (64 -> 256) on 13x13, pool -> 11x11, written with a zero border for conv3
        static const bool pair2 = !getenv("GN_CONV2_SINGLE");      // development switch: the cta_group::1 kernel for A/B runs
        if (pair2)
            rc = run_conv_pair<2, 128, 4, 13, 128>(m, n, ws.a1p, 64, m->conv2_wk, 256, m->conv2_b, m->bn_scale[1],
                                                   m->bn_shift[1], ws.a2p, 1, st);
        else
        rc = run_conv<2, 128, 3, 13>(m, n, ws.a1p, 64, m->conv2_wk, 256, m->conv2_b, m->bn_scale[1], m->bn_shift[1],
                                     ws.a2p, 1, st);
        if (rc) return rc;
        GN_STAGE_MARK(m, 3, st);
        // block 3: conv3 (256 -> 512) on 11x11, pool -> 9x9, dense (p, c) rows for linear5
        // (4 operand stages: with 3 the TMA round trip was exposed -- tensor pipe 83 % busy; the pooling stage shrinks
        //  to 64 channels per pass to make room)
        static const bool pair3 = !getenv("GN_CONV3_SINGLE");      // development switch: the cta_group::1 kernel for A/B runs
        if (pair3)
            rc = run_conv_pair<1, 256, 6, 11, 64>(m, n, ws.a2p, 256, m->conv3_wk, 512, m->conv3_b, m->bn_scale[2],
                                               m->bn_shift[2], ws.a3, 0, st);
        else
        rc = run_conv<1, 256, 4, 11, 64>(m, n, ws.a2p, 256, m->conv3_wk, 512, m->conv3_b, m->bn_scale[2], m->bn_shift[2],
                                         ws.a3, 0, st);
    } else {
        // any other frame size (e.g. 224x224: conv2 on 74x74, conv3 on 72x72): overlapping 16x16 spatial tiles
        rc = run_conv_tiled(m, n, ws.a1p, m->P1, 64, m->conv2_wk, 256, m->conv2_b, m->bn_scale[1], m->bn_shift[1],
                            ws.a2p, 1, st);
        if (rc) return rc;
        GN_STAGE_MARK(m, 3, st);
        rc = run_conv_tiled(m, n, ws.a2p, m->P2, 256, m->conv3_wk, 512, m->conv3_b, m->bn_scale[2], m->bn_shift[2],
                            ws.a3, 0, st);
    }
    if (rc) return rc;
    GN_STAGE_MARK(m, 4, st);
    const int64_t K5 = (int64_t)m->P3 * m->P3 * 512;
    const int aoff = m->audio_included ? 128 : 0;
    // linear5 + ReLU -> columns [aoff, aoff+512) of the fusion input (audio features occupy [0,128))
    rc = run_linear_splitk(m, n, ws.a3, K5, m->linear5_wk, 512, m->linear5_b, gn_linear5_splits(m), ws.part, ws.h0,
                           m->fin, aoff, d_visfeat, st);
    if (rc) return rc;
    GN_STAGE_MARK(m, 5, st);
    rc = run_linear<256, 4, kLinear>(m, n, ws.h0, m->fin, m->fin, m->f_wk[0], 512, m->f_b[0], ws.h1, 512, 0, nullptr,
                                     nullptr, nullptr, st);
    if (rc) return rc;
    rc = run_linear<256, 4, kLinear>(m, n, ws.h1, 512, 512, m->f_wk[1], 512, m->f_b[1], ws.h2, 512, 0, nullptr,
                                     nullptr, nullptr, st);
    if (rc) return rc;
    rc = run_linear<256, 4, kLinear>(m, n, ws.h2, 512, 512, m->f_wk[2], 256, m->f_b[2], ws.h3, 256, 0, nullptr,
                                     nullptr, nullptr, st);
    if (rc) return rc;
    // fusion.9 + ReLU + fusion.12 + sigmoid + 4y+1 -> scores
    // (classifier head, gn_model_set_classifier_head: fusion.12 is 128 -> 5 and softmax replaces sigmoid; d_scores is [n][5])
    rc = run_linear<128, 4, kScore>(m, n, ws.h3, 256, 256, m->f_wk[3], 128, m->f_b[3], nullptr, 0, 0,
                                    m->n_cls > 1 ? m->cls_w : m->f12_w, m->n_cls > 1 ? m->cls_b : m->f12_b, d_scores, st,
                                    m->n_cls > 1 ? m->n_cls : 1, sink);
    GN_STAGE_MARK(m, 6, st);
    return rc;
}

// Training GEMMs of the fusion MLP (gn_train.cu): C = A[M][K] . B[N][K]^T, bf16 operands, fp32 accumulation.
//   mode 3 (forward)  : out = dropout(relu(C + bias)) as bf16 [M][N] and transposed [N][ld_t]
//   mode 4 (dgrad)    : out = C * scale * [gate > 0]  as bf16 [M][N] and transposed [N][ld_t]
//   mode 5 (wgrad)    : out_f32 [M][N] = C
int gn_tc_train_gemm(gn_handle* h, int mode, const __nv_bfloat16* A, int64_t M, int K, int64_t lda,
                     const __nv_bfloat16* B, int N, int64_t ldb, const float* bias, __nv_bfloat16* out,
                     __nv_bfloat16* out_t, int64_t ld_t, float* out_f32, const __nv_bfloat16* gate,
                     const uint8_t* mask, float scale, float keep_prob, unsigned long long seed, cudaStream_t st,
                     const gn_step_state* step_state) {
    constexpr int BN = 128;
    if (K % kBK || N % BN || M <= 0) return GN_E_SHAPE;
    CUtensorMap ta, tb;
    int rc = make_tmap_2d(h, &ta, A, M, K, lda, 128);
    if (rc) return rc;
    rc = make_tmap_2d(h, &tb, B, N, K, ldb, BN);
    if (rc) return rc;
    GemmArgs a{};
    a.M = M; a.N = N; a.K = K; a.bias = bias; a.out = out; a.ld_out = N; a.col_off = 0; a.out_f32 = out_f32;
    a.out_t = out_t; a.ld_t = ld_t; a.gate = gate; a.mask = mask; a.scale = scale; a.keep_prob = keep_prob; a.seed = seed;
    a.step_state = step_state;
    const int64_t tiles = ((M + 127) / 128) * (N / BN);
    switch (mode) {
        case kTrainFwd: return launch<1, BN, 4, kTrainFwd, 0>(h, ta, tb, a, tiles, st);
        case kDgrad: return launch<1, BN, 4, kDgrad, 0>(h, ta, tb, a, tiles, st);
        case kWgrad: return launch<1, BN, 4, kWgrad, 0>(h, ta, tb, a, tiles, st);
        default: return GN_E_ARG;
    }
}

// C[M][N] (fp32, row-major) = A[M][K] . B[N][K]^T with fp32 operands multiplied as TF32 on the tensor cores; K % 32 == 0,
// N % 128 == 0, lda/ldb in elements (multiples of 4).  Rows of A beyond M are zero-filled by TMA and never stored.
int gn_tc_gemm_tf32(gn_handle* h, const float* A, int64_t M, int K, int64_t lda, const float* B, int N, int64_t ldb,
                    float* C, cudaStream_t st, int k_splits) {
    constexpr int BN = 128;
    if (K % 32 || N % BN || M <= 0 || (lda & 3) || (ldb & 3) || k_splits < 1) return GN_E_SHAPE;
    {   // every K range non-empty
        const int ksteps = K / 32, per = (ksteps + k_splits - 1) / k_splits;
        k_splits = (ksteps + per - 1) / per;
    }
    if (!h->encode_tiled) return GN_E_NODEVICE;
    CUtensorMap ta, tb;
    for (int which = 0; which < 2; ++which) {
        const cuuint64_t gdim[2] = {(cuuint64_t)K, (cuuint64_t)(which ? N : M)};
        const cuuint64_t gstr[1] = {(cuuint64_t)((which ? ldb : lda) * 4)};
        const cuuint32_t bx[2] = {32u, (cuuint32_t)(which ? BN : 128)};
        const cuuint32_t es[2] = {1u, 1u};
        CUresult r = reinterpret_cast<EncodeTiledFn>(h->encode_tiled)(
            which ? &tb : &ta, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(which ? B : A), gdim, gstr, bx, es,
            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) return (int)r;
    }
    GemmArgs a{};
    a.M = M; a.N = N; a.K = K; a.out_f32 = C; a.k_splits = k_splits; a.split_stride = M * N;
    return launch<1, BN, 4, kGemmTF32, 0>(h, ta, tb, a, ((M + 127) / 128) * (N / BN) * k_splits, st);
}

extern "C" int gn_dbg_gemm_tf32(gn_handle* h, const float* d_a, int64_t M, int K, const float* d_b, int N, float* d_c,
                                void* stream) {
    if (!h || !d_a || !d_b || !d_c) return GN_E_ARG;
    return gn_tc_gemm_tf32(h, d_a, M, K, K, d_b, N, K, d_c, gn_stream(stream), 1);
}
