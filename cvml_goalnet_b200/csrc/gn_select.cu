// (a6-a10) keyshot selection and evaluation counts: integer work, bit-exact with the reference.
//   gn_clip_information  <- utils.py:611 (round half-even -> int8), 396-410 (expand_array), 445-464
//   gn_knapsack          <- utils.py:466-510 (0/1 DP + backtrack)
//   gn_summary_indicator <- utils.py:638-641
//   gn_fscore_counts     <- utils.py:571-573 (the three sums per annotator)
#include <stdio.h>
#include <stdlib.h>
#include "gn_common.cuh"

namespace {

// ---------------------------------------------------------------------------------------------
// clip statistics
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ int round_importance(float s) {
    // torch.round (half to even) then .type(torch.int8)
    int v = (int)rintf(s);
    return (int)(signed char)v;
}

__global__ void __launch_bounds__(256)
clip_info_kernel(const float* __restrict__ scores, int64_t n, int rate, int identity, int64_t n_raw,
                 const int32_t* __restrict__ clips, int n_clips, int32_t* __restrict__ values,
                 int32_t* __restrict__ lengths) {
    const int lane = threadIdx.x & 31;
    const int warp_global = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int nwarps = (gridDim.x * blockDim.x) >> 5;
    for (int c = warp_global; c < n_clips; c += nwarps) {
        // Python slice imp[a:b]: clamp to [0, n_raw]; empty when b <= a
        int64_t a = clips[2 * c], b = clips[2 * c + 1];
        a = a < 0 ? 0 : (a > n_raw ? n_raw : a);
        b = b < 0 ? 0 : (b > n_raw ? n_raw : b);
        int sum = 0;
        for (int64_t f = a + lane; f < b; f += 32) {
            // expand_array: repeat each score `rate` times, truncate, or pad with the last one
            int64_t i = identity ? f : f / rate;
            if (i > n - 1) i = n - 1;
            sum += round_importance(scores[i]);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
        if (lane == 0) {
            values[c] = sum;
            lengths[c] = (int32_t)(b > a ? b - a : 0);
        }
    }
}

// ---------------------------------------------------------------------------------------------
// 0/1 knapsack: one CTA per problem.  The table row lives in shared memory (ping-pong) when two rows
// fit, else in the caller's workspace; one decision bit per cell (K[i][w] != K[i-1][w]) is all the
// backtrack of utils.py:495-509 needs, because its running `res` always equals K[i][w].
// ---------------------------------------------------------------------------------------------
constexpr int kKnapThreads = 1024;

__global__ void __launch_bounds__(kKnapThreads, 1)
knapsack_kernel(const int32_t* __restrict__ values_all, const int32_t* __restrict__ weights_all, int n_items,
                int64_t capacity, int zero_col_forced, uint8_t* __restrict__ selected_all,
                int32_t* __restrict__ best_all, uint32_t* __restrict__ bits_all, int64_t words,
                int32_t* __restrict__ rows_global, int use_smem) {
    extern __shared__ int32_t row_smem[];
    const int prob = blockIdx.x;
    const int32_t* values = values_all + (size_t)prob * n_items;
    const int32_t* weights = weights_all + (size_t)prob * n_items;
    uint8_t* selected = selected_all + (size_t)prob * n_items;
    uint32_t* bits = bits_all + (size_t)prob * n_items * words;
    const int64_t C = capacity + 1;
    const int64_t Cpad = words * 32;
    int32_t* r0 = use_smem ? row_smem : rows_global + (size_t)prob * 2 * Cpad;
    int32_t* r1 = r0 + Cpad;
    const int tid = threadIdx.x, lane = tid & 31;

    for (int64_t w = tid; w < Cpad; w += kKnapThreads) r0[w] = 0;      // K[0][*] = 0 (utils.py:487)
    for (int i = tid; i < n_items; i += kKnapThreads) selected[i] = 0;
    __syncthreads();

    int32_t* prev = r0;
    int32_t* cur = r1;
    for (int i = 0; i < n_items; ++i) {
        const int32_t wt = weights[i], v = values[i];
        uint32_t* brow = bits + (size_t)i * words;
        for (int64_t w = tid; w < Cpad; w += kKnapThreads) {
            int32_t old = 0, nv = 0;
            if (w < C) {
                old = prev[w];
                nv = old;
                // utils.py:487-492; column 0 of the reference table is forced to 0
                if (wt <= w && !(w == 0 && zero_col_forced)) {
                    const int32_t cand = v + prev[w - wt];
                    nv = cand > old ? cand : old;
                }
                cur[w] = nv;
            }
            const uint32_t word = __ballot_sync(0xffffffffu, nv != old);
            if (lane == 0) brow[w >> 5] = word;
        }
        __syncthreads();
        int32_t* t = prev; prev = cur; cur = t;
    }
    // backtrack (utils.py:495-509): skip item i when K[i][w] == K[i-1][w]
    if (tid == 0) {
        best_all[prob] = prev[capacity];
        int64_t w = capacity;
        for (int i = n_items; i >= 1; --i) {
            const uint32_t word = bits[(size_t)(i - 1) * words + (w >> 5)];
            if ((word >> (w & 31)) & 1u) {
                selected[i - 1] = 1;
                w -= weights[i - 1];
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------
// 0/1 knapsack, thread-block-cluster version: the table columns are split over the CTAs of one
// cluster (one cluster per problem); every CTA keeps its slice of the previous/current row in its
// own shared memory and reads K[i-1][w - wt] from the owning CTA through distributed shared memory.
// One barrier.cluster per item replaces the single-CTA __syncthreads() and divides the per-row work
// by the cluster size.  The backtrack is a single warp that tests 32 rows per L2 round trip (rows
// whose decision bit at the current w is clear leave w unchanged, so they can be skipped together).
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ uint32_t cluster_id_x() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%clusterid.x;" : "=r"(r));
    return r;
}
__device__ __forceinline__ uint32_t map_to_cta(uint32_t local_smem_addr, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local_smem_addr), "r"(rank));
    return r;
}
__device__ __forceinline__ int32_t ld_cluster_s32(uint32_t addr) {
    int32_t v;
    asm volatile("ld.shared::cluster.s32 %0, [%1];" : "=r"(v) : "r"(addr) : "memory");
    return v;
}
__device__ __forceinline__ void cluster_barrier() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}

constexpr int kKnapRing = 32;     // rows of decision bits buffered in shared memory between flushes

template <int THREADS>
__global__ void __launch_bounds__(THREADS, 1)
knapsack_cluster_kernel(const int32_t* __restrict__ values_all, const int32_t* __restrict__ weights_all, int n_items,
                        int capacity, int zero_col_forced, uint8_t* __restrict__ selected_all,
                        int32_t* __restrict__ best_all, uint32_t* bits_all, int words, int slice_w) {
    extern __shared__ int32_t row_smem[];            // [2][slice_w] rows, then [kKnapRing][slice_w/32] decision bits
    const uint32_t rank = cluster_ctarank();
    const int prob = (int)cluster_id_x();
    const int32_t* values = values_all + (size_t)prob * n_items;
    const int32_t* weights = weights_all + (size_t)prob * n_items;
    uint8_t* selected = selected_all + (size_t)prob * n_items;
    uint32_t* bits = bits_all + (size_t)prob * n_items * words;
    const int C = capacity + 1;
    const int c0 = (int)rank * slice_w;              // first table column of this CTA's slice
    const int sw = slice_w >> 5;                     // decision words per row in this slice
    const int w0 = c0 >> 5;                          // first decision word of this slice
    const int nw = (w0 + sw <= words) ? sw : (words > w0 ? words - w0 : 0);   // words of the slice that exist
    uint32_t* ring = reinterpret_cast<uint32_t*>(row_smem + 2 * slice_w);
    const int tid = threadIdx.x, lane = tid & 31;
    const uint32_t smem_base = (uint32_t)__cvta_generic_to_shared(row_smem);

    for (int l = tid; l < 2 * slice_w; l += THREADS) row_smem[l] = 0;      // K[0][*] = 0 (utils.py:487)
    if (rank == 0)
        for (int i = tid; i < n_items; i += THREADS) selected[i] = 0;
    cluster_barrier();

    int pb = 0;                                       // buffer holding row i-1
    for (int i = 0; i < n_items; ++i) {
        const int32_t wt = __ldg(weights + i), v = __ldg(values + i);
        const int q = wt / slice_w, rem = wt - q * slice_w;
        const int own_a = (int)rank - q, own_b = own_a - 1;         // CTAs holding K[i-1][w - wt]
        const int32_t* prev = row_smem + pb * slice_w;
        int32_t* cur = row_smem + (pb ^ 1) * slice_w;
        const uint32_t prev_addr = smem_base + (uint32_t)(pb * slice_w) * 4u;
        // shared::cluster addresses of row i-1 in the two CTAs that can hold column w - wt
        const uint32_t rem_a = map_to_cta(prev_addr, (uint32_t)(own_a >= 0 ? own_a : 0));
        const uint32_t rem_b = map_to_cta(prev_addr, (uint32_t)(own_b >= 0 ? own_b : 0));
        uint32_t* rrow = ring + (i & (kKnapRing - 1)) * sw;
#pragma unroll 1
        for (int l = tid; l < slice_w; l += THREADS) {
            const int w = c0 + l;
            const int32_t old = prev[l];
            int32_t nv = old;
            // utils.py:487-492; column 0 of the reference table is forced to 0
            if (w < C && wt <= w && !(w == 0 && zero_col_forced)) {
                const int sl = l - rem;
                int32_t src;
                if (q == 0 && sl >= 0) {
                    src = prev[sl];                                   // same CTA: plain shared load
                } else {
                    const uint32_t addr = sl >= 0 ? rem_a + (uint32_t)sl * 4u : rem_b + (uint32_t)(sl + slice_w) * 4u;
                    src = ld_cluster_s32(addr);
                }
                const int32_t cand = v + src;
                nv = cand > old ? cand : old;
            }
            cur[l] = nv;
            const uint32_t word = __ballot_sync(0xffffffffu, nv != old);
            if (lane == 0) rrow[l >> 5] = word;
        }
        // Decision bits leave shared memory once per kKnapRing rows, so that the release fence of the
        // per-row cluster barrier normally has no global stores to wait for.
        if ((i & (kKnapRing - 1)) == kKnapRing - 1 || i == n_items - 1) {
            __syncthreads();
            const int r0 = i & ~(kKnapRing - 1);
            const int nr = i - r0 + 1;
            for (int e = tid; e < nr * sw; e += THREADS) {
                const int r = e / sw, wi = e - r * sw;
                if (wi < nw) bits[(size_t)(r0 + r) * words + w0 + wi] = ring[e];
            }
        }
        cluster_barrier();       // row i complete everywhere; row i-1 no longer read by anyone
        pb ^= 1;
    }
    // K[n][capacity]
    if (tid == 0 && capacity >= c0 && capacity < c0 + slice_w) best_all[prob] = row_smem[pb * slice_w + capacity - c0];
    // backtrack (utils.py:495-509): skip item i when K[i][w] == K[i-1][w]
    if (rank == 0 && tid < 32) {
        int w = capacity;
        int i = n_items;
        while (i > 0) {
            const int row = i - 1 - lane;
            uint32_t bit = 0;
            if (row >= 0) {
                const volatile uint32_t* bp = bits + (size_t)row * words + (w >> 5);
                bit = (*bp >> (w & 31)) & 1u;
            }
            const uint32_t mask = __ballot_sync(0xffffffffu, bit != 0);
            if (mask == 0) { i -= 32; continue; }
            const int r = i - 1 - (__ffs(mask) - 1);          // nearest selected row below i
            if (lane == 0) selected[r] = 1;
            w -= __ldg(weights + r);
            i = r;
        }
    }
}

// ---------------------------------------------------------------------------------------------
// summary indicator
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
indicator_kernel(const int32_t* __restrict__ clips, const uint8_t* __restrict__ selected, int n_clips, int64_t n_raw,
                 uint8_t* __restrict__ ind) {
    for (int c = blockIdx.x; c < n_clips; c += gridDim.x) {
        if (!selected[c]) continue;
        int64_t a = clips[2 * c], b = (int64_t)clips[2 * c + 1] + 1;    // inclusive end (utils.py:640)
        if (a < 0) a = 0;
        if (b > n_raw) b = n_raw;
        for (int64_t f = a + threadIdx.x; f < b; f += blockDim.x) ind[f] = 1;
    }
}

// ---------------------------------------------------------------------------------------------
// F-score counts
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
fscore_kernel(const uint8_t* __restrict__ S, const uint8_t* __restrict__ G, int n_users, int64_t n_raw,
              unsigned long long* __restrict__ counts, int vec_ok) {
    const int u = blockIdx.y;                     // n_users rows of G, plus one extra row for sum(S)
    const uint8_t* g = (u < n_users) ? G + (size_t)u * n_raw : S;
    unsigned long long overlap = 0, gsum = 0;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    const int64_t t0 = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (vec_ok) {
        const int64_t nv = n_raw / 16;
        const uint4* gv = reinterpret_cast<const uint4*>(g);
        const uint4* sv = reinterpret_cast<const uint4*>(S);
        for (int64_t i = t0; i < nv; i += stride) {
            const uint4 a = gv[i], s = sv[i];
            const unsigned int ga[4] = {a.x, a.y, a.z, a.w}, sa[4] = {s.x, s.y, s.z, s.w};
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                gsum += __vsadu4(ga[k], 0u);
                overlap += __popc(__vcmpne4(ga[k], 0u) & __vcmpne4(sa[k], 0u)) >> 3;
            }
        }
        for (int64_t f = nv * 16 + t0; f < n_raw; f += stride) {
            gsum += g[f];
            overlap += (g[f] != 0 && S[f] != 0) ? 1 : 0;
        }
    } else {
        for (int64_t f = t0; f < n_raw; f += stride) {
            gsum += g[f];
            overlap += (g[f] != 0 && S[f] != 0) ? 1 : 0;
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        overlap += __shfl_xor_sync(0xffffffffu, overlap, o);
        gsum += __shfl_xor_sync(0xffffffffu, gsum, o);
    }
    if ((threadIdx.x & 31) == 0) {
        if (u < n_users) {
            atomicAdd(counts + u, overlap);
            atomicAdd(counts + n_users + u, gsum);
        } else {
            atomicAdd(counts + 2 * n_users, gsum);
        }
    }
}

constexpr int64_t kKnapSmemLimit = 200 * 1024;

}  // namespace

extern "C" int gn_clip_information(gn_handle* h, const float* d_scores, int64_t n, int skip_frames, int64_t n_raw,
                                   const int32_t* d_clips, int n_clips, int32_t* d_values, int32_t* d_lengths,
                                   void* stream) {
    if (!h || n <= 0 || n_raw < 0 || n_clips < 0 || skip_frames <= 0) return GN_E_ARG;
    if (n_clips == 0) return GN_OK;
    if (!d_scores || !d_clips || !d_values || !d_lengths) return GN_E_ARG;
    if (n_raw > (1LL << 31) / 128) return GN_E_RANGE;      // |importance| <= 128 per frame, int32 sums
    const int grid = (n_clips + 7) / 8;
    clip_info_kernel<<<grid, 256, 0, gn_stream(stream)>>>(d_scores, n, skip_frames, n == n_raw ? 1 : 0, n_raw, d_clips,
                                                          n_clips, d_values, d_lengths);
    GN_LAUNCHED(h);
    return GN_OK;
}

static inline int64_t knap_words(int64_t capacity) { return (capacity + 1 + 31) / 32; }

extern "C" int64_t gn_knapsack_workspace_bytes(int n_items, int64_t capacity_columns) {
    if (n_items < 0 || capacity_columns < 0) return GN_E_ARG;
    const int64_t words = knap_words(capacity_columns);
    int64_t bytes = (int64_t)n_items * words * 4;                 // decision bits
    if (2 * words * 32 * 4 > kKnapSmemLimit) bytes += 2 * words * 32 * 4;   // rows do not fit shared memory
    return (bytes + 255) & ~255LL;
}

extern "C" int gn_knapsack(gn_handle* h, const int32_t* d_values, const int32_t* d_weights, int n_items,
                           int64_t capacity, int capacity_rem, int batch, uint8_t* d_selected, int32_t* d_best,
                           void* d_workspace, int64_t workspace_bytes, void* stream) {
    if (!h || n_items < 0 || capacity < 0 || batch < 0 || capacity_rem < 0) return GN_E_ARG;
    if (batch == 0) return GN_OK;
    if (!d_best || (n_items > 0 && (!d_values || !d_weights || !d_selected))) return GN_E_ARG;
    if (capacity > (1LL << 31) - 64) return GN_E_RANGE;
    const int64_t words = knap_words(capacity);
    const int64_t per = gn_knapsack_workspace_bytes(n_items, capacity);
    if (per * batch > workspace_bytes || (!d_workspace && per > 0)) return GN_E_WORKSPACE;
    // cluster path: columns split over up to 8 CTAs (portable cluster size), at least 1024 columns per CTA
    int cl = 8;
    while (cl > 1 && (capacity + 1) < (int64_t)cl * 1024) cl >>= 1;
    if (const char* e = getenv("GN_KNAP_CL")) { const int v = atoi(e); if (v == 1 || v == 2 || v == 4 || v == 8) cl = v; }
    int64_t slice = (capacity + 1 + cl - 1) / cl;
    slice = (slice + 31) & ~31LL;
    const int64_t cl_smem = 2 * slice * 4 + kKnapRing * (slice / 32) * 4;
    const int use_smem = cl_smem <= kKnapSmemLimit ? 1 : 0;
    // workspace: [batch][n_items][words] bits, then (if needed) [batch][2][words*32] rows
    uint32_t* bits = reinterpret_cast<uint32_t*>(d_workspace);
    if (use_smem) {
        constexpr int kT = 1024;
        auto kern = knapsack_cluster_kernel<kT>;
        static bool cattr = false;
        if (!cattr) {
            GN_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kKnapSmemLimit));
            cattr = true;
        }
        cudaLaunchConfig_t cfg{};
        cfg.gridDim = dim3((unsigned)(batch * cl));
        cfg.blockDim = dim3(kT);
        cfg.dynamicSmemBytes = (size_t)cl_smem;
        cfg.stream = gn_stream(stream);
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = (unsigned)cl;
        at[0].val.clusterDim.y = 1;
        at[0].val.clusterDim.z = 1;
        cfg.attrs = at;
        cfg.numAttrs = 1;
        GN_CUDA(cudaLaunchKernelEx(&cfg, kern, d_values, d_weights, n_items, (int)capacity,
                                   capacity_rem == 0 ? 1 : 0, d_selected, d_best, bits, (int)words, (int)slice));
        GN_LAUNCHED(h);
        return GN_OK;
    }
    int32_t* rows = reinterpret_cast<int32_t*>(bits + (size_t)batch * n_items * words);
    knapsack_kernel<<<batch, kKnapThreads, 0, gn_stream(stream)>>>(
        d_values, d_weights, n_items, capacity, capacity_rem == 0 ? 1 : 0, d_selected, d_best, bits, words, rows,
        0);
    GN_LAUNCHED(h);
    return GN_OK;
}

extern "C" int gn_summary_indicator(gn_handle* h, const int32_t* d_clips, const uint8_t* d_selected, int n_clips,
                                    int64_t n_raw, uint8_t* d_indicator, void* stream) {
    if (!h || n_clips < 0 || n_raw < 0) return GN_E_ARG;
    if (n_raw == 0) return GN_OK;
    if (!d_indicator) return GN_E_ARG;
    GN_CUDA(cudaMemsetAsync(d_indicator, 0, (size_t)n_raw, gn_stream(stream)));
    if (n_clips == 0) return GN_OK;
    if (!d_clips || !d_selected) return GN_E_ARG;
    const int grid = n_clips < 4 * h->num_sms ? n_clips : 4 * h->num_sms;
    indicator_kernel<<<grid, 256, 0, gn_stream(stream)>>>(d_clips, d_selected, n_clips, n_raw, d_indicator);
    GN_LAUNCHED(h);
    return GN_OK;
}

extern "C" int gn_fscore_counts(gn_handle* h, const uint8_t* d_pred, const uint8_t* d_gd, int n_users, int64_t n_raw,
                                int64_t* d_counts, void* stream) {
    if (!h || n_users < 0 || n_raw < 0 || !d_counts) return GN_E_ARG;
    GN_CUDA(cudaMemsetAsync(d_counts, 0, sizeof(int64_t) * (2 * n_users + 1), gn_stream(stream)));
    if (n_raw == 0) return GN_OK;
    if (!d_pred || (n_users > 0 && !d_gd)) return GN_E_ARG;
    const int vec_ok = (n_raw % 16 == 0) && ((reinterpret_cast<uintptr_t>(d_pred) & 15) == 0) &&
                       ((reinterpret_cast<uintptr_t>(d_gd) & 15) == 0);
    int64_t bx = (n_raw / 16 + 255) / 256;
    if (bx < 1) bx = 1;
    if (bx > 2 * h->num_sms) bx = 2 * h->num_sms;
    fscore_kernel<<<dim3((unsigned)bx, (unsigned)(n_users + 1)), 256, 0, gn_stream(stream)>>>(
        d_pred, d_gd, n_users, n_raw, reinterpret_cast<unsigned long long*>(d_counts), vec_ok);
    GN_LAUNCHED(h);
    return GN_OK;
}
