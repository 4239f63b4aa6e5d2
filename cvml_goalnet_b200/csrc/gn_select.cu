// (a6-a10) keyshot selection and evaluation counts: integer work, bit-exact with the reference.
//   gn_clip_information  <- utils.py:611 (round half-even -> int8), 396-410 (expand_array), 445-464
//   gn_knapsack          <- utils.py:466-510 (0/1 DP + backtrack)
//   gn_summary_indicator <- utils.py:638-641
//   gn_fscore_counts     <- utils.py:571-573 (the three sums per annotator)
#include <mutex>
#include <stdio.h>
#include <stdlib.h>
#include "gn_common.cuh"
#include "gn_tc_ptx.cuh"

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
                        int32_t* __restrict__ best_all, uint32_t* bits_all, int words, int slice_w,
                        const int* __restrict__ status_all) {
    extern __shared__ int32_t row_smem[];            // [2][slice_w] rows, [kKnapRing][slice_w/32] decision bits, items
    const uint32_t rank = cluster_ctarank();
    const int prob = (int)cluster_id_x();
    // launched behind the multi-row kernel: nothing to do unless that one declined the problem (status 1)
    if (status_all != nullptr && status_all[prob] == 0) return;
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
    // (weight, value) of every item in shared memory: the per-row cluster barrier invalidates L1, so a global
    // load at the top of each row would pay an L2 round trip on the critical path
    int2* items = reinterpret_cast<int2*>(ring + kKnapRing * sw);
    const int tid = threadIdx.x, lane = tid & 31;
    const uint32_t smem_base = (uint32_t)__cvta_generic_to_shared(row_smem);

    for (int i = tid; i < n_items; i += THREADS) items[i] = make_int2(weights[i], values[i]);
    for (int l = tid; l < 2 * slice_w; l += THREADS) row_smem[l] = 0;      // K[0][*] = 0 (utils.py:487)
    if (rank == 0)
        for (int i = tid; i < n_items; i += THREADS) selected[i] = 0;
    cluster_barrier();

    int pb = 0;                                       // buffer holding row i-1
    for (int i = 0; i < n_items; ++i) {
        const int2 it = items[i];
        const int32_t wt = it.x, v = it.y;
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
// 0/1 knapsack, T rows per step, neighbour-to-neighbour pipeline.  Row i only depends on row i-1 at columns
// w and w - wt_i, so rows i0 .. i0+T-1 at column w are a function of row i0-1 at the 2^T columns
// w - (sum of a subset of the T weights).  Every thread loads those 2^T values and evaluates the T levels in
// registers (one add-max per cell).  Data only ever flows towards higher columns, so there is no cluster-wide
// barrier: after a step, a CTA PUSHES the tail of its new row into the halo of its right neighbour through
// distributed shared memory and arrives on that neighbour's mbarrier; the neighbour returns a credit once it
// has finished reading a halo buffer.  CTA r simply runs a little behind CTA r-1.
// Columns to the left of column 0 hold a large negative sentinel, so "item does not fit" needs no test.
// Declines (status 1; the one-row kernel behind it then runs) when T * max weight exceeds the halo, when a
// zero weight meets the forced-zero column, or when values are large enough to threaten the sentinel.
// ---------------------------------------------------------------------------------------------
constexpr int32_t kKnapNeg = -(1 << 29);

template <int THREADS, int T, int CPT>
__global__ void __launch_bounds__(THREADS, 1)
knapsack_multirow_kernel(const int32_t* __restrict__ values_all, const int32_t* __restrict__ weights_all, int n_items,
                         int capacity, int zero_col_forced, uint8_t* __restrict__ selected_all,
                         int32_t* __restrict__ best_all, uint32_t* bits_all, int words, int slice_w, int halo,
                         int* __restrict__ status_all, int cl) {
    constexpr int NS = 1 << T;
    constexpr int NWARPS = THREADS / 32;
    extern __shared__ int32_t row_smem[];            // [2][halo + slice_w] rows, [kKnapRing][slice_w/32] bits, items
    __shared__ int s_maxw, s_minw, s_maxv;
    __shared__ __align__(8) uint64_t bars[4];        // full[2] (halo of buffer b has arrived), credit[2] (right
                                                     // neighbour has finished reading its halo of buffer b)
    const uint32_t rank = tcx::cluster_ctarank();
    const int prob = (int)tcx::cluster_id_x();
    const int32_t* values = values_all + (size_t)prob * n_items;
    const int32_t* weights = weights_all + (size_t)prob * n_items;
    uint8_t* selected = selected_all + (size_t)prob * n_items;
    uint32_t* bits = bits_all + (size_t)prob * n_items * words;
    const int c0 = (int)rank * slice_w;
    const int sw = slice_w >> 5;
    const int w0 = c0 >> 5;
    const int nw = (w0 + sw <= words) ? sw : (words > w0 ? words - w0 : 0);
    const int rowlen = halo + slice_w;
    uint32_t* ring = reinterpret_cast<uint32_t*>(row_smem + 2 * rowlen);
    int2* items = reinterpret_cast<int2*>(ring + kKnapRing * sw);
    const int tid = threadIdx.x, lane = tid & 31;
    const uint32_t smem_base = tcx::smem_u32(row_smem);
    const uint32_t bar0 = tcx::smem_u32(bars);
    const bool has_left = rank > 0, has_right = (int)rank + 1 < cl;

    if (tid == 0) {
        s_maxw = 0; s_minw = 0x7fffffff; s_maxv = 0;
        tcx::mbar_init(bar0, NWARPS); tcx::mbar_init(bar0 + 8, NWARPS);      // full[0], full[1]
        tcx::mbar_init(bar0 + 16, 1); tcx::mbar_init(bar0 + 24, 1);          // credit[0], credit[1]
        tcx::fence_barrier_init();
    }
    __syncthreads();
    int mw = 0, mnw = 0x7fffffff, mv = 0;
    for (int i = tid; i < n_items; i += THREADS) {
        const int2 it = make_int2(weights[i], values[i]);
        items[i] = it;
        mw = max(mw, it.x); mnw = min(mnw, it.x); mv = max(mv, abs(it.y));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        mw = max(mw, __shfl_xor_sync(0xffffffffu, mw, o));
        mnw = min(mnw, __shfl_xor_sync(0xffffffffu, mnw, o));
        mv = max(mv, __shfl_xor_sync(0xffffffffu, mv, o));
    }
    if (lane == 0) { atomicMax(&s_maxw, mw); atomicMin(&s_minw, mnw); atomicMax(&s_maxv, mv); }
    // K[0][*] = 0 (utils.py:487); columns left of column 0 (the halo of rank 0) = sentinel
    for (int l = tid; l < 2 * rowlen; l += THREADS) {
        const int p = l < rowlen ? l : l - rowlen;
        row_smem[l] = (rank == 0 && p < halo) ? kKnapNeg : 0;
    }
    if (rank == 0)
        for (int i = tid; i < n_items; i += THREADS) selected[i] = 0;
    __syncthreads();
    // the host picked this kernel without knowing the items: tell it to use the one-row kernel instead
    const bool decline = (int64_t)s_maxw * T > halo || s_minw < 0 || (s_minw == 0 && zero_col_forced) ||
                         (int64_t)s_maxv * (int64_t)(n_items + T) >= (1 << 28);
    if (rank == 0 && tid == 0) status_all[prob] = decline ? 1 : 0;
    if (decline) return;                               // uniform over the whole cluster: same data everywhere
    tcx::cluster_sync_all();                           // every CTA's mbarriers exist before anyone arrives remotely

    const int nsteps = (n_items + T - 1) / T;
    for (int k = 0; k < nsteps; ++k) {
        const int i0 = k * T;
        const int pb = k & 1;                          // buffer holding row i0 - 1
        // this step's items (missing ones at the tail: weight 0, value 0 = no-op rows) and the 2^T subset offsets
        int val[T], off[NS];
        int need_next = 0;
        off[0] = 0;
#pragma unroll
        for (int j = 0; j < T; ++j) {
            const int2 it = (i0 + j < n_items) ? items[i0 + j] : make_int2(0, 0);
            val[j] = it.y;
#pragma unroll
            for (int sidx = 0; sidx < (1 << j); ++sidx) off[sidx | (1 << j)] = off[sidx] + it.x;
            if (i0 + T + j < n_items) need_next += items[i0 + T + j].x;
        }
        int32_t* prev = row_smem + pb * rowlen;
        int32_t* cur = row_smem + (pb ^ 1) * rowlen;
        // the halo of row i0-1 was pushed by the left neighbour after its step k-1
        if (has_left && k > 0) tcx::mbar_wait_acq_cluster(bar0 + 8 * pb, ((k - 1) >> 1) & 1);
        const int rbase = i0 & (kKnapRing - 1);        // kKnapRing is a multiple of every T used
        // CPT columns per thread and pass, THREADS apart: the 2^T load addresses are formed once and the other
        // columns of the pass reach theirs through immediate offsets
#pragma unroll 1
        for (int lb = tid; lb < slice_w; lb += CPT * THREADS) {
            const int32_t* pp = prev + halo + lb;
            int32_t a[CPT][NS];
#pragma unroll
            for (int sidx = 0; sidx < NS; ++sidx) {
                const int32_t* ps = pp - off[sidx];
#pragma unroll
                for (int c = 0; c < CPT; ++c) a[c][sidx] = ps[c * THREADS];
            }
#pragma unroll
            for (int c = 0; c < CPT; ++c) {
                const int l = lb + c * THREADS;
                {                                       // slice_w is a multiple of CPT * THREADS (host)
                    uint32_t dec[T];
                    // K = max(K[w], v + K[w - wt]) (utils.py:489-490); a column left of 0 holds the sentinel and loses
#pragma unroll
                    for (int j = 0; j < T; ++j) {
                        const int32_t old = a[c][0];
#pragma unroll
                        for (int sidx = 0; sidx < NS; ++sidx)
                            if ((sidx & ((2 << j) - 1)) == 0)
                                a[c][sidx] = max(a[c][sidx], val[j] + a[c][sidx | (1 << j)]);
                        dec[j] = a[c][0] != old;
                    }
                    cur[halo + l] = a[c][0];
#pragma unroll
                    for (int j = 0; j < T; ++j) {
                        const uint32_t word = __ballot_sync(0xffffffffu, dec[j] != 0);
                        if (lane == 0) ring[(rbase + j) * sw + (l >> 5)] = word;
                    }
                }
            }
        }
        __syncthreads();                               // row i0+T-1 complete in this CTA; nobody reads `prev` any more
        // decision bits leave shared memory once per kKnapRing rows
        const int last = (i0 + T < n_items) ? i0 + T - 1 : n_items - 1;
        if (((i0 + T) & (kKnapRing - 1)) == 0 || i0 + T >= n_items) {
            const int r0 = i0 & ~(kKnapRing - 1);
            const int nr = last - r0 + 1;
            for (int e = tid; e < nr * sw; e += THREADS) {
                const int r = e / sw, wi = e - r * sw;
                if (wi < nw) bits[(size_t)(r0 + r) * words + w0 + wi] = ring[e];
            }
            __syncthreads();                           // the ring is rewritten by the next step
        }
        if (k + 1 < nsteps) {
            // credit to the left neighbour: this CTA no longer reads its halo of buffer pb
            if (has_left && tid == 0) tcx::mbar_arrive_cluster(tcx::mapa(bar0 + 16 + 8 * pb, rank - 1));
            if (has_right) {
                // the right neighbour reads its halo of buffer pb^1 at step k+1 and read it last at step k-1
                if (k > 0) tcx::mbar_wait_acq_cluster(bar0 + 16 + 8 * (pb ^ 1), ((k - 1) >> 1) & 1);
                const uint32_t dst = tcx::mapa(smem_base + (uint32_t)((pb ^ 1) * rowlen) * 4u, rank + 1);
                for (int h = tid; h < need_next; h += THREADS)
                    tcx::st_cluster_u32(dst + (uint32_t)(halo - 1 - h) * 4u, (uint32_t)cur[rowlen - 1 - h]);
                __syncwarp();
                if (lane == 0) tcx::mbar_arrive_cluster(tcx::mapa(bar0 + 8 * (pb ^ 1), rank + 1));
            }
        }
    }
    tcx::cluster_sync_all();                           // all decision bits are in global memory; no DSMEM traffic left
    const int pbl = nsteps & 1;
    if (tid == 0 && capacity >= c0 && capacity < c0 + slice_w)
        best_all[prob] = row_smem[pbl * rowlen + halo + capacity - c0];
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
            w -= items[r].x;
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
    // two table rows for knapsack_kernel (the fallback when the cluster kernels' shared-memory need -- which also grows
    // with n_items -- exceeds the limit): always reserved, so that no item count can make the fallback write past the end
    bytes = ((bytes + 15) & ~15LL) + 2 * words * 32 * 4;
    return ((bytes + 255) & ~255LL) + 256;                        // + status word of the multi-row kernel
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
    // cluster path: columns split over the CTAs of one cluster, at least 1024 columns per CTA.  A single wide
    // problem takes a 16-CTA cluster (non-portable size, one GPC); batches keep the portable 8 so that more
    // problems are resident at once.
    int cl = batch <= 4 ? 16 : 8;
    while (cl > 1 && (capacity + 1) < (int64_t)cl * 1024) cl >>= 1;
    if (const char* e = getenv("GN_KNAP_CL")) { const int v = atoi(e); if (v == 1 || v == 2 || v == 4 || v == 8 || v == 16) cl = v; }
    int64_t slice = (capacity + 1 + cl - 1) / cl;
    slice = (slice + 31) & ~31LL;
    constexpr int kMT = 512, kMCPT = 3;            // multi-row kernel: threads, columns per thread and pass
    {
        // whole passes for the multi-row kernel when that costs little padding
        const int64_t padded = (slice + kMT * kMCPT - 1) / (kMT * kMCPT) * (kMT * kMCPT);
        if (padded * 4 <= slice * 5) slice = padded;
    }
    const int64_t cl_smem = 2 * slice * 4 + kKnapRing * (slice / 32) * 4 + 8 + (int64_t)n_items * 8;
    const int use_smem = cl_smem <= kKnapSmemLimit ? 1 : 0;
    // workspace: [batch][n_items][words] bits, then (if needed) [batch][2][words*32] rows
    uint32_t* bits = reinterpret_cast<uint32_t*>(d_workspace);
    int* status = reinterpret_cast<int*>(reinterpret_cast<char*>(d_workspace) + per * batch - 256 * (int64_t)batch);
    if (use_smem) {
        constexpr int kT = 1024;
        cudaLaunchConfig_t cfg{};
        cfg.gridDim = dim3((unsigned)(batch * cl));
        cfg.blockDim = dim3(kT);
        cfg.stream = gn_stream(stream);
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = (unsigned)cl;
        at[0].val.clusterDim.y = 1;
        at[0].val.clusterDim.z = 1;
        cfg.attrs = at;
        cfg.numAttrs = 1;
        // multi-row kernel first (T rows per cluster barrier); it declines, per problem, when T * max weight
        // does not fit its halo, and the one-row kernel behind it then does that problem
        int rows_per_step = 4;
        if (const char* e = getenv("GN_KNAP_T")) rows_per_step = atoi(e);
        int64_t halo = slice < 4096 ? slice : 4096;
        const int64_t mr_smem = 2 * (halo + slice) * 4 + kKnapRing * (slice / 32) * 4 + 8 + (int64_t)n_items * 8;
        const int* skip = nullptr;
        if ((rows_per_step == 2 || rows_per_step == 4) && mr_smem <= kKnapSmemLimit && n_items > 0 &&
            slice % (kMT * kMCPT) == 0) {
            // 512 threads -> 128 registers per thread: 3 columns x 16 cells in flight
            auto k4 = knapsack_multirow_kernel<kMT, 4, kMCPT>;
            auto k2 = knapsack_multirow_kernel<kMT, 2, kMCPT>;
            cfg.blockDim = dim3(kMT);
            if (gn_attr_first(h, reinterpret_cast<const void*>(k4))) {
                GN_CUDA(cudaFuncSetAttribute(k4, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kKnapSmemLimit));
                GN_CUDA(cudaFuncSetAttribute(k2, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kKnapSmemLimit));
                GN_CUDA(cudaFuncSetAttribute(k4, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
                GN_CUDA(cudaFuncSetAttribute(k2, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
            }
            cfg.dynamicSmemBytes = (size_t)mr_smem;
            GN_CUDA(cudaLaunchKernelEx(&cfg, rows_per_step == 4 ? k4 : k2, d_values, d_weights, n_items, (int)capacity,
                                       capacity_rem == 0 ? 1 : 0, d_selected, d_best, bits, (int)words, (int)slice,
                                       (int)halo, status, cl));
            GN_LAUNCHED(h);
            skip = status;
            cfg.blockDim = dim3(kT);
        }
        auto kern = knapsack_cluster_kernel<kT>;
        if (gn_attr_first(h, reinterpret_cast<const void*>(kern))) {
            GN_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kKnapSmemLimit));
            GN_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
        }
        cfg.dynamicSmemBytes = (size_t)cl_smem;
        GN_CUDA(cudaLaunchKernelEx(&cfg, kern, d_values, d_weights, n_items, (int)capacity,
                                   capacity_rem == 0 ? 1 : 0, d_selected, d_best, bits, (int)words, (int)slice, skip));
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

// ---- gn_knapsack without a caller workspace and without any allocation ----------------------------------------------
// The scratch (decision bits, fallback table rows, status words) lives in a file-scope device array: every GPU that
// loads the library holds GN_KNAPSACK_STATIC_MAX_BYTES of it until exit.  In host code `g_knap_scratch` is only a
// stand-in for the symbol; its device address is looked up once per device with cudaGetSymbolAddress (that device
// current) and it is that address the kernels receive.
static __device__ __align__(256) unsigned char g_knap_scratch[GN_KNAPSACK_STATIC_MAX_BYTES];

namespace {
constexpr int kMaxStaticDevices = 64;
void* g_knap_scratch_ptr[kMaxStaticDevices];       // device address of g_knap_scratch on each device, 0 = not looked up
std::mutex g_knap_scratch_mutex;

int knap_scratch_address(int device, void** out) {
    if (device < 0 || device >= kMaxStaticDevices) return GN_E_ARG;
    std::lock_guard<std::mutex> lock(g_knap_scratch_mutex);
    if (!g_knap_scratch_ptr[device]) {
        int cur = -1;
        GN_CUDA(cudaGetDevice(&cur));
        if (cur != device) GN_CUDA(cudaSetDevice(device));
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_knap_scratch);
        if (cur != device) cudaSetDevice(cur);
        if (e != cudaSuccess) { cudaGetLastError(); return (int)e; }
        if (!p) return GN_E_NODEVICE;
        g_knap_scratch_ptr[device] = p;
    }
    *out = g_knap_scratch_ptr[device];
    return GN_OK;
}
}  // namespace

extern "C" int gn_knapsack_static(gn_handle* h, const int32_t* d_values, const int32_t* d_weights, int n_items,
                                  int64_t capacity, int capacity_rem, int batch, uint8_t* d_selected,
                                  int32_t* d_best, void* stream) {
    if (!h || n_items < 0 || capacity < 0 || batch < 0 || capacity_rem < 0) return GN_E_ARG;
    if (batch == 0) return GN_OK;
    if (capacity > (1LL << 31) - 64) return GN_E_RANGE;
    const int64_t per = gn_knapsack_workspace_bytes(n_items, capacity);
    if (per < 0) return (int)per;
    if (per > (int64_t)GN_KNAPSACK_STATIC_MAX_BYTES / batch) return GN_E_STATIC_LIMIT;
    void* scratch = nullptr;
    const int rc = knap_scratch_address(h->device, &scratch);
    if (rc != GN_OK) return rc;
    // Nothing in the scratch has to start at zero: every word a kernel reads (decision bits, table rows, status) is
    // written earlier in the same call.  The status words are cleared all the same, so that a call never sees the
    // previous call's "declined" marks.
    GN_CUDA(cudaMemsetAsync(reinterpret_cast<char*>(scratch) + per * batch - 256 * (int64_t)batch, 0,
                            256 * (size_t)batch, gn_stream(stream)));
    return gn_knapsack(h, d_values, d_weights, n_items, capacity, capacity_rem, batch, d_selected, d_best, scratch,
                       per * batch, stream);
}

// d_dst[m][r * per + j] = d_src[r][m][j] (see goalnet_b200.h): the layout change after an all-gather of frame shards
__global__ void __launch_bounds__(256)
unshard_rows_kernel(const float* __restrict__ src, int world, int64_t n_matches, int64_t per, int64_t n,
                    float* __restrict__ dst) {
    const int64_t total = n_matches * n;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t mt = i / n, f = i - mt * n;
        const int64_t r = f / per, j = f - r * per;
        dst[i] = src[(r * n_matches + mt) * per + j];
    }
}

extern "C" int gn_unshard_rows(gn_handle* h, const float* d_src, int world, int64_t n_matches, int64_t per, int64_t n,
                               float* d_dst, void* stream) {
    if (!h || world < 1 || n_matches < 0 || per < 0 || n < 0 || n > (int64_t)world * per) return GN_E_ARG;
    if (n_matches == 0 || n == 0) return GN_OK;
    if (!d_src || !d_dst) return GN_E_ARG;
    const int64_t blocks = (n_matches * n + 255) / 256, cap = (int64_t)h->num_sms * 8;
    unshard_rows_kernel<<<(int)(blocks < cap ? blocks : cap), 256, 0, gn_stream(stream)>>>(d_src, world, n_matches, per, n,
                                                                                          d_dst);
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
