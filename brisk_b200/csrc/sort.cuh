// sort.cuh — stable LSD radix sort (8 bits per pass) used by the ordered dump and the KFF writer (SURVEY.md §8f rows
// 1 and 3: DenseMenuYo::next walks the buckets in increasing id, brisk/DenseMenuYo.hpp:476-521).
//
// What is sorted is a permutation (u32 values); the key of element i is produced by a functor from the value, so the
// keys never have to be materialised or moved: temporaries are the two permutation buffers plus the per-tile digit
// histograms.  One pass = sort_hist_kernel (digit histogram of every tile) -> sort_scan_*_kernel (exclusive scan in
// digit-major order: where every (digit, tile) run starts in the output) -> sort_scatter_kernel (every warp walks its
// rows in input order and ranks equal digits with match.any, which keeps the pass stable).
#pragma once
#include "common.cuh"

#define SORT_THREADS 256
#define SORT_ROWS 8                                   // rows of 32 elements per warp
#define SORT_TILE (SORT_THREADS * SORT_ROWS)          // 2048 elements per CTA

template <class KeyOf>
__global__ void __launch_bounds__(SORT_THREADS) sort_hist_kernel(KeyOf key_of, const u32 *__restrict__ perm, u64 n, int shift,
                                                                 u32 *__restrict__ hist, u32 n_tiles) {
    __shared__ u32 h[256];
    h[threadIdx.x] = 0;
    __syncthreads();
    const u64 base = (u64)blockIdx.x * SORT_TILE;
#pragma unroll
    for (int r = 0; r < SORT_ROWS; r++) {
        const u64 i = base + (u64)r * SORT_THREADS + threadIdx.x;
        if (i < n) atomicAdd(&h[(key_of(perm[i]) >> shift) & 255u], 1u);
    }
    __syncthreads();
    hist[(u64)threadIdx.x * n_tiles + blockIdx.x] = h[threadIdx.x];
}

// exclusive scan of hist[0, total) in place (the totals stay below 2^32), in three steps: every CTA scans its chunk of
// SCAN_CHUNK entries and leaves the chunk's sum in partials[], one CTA scans partials[], every CTA adds its offset
#define SCAN_THREADS 1024
#define SCAN_PER_THREAD 8
#define SCAN_CHUNK (SCAN_THREADS * SCAN_PER_THREAD)

// exclusive scan of one value per thread over the CTA (1024 threads); *total = the CTA's sum
__device__ __forceinline__ u32 scan_block_exclusive(u32 v, u32 *total) {
    __shared__ u32 wsum[SCAN_THREADS / 32];
    __shared__ u32 wtot;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    u32 inc = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const u32 t = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= d) inc += t;
    }
    if (lane == 31) wsum[w] = inc;
    __syncthreads();
    if (w == 0) {
        const u32 x = wsum[lane];
        u32 xi = x;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const u32 t = __shfl_up_sync(0xffffffffu, xi, d);
            if (lane >= d) xi += t;
        }
        wsum[lane] = xi - x;
        if (lane == 31) wtot = xi;
    }
    __syncthreads();
    const u32 r = wsum[w] + inc - v;
    *total = wtot;
    __syncthreads();  // the shared words may be reused by the caller's next scan
    return r;
}

__global__ void __launch_bounds__(SCAN_THREADS) sort_scan_chunks_kernel(u32 *hist, u64 total, u32 *partials) {
    const u64 base = (u64)blockIdx.x * SCAN_CHUNK + (u64)threadIdx.x * SCAN_PER_THREAD;
    u32 v[SCAN_PER_THREAD], s = 0;
#pragma unroll
    for (int i = 0; i < SCAN_PER_THREAD; i++) {
        v[i] = base + i < total ? hist[base + i] : 0u;
        s += v[i];
    }
    u32 chunk_total;
    u32 run = scan_block_exclusive(s, &chunk_total);
#pragma unroll
    for (int i = 0; i < SCAN_PER_THREAD; i++) {
        if (base + i < total) hist[base + i] = run;
        run += v[i];
    }
    if (threadIdx.x == 0) partials[blockIdx.x] = chunk_total;
}

// exclusive scan of partials[0, n) in place by one CTA (n = number of chunks: a few thousand at most)
__global__ void __launch_bounds__(SCAN_THREADS) sort_scan_kernel(u32 *partials, u64 n) {
    __shared__ u32 carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (u64 base = 0; base < n; base += SCAN_THREADS) {
        const u64 i = base + threadIdx.x;
        const u32 v = i < n ? partials[i] : 0u;
        u32 tot;
        const u32 ex = scan_block_exclusive(v, &tot);
        if (i < n) partials[i] = carry + ex;
        __syncthreads();
        if (threadIdx.x == 0) carry += tot;
        __syncthreads();
    }
}

__global__ void __launch_bounds__(SCAN_THREADS) sort_scan_add_kernel(u32 *hist, u64 total, const u32 *__restrict__ partials) {
    const u32 off = partials[blockIdx.x];
    const u64 base = (u64)blockIdx.x * SCAN_CHUNK + (u64)threadIdx.x * SCAN_PER_THREAD;
#pragma unroll
    for (int i = 0; i < SCAN_PER_THREAD; i++)
        if (base + i < total) hist[base + i] += off;
}

template <class KeyOf>
__global__ void __launch_bounds__(SORT_THREADS) sort_scatter_kernel(KeyOf key_of, const u32 *__restrict__ perm, u64 n, int shift,
                                                                    const u32 *__restrict__ hist, u32 n_tiles, u32 *__restrict__ out) {
    __shared__ u32 wh[SORT_THREADS / 32][256];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const u32 lane_lt = (1u << lane) - 1u;
    for (int i = threadIdx.x; i < (SORT_THREADS / 32) * 256; i += SORT_THREADS) (&wh[0][0])[i] = 0;
    __syncthreads();
    // warp w owns the elements [base + w * 256, base + (w + 1) * 256) of the tile, row after row
    const u64 base = (u64)blockIdx.x * SORT_TILE + (u64)w * (32 * SORT_ROWS);
    u32 v[SORT_ROWS], d[SORT_ROWS];
#pragma unroll
    for (int r = 0; r < SORT_ROWS; r++) {
        const u64 i = base + (u64)r * 32 + lane;
        const bool ok = i < n;
        v[r] = ok ? perm[i] : 0u;
        d[r] = ok ? (u32)((key_of(v[r]) >> shift) & 255u) : 256u;
        const unsigned act = __ballot_sync(0xffffffffu, ok);
        if (ok) {
            const unsigned peers = __match_any_sync(act, d[r]);
            if ((peers & lane_lt) == 0) wh[w][d[r]] += __popc(peers);  // one leader per digit value
        }
        __syncwarp();
    }
    __syncthreads();
    {   // digit t: where its run of this tile starts in the output, split over the warps in order
        const int t = threadIdx.x;
        u32 run = hist[(u64)t * n_tiles + blockIdx.x];
        for (int ww = 0; ww < SORT_THREADS / 32; ww++) { const u32 c = wh[ww][t]; wh[ww][t] = run; run += c; }
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < SORT_ROWS; r++) {
        const bool ok = d[r] < 256u;
        const unsigned act = __ballot_sync(0xffffffffu, ok);
        if (ok) {
            const unsigned peers = __match_any_sync(act, d[r]);
            const int leader = __ffs(peers) - 1;
            u32 start = 0;
            if (lane == leader) { start = wh[w][d[r]]; wh[w][d[r]] = start + __popc(peers); }
            start = __shfl_sync(peers, start, leader);
            out[start + __popc(peers & lane_lt)] = v[r];
        }
        __syncwarp();
    }
}

__global__ void __launch_bounds__(256) sort_iota_kernel(u32 *perm, u64 n) {
    const u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) perm[i] = (u32)i;
}

// sorts perm_a[0, n) by the bits [bit0, bit1) of key_of(value), least significant digit first; the result ends in
// *result (perm_a or perm_b).  hist must hold radix_sort_hist_bytes(n): 256 * ceil(n / SORT_TILE) u32 plus the scan's partials.
template <class KeyOf>
static cudaError_t radix_sort_perm(KeyOf key_of, u32 *perm_a, u32 *perm_b, u64 n, int bit0, int bit1, u32 *hist, cudaStream_t stream,
                                   u32 **result) {
    *result = perm_a;
    if (n == 0) return cudaSuccess;
    const u32 n_tiles = (u32)((n + SORT_TILE - 1) / SORT_TILE);
    u32 *in = perm_a, *out = perm_b;
    const u64 total = (u64)256 * n_tiles;
    const u32 n_chunks = (u32)((total + SCAN_CHUNK - 1) / SCAN_CHUNK);
    u32 *partials = hist + total;  // behind the histograms (radix_sort_hist_bytes)
    for (int shift = bit0; shift < bit1; shift += 8) {
        sort_hist_kernel<KeyOf><<<n_tiles, SORT_THREADS, 0, stream>>>(key_of, in, n, shift, hist, n_tiles);
        sort_scan_chunks_kernel<<<n_chunks, SCAN_THREADS, 0, stream>>>(hist, total, partials);
        sort_scan_kernel<<<1, SCAN_THREADS, 0, stream>>>(partials, n_chunks);
        sort_scan_add_kernel<<<n_chunks, SCAN_THREADS, 0, stream>>>(hist, total, partials);
        sort_scatter_kernel<KeyOf><<<n_tiles, SORT_THREADS, 0, stream>>>(key_of, in, n, shift, hist, n_tiles, out);
        u32 *t = in; in = out; out = t;
    }
    *result = in;
    return cudaGetLastError();
}
static inline size_t radix_sort_hist_bytes(u64 n) {
    const size_t total = (size_t)256 * ((n + SORT_TILE - 1) / SORT_TILE);
    return (total + (total + SCAN_CHUNK - 1) / SCAN_CHUNK + 1) * sizeof(u32);
}
