// sort.cuh — stable LSD radix sort (8 bits per pass) used by the ordered dump and the KFF writer (SURVEY.md §8f rows
// 1 and 3: DenseMenuYo::next walks the buckets in increasing id, brisk/DenseMenuYo.hpp:476-521).
//
// What is sorted is a permutation (u32 values); the key of element i is produced by a functor from the value, so the
// keys never have to be materialised or moved: temporaries are the two permutation buffers plus the per-tile digit
// histograms.  One pass = sort_hist_kernel (digit histogram of every tile) -> sort_scan_kernel (exclusive scan in
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

// exclusive scan of hist[0, total) in place (one CTA; the totals stay below 2^32)
__global__ void __launch_bounds__(1024) sort_scan_kernel(u32 *hist, u64 total) {
    __shared__ u32 part[1024];
    const u64 per = (total + 1023) / 1024;
    const u64 a = (u64)threadIdx.x * per, b = a + per < total ? a + per : total;
    u32 s = 0;
    for (u64 i = a; i < b; i++) s += hist[i];
    part[threadIdx.x] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        u32 acc = 0;
        for (int i = 0; i < 1024; i++) { u32 v = part[i]; part[i] = acc; acc += v; }
    }
    __syncthreads();
    u32 run = part[threadIdx.x];
    for (u64 i = a; i < b; i++) { u32 v = hist[i]; hist[i] = run; run += v; }
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
// *result (perm_a or perm_b).  hist must hold 256 * ceil(n / SORT_TILE) u32.
template <class KeyOf>
static cudaError_t radix_sort_perm(KeyOf key_of, u32 *perm_a, u32 *perm_b, u64 n, int bit0, int bit1, u32 *hist, cudaStream_t stream,
                                   u32 **result) {
    *result = perm_a;
    if (n == 0) return cudaSuccess;
    const u32 n_tiles = (u32)((n + SORT_TILE - 1) / SORT_TILE);
    u32 *in = perm_a, *out = perm_b;
    for (int shift = bit0; shift < bit1; shift += 8) {
        sort_hist_kernel<KeyOf><<<n_tiles, SORT_THREADS, 0, stream>>>(key_of, in, n, shift, hist, n_tiles);
        sort_scan_kernel<<<1, 1024, 0, stream>>>(hist, (u64)256 * n_tiles);
        sort_scatter_kernel<KeyOf><<<n_tiles, SORT_THREADS, 0, stream>>>(key_of, in, n, shift, hist, n_tiles, out);
        u32 *t = in; in = out; out = t;
    }
    *result = in;
    return cudaGetLastError();
}
static inline size_t radix_sort_hist_bytes(u64 n) { return (size_t)256 * ((n + SORT_TILE - 1) / SORT_TILE) * sizeof(u32); }
