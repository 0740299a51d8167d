// partition.cuh — radix partition of superkmer records by the owner of their bucket (north-star item 5).
//
// In the reference every superkmer is routed to its bucket through the lock stripes of
// DenseMenuYo::insert_kmer_vector (brisk/DenseMenuYo.hpp:248-308).  Here the bucket id sits in the top 2b bits of
// the record header; the 4^b bucket ids are split over the ranks by brisk_owner_of_bucket and the records of a
// batch are partitioned into one contiguous region per destination rank before the all-to-all (comm.cuh).
//
//   partition_count_kernel   : per-destination totals (shared-memory histogram, one global atomic per CTA and
//                              destination)
//   partition_scatter_kernel : every CTA takes a tile of records, ranks them per destination in shared memory,
//                              reserves one contiguous run per destination with a single atomic, stages the tile
//                              in shared memory ordered by destination and writes it out with 128-bit stores, so
//                              that the lanes of a warp write consecutive 16-byte pieces of a run.
// Both passes stream the records once: algorithmic bytes = read 2x + write 1x the record bytes (pass 1 only needs
// the header word but a 32-byte sector is the unit HBM moves).
#pragma once
#include "common.cuh"

#define PART_MAX_RANKS 64
#define PART_THREADS 256
#define PART_TILE 512  // records per CTA iteration (2 per thread; 24 KB of staging at 48-byte records)

// owner rank of a bucket id — the device twin of brisk_owner_of_bucket (brisk_capi.cu).  Bucket ids are the middle
// bits of the hashed minimizer; the minimum of a window is biased towards small hashes, so the id is scrambled
// before the range split to keep the shards even.
__host__ __device__ __forceinline__ u32 owner_of_bucket(u64 bucket_id, u32 n_ranks) {
    u64 x = bucket_id * 0x9E3779B97F4A7C15ull;
    x ^= x >> 32;
#ifdef __CUDA_ARCH__
    // explicit mul-hi: nvcc 12.9 folded the 64-bit form ((x & 0xffffffff) * n_ranks) >> 32, scaled by the histogram's
    // element size, into x * ((n_ranks << 2) >> 32) = 0 inside partition_count_kernel (seen in the SASS and on the GPU)
    return __umulhi((u32)x, n_ranks);
#else
    return (u32)(((x & 0xffffffffull) * n_ranks) >> 32);
#endif
}

// destination of a record: mode 0 = owner rank of its bucket (sharded index), mode 1 = table region of its bucket
// (n_ranks = STORE_REGIONS; common.cuh)
__device__ __forceinline__ u32 partition_dest(u64 bucket, u32 n_ranks, int mode) {
    return mode == 0 ? owner_of_bucket(bucket, n_ranks) : region_of_bucket(bucket);
}

template <int RW>
__global__ void __launch_bounds__(PART_THREADS) partition_count_kernel(const u64 *__restrict__ recs, u64 n_recs, int bucket_shift,
                                                                       u32 n_ranks, int mode, u64 *__restrict__ totals) {
    __shared__ u32 hist[PART_MAX_RANKS];
    if (threadIdx.x < PART_MAX_RANKS) hist[threadIdx.x] = 0;
    __syncthreads();
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n_recs; i += (u64)gridDim.x * blockDim.x) {
        u64 top = __ldg(recs + i * RW + (RW - 1));
        atomicAdd(&hist[partition_dest(top >> bucket_shift, n_ranks, mode)], 1u);
    }
    __syncthreads();
    if (threadIdx.x < n_ranks && hist[threadIdx.x]) atomicAdd(totals + threadIdx.x, (u64)hist[threadIdx.x]);
}

// totals[n_ranks] -> exclusive prefix in starts[n_ranks + 1]; cursors[] reset to the region starts
__global__ void partition_starts_kernel(const u64 *totals, u32 n_ranks, u64 *starts, u64 *cursors) {
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        u64 acc = 0;
        for (u32 d = 0; d < n_ranks; d++) {
            starts[d] = acc;
            cursors[d] = acc;
            acc += totals[d];
        }
        starts[n_ranks] = acc;
    }
}

// rec_occ_in / rec_occ_out (optional): the per-record read-order occurrence tag travels with its record
template <int RW>
__global__ void __launch_bounds__(PART_THREADS) partition_scatter_kernel(const u64 *__restrict__ recs, u64 n_recs, int bucket_shift,
                                                                         u32 n_ranks, int mode, u64 *__restrict__ cursors,
                                                                         u64 *__restrict__ out, const u64 *__restrict__ occ_in,
                                                                         u64 *__restrict__ occ_out) {
    __shared__ u32 hist[PART_MAX_RANKS];       // records of the tile per destination
    __shared__ u32 lstart[PART_MAX_RANKS + 1]; // exclusive prefix of hist: where a destination's run starts in the tile
    __shared__ u64 gstart[PART_MAX_RANKS];     // where that run goes in the output
    __shared__ __align__(16) u64 stage[PART_TILE * RW];
    __shared__ u64 stage_occ[PART_TILE];
    constexpr int PER = PART_TILE / PART_THREADS;
    const u64 n_tiles = (n_recs + PART_TILE - 1) / PART_TILE;
    for (u64 tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const u64 base = tile * PART_TILE;
        const u32 cnt = (u32)((n_recs - base) < (u64)PART_TILE ? (n_recs - base) : (u64)PART_TILE);
        if (threadIdx.x < PART_MAX_RANKS) hist[threadIdx.x] = 0;
        __syncthreads();
        u64 rec[PER][RW];
        u64 occ[PER];
        u32 dest[PER], rank[PER];
#pragma unroll
        for (int q = 0; q < PER; q++) {
            const u32 e = q * PART_THREADS + threadIdx.x;  // consecutive threads read consecutive records
            dest[q] = 0xffffffffu;
            if (e < cnt) {
                const ulonglong2 *src = reinterpret_cast<const ulonglong2 *>(recs + (base + e) * RW);
#pragma unroll
                for (int i = 0; i < RW; i += 2) {
                    ulonglong2 v = __ldcs(src + (i >> 1));
                    rec[q][i] = v.x;
                    rec[q][i + 1] = v.y;
                }
                occ[q] = occ_in ? __ldcs(occ_in + base + e) : 0ull;
                dest[q] = partition_dest(rec[q][RW - 1] >> bucket_shift, n_ranks, mode);
                rank[q] = atomicAdd(&hist[dest[q]], 1u);
            }
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            u32 acc = 0;
            for (u32 d = 0; d < n_ranks; d++) {
                lstart[d] = acc;
                acc += hist[d];
            }
            lstart[n_ranks] = acc;
        }
        if (threadIdx.x < n_ranks && hist[threadIdx.x]) gstart[threadIdx.x] = atomicAdd(cursors + threadIdx.x, (u64)hist[threadIdx.x]);
        __syncthreads();
#pragma unroll
        for (int q = 0; q < PER; q++) {
            if (dest[q] != 0xffffffffu) {
                const u32 p = lstart[dest[q]] + rank[q];
                ulonglong2 *dst = reinterpret_cast<ulonglong2 *>(stage + (size_t)p * RW);
#pragma unroll
                for (int i = 0; i < RW; i += 2) dst[i >> 1] = make_ulonglong2(rec[q][i], rec[q][i + 1]);
                if (occ_in) stage_occ[p] = occ[q];
            }
        }
        __syncthreads();
        // write the tile out in 16-byte pieces: piece x of the tile belongs to staged record x / (RW/2)
        constexpr int PIECES = RW / 2;
        const ulonglong2 *st = reinterpret_cast<const ulonglong2 *>(stage);
        for (u32 x = threadIdx.x; x < cnt * PIECES; x += PART_THREADS) {
            const u32 p = x / PIECES, piece = x % PIECES;
            // destination of staged record p: the last d with lstart[d] <= p
            u32 d = 0;
            for (u32 step = PART_MAX_RANKS / 2; step > 0; step >>= 1)
                if (d + step < n_ranks && lstart[d + step] <= p) d += step;
            const u64 g = gstart[d] + (p - lstart[d]);
            reinterpret_cast<ulonglong2 *>(out + g * RW)[piece] = st[x];
            if (occ_out && piece == 0) occ_out[g] = stage_occ[p];
        }
        __syncthreads();
    }
}
