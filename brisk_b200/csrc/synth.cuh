// synth.cuh — the synthetic read generator of SURVEY.md §8(d) on the device (H6: 30 GB of ASCII reads cannot come over
// PCIe at kernel speed), bit-identical to brisk_b200/synth.py:
//     genome[i] = "ACGT"[splitmix64(seed ^ (0x9E3779B97F4A7C15 * (i + 1))) & 3]
//     read r    : start  = splitmix64(seed + 1 + 2r) mod (G - L + 1)
//                 strand = splitmix64(seed + 1 + 2r + 1) & 1
//                 bases  = genome[start : start + L], reverse-complemented when strand = 1
// Benchmark / test support, not part of the insertion path; the reference's counterpart is the FASTA file on disk.
#pragma once
#include "common.cuh"

__host__ __device__ __forceinline__ u64 splitmix64(u64 x) {
    u64 z = x + 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

// 16 genome bases per thread, one 128-bit store
__global__ void __launch_bounds__(256) synth_genome_kernel(unsigned char *__restrict__ genome, u64 G, u64 seed) {
    const u64 i0 = ((u64)blockIdx.x * blockDim.x + threadIdx.x) * 16;
    if (i0 >= G) return;
    unsigned char b[16];
#pragma unroll
    for (int j = 0; j < 16; j++) b[j] = "ACGT"[splitmix64(seed ^ (0x9E3779B97F4A7C15ull * (i0 + j + 1))) & 3];
    if (i0 + 16 <= G && (reinterpret_cast<size_t>(genome + i0) & 15) == 0) {
        *reinterpret_cast<uint4 *>(genome + i0) = *reinterpret_cast<const uint4 *>(b);
    } else {
        for (int j = 0; j < 16 && i0 + j < G; j++) genome[i0 + j] = b[j];
    }
}

// one warp per read: lanes copy (or reverse-complement) the read's bases
__global__ void __launch_bounds__(256) synth_reads_kernel(const unsigned char *__restrict__ genome, u64 G, u32 L, u64 seed, u64 first,
                                                          u64 count, unsigned char *__restrict__ out) {
    const u64 r = ((u64)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (r >= count) return;
    const u64 base = seed + 1 + 2 * (first + r);
    const u64 start = splitmix64(base) % (G - L + 1);
    const bool strand = splitmix64(base + 1) & 1;
    unsigned char *dst = out + r * L;
    for (u32 j = lane; j < L; j += 32) {
        unsigned char c;
        if (!strand) {
            c = genome[start + j];
        } else {
            const unsigned char g = genome[start + (L - 1 - j)];
            c = g == 'A' ? 'T' : g == 'C' ? 'G' : g == 'G' ? 'C' : 'A';
        }
        dst[j] = c;
    }
}

// ---- order-independent checksum of the stored set: sum over the stored (kmer_s, minimizer_idx, DATA byte) of a 64-bit
// mix, and the sum of the DATA bytes.  The sums of the shards of a sharded index add up to the sums of the unsharded
// index over the same reads (bench.py's parity_check, "a checksum of checksums").
__device__ __forceinline__ u64 checksum_mix(u64 lo, u64 hi, u32 idx, u32 data) {
    u64 x = lo * 0x9E3779B97F4A7C15ull ^ (hi + 0x632BE59BD9B4E019ull) * 0xD6E8FEB86659FD93ull ^ ((u64)idx << 8 | data) * 0xff51afd7ed558ccdull;
    x ^= x >> 32;
    x *= 0xc4ceb9fe1a85ec53ull;
    x ^= x >> 29;
    return x;
}
