// fasta.cuh — FASTA ingest on the device: getLineFasta + clean_dna of the reference's demo app
// (apps/counter.cpp:128-190) for a whole text buffer at once.
//
// What the reference's reader loop hands to count_sequence (one reader thread): for every record — the lines between
// two header lines, a header being the first line of the file or any line starting with '>' — the maximal runs of
// [ACGTacgt] of its concatenated lines ('\n' is transparent, every other byte cuts, '\r' included), upper-cased, in
// file order.  Three passes over the text, all HBM-bound:
//   1. fasta_linestart_kernel + inclusive max-scan : the start of the line every byte belongs to
//   2. fasta_classify_kernel + exclusive max-scan  : BASE / SKIP ('\n' of a sequence line) / BREAK per byte and, for
//                                                    every byte, the nearest preceding non-SKIP byte's class
//   3. exclusive sum of (is_base, is_run_start) packed in one u64, then fasta_emit_kernel writes the upper-cased
//      bases densely and one offset per run
// Texts are limited to 2^32 - 2 bytes per call (callers split big files at record boundaries).
#pragma once
#include "common.cuh"

__device__ __forceinline__ bool fasta_is_base(unsigned char c) {
    c &= 0xDF;  // upper-case
    return c == 'A' || c == 'C' || c == 'G' || c == 'T';
}

// marker[i] = i + 1 when byte i starts a line, else 0
__global__ void __launch_bounds__(256) fasta_linestart_kernel(const unsigned char *__restrict__ text, u64 n, u32 *__restrict__ marker) {
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x)
        marker[i] = (i == 0 || text[i - 1] == '\n') ? (u32)(i + 1) : 0u;
}

// linestart[i] (1-based, from the scan) -> cls[i] = (i + 1) << 1 | is_break for non-SKIP bytes, 0 for SKIP bytes;
// kind[i] = 1 for a BASE byte, 0 otherwise
__global__ void __launch_bounds__(256) fasta_classify_kernel(const unsigned char *__restrict__ text, u64 n, const u32 *__restrict__ linestart,
                                                             u64 *__restrict__ cls, unsigned char *__restrict__ kind) {
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x) {
        const u32 ls = linestart[i] - 1;
        const unsigned char c = text[i];
        const bool header = ls == 0 || text[ls] == '>';  // the first line is consumed as a header whatever it holds
        const bool base = !header && fasta_is_base(c);
        const bool skip = !header && c == '\n';
        kind[i] = base ? 1 : 0;
        cls[i] = skip ? 0ull : (((i + 1) << 1) | (base ? 0ull : 1ull));
    }
}

// prev[i] = exclusive max-scan of cls: the nearest preceding non-SKIP byte.  flags[i] = is_base | is_run_start << 32
__global__ void __launch_bounds__(256) fasta_flags_kernel(const unsigned char *__restrict__ kind, const u64 *__restrict__ prev, u64 n,
                                                          u64 *__restrict__ flags) {
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x) {
        const u64 base = kind[i];
        const u64 start = base && (prev[i] == 0 || (prev[i] & 1)) ? 1ull : 0ull;
        flags[i] = base | (start << 32);
    }
}

// pos[i] = exclusive sum of flags: low word = index among the bases, high word = index among the runs
__global__ void __launch_bounds__(256) fasta_emit_kernel(const unsigned char *__restrict__ text, const u64 *__restrict__ flags,
                                                         const u64 *__restrict__ pos, u64 n, unsigned char *__restrict__ bases,
                                                         u64 *__restrict__ offsets) {
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x) {
        const u64 f = flags[i];
        if (f & 1) {
            const u64 p = pos[i];
            bases[p & 0xffffffffull] = text[i] & 0xDF;
            if (f >> 32) offsets[p >> 32] = p & 0xffffffffull;
        }
    }
}

struct MaxOp64 {
    __device__ __forceinline__ u64 operator()(u64 a, u64 b) const { return a > b ? a : b; }
};
struct MaxOp32 {
    __device__ __forceinline__ u32 operator()(u32 a, u32 b) const { return a > b ? a : b; }
};
