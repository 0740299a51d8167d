// common.cuh — parameters and device helpers shared by the kernels of libbrisk_b200.
//
// Bit conventions follow the reference: nucleotide code (c>>1)&3 (brisk/Kmers.cpp:442-444),
// a k-mer is an integer whose least-significant 2 bits are its LAST base, complement = ^2.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

typedef unsigned long long u64;
typedef unsigned int u32;

#define BRISK_MAX_M 31
#define BRISK_MAX_RW 6

// Parameters(k,m,b) (brisk/parameters.hpp:24-34) plus everything derived from them that the
// kernels need.  Passed by value as a kernel argument.
struct BriskParams {
    int k, m, b;
    int w;   // k - m + 1: number of m-mers in a k-mer = largest superkmer
    int s;   // (m - b + 1) / 2: low nucleotides of the hashed minimizer that stay in the key (Brisk.hpp:135)
    int wb;  // bits of the (n-1) and idx0 record fields
    int rec_words;     // u64 words per superkmer record (2, 4 or 6)
    int payload_bits;  // 2 * (2k - m)
    int key_words;     // 1 when 2k <= 64, else 2
    int literal_decycle;
    u64 m_mask;        // 4^m - 1
    u64 kmask_lo, kmask_hi;
    u64 bucket_mask;   // 4^b - 1
    u64 polyA_hash;    // bfc_hash_64 of the all-A m-mer (zero-padded pseudo m-mers)
    double eps;        // Decycling.h:19
    double coef[4 * 32];    // DecyclingSet::coef, computed on the host with the same libm (Decycling.cpp:7-13)
    // decycling class in fixed point (enumerate.cuh): w = exp(2 pi i / m)
    int2 dec_tw[32];        // (cos, sin)(2 pi q / m) * 2^24: weight of a base at read position u, q = u mod m
    int2 dec_rot[32];       // (cos, sin)(2 pi q / m) * 2^30: rotation back to the m-mer's own frame
    int rot1_c30, rot1_s30; // = dec_rot[1]
    int dec_th;             // |R|, |R(rot)| (units of 2^-22) at or below this -> literal recomputation
    // [dec_groups][64] fixed-point (2^24) complex partial sums of the decycling weights over groups of 3 nucleotides
    // (device memory, copied to shared memory by the enumeration kernel): entry (g, v) = sum_{j<3} code(nucleotide j
    // of v) * w^(m-1-(3g+j)); built in brisk_create
    const int2 *dec_tab;
    int dec_groups;  // ceil(m / 3)
};

// ---------------------------------------------------------------- hashing.cpp
// integer part of bfc_hash_64 (brisk/hashing.cpp:10-16)
__host__ __device__ __forceinline__ u64 bfc_mix(u64 key, u64 mask) {
    key = (~key + (key << 21)) & mask;
    key = key ^ key >> 24;
    key = ((key + (key << 3)) + (key << 8)) & mask;
    key = key ^ key >> 14;
    key = ((key + (key << 2)) + (key << 4)) & mask;
    key = key ^ key >> 28;
    key = (key + (key << 31)) & mask;
    return key;
}
// the same on 32-bit registers, valid when mask < 2^32 (m <= 16): every step is masked, left
// shifts and additions only carry upwards, right shifts act on already-masked values.
__device__ __forceinline__ u32 bfc_mix32(u32 key, u32 mask) {
    key = (key * 0x1fffffu - 1u) & mask;  // ~key + (key << 21) = key * (2^21 - 1) - 1
    key = key ^ key >> 24;
    key = (key * 265u) & mask;            // key + (key << 3) + (key << 8)
    key = key ^ key >> 14;
    key = (key * 21u) & mask;             // key + (key << 2) + (key << 4)
    key = key ^ key >> 28;
    key = (key * 0x80000001u) & mask;     // key + (key << 31)
    return key;
}
// bfc_mix with its shift-and-add rounds written as the multiplications they are (identical modulo 2^64): the
// enumeration kernel is bound by the integer ALU pipe (shifts, logic ops, adds), multiplies run on the FMA pipe
__device__ __forceinline__ u64 bfc_mix_mul(u64 key, u64 mask) {
    key = (key * 0x1fffffull - 1ull) & mask;
    key = key ^ key >> 24;
    key = (key * 265ull) & mask;
    key = key ^ key >> 14;
    key = (key * 21ull) & mask;
    key = key ^ key >> 28;
    key = (key * 0x80000001ull) & mask;
    return key;
}

// ---------------------------------------------------------------- Kmers.cpp
// reverse the order of the 32 two-bit groups of a word
__host__ __device__ __forceinline__ u64 rev2(u64 x) {
#ifdef __CUDA_ARCH__
    x = __brevll(x);  // all 64 bits reversed; swapping neighbours restores the order inside each pair
    return ((x & 0x5555555555555555ull) << 1) | ((x >> 1) & 0x5555555555555555ull);
#else
    x = __builtin_bswap64(x);
    x = ((x & 0x0f0f0f0f0f0f0f0full) << 4) | ((x >> 4) & 0x0f0f0f0f0f0f0f0full);
    return ((x & 0x3333333333333333ull) << 2) | ((x >> 2) & 0x3333333333333333ull);
#endif
}
// rcbc (brisk/Kmers.cpp:320-332): reverse complement of an n-mer, n <= 32
__host__ __device__ __forceinline__ u64 rcbc(u64 in, int n) { return rev2(in ^ 0xaaaaaaaaaaaaaaaaull) >> (64 - 2 * n); }

// rcb (brisk/Kmers.cpp:293-316) AS WRITTEN: the byte shuffle's result is discarded there, so only the
// four nucleotides inside every byte are reversed, then complemented, then the 128-bit value is
// shifted right by 128 - 2n.
__host__ __device__ __forceinline__ void rcb128(u64 lo, u64 hi, int n, u64 &olo, u64 &ohi) {
    const u64 c1 = 0x0f0f0f0f0f0f0f0full, c2 = 0x3333333333333333ull;
    lo = ((lo & c1) << 4) | ((lo & (c1 << 4)) >> 4);
    lo = ((lo & c2) << 2) | ((lo & (c2 << 2)) >> 2);
    lo ^= 0xaaaaaaaaaaaaaaaaull;
    hi = ((hi & c1) << 4) | ((hi & (c1 << 4)) >> 4);
    hi = ((hi & c2) << 2) | ((hi & (c2 << 2)) >> 2);
    hi ^= 0xaaaaaaaaaaaaaaaaull;
    int count = 128 - 2 * n;
    if (count >= 128) { olo = 0; ohi = 0; }
    else if (count == 0) { olo = lo; ohi = hi; }
    else if (count >= 64) { olo = hi >> (count - 64); ohi = 0; }
    else { olo = (lo >> count) | (hi << (64 - count)); ohi = hi >> count; }
}
// canonized(kint) (brisk/Kmers.cpp:342-353)
__host__ __device__ __forceinline__ bool canonized128(u64 lo, u64 hi, int n) {
    u64 rlo, rhi;
    rcb128(lo, hi, n, rlo, rhi);
    bool r_smaller = (rhi < hi) || (rhi == hi && rlo < lo);
    return !r_smaller;  // x == min(x, rcb(x))
}

// ---------------------------------------------------------------- Decycling.cpp
// computeR (brisk/Decycling.cpp:17-24), literal summation order
__host__ __device__ __forceinline__ double dec_computeR(const double *coef, u64 seq, int m) {
    double R = 0;
    for (int i = 4 * (m - 1); i > 0; i -= 4) {
        R += coef[i + (int)(seq & 3)];
        seq >>= 2;
    }
    return R;
}
// memDouble (brisk/Decycling.cpp:38-52)
__host__ __device__ __forceinline__ u32 dec_mem_double(const double *coef, double eps, u64 seq, int m) {
    double R = dec_computeR(coef, seq, m);
    if (R > eps) {
        u64 rot = ((seq & 3) << (2 * (m - 1))) + (seq >> 2);
        if (dec_computeR(coef, rot, m) < eps) return 0;
    } else if (R < -eps) {
        u64 rot = ((seq & 3) << (2 * (m - 1))) + (seq >> 2);
        if (dec_computeR(coef, rot, m) > -eps) return 1;
    }
    return 2;
}
// bfc_hash_64 (brisk/hashing.cpp:8-19), literal
__host__ __device__ __forceinline__ u64 bfc_hash_literal(const double *coef, double eps, u64 key, u64 mask, int m) {
    return ((u64)dec_mem_double(coef, eps, key, m) << 62) + bfc_mix(key, mask);
}

// ---------------------------------------------------------------- packed reads
// Reads are packed 32 bases per u64, FIRST base in the two most significant bits, each read starting
// on a word boundary: first word of read r = offsets[r]/32 + r (no scan needed, never overlaps).
__host__ __device__ __forceinline__ u64 packed_start(u64 byte_off, u64 r) { return (byte_off >> 5) + r; }

// the 32 bases [a, a+32) of a packed read, base a in the top bits (bases past the end are garbage)
__device__ __forceinline__ u64 get_bits64(const u64 *pk, u64 a) {
    u64 q = a >> 5;
    int r = 2 * (int)(a & 31);
    u64 x = __ldg(pk + q);
    if (r) x = (x << r) | (__ldg(pk + q + 1) >> (64 - r));
    return x;
}
// bases [a, a+len) right-aligned in one word, len <= 32
__device__ __forceinline__ u64 get_bases(const u64 *pk, u64 a, int len) { return len ? get_bits64(pk, a) >> (64 - 2 * len) : 0; }

// ---------------------------------------------------------------- multiword helpers (registers)
// Shifts the little-endian multiword value v[NW] right by sh bits, in place.  Whole-word moves are done as
// log-steps with static indices so that v[] stays in registers (a dynamically indexed v[q] would be spilled
// to local memory, which the first ncu capture showed as one extra DRAM sector write per k-mer).
template <int NW, int MAXQ>
__device__ __forceinline__ void mw_shr(u64 (&v)[NW], int sh) {
    const int q = sh >> 6, r = sh & 63;
    if (MAXQ >= 1 && (q & 1)) {
#pragma unroll
        for (int i = 0; i < NW; i++) v[i] = i + 1 < NW ? v[i + 1] : 0ull;
    }
    if (MAXQ >= 2 && (q & 2)) {
#pragma unroll
        for (int i = 0; i < NW; i++) v[i] = i + 2 < NW ? v[i + 2] : 0ull;
    }
    if (MAXQ >= 4 && (q & 4)) {
#pragma unroll
        for (int i = 0; i < NW; i++) v[i] = i + 4 < NW ? v[i + 4] : 0ull;
    }
    if (r) {
#pragma unroll
        for (int i = 0; i < NW; i++) v[i] = (v[i] >> r) | (i + 1 < NW ? v[i + 1] << (64 - r) : 0ull);
    }
}
// 64 bits of v[NW] starting at bit sh (v is not modified)
template <int NW, int MAXQ>
__device__ __forceinline__ u64 mw_extract64(const u64 (&v)[NW], int sh) {
    u64 t[NW];
#pragma unroll
    for (int i = 0; i < NW; i++) t[i] = v[i];
    mw_shr<NW, MAXQ>(t, sh);
    return t[0];
}

// ---------------------------------------------------------------- record layout
// record = rec_words u64, little-endian multiword:
//   bits [0, 2*len)                  the superkmer's bases in its orientation, last base at bits 1..0
//   top word: bucket id in the top 2b bits, then (n-1) in wb bits, then idx0 in wb bits
__host__ __device__ __forceinline__ u64 rec_header(const BriskParams &P, u64 bucket, int n, int idx0) {
    int sb = 64 - 2 * P.b;
    return (bucket << sb) | ((u64)(n - 1) << (sb - P.wb)) | ((u64)idx0 << (sb - 2 * P.wb));
}
__host__ __device__ __forceinline__ void rec_unpack_header(const BriskParams &P, u64 top, u64 &bucket, int &n, int &idx0) {
    int sb = 64 - 2 * P.b;
    u64 fm = ((u64)1 << P.wb) - 1;
    bucket = top >> sb;
    n = (int)((top >> (sb - P.wb)) & fm) + 1;
    idx0 = (int)((top >> (sb - 2 * P.wb)) & fm);
}
__host__ __device__ __forceinline__ u64 rec_payload_mask_top(const BriskParams &P) {
    int hb = 2 * P.b + 2 * P.wb;  // header bits in the top word
    return hb >= 64 ? 0 : (~(u64)0 >> hb);
}

// scramble of the table position: murmur3 finaliser over (lo, hi, idx)
__host__ __device__ __forceinline__ u64 slot_hash(u64 lo, u64 hi, u32 idx) {
    u64 x = lo ^ (hi * 0x9E3779B97F4A7C15ull) ^ ((u64)(idx + 1) * 0xD6E8FEB86659FD93ull);
    x ^= x >> 33;
    x *= 0xff51afd7ed558ccdull;
    x ^= x >> 33;
    x *= 0xc4ceb9fe1a85ec53ull;
    x ^= x >> 33;
    return x;
}

// ---- placement in the bucket store -----------------------------------------------------------------
// All k-mers of a superkmer share the m-mer found at minimizer_idx (the minimizer, as it is written in the
// superkmer's orientation) and have consecutive minimizer_idx.  The home slot of (kmer_s, idx) is therefore
//     group_home(minimizer m-mer, its bucket id) + idx
// so that the n k-mers of a superkmer sit in n CONSECUTIVE slots: the lanes of a warp that walk a superkmer
// touch whole 128-byte lines instead of n random 32-byte sectors.  Collisions jump by STORE_STRIDE (>= the
// largest superkmer), which keeps a displaced piece of a superkmer contiguous; after STORE_GROUP_PROBES jumps
// the probe sequence switches to a per-key hash so that a minimizer shared by very many k-mers (low-complexity
// sequence) cannot build a long chain.
#define STORE_STRIDE 64u
#define STORE_GROUP_PROBES 4u
__host__ __device__ __forceinline__ u64 minimizer_word(const BriskParams &P, u64 lo, u64 hi, int idx) {
    int sh = 2 * idx;
    u64 W = sh == 0 ? lo : (sh >= 64 ? hi >> (sh - 64) : (lo >> sh) | (hi << (64 - sh)));
    return W & P.m_mask;
}
__host__ __device__ __forceinline__ u64 group_base(u64 W) {
    u64 x = (W + 0x9E3779B97F4A7C15ull) * 0xD6E8FEB86659FD93ull;
    x ^= x >> 32;
    x *= 0xD6E8FEB86659FD93ull;
    x ^= x >> 29;
    return x;
}
// Table regions.  The table is cut into STORE_REGIONS equal regions and a minimizer's group lives in the region chosen
// by a hash of its BUCKET id (the record header carries the bucket id), so a batch whose records have been partitioned
// by region (partition.cuh, mode 1) walks the table region by region: the slots a batch touches in one region (a few
// tens of MB) stay in the 126 MB L2 while the batch's repeated k-mers come back to them, and the TLB sees one region at
// a time.  The region hash is independent of the rank split of the sharded index.
#define STORE_REGIONS 64u
__host__ __device__ __forceinline__ u32 region_of_bucket(u64 bucket_id) {
    u64 x = (bucket_id + 0x51ED270B27B4F3CFull) * 0xD6E8FEB86659FD93ull;
    x ^= x >> 29;
    return (u32)(x >> 40) & (STORE_REGIONS - 1);
}
// first slot of the group of minimizer m-mer W of bucket `bucket` in a table of mask + 1 slots (mask + 1 >= 1024)
__host__ __device__ __forceinline__ u64 group_home(u64 W, u64 bucket, u64 mask) {
    const u64 region_slots = (mask + 1) / STORE_REGIONS;
    return (u64)region_of_bucket(bucket) * region_slots + (group_base(W) & (region_slots - 1));
}
// slot visited by probe number i (i = 0 is the home slot)
__host__ __device__ __forceinline__ u64 probe_slot(u64 home, u64 lo, u64 hi, u32 idx, u32 i, u64 mask) {
    if (i < STORE_GROUP_PROBES) return (home + (u64)i * STORE_STRIDE) & mask;
    return (slot_hash(lo, hi, idx) + (i - STORE_GROUP_PROBES)) & mask;
}

// bucket id of a k-mer (brisk/Brisk.hpp:135-137 after hash_kmer_minimizer_inplace, Kmers.cpp:191-200):
// middle 2b bits of the hashed m-mer found at minimizer_idx
__host__ __device__ __forceinline__ u64 bucket_of_kmer(const BriskParams &P, u64 lo, u64 hi, int idx) {
    int sh = 2 * idx;
    u64 W = sh == 0 ? lo : (sh >= 64 ? hi >> (sh - 64) : (lo >> sh) | (hi << (64 - sh)));
    W &= P.m_mask;
    return (bfc_mix(W, P.m_mask) >> (2 * P.s)) & P.bucket_mask;
}
