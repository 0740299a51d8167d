// kff.cuh — device side of the KFF writer (the reference: BriskWriter::write, brisk/writer.hpp:75-179,
// over brisk/lib/kff).  The reference walks its buckets in id order and emits, per bucket and per run of equal
// minimizers, one Section_Minimizer whose blocks are its stored superkmers (SKL records): the sequence without the
// minimizer, the minimizer's position, the DATA of its k-mers.  Its SKL records are what insertion order happened to
// compact; here the stored k-mers are re-assembled into superkmers from the table itself:
//
//   kff_links_kernel   every stored (kmer_s, idx) looks up its <= 4 possible successors ((kmer_s << 2 | c), idx + 1) and
//                      <= 4 possible predecessors in the table; an edge A -> B exists when B is A's only successor and A
//                      is B's only predecessor, so the chains are disjoint paths (idx grows by one per step: no cycles)
//   kff_edges_kernel   marks the edges, kff_chains_kernel emits one chain record per path start: (sort key = bucket id,
//                      then the rest of the hashed minimizer; first slot; number of k-mers)
//   -> sort.cuh orders the chains by bucket id / minimizer, a scan lays out sections and blocks,
//   kff_emit_kernel    writes the section headers and the blocks (big-endian, right-aligned 2-bit sequences exactly as
//                      to_big_endian_compact / Section_Minimizer::write_compacted_sequence_without_mini produce them).
#pragma once
#include "store.cuh"

#define KFF_SUCC_MASK ((1ull << 40) - 1)
#define KFF_NSUCC_SHIFT 40
#define KFF_NPRED_SHIFT 44
#define KFF_EDGE (1ull << 48)       // this key's only successor takes it as its only predecessor
#define KFF_HAS_PRED (1ull << 49)   // this key is the target of an edge
#define KFF_VALID (1ull << 63)

struct KffChain { u64 sortkey; u64 slot; u32 n; u32 pad; };

struct KffGeom {
    int k, m, data_size;
    int nb_kmers_bytes, mini_pos_bytes, mini_bytes;
    __host__ __device__ u64 block_bytes(u32 n) const {
        return (u64)nb_kmers_bytes + mini_pos_bytes + ((u64)(n + k - 1 - m) + 3) / 4 + (u64)n * data_size;
    }
    __host__ __device__ u64 section_header_bytes() const { return 1 + (u64)mini_bytes + 8; }
};

template <int KW>
__device__ __forceinline__ bool kff_load(const StoreView &T, u64 slot, u64 &lo, u64 &hi, u32 &idx) {
    typedef typename SlotOf<KW>::type S;
    const S *p = reinterpret_cast<const S *>(T.slots) + slot;
    lo = *reinterpret_cast<const u64 *>(p);
    hi = KW == 2 ? reinterpret_cast<const u64 *>(p)[1] : 0;
    if (lo == STORE_EMPTY && (KW == 1 || hi == STORE_EMPTY)) return false;
    idx = (p->meta & META_TAG_MASK) - 1;
    return true;
}

template <int KW>
__global__ void __launch_bounds__(256) kff_links_kernel(const BriskParams P, StoreView T, u64 *__restrict__ link) {
    const int k = P.k, kmax = P.k - P.m;
    for (u64 s = (u64)blockIdx.x * blockDim.x + threadIdx.x; s <= T.mask; s += (u64)gridDim.x * blockDim.x) {
        u64 lo, hi; u32 idx;
        if (!kff_load<KW>(T, s, lo, hi, idx)) { link[s] = 0; continue; }
        const u64 gbase = group_home(minimizer_word(P, lo, hi, (int)idx), bucket_of_kmer(P, lo, hi, (int)idx), T.mask);
        u64 nsucc = 0, npred = 0, succ = 0;
        if ((int)idx + 1 <= kmax) {
            // successors: kmer_s shifted left by one base, a new last base, the minimizer one position further from the end
            u64 blo = lo << 2, bhi = KW == 2 ? ((hi << 2) | (lo >> 62)) : 0;
            blo &= P.kmask_lo; bhi &= P.kmask_hi;
            for (u64 c = 0; c < 4; c++) {
                u64 fs; u32 mt;
                const u64 qlo = blo | c;
                if (qlo == STORE_EMPTY && (KW == 1 || bhi == STORE_EMPTY)) continue;  // the key outside the table: its own chain
                if (store_find<KW>(T, qlo, bhi, idx + 1, (gbase + idx + 1) & T.mask, fs, mt)) { nsucc++; succ = fs; }
            }
        }
        if (idx >= 1) {
            // predecessors: shifted right by one base, a new first base
            const int top = 2 * (k - 1);
            u64 blo = (lo >> 2) | (KW == 2 ? hi << 62 : 0), bhi = KW == 2 ? hi >> 2 : 0;
            for (u64 c = 0; c < 4; c++) {
                u64 qlo = blo, qhi = bhi;
                if (top < 64) qlo |= c << top; else qhi |= c << (top - 64);
                if (qlo == STORE_EMPTY && (KW == 1 || qhi == STORE_EMPTY)) continue;
                u64 fs; u32 mt;
                if (store_find<KW>(T, qlo, qhi, idx - 1, (gbase + idx - 1) & T.mask, fs, mt)) npred++;
            }
        }
        link[s] = KFF_VALID | (succ & KFF_SUCC_MASK) | (nsucc << KFF_NSUCC_SHIFT) | (npred << KFF_NPRED_SHIFT);
    }
}

__global__ void __launch_bounds__(256) kff_edges_kernel(u64 *link, u64 n_slots) {
    for (u64 s = (u64)blockIdx.x * blockDim.x + threadIdx.x; s < n_slots; s += (u64)gridDim.x * blockDim.x) {
        const u64 a = link[s];
        if (!(a & KFF_VALID) || ((a >> KFF_NSUCC_SHIFT) & 15) != 1) continue;
        const u64 b = a & KFF_SUCC_MASK;
        if (((link[b] >> KFF_NPRED_SHIFT) & 15) != 1) continue;
        atomicOr((unsigned long long *)link + s, KFF_EDGE);
        atomicOr((unsigned long long *)link + b, KFF_HAS_PRED);
    }
}

// hashed minimizer with the bucket id moved to the top: buckets in increasing id (the reference's outer loop), equal
// minimizers adjacent
__device__ __forceinline__ u64 kff_sortkey(const BriskParams &P, u64 W) {
    const u64 H = bfc_mix(W, P.m_mask);
    const int s2 = 2 * P.s, b2 = 2 * P.b;
    const u64 low = H & ((1ull << s2) - 1), bucket = (H >> s2) & P.bucket_mask, high = H >> (s2 + b2);
    return (bucket << (2 * P.m - b2)) | (high << s2) | low;
}

template <int KW>
__global__ void __launch_bounds__(256) kff_chains_kernel(const BriskParams P, StoreView T, const u64 *__restrict__ link, int n_special,
                                                         KffChain *chains, u64 cap_chains, u64 *counter) {
    const u64 total = T.mask + 1 + (u64)n_special;
    for (u64 s = (u64)blockIdx.x * blockDim.x + threadIdx.x; s < total; s += (u64)gridDim.x * blockDim.x) {
        u64 lo, hi; u32 idx, n = 1;
        if (s <= T.mask) {
            const u64 a = link[s];
            if (!(a & KFF_VALID) || (a & KFF_HAS_PRED)) continue;
            kff_load<KW>(T, s, lo, hi, idx);
            u64 cur = a;
            while (cur & KFF_EDGE) { cur = link[cur & KFF_SUCC_MASK]; n++; }
        } else {  // the all-ones k-mer under minimizer_idx s - cap: a block of its own
            const u32 mt = T.special[s - T.mask - 1];
            if ((mt & META_TAG_MASK) == 0) continue;
            lo = STORE_EMPTY; hi = KW == 2 ? STORE_EMPTY : 0;
            lo &= P.kmask_lo; hi &= P.kmask_hi;
            idx = (mt & META_TAG_MASK) - 1;
        }
        const u64 o = atomicAdd((unsigned long long *)counter, 1ull);
        if (o < cap_chains) {
            KffChain c;
            c.sortkey = kff_sortkey(P, minimizer_word(P, lo, hi, (int)idx));
            c.slot = s; c.n = n; c.pad = 0;
            chains[o] = c;
        }
    }
}

struct KffChainKey {
    const KffChain *chains;
    __device__ __forceinline__ u64 operator()(u32 i) const { return chains[i].sortkey; }
};

// sizes[i] = bytes chain perm[i] contributes (its block + a section header when it opens a section); first[i] = 1 when
// it opens a section
__global__ void __launch_bounds__(256) kff_sizes_kernel(KffGeom G, const KffChain *__restrict__ chains, const u32 *__restrict__ perm, u64 n,
                                                        u64 *sizes, u64 *first) {
    u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const KffChain c = chains[perm[i]];
    const bool open = i == 0 || chains[perm[i - 1]].sortkey != c.sortkey;
    sizes[i] = G.block_bytes(c.n) + (open ? G.section_header_bytes() : 0);
    first[i] = open ? 1 : 0;
}

__device__ __forceinline__ void kff_put_be(unsigned char *dst, int bytes, u64 v) {
    for (int i = 0; i < bytes; i++) dst[i] = (unsigned char)(v >> (8 * (bytes - 1 - i)));
}

// nucleotide `pos` (0 = leftmost) of a sequence of `len` nucleotides held right-aligned in w[] (little-endian words)
__device__ __forceinline__ u32 kff_nt(const u64 *w, int len, int pos) {
    const int bit = 2 * (len - 1 - pos);
    return (u32)(w[bit >> 6] >> (bit & 63)) & 3u;
}

template <int KW>
__global__ void __launch_bounds__(128) kff_emit_kernel(const BriskParams P, KffGeom G, StoreView T, const u64 *__restrict__ link,
                                                       const KffChain *__restrict__ chains, const u32 *__restrict__ perm, u64 n,
                                                       const u64 *__restrict__ offs, const u64 *__restrict__ sec_ord,
                                                       unsigned char *__restrict__ out, u64 *__restrict__ sec_pos) {
    u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const KffChain c = chains[perm[i]];
    const bool open = i == 0 || chains[perm[i - 1]].sortkey != c.sortkey;
    unsigned char *dst = out + offs[i];
    // the superkmer: first k-mer, then the last base of every following k-mer
    u64 S[4] = {0, 0, 0, 0};
    u64 lo, hi; u32 idx0;
    if (c.slot <= T.mask) kff_load<KW>(T, c.slot, lo, hi, idx0);
    else {
        lo = STORE_EMPTY & P.kmask_lo; hi = KW == 2 ? (STORE_EMPTY & P.kmask_hi) : 0;
        idx0 = (T.special[c.slot - T.mask - 1] & META_TAG_MASK) - 1;
    }
    if (open) {
        const u64 W = minimizer_word(P, lo, hi, (int)idx0);
        u64 cnt = 1;
        while (i + cnt < n && chains[perm[i + cnt]].sortkey == c.sortkey) cnt++;
        sec_pos[sec_ord[i]] = offs[i];
        dst[0] = 'm';
        kff_put_be(dst + 1, G.mini_bytes, W);
        kff_put_be(dst + 1 + G.mini_bytes, 8, cnt);
        dst += G.section_header_bytes();
    }
    const int n_k = (int)c.n, len = n_k + G.k - 1, seq_len = len - G.m;
    const int mini_pos = G.k - G.m - (int)idx0;  // nucleotides of the superkmer before its minimizer
    kff_put_be(dst, G.nb_kmers_bytes, (u64)n_k);
    kff_put_be(dst + G.nb_kmers_bytes, G.mini_pos_bytes, (u64)mini_pos);
    unsigned char *seq = dst + G.nb_kmers_bytes + G.mini_pos_bytes;
    const int seq_bytes = (seq_len + 3) / 4;
    unsigned char *dat = seq + seq_bytes;
    S[0] = lo; S[1] = hi;
    u64 slot = c.slot;
    for (int j = 0; j < n_k; j++) {
        if (j > 0) {
            slot = link[slot] & KFF_SUCC_MASK;
            u64 l2, h2; u32 i2;
            kff_load<KW>(T, slot, l2, h2, i2);
            // S = S << 2 | last base
            S[3] = (S[3] << 2) | (S[2] >> 62); S[2] = (S[2] << 2) | (S[1] >> 62); S[1] = (S[1] << 2) | (S[0] >> 62);
            S[0] = (S[0] << 2) | (l2 & 3);
        }
        if (T.data_size == 1) dat[j] = (unsigned char)T.cnt[slot];
        else for (u32 q = 0; q < T.data_size; q++) dat[(u64)j * T.data_size + q] = T.data[slot * T.data_size + q];
    }
    // sequence without the minimizer, right-aligned, first nucleotide in the most significant bits
    for (int b = 0; b < seq_bytes; b++) seq[b] = 0;
    const int pad = 4 * seq_bytes - seq_len;
    for (int q = 0; q < seq_len; q++) {
        const int pos = q < mini_pos ? q : q + G.m;
        const int at = pad + q;
        seq[at >> 2] |= (unsigned char)(kff_nt(S, len, pos) << (2 * (3 - (at & 3))));
    }
}
