// enumerate.cuh — 2-bit packing and superkmer enumeration (north-star items 1-4).
//
// pack_reads_kernel        : ASCII -> 2-bit, one warp per read                         (Kmers.cpp:442-444)
// enumerate_kernel<KT,MT,…>: one THREAD per read (or per segment of a long read, see EnumSeg) streams its bases once,
//                            rolling the forward / reverse m-mer, its order hash (hashing.cpp:8-19) and decycling
//                            class (Decycling.cpp:38-52), and replays SuperKmerEnumerator::next (Kmers.cpp:522-603)
//                            literally: seed get_minimizer on the (k-1)-mer, strict "smaller hash" takeover, full
//                            rescan when the minimizer leaves the window, the 64-bit truncation of get_minimizer for
//                            k > 32 (Q3) and its tie rules (Q5).  Rescans are done by the whole warp for one requesting
//                            lane at a time.  Superkmer descriptors are staged in shared memory and expanded into
//                            packed records by the whole CTA with coalesced 128-bit stores.
//
// KT, MT > 0: k and m are compile-time constants (the five configurations of BASELINE.json); every shift, mask and
// window length is an immediate, m-mers of m <= 16 live in 32-bit registers, order hashes of m <= 15 too.
// KT = MT = 0: the same code with k, m read from BriskParams (any other parameter set).
#pragma once
#include <type_traits>

#include "common.cuh"

#define ENUM_THREADS 128
#define ENUM_WPOOL 128   // superkmer descriptors staged per warp
#define ENUM_WFLUSH 64   // a warp expands its staged descriptors into records once it holds this many
#define ENUM_SYNC_STEPS 32

// ---------------------------------------------------------------------------------------------
// 2 bits per base, first base on top: the 4 bases in the bytes of h (byte 0 first) -> 8 bits
__device__ __forceinline__ u32 pack4(u32 h) {
    u32 t = (__byte_perm(h, 0, 0x0123) >> 1) & 0x03030303u;  // byte 0 on top, (c >> 1) & 3 (Kmers.cpp:442-444)
    t = (t | (t >> 6)) & 0x000F000Fu;
    return (t | (t >> 12)) & 0xFFu;
}
__device__ __forceinline__ u32 pack8(u64 x) { return (pack4((u32)x) << 8) | pack4((u32)(x >> 32)); }

#define PACK_READS_PER_WARP 4
#define PACK_READS_PER_CTA (8 * PACK_READS_PER_WARP)  // 256 threads
// A warp packs PACK_READS_PER_WARP consecutive reads: 8 lanes per read, one lane per output word (32 bases), so a read
// of up to 256 bases takes one pass and no exchange between lanes.  The offsets of the warp's reads come with one
// request.  A lane fetches its 32 bases as the (at most five) ALIGNED 8-byte words that hold them (a read starts at
// any byte; only words with at least one byte of the read are touched) and funnel-shifts them into place.
__global__ void __launch_bounds__(256) pack_reads_kernel(const unsigned char *__restrict__ bases,
                                                         const u64 *__restrict__ offsets, u64 read0, u64 read_end,
                                                         u64 *__restrict__ packed) {
    constexpr int R = PACK_READS_PER_WARP;
    const int lane = threadIdx.x & 31, sub = lane >> 3, wl = lane & 7;
    const u64 r0 = read0 + ((u64)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) * R;  // reads [read0, read_end) of the batch
    if (r0 >= read_end) return;
    const int nr = (int)(read_end - r0 < (u64)R ? read_end - r0 : (u64)R);
    const u64 my_off = lane <= nr ? offsets[r0 + lane] : 0;
    const u64 beg = __shfl_sync(0xffffffffu, my_off, sub), end = __shfl_sync(0xffffffffu, my_off, sub + 1);
    if (sub >= nr) return;
    const u64 len = end - beg, n_words = (len + 31) >> 5;
    const unsigned char *src = bases + beg;
    const unsigned mis = (unsigned)((uintptr_t)src & 7u), sh = mis * 8u;
    const u64 *asrc = reinterpret_cast<const u64 *>((uintptr_t)src - mis);
    const u64 last_word = len ? (mis + len - 1) >> 3 : 0;  // the aligned word (from asrc) with the read's last base
    u64 *out = packed + packed_start(beg, r0 + sub);
    for (u64 w = wl; w < n_words; w += 8) {
        const u64 wi = 4 * w;  // bases [32w, 32w + 32) start in aligned word wi at bit sh
        u64 q[5];
#pragma unroll
        for (int j = 0; j < 5; j++) q[j] = (wi + j <= last_word && (j < 4 || sh)) ? __ldg(asrc + wi + j) : 0ull;
        const u64 left = len - 32 * w;  // >= 1 bases from here to the read's end
        u64 word = 0;
#pragma unroll
        for (int j = 0; j < 4; j++) {
            u64 x = sh ? (q[j] >> sh) | (q[j + 1] << (64 - sh)) : q[j];  // byte i = base 32w + 8j + i
            if (left < 8u * (j + 1)) x = left > 8u * j ? x & (~0ull >> (64 - 8 * (left - 8u * j))) : 0ull;
            word = (word << 16) | pack8(x);
        }
        out[w] = word;
    }
}

// ---------------------------------------------------------------------------------------------
// compile-time / run-time view of (k, m)
template <int KT, int MT>
struct KM {
    typedef typename std::conditional<(MT > 0 && MT <= 16), u32, u64>::type mer_t;    // an m-mer
    typedef typename std::conditional<(MT > 0 && MT <= 15), u32, u64>::type hash_t;   // class bits + mix of an m-mer
    static constexpr bool HASH32 = (MT > 0 && MT <= 15);
    __device__ __forceinline__ static int k(const BriskParams &P) { return KT ? KT : P.k; }
    __device__ __forceinline__ static int m(const BriskParams &P) { return MT ? MT : P.m; }
    __device__ __forceinline__ static mer_t m_mask(const BriskParams &P) {
        return MT ? (mer_t)((MT >= 32) ? ~0ull : ((1ull << (2 * (MT & 31))) - 1)) : (mer_t)P.m_mask;
    }
    __device__ __forceinline__ static int s(const BriskParams &P) { return P.s; }
    // order hash = (class << 62) + mix (hashing.cpp:18); in 32 bits the class sits right above the 2m <= 30 mix bits
    __device__ __forceinline__ static hash_t make_hash(u32 cls, mer_t mix) {
        if (HASH32) return (hash_t)((cls << 30) | (u32)mix);
        return (hash_t)(((u64)cls << 62) + (u64)mix);
    }
    __device__ __forceinline__ static mer_t mix(mer_t key, mer_t mask) {
        if (sizeof(mer_t) == 4) return (mer_t)bfc_mix32((u32)key, (u32)mask);
        return (mer_t)bfc_mix_mul((u64)key, (u64)mask);
    }
    // rcbc (brisk/Kmers.cpp:320-332) on an m-mer
    __device__ __forceinline__ static mer_t rc(mer_t v, int m) {
        if (sizeof(mer_t) == 4) {
            u32 x = __brev((u32)v ^ 0xaaaaaaaau);
            x = ((x & 0x55555555u) << 1) | ((x >> 1) & 0x55555555u);
            return (mer_t)(x >> (32 - 2 * m));
        }
        return (mer_t)rcbc((u64)v, m);
    }
};

// minimum over the warp
__device__ __forceinline__ u64 warp_min(u64 v) {
    u32 hi = (u32)(v >> 32), lo = (u32)v;
    u32 mhi = __reduce_min_sync(0xffffffffu, hi);
    u32 mlo = __reduce_min_sync(0xffffffffu, hi == mhi ? lo : 0xffffffffu);
    return ((u64)mhi << 32) | mlo;
}
__device__ __forceinline__ u32 warp_min(u32 v) { return __reduce_min_sync(0xffffffffu, v); }

// ---- decycling class in fixed point -------------------------------------------------------------
// DecyclingSet::memDouble (Decycling.cpp:38-52) classifies an m-mer by R = Im(S) and R(rot) = Im(w S),
// S = sum_j code_j w^j over the LEFT positions j of the m-mer, w = exp(2 pi i / m), against +-1e-6.
// Rolled: a base at read position u weighs w^(u mod m) (exact integer table, 2^24 fixed point), so the window sum
// T = sum_u code_u w^(u mod m) rolls with one table entry per step and never drifts; S = w^(-p) T for the m-mer
// starting at p.  The two code bit planes are kept apart (code = a + 2b) because the reverse strand's sum is
// w^(-1) conj(A - 2B).  Whenever |R| or |R(rot)| is inside the threshold band (eps + the worst-case fixed-point
// error, P.dec_th) the literal double summation decides: the class is always the reference's.
__device__ __forceinline__ int dec_im(int Xr, int Xi, int c30, int s30) { return __mulhi(Xi, c30) - __mulhi(Xr, s30); }

// Class from D0 = Im(w^-p X), D1 = Im(w^(1-p) X), X = A + 2B (forward strand canonical) or A - 2B (reverse strand).
// Forward: R = D0, R(rot) = D1.  Reverse: R = -D1, R(rot) = -D0.  Either way the class is 0 when D0 > 0 > D1, 1 when
// D0 < 0 < D1 and 2 otherwise, and the band test is symmetric too.
__device__ __forceinline__ u32 dec_class_signs(int D0, int D1, int th, bool &ambiguous) {
    ambiguous = (u32)(D0 + th) <= 2u * (u32)th || (u32)(D1 + th) <= 2u * (u32)th;
    return ((D0 ^ D1) < 0) ? (D1 > 0 ? 1u : 0u) : 2u;
}

// class of an arbitrary m-mer from the grouped tables: tab[g][v] = fixed-point sum of code * w^(left position) over the
// three nucleotides 3g .. 3g+2 (from the right end) of the m-mer
template <int MT>
__device__ __forceinline__ u32 dec_class_tab(const BriskParams &P, const double *coef, const int2 *tab, u64 seq, int m) {
    // Only the last nucleotide set (the zero-padded tops of a truncated k-mer are mostly of this kind, Q3): R(rot) is the
    // empty sum, exactly 0, which no fixed-point band can tell from +-1e-6.  memDouble gives R = -code sin(2 pi / m) <
    // -eps and R(rot) = 0 > -eps: class 1; the all-A m-mer has R = 0: class 2.
    if (seq < 4) return seq ? 1u : 2u;
    int Sr = 0, Si = 0;
    constexpr int G = MT ? (MT + 2) / 3 : 11;
#pragma unroll
    for (int g = 0; g < G; g++) {
        if (MT || g < P.dec_groups) {
            const int2 e = tab[g * 64 + (int)((seq >> (6 * g)) & 63)];
            Sr += e.x;
            Si += e.y;
        }
    }
    bool amb;
    const u32 cls = dec_class_signs(Si >> 2, __mulhi(Sr, P.rot1_s30) + __mulhi(Si, P.rot1_c30), P.dec_th, amb);
    return amb ? dec_mem_double(coef, P.eps, seq, m) : cls;
}

// ---------------------------------------------------------------------------------------------
// superkmer descriptor staged in shared memory
__device__ __forceinline__ u64 make_desc(u32 i0, int n, int pos0, int rev, int lane) {
    return (u64)i0 | ((u64)(u32)n << 32) | ((u64)(u32)pos0 << 40) | ((u64)(u32)rev << 48) | ((u64)(u32)lane << 49);
}

// Builds the packed record of one superkmer: k-mers i0 .. i0+n-1 of the read (read order), whose first
// k-mer had mini_pos = pos0, minimizer on the reverse strand when rev.  Emission order and minimizer_idx
// follow Kmers.cpp:578-584 and the vector reversal at :554-556/:567-569/:597-599.
template <int KT, int MT, int RW>
__device__ __forceinline__ void build_record(const BriskParams &P, const u64 *pk, u64 i0, int n, int pos0, bool rev,
                                             u64 (&rec)[RW]) {
    const int k = KM<KT, MT>::k(P), m = KM<KT, MT>::m(P);
    const u64 m_mask = (u64)KM<KT, MT>::m_mask(P);
    const int len = n + k - 1;
    // bases [i0, i0+len) right-aligned in rec[]
#pragma unroll
    for (int wi = 0; wi < RW; wi++) {
        int endb = len - 32 * wi;  // bases [endb-32, endb) of the segment go to word wi
        u64 x = 0;
        if (endb > 0) {
            if (endb >= 32) x = get_bits64(pk, i0 + endb - 32);
            else x = get_bits64(pk, i0) >> (64 - 2 * endb);
        }
        rec[wi] = x;
    }
    int idx0 = pos0;
    if (rev) {
        // reverse complement of the whole segment: reverse all 2-bit groups of the RW-word value,
        // complement, shift right by (64*RW - 2*len)
        u64 t[RW];
#pragma unroll
        for (int wi = 0; wi < RW; wi++) t[wi] = rev2(rec[RW - 1 - wi] ^ 0xaaaaaaaaaaaaaaaaull);
#pragma unroll
        for (int wi = 0; wi < RW; wi++) rec[wi] = t[wi];
        mw_shr<RW, 7>(rec, 64 * RW - 2 * len);
        idx0 = k - m - pos0 - n + 1;
    }
    // minimizer slice of the first emitted k-mer: nucleotides [idx0, idx0+m) from its right end
    u64 W = mw_extract64<RW, 3>(rec, 2 * (n - 1 + idx0)) & m_mask;
    u64 bucket = (bfc_mix(W, m_mask) >> (2 * P.s)) & P.bucket_mask;
    rec[RW - 1] |= rec_header(P, bucket, n, idx0);
}

template <int RW>
__device__ __forceinline__ void store_record(u64 *recs, u64 g, const u64 (&rec)[RW]) {
    ulonglong2 *dst = reinterpret_cast<ulonglong2 *>(recs + g * RW);
#pragma unroll
    for (int wi = 0; wi < RW; wi += 2) dst[wi >> 1] = make_ulonglong2(rec[wi], rec[wi + 1]);
}

struct EnumOut {
    u64 *recs;       // [rec_cap][RW] or null
    u64 rec_cap;
    u64 *rec_count;  // device counter (may exceed rec_cap: caller retries with a larger buffer)
    u64 *rec_occ;    // optional [rec_cap]: (index of the first read-order k-mer occurrence << 1) | rev
    const u64 *kmer_base;  // optional [n_reads]: occurrence index of k-mer 0 of each read
};

// Long reads.  One thread per read is the right grain for sequencing reads (hundreds of bases), not for a contig or a
// chromosome handed over by the FASTA entry points: a thread of the plain launch skips any read longer than
// EnumLong::limit bases and appends its index to a list instead.  Those reads are then cut into segments of
// ENUM_SEG_KMERS k-mers that are enumerated by different threads (SEG = true).  The state machine of
// SuperKmerEnumerator::next depends on history (Q4), so a segment that does not start at the read's first base starts
// `warm` k-mers early from a full rescan (get_minimizer on that k-mer, the state the reference is in right after an
// "outdated" event) and only emits the superkmers that START inside [kmer0, kmer1), running on until the superkmer that
// is open at kmer1 has closed.  Every segment writes down its state after the k-mer kmer0 - 1 (its own warm-up) and
// after kmer1 - 1; the output is exact if, for every pair of neighbours, the state segment s + 1 reached by warming up
// equals the state segment s arrived at from the read's start (equal states have equal futures; by induction from
// segment 0, which is exact).  enumerate_seg_verify_kernel counts the pairs that differ; the host then repeats the long
// reads with a longer warm-up and finally with one thread per read.
#define ENUM_SEG_KMERS 1024
struct EnumLong {
    u64 *list;    // [cap] indices of the reads the plain launch skipped (null: nothing is skipped)
    u64 cap;
    u64 *count;   // device counter
    u32 limit;    // reads longer than this many bases are skipped
};
struct EnumSeg {
    u64 read;       // index of the read in the batch
    u64 kmer0;      // first k-mer whose superkmer starts belong to this segment
    u64 kmer1;      // one past the last
    u64 start;      // k-mer at which the thread starts rolling (kmer0 - warm, or 0)
};
struct EnumSegState { u64 H; int pos; int rev; };
struct EnumSegIn {
    const EnumSeg *segs;   // null: one thread per read
    u64 n_segs;
    EnumSegState *at0, *at1;  // [n_segs] state after k-mer kmer0 - 1 / kmer1 - 1
};

// segments of the long reads: read j of the list gets n_seg[j] = ceil(n_kmers / ENUM_SEG_KMERS) segments
__global__ void __launch_bounds__(256) enumerate_seg_count_kernel(const u64 *__restrict__ offsets, const u64 *__restrict__ list, u64 n_long,
                                                                  int k, u64 *__restrict__ n_seg) {
    const u64 j = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (j > n_long) return;
    u64 v = 0;
    if (j < n_long) {
        const u64 r = list[j], len = offsets[r + 1] - offsets[r];
        v = len >= (u64)k ? (len - k + 1 + ENUM_SEG_KMERS - 1) / ENUM_SEG_KMERS : 0;
    }
    n_seg[j] = v;
}
__global__ void __launch_bounds__(256) enumerate_seg_fill_kernel(const u64 *__restrict__ offsets, const u64 *__restrict__ list, u64 n_long,
                                                                 int k, u64 warm, const u64 *__restrict__ seg_start, EnumSeg *__restrict__ segs) {
    const u64 j = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n_long) return;
    const u64 r = list[j], len = offsets[r + 1] - offsets[r];
    if (len < (u64)k) return;
    const u64 nk = len - k + 1;
    u64 o = seg_start[j];
    for (u64 a = 0; a < nk; a += ENUM_SEG_KMERS, o++) {
        EnumSeg g;
        g.read = r; g.kmer0 = a; g.kmer1 = a + ENUM_SEG_KMERS < nk ? a + ENUM_SEG_KMERS : nk;
        g.start = a > warm ? a - warm : 0;
        segs[o] = g;
    }
}
// neighbours of one read must agree on the state at their common boundary
__global__ void __launch_bounds__(256) enumerate_seg_verify_kernel(const EnumSeg *__restrict__ segs, const EnumSegState *__restrict__ at0,
                                                                   const EnumSegState *__restrict__ at1, u64 n_segs, u64 *mismatches) {
    const u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i + 1 >= n_segs) return;
    if (segs[i].read != segs[i + 1].read) return;
    const EnumSegState a = at1[i], b = at0[i + 1];
    if (a.H != b.H || a.pos != b.pos || a.rev != b.rev) atomicAdd((unsigned long long *)mismatches, 1ull);
}

// ring of order hashes: slot-major, the column is XOR-swizzled with the slot so that a thread writing its own
// column and a warp reading one thread's column across slots are both free of bank conflicts
__device__ __forceinline__ int ring_at(int slot, int tid) { return slot * ENUM_THREADS + (tid ^ (slot & 31)); }

// RING = true : every thread keeps the order hashes of its last w m-mers in a shared-memory ring; a rescan reads them.
// RING = false: no ring (shared memory: staging pool + tables only); a rescan recomputes the hash of every window
//               position from the requester's last 32 bases, which k > 32 has to do anyway for most positions (the
//               64-bit truncation of get_minimizer, Q3); no order hash is computed before the first k-mer is complete.
// LIT  = true : DecyclingSet::memDouble evaluated literally for every m-mer (BRISK_F_LITERAL_DECYCLE).
//
// The warps of a CTA only share the read-only tables: every warp stages the descriptors of its own superkmers (slots
// handed out with a ballot, no atomics) and expands them into records on its own, so there is no CTA barrier after
// the tables are loaded.
// SEG  = true : one thread per SEGMENT of a long read (EnumSegIn) instead of one per read.
template <int KT, int MT, int RW, bool RING, bool LIT, bool SEG>
__global__ void __launch_bounds__(ENUM_THREADS, 8) enumerate_kernel(const BriskParams P, const u64 *__restrict__ packed,
                                                                 const u64 *__restrict__ offsets, u64 read0, u64 n_reads,
                                                                 EnumOut out, EnumLong lg, EnumSegIn sg) {
    typedef KM<KT, MT> C;
    typedef typename C::mer_t mer_t;
    typedef typename C::hash_t hash_t;
    extern __shared__ u64 smem[];
    const int k = C::k(P), m = C::m(P);
    const int w = k - m + 1;
    const mer_t m_mask = C::m_mask(P);
    hash_t *ring = reinterpret_cast<hash_t *>(smem);   // [w][ENUM_THREADS] order hashes of the last w m-mers (RING only)
    u64 *pool = smem + (RING ? ((size_t)w * ENUM_THREADS * sizeof(hash_t) + 7) / 8 : 0);  // [warps][ENUM_WPOOL] superkmer descriptors
    double *coef = (double *)(pool + (ENUM_THREADS / 32) * ENUM_WPOOL);  // [4m]
    int2 *dtab = (int2 *)(coef + 4 * m);               // [dec_groups][64] grouped decycling weights

    const unsigned FULL = 0xffffffffu;
    const int tid = threadIdx.x;
    const int lane = tid & 31;
    const u32 lane_lt = (1u << lane) - 1u;
    u64 *wpool = pool + (tid >> 5) * ENUM_WPOOL;
    const u64 gid = read0 + (u64)blockIdx.x * ENUM_THREADS + tid;  // this launch covers reads [read0, n_reads) of the batch
    u64 rid = gid;          // index of the read in the batch
    u32 L = 0;              // bases this thread consumes
    u64 wstart = 0;
    u64 boff = 0;           // SEG: base of the read at which this thread starts rolling
    u32 seg_k0 = 0, seg_k1 = 0xffffffffu;  // SEG: k-mers (relative to boff) whose superkmer starts this thread emits
    bool mid = false;       // SEG: the thread does not start at the read's first base
    if (!SEG) {
        if (rid < n_reads) {
            u64 beg = offsets[rid], end = offsets[rid + 1];
            u64 len = end - beg;
            wstart = packed_start(beg, rid);
            if (len >= (u64)k) {  // apps/counter.cpp:233
                if (lg.list && len > (u64)lg.limit) {  // a long read: left to the segmented launch
                    const u64 o = atomicAdd((unsigned long long *)lg.count, 1ull);
                    if (o < lg.cap) lg.list[o] = rid;
                } else if (len <= 0xffffffffull) {
                    L = (u32)len;
                }
            }
        }
    } else if (gid < sg.n_segs) {
        const EnumSeg g = sg.segs[gid];
        rid = g.read;
        const u64 beg = offsets[rid], len = offsets[rid + 1] - beg;
        wstart = packed_start(beg, rid);
        boff = g.start;
        mid = g.start > 0;
        // run on until the superkmer that is open at kmer1 has closed: at most w - 1 k-mers further
        const u64 last = g.kmer1 + (u64)(k - m + 1) + (u64)k - 1;
        L = (u32)((last < len ? last : len) - boff);
        seg_k0 = (u32)(g.kmer0 - boff);
        seg_k1 = (u32)(g.kmer1 - boff);
    }
    const u64 *pk = packed + wstart;
    for (int i = tid; i < 4 * m; i += ENUM_THREADS) coef[i] = P.coef[i];
    for (int i = tid; i < 64 * P.dec_groups; i += ENUM_THREADS) dtab[i] = P.dec_tab[i];
    __syncthreads();
    const u32 maxL = __reduce_max_sync(FULL, L);
    const u32 minL = __reduce_min_sync(FULL, L ? L : 0xffffffffu);  // first step at which a read of the warp ends
    if (maxL == 0) return;

    // rolling state
    mer_t f = 0, r = 0;
    u64 word = 0, revbits = 0, hist = 0;  // hist: the last 32 bases (RING = false)
    int Pr = 0, Pi = 0, Mr = 0, Mi = 0;   // window sums of code * w^(u mod m) and of (a - 2b) * w^(u mod m), fixed point
    int rslot = -1;           // ring slot of the newest m-mer (the same for every thread: all reads start at t = 0)
    int q = 0;                // t mod m (the same for every thread)
    const int DEAD_POS = -(1 << 30);
    hash_t H = 0;             // mini_hash
    int pos = L ? (SEG && mid ? k - m : 0) : DEAD_POS;  // mini_pos (a segment that starts inside its read: rescan at its first k-mer)
    int rev = 0;              // reversed
    u32 i0 = 0;               // first k-mer of the open superkmer
    int pos0 = 0;             // mini_pos of that k-mer
    int pcount = 0;           // descriptors staged by the warp (the same in every lane)
    const int rshift = 2 * m - 2;
    const hash_t polyA = C::make_hash((u32)(P.polyA_hash >> 62), (mer_t)(P.polyA_hash & (u64)m_mask));

    auto emit_record = [&](const u64 *rpk, u64 roff, u64 g, u64 d, u64 read_id) {
        if (out.recs && g < out.rec_cap) {
            const int rvd = (int)(d >> 48) & 1;
            u64 rec[RW];
            build_record<KT, MT, RW>(P, rpk, roff + (u32)d, (int)(d >> 32) & 255, (int)(d >> 40) & 255, rvd != 0, rec);
            store_record<RW>(out.recs, g, rec);
            if (out.rec_occ) out.rec_occ[g] = ((out.kmer_base[read_id] + roff + (u32)d) << 1) | (u64)rvd;
        }
    };
    // stages the descriptor of every lane that wants one
    auto append = [&](bool want, u32 a, int n, int p0, int rv) {
        if (SEG) want = want && a >= seg_k0 && a < seg_k1;  // only the superkmers that start inside the segment
        const unsigned mk = __ballot_sync(FULL, want);
        if (mk) {
            if (want) {
                const u64 desc = make_desc(a, n, p0, rv, lane);
                const int slot = pcount + __popc(mk & lane_lt);
                if (slot < ENUM_WPOOL) wpool[slot] = desc;
                else emit_record(pk, boff, atomicAdd(out.rec_count, 1ull), desc, rid);  // staging full: emit directly
            }
            pcount += __popc(mk);
        }
    };
    // the warp expands its staged descriptors into packed records
    auto flush = [&]() {
        const u32 cnt = pcount < ENUM_WPOOL ? (u32)pcount : (u32)ENUM_WPOOL;
        u64 base = 0;
        if (lane == 0) base = atomicAdd(out.rec_count, (u64)cnt);
        base = __shfl_sync(FULL, base, 0);
        __syncwarp();
        for (u32 e0 = 0; e0 < cnt; e0 += 32) {
            const u32 e = e0 + lane;
            const u64 d = e < cnt ? wpool[e] : 0ull;
            const int owner = (int)(d >> 49) & 31;
            const u64 o_start = __shfl_sync(FULL, wstart, owner);
            const u64 o_rid = __shfl_sync(FULL, rid, owner);
            const u64 o_off = SEG ? __shfl_sync(FULL, boff, owner) : 0ull;
            if (e < cnt) emit_record(packed + o_start, o_off, base + e, d, o_rid);
        }
        __syncwarp();
        pcount = 0;
    };

    // one base: forward / reverse m-mer, last 32 bases, window sums
    auto roll = [&]() {
        const u32 code = (u32)(word >> 62);
        word <<= 2;
        const u32 leave = (u32)(f >> rshift) & 3u;
        const u32 ccode = code ^ 2u;
        f = (mer_t)((f * 4 + code) & m_mask);
        r = (mer_t)((r >> 2) | ((mer_t)ccode << rshift));
        if (!RING) hist = hist * 4 + code;
        if (!LIT) {
            // code = a + 2b: X+ = sum of code * tw, X- = sum of (a - 2b) * tw, and a - 2b = (code ^ 2) - 2
            const int dp = (int)code - (int)leave, dm = (int)ccode - (int)(leave ^ 2u);
            const int tc = P.dec_tw[q].x, ts = P.dec_tw[q].y;
            Pr += dp * tc; Pi += dp * ts;
            Mr += dm * tc; Mi += dm * ts;
        }
    };
    // order hash of the newest m-mer (hashing.cpp:8-19); rv = it is the reverse strand that is canonical
    auto order_hash = [&](int &rv) -> hash_t {
        const bool rvb = r < f;
        rv = rvb;
        const mer_t canon = rvb ? r : f;
        u32 cls;
        if (LIT) {
            cls = dec_mem_double(coef, P.eps, (u64)canon, m);
        } else {
            // the m-mer ending at base t starts at phase p = (t + 1) mod m = q + 1
            const int p = q + 1 == m ? 0 : q + 1;
            const int Xr = rvb ? Mr : Pr, Xi = rvb ? Mi : Pi;
            bool amb;
            cls = dec_class_signs(dec_im(Xr, Xi, P.dec_rot[p].x, P.dec_rot[p].y), dec_im(Xr, Xi, P.dec_rot[q].x, P.dec_rot[q].y),
                                  P.dec_th, amb);
            if (amb && pos > DEAD_POS / 2) cls = dec_mem_double(coef, P.eps, (u64)canon, m);  // (not for a lane whose read has ended)
        }
        return C::make_hash(cls, C::mix(canon, m_mask));
    };
    // get_minimizer (Kmers.cpp:367-408) on the kk-mer ending at base t for the lanes that asked, one requester at a
    // time, the warp's lanes taking one window position each
    auto rescans = [&](unsigned req, const int kk, const u32 t) {
        const int nmax = kk - m;
        const int real_max = kk > 32 ? 32 - m : nmax;
        const bool has_hi = nmax >= 32;  // positions >= 32 hold the all-A pseudo m-mer (Q3)
        u64 cur64 = 0;
        if (!RING) cur64 = hist;
        else if (kk > 32 && ((req >> lane) & 1u)) cur64 = get_bits64(pk, boff + t - 31);
        while (req) {
            const int src = __ffs(req) - 1;
            req &= req - 1;
            const u64 s_cur = (!RING || kk > 32) ? __shfl_sync(FULL, cur64, src) : 0ull;
            const u64 s_revbits = RING ? __shfl_sync(FULL, revbits, src) : 0ull;
            const int s_tid = (tid & ~31) | src;
            hash_t nh = (hash_t)~(hash_t)0;
            bool prv = false;
            if (lane <= nmax) {
                if (RING && lane <= real_max) {
                    int sl = rslot - lane;
                    if (sl < 0) sl += w;
                    nh = ring[ring_at(sl, s_tid)];
                    prv = (s_revbits >> lane) & 1;
                } else {
                    // real m-mer (masked) or the zero-padded top of the truncated k-mer (the mask is a no-op)
                    const mer_t v = (mer_t)((s_cur >> (2 * lane)) & (u64)m_mask);
                    const mer_t rc = C::rc(v, m);
                    const mer_t c = v < rc ? v : rc;
                    const u32 pc = LIT ? dec_mem_double(coef, P.eps, (u64)c, m) : dec_class_tab<MT>(P, coef, dtab, (u64)c, m);
                    nh = C::make_hash(pc, C::mix(c, m_mask));
                    prv = c != v;
                }
            }
            hash_t hm = warp_min(nh);
            if (has_hi && polyA < hm) hm = polyA;
            u64 tmask = __ballot_sync(FULL, nh == hm);
            if (has_hi && polyA == hm) tmask |= ((((u64)1) << (nmax - 31)) - 1) << 32;
            const unsigned rvlo = __ballot_sync(FULL, prv);
            int mp = __ffsll((long long)tmask) - 1;
            int rv0 = mp < 32 ? (int)((rvlo >> mp) & 1u) : 0;
            u64 rest = tmask & (tmask - 1);
            if (rest) {  // equal hashes in the window: the tie rules of Kmers.cpp:389-404, in scan order
                int cs = 0;
                if (lane == src) {
                    int nlo = kk < 32 ? kk : 32;
                    u64 lo = get_bases(pk, boff + t + 1 - nlo, nlo);
                    u64 hi = kk > 32 ? get_bases(pk, boff + t + 1 - kk, kk - 32) : 0;
                    cs = canonized128(lo, hi, kk) ? 1 : 0;
                }
                cs = __shfl_sync(FULL, cs, src);
                while (rest) {
                    const int i2 = __ffsll((long long)rest) - 1;
                    rest &= rest - 1;
                    const int rvi = i2 < 32 ? (int)((rvlo >> i2) & 1u) : 0;
                    if (nmax - i2 < mp) { mp = nmax - i2; rv0 = rvi; }
                    else if (nmax - i2 == mp) { if (!cs) rv0 = 0; }
                }
            }
            if (lane == src) { H = hm; pos = mp; rev = rv0; }
        }
    };
    auto next_q = [&]() { q = q + 1 == m ? 0 : q + 1; };

    static_assert(ENUM_SYNC_STEPS == 32, "one packed word per block of steps");
    // phases of a read (t = index of the base being consumed; the same t in every thread):
    //   t <  e_roll          : the window fills (RING: until the first m-mer is complete; else: until the seed step)
    //   e_roll <= t < k - 2  : RING only: order hashes go to the ring
    //   t == k - 2           : seed get_minimizer on the (k-1)-mer (Kmers.cpp:533)
    //   t >= k - 1           : one k-mer per step
    const u32 e_roll = RING ? (u32)(m - 1) : (u32)(k - 2);
    for (u32 t0 = 0; t0 < maxL; t0 += ENUM_SYNC_STEPS) {
        if (t0 < L) word = SEG ? get_bits64(pk, boff + t0) : __ldg(pk + (t0 >> 5));  // the 32 bases of this block of steps
        const u32 t1 = t0 + ENUM_SYNC_STEPS;
        u32 t = t0;
        for (; t < t1 && t < e_roll; t++) {
            if (L) roll();
            next_q();
        }
        if (RING) {
            for (; t < t1 && t + 2 < (u32)k; t++) {
                rslot = rslot + 1 == w ? 0 : rslot + 1;
                if (L) {
                    roll();
                    int rv;
                    ring[ring_at(rslot, tid)] = order_hash(rv);
                    revbits = (revbits << 1) | (u64)rv;
                }
                next_q();
            }
        }
        if (t < t1 && t + 2 == (u32)k) {
            if (RING && t + 1 >= (u32)m) rslot = rslot + 1 == w ? 0 : rslot + 1;
            if (L) {
                roll();
                if (RING && t + 1 >= (u32)m) {
                    int rv;
                    ring[ring_at(rslot, tid)] = order_hash(rv);
                    revbits = (revbits << 1) | (u64)rv;
                }
            }
            __syncwarp();
            const unsigned req = __ballot_sync(FULL, L != 0 && !(SEG && mid));
            if (req) rescans(req, k - 1, t);
            next_q();
            t++;
        }
        for (; t < t1; t++) {
            // a lane whose read has ended (or is too short) keeps rolling zeros with H = 0 and mini_pos far below zero: it
            // can neither take over (h < 0) nor ask for a rescan, so the steady state needs no "active" predicate
            const bool first = t + 1 == (u32)k;  // the read's first k-mer never closes a superkmer (Kmers.cpp:585-593)
            const int old_rev = rev;
            if (RING) rslot = rslot + 1 == w ? 0 : rslot + 1;
            roll();
            int rv;
            const hash_t h = order_hash(rv);
            if (RING) {
                ring[ring_at(rslot, tid)] = h;
                revbits = (revbits << 1) | (u64)rv;
            }
            pos += 1;
            const unsigned req = __ballot_sync(FULL, pos > k - m);  // the minimizer left the window (Kmers.cpp:551-562)
            const bool need = (req >> lane) & 1u;
            if (req) rescans(req, k, t);
            const u32 i = t + 1 - k;
            bool boundary = need;
            if (!need && h < H) {
                boundary = true;
                H = h; pos = 0; rev = rv;
            }
            append(boundary && !first, i0, (int)(i - i0), pos0, old_rev);
            if (boundary || first) { i0 = i; pos0 = pos; }
            if (SEG && L) {  // the state after this k-mer, where the neighbours of the segment compare theirs
                if (i + 1 == seg_k0) { EnumSegState z; z.H = (u64)H; z.pos = pos; z.rev = rev; sg.at0[gid] = z; }
                if (i + 1 == seg_k1) { EnumSegState z; z.H = (u64)H; z.pos = pos; z.rev = rev; sg.at1[gid] = z; }
            }
            if (t + 1 >= minL) {  // some read of the warp may end here: its open superkmer is the last one
                const bool fin = t + 1 == L;
                append(fin, i0, (int)(i + 1 - i0), pos0, rev);
                if (fin) { H = 0; pos = DEAD_POS; }
            }
            next_q();
        }
        if (pcount >= ENUM_WFLUSH || (t1 >= maxL && pcount > 0)) flush();
    }
}
