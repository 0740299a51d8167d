// enumerate.cuh — 2-bit packing and superkmer enumeration (north-star items 1-4).
//
// pack_reads_kernel     : ASCII -> 2-bit, one warp per read                         (Kmers.cpp:442-444)
// enumerate_kernel<RW>  : one THREAD per read streams the read once, rolling the forward / reverse
//                         m-mer, its order hash (hashing.cpp:8-19) and decycling class
//                         (Decycling.cpp:38-52), and replays SuperKmerEnumerator::next
//                         (Kmers.cpp:522-603) literally: seed get_minimizer on the (k-1)-mer, strict
//                         "smaller hash" takeover, full rescan when the minimizer leaves the window,
//                         the 64-bit truncation of get_minimizer for k > 32 (Q3) and its tie rules (Q5).
//                         Superkmer descriptors are staged in shared memory and expanded into packed
//                         records by the whole CTA with coalesced 128-bit stores.
#pragma once
#include "common.cuh"

#define ENUM_THREADS 128
#define ENUM_POOL 384
#define ENUM_SYNC_STEPS 32
#define DEC_RESYNC 128

// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) pack_reads_kernel(const unsigned char *__restrict__ bases,
                                                         const u64 *__restrict__ offsets, u64 read0, u64 read_end,
                                                         u64 *__restrict__ packed) {
    const int lane = threadIdx.x & 31;
    const u64 r = read0 + (u64)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);  // reads [read0, read_end) of the batch
    if (r >= read_end) return;
    const u64 beg = offsets[r], end = offsets[r + 1];
    const u64 len = end - beg;
    u64 *out = packed + packed_start(beg, r);
    const unsigned char *src = bases + beg;
    // a warp packs 128 bases (4 words) per iteration: lane -> 4 consecutive bases
    for (u64 base0 = 0; base0 < len; base0 += 128) {
        u64 p = base0 + 4 * (u64)lane;
        u32 v = 0;
#pragma unroll
        for (int j = 0; j < 4; j++) {
            u32 c = (p + j < len) ? (u32)src[p + j] : 0u;
            v = (v << 2) | ((c >> 1) & 3u);
        }
        // lane g*8+i holds bases [32g+4i, 32g+4i+4): place at bits (7-i)*8 of word g
        u64 x = (u64)v << (8 * (7 - (lane & 7)));
        x |= __shfl_xor_sync(0xffffffffu, x, 1);
        x |= __shfl_xor_sync(0xffffffffu, x, 2);
        x |= __shfl_xor_sync(0xffffffffu, x, 4);
        if ((lane & 7) == 0 && base0 + 32 * (u64)(lane >> 3) < len) out[(base0 >> 5) + (lane >> 3)] = x;
    }
}

// ---------------------------------------------------------------------------------------------
// superkmer descriptor staged in shared memory
__device__ __forceinline__ u64 make_desc(u32 i0, int n, int pos0, bool rev, int tid) {
    return (u64)i0 | ((u64)n << 32) | ((u64)pos0 << 40) | ((u64)rev << 48) | ((u64)tid << 49);
}

// Builds the packed record of one superkmer: k-mers i0 .. i0+n-1 of the read (read order), whose first
// k-mer had mini_pos = pos0, minimizer on the reverse strand when rev.  Emission order and minimizer_idx
// follow Kmers.cpp:578-584 and the vector reversal at :554-556/:567-569/:597-599.
template <int RW>
__device__ __forceinline__ void build_record(const BriskParams &P, const u64 *pk, u32 i0, int n, int pos0, bool rev,
                                             u64 (&rec)[RW]) {
    const int len = n + P.k - 1;
    // bases [i0, i0+len) right-aligned in rec[]
#pragma unroll
    for (int wi = 0; wi < RW; wi++) {
        int endb = len - 32 * wi;  // bases [endb-32, endb) of the segment go to word wi
        u64 x = 0;
        if (endb > 0) {
            if (endb >= 32) x = get_bits64(pk, (u64)i0 + endb - 32);
            else x = get_bits64(pk, (u64)i0) >> (64 - 2 * endb);
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
        idx0 = P.k - P.m - pos0 - n + 1;
    }
    // minimizer slice of the first emitted k-mer: nucleotides [idx0, idx0+m) from its right end
    u64 W = mw_extract64<RW, 3>(rec, 2 * (n - 1 + idx0)) & P.m_mask;
    u64 bucket = (bfc_mix(W, P.m_mask) >> (2 * P.s)) & P.bucket_mask;
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

// 64-bit minimum over the warp
__device__ __forceinline__ u64 warp_min64(u64 v) {
    u32 hi = (u32)(v >> 32), lo = (u32)v;
    u32 mhi = __reduce_min_sync(0xffffffffu, hi);
    u32 mlo = __reduce_min_sync(0xffffffffu, hi == mhi ? lo : 0xffffffffu);
    return ((u64)mhi << 32) | mlo;
}

// ring of order hashes: slot-major, the column is XOR-swizzled with the slot so that a thread writing its own
// column and a warp reading one thread's column across slots are both free of bank conflicts
__device__ __forceinline__ int ring_at(int slot, int tid) { return slot * ENUM_THREADS + (tid ^ (slot & 31)); }

// RING = true : every thread keeps the order hashes of its last w m-mers in a shared-memory ring; a rescan reads them.
// RING = false: no ring (shared memory: staging pool + tables only, twice the resident warps at k = 63); a rescan
//               recomputes the hash of every window position from the requester's last 32 bases, which k > 32 has to
//               do anyway for most positions (the 64-bit truncation of get_minimizer, Q3).
template <int RW, bool RING>
__global__ void __launch_bounds__(ENUM_THREADS) enumerate_kernel(const BriskParams P, const u64 *__restrict__ packed,
                                                                 const u64 *__restrict__ offsets, u64 read0, u64 n_reads,
                                                                 EnumOut out) {
    extern __shared__ u64 smem[];
    u64 *ring = smem;                                  // [w][ENUM_THREADS] order hashes of the last w m-mers (RING only)
    u64 *pool = ring + (RING ? (size_t)P.w * ENUM_THREADS : 0);  // [ENUM_POOL] superkmer descriptors
    double *coef = (double *)(pool + ENUM_POOL);       // [4m]
    double2 *dtab = (double2 *)(coef + 4 * P.m);  // [dec_groups][64] grouped decycling weights (16-byte aligned: 4m doubles)
    __shared__ u32 pool_n;
    __shared__ u32 s_maxL;
    __shared__ u64 s_base;
    __shared__ u64 rd_start[ENUM_THREADS];
    __shared__ u64 rd_id[ENUM_THREADS];

    const int tid = threadIdx.x;
    const int lane = tid & 31;
    const int k = P.k, m = P.m;
    const u64 rid = read0 + (u64)blockIdx.x * ENUM_THREADS + tid;  // this launch covers reads [read0, n_reads) of the batch
    u32 L = 0;
    u64 wstart = 0;
    if (rid < n_reads) {
        u64 beg = offsets[rid], end = offsets[rid + 1];
        u64 len = end - beg;
        wstart = packed_start(beg, rid);
        L = len >= (u64)k ? (u32)len : 0u;  // apps/counter.cpp:233
    }
    const u64 *pk = packed + wstart;
    rd_start[tid] = wstart;
    rd_id[tid] = rid;
    if (tid == 0) { pool_n = 0; s_maxL = 0; }
    for (int i = tid; i < 4 * m; i += ENUM_THREADS) coef[i] = P.coef[i];
    for (int i = tid; i < 64 * P.dec_groups; i += ENUM_THREADS) dtab[i] = P.dec_tab[i];
    __syncthreads();
    atomicMax(&s_maxL, L);
    __syncthreads();
    const u32 maxL = s_maxL;

    // rolling state
    u64 f = 0, r = 0, word = 0, revbits = 0, hist = 0;  // hist: the last 32 bases (RING = false)
    double Ar = 0, Ai = 0, Br = 0, Bi = 0;
    int rslot = -1;           // ring slot of the newest m-mer (the same for every thread: all reads start at t = 0)
    u64 H = 0;                // mini_hash
    int pos = 0;              // mini_pos
    bool rev = false;         // reversed
    u32 i0 = 0;               // first k-mer of the open superkmer
    int pos0 = 0;             // mini_pos of that k-mer
    const int rshift = 2 * m - 2;

    auto append = [&](u32 a, int n, int p0, bool rv) {
        u32 slot = atomicAdd(&pool_n, 1u);
        if (slot < ENUM_POOL) {
            pool[slot] = make_desc(a, n, p0, rv, tid);
        } else {  // staging full: emit directly
            u64 g = atomicAdd(out.rec_count, 1ull);
            if (out.recs && g < out.rec_cap) {
                u64 rec[RW];
                build_record<RW>(P, pk, a, n, p0, rv, rec);
                store_record<RW>(out.recs, g, rec);
                if (out.rec_occ) out.rec_occ[g] = ((out.kmer_base[rid] + a) << 1) | (u64)rv;
            }
        }
    };

    static_assert(ENUM_SYNC_STEPS == 32, "one packed word per block of steps");
    for (u32 t0 = 0; t0 < maxL; t0 += ENUM_SYNC_STEPS) {
        if (t0 < L) word = __ldg(pk + (t0 >> 5));  // the 32 bases of this block of steps
        for (u32 t = t0; t < t0 + ENUM_SYNC_STEPS; t++) {
            const bool active = t < L;
            u64 h = 0;
            bool rv = false;
            int need_kk = 0;  // k-1: seed get_minimizer (Kmers.cpp:533), k: the minimizer left the window (:551-562)
            const bool old_rev = rev;
            if (RING && t + 1 >= (u32)m) rslot = rslot + 1 == P.w ? 0 : rslot + 1;  // same t in every thread: uniform
            if (active) {
                const u32 code = (u32)(word >> 62);
                word <<= 2;
                const u32 leave = (u32)(f >> rshift) & 3u;
                f = ((f << 2) | code) & P.m_mask;
                r = (r >> 2) | ((u64)(code ^ 2u) << rshift);
                if (!RING) hist = (hist << 2) | code;
                if (!P.literal_decycle) {
                    // A = sum of low code bits * w^q, B = sum of high code bits * w^q over the window
                    double la = (double)(leave & 1u), lb = (double)(leave >> 1);
                    double ea = (double)(code & 1u), eb = (double)(code >> 1);
                    double ar = Ar - la, br = Br - lb;
                    double nAr = fma(Ai, P.rot_s, ar * P.rot_c), nAi = fma(-ar, P.rot_s, Ai * P.rot_c);
                    double nBr = fma(Bi, P.rot_s, br * P.rot_c), nBi = fma(-br, P.rot_s, Bi * P.rot_c);
                    Ar = fma(ea, P.in_c, nAr); Ai = fma(ea, P.in_s, nAi);
                    Br = fma(eb, P.in_c, nBr); Bi = fma(eb, P.in_s, nBi);
                    if ((t & (DEC_RESYNC - 1)) == DEC_RESYNC - 1) {
                        // bound the rounding drift on long reads: recompute the sums from the window
                        double sAr = 0, sAi = 0, sBr = 0, sBi = 0;
                        for (int q = 0; q < m; q++) {
                            u32 c = (u32)(f >> (2 * (m - 1 - q))) & 3u;
                            double tc = P.tw_c[q], ts = P.tw_s[q];
                            if (c & 1u) { sAr += tc; sAi += ts; }
                            if (c & 2u) { sBr += tc; sBi += ts; }
                        }
                        Ar = sAr; Ai = sAi; Br = sBr; Bi = sBi;
                    }
                }
                if (t + 1 >= (u32)m) {
                    rv = r < f;
                    const u64 canon = rv ? r : f;
                    u32 cls;
                    if (P.literal_decycle) {
                        cls = dec_mem_double(coef, P.eps, canon, m);
                    } else {
                        // R = Im(S), R(rot) = Im(w S) for the canonical strand (decycling)
                        double z = rv ? -2.0 : 2.0;
                        double Xr = fma(z, Br, Ar), Xi = fma(z, Bi, Ai);
                        double T = fma(Xr, P.rot_s, Xi * P.rot_c);
                        double R = rv ? -T : Xi, R2 = rv ? -Xi : T;
                        cls = 2;
                        if (R > P.eps) { if (R2 < P.eps) cls = 0; }
                        else if (R < -P.eps) { if (R2 > -P.eps) cls = 1; }
                        if (fabs(fabs(R) - P.eps) < P.guard || fabs(fabs(R2) - P.eps) < P.guard)
                            cls = dec_mem_double(coef, P.eps, canon, m);
                    }
                    h = ((u64)cls << 62) + bfc_mix(canon, P.m_mask);
                    if (RING) {
                        ring[ring_at(rslot, tid)] = h;
                        revbits = (revbits << 1) | (u64)rv;
                    }
                    if (t + 2 == (u32)k) {
                        need_kk = k - 1;
                    } else if (t + 1 >= (u32)k) {
                        pos += 1;
                        if (pos > k - m) need_kk = k;
                    }
                }
            }
            __syncwarp();
            // get_minimizer (Kmers.cpp:367-408) for the threads that asked, one requester at a time, the warp's
            // lanes taking one window position each
            unsigned req = __ballot_sync(0xffffffffu, need_kk != 0);
            if (req) {
                const int kk = (t + 2 == (u32)k) ? k - 1 : k;
                const int nmax = kk - m;
                const int real_max = kk > 32 ? 32 - m : nmax;
                const bool has_hi = nmax >= 32;  // positions >= 32 hold the all-A pseudo m-mer (Q3)
                u64 cur64 = 0;
                if (!RING) cur64 = hist;
                else if (need_kk && kk > 32) cur64 = get_bits64(pk, (u64)t - 31);
                while (req) {
                    const int src = __ffs(req) - 1;
                    req &= req - 1;
                    const u64 s_cur = __shfl_sync(0xffffffffu, cur64, src);
                    const u64 s_revbits = __shfl_sync(0xffffffffu, revbits, src);
                    const int s_tid = (tid & ~31) | src;
                    u64 nh = ~0ull;
                    bool prv = false;
                    if (lane <= nmax) {
                        if (RING && lane <= real_max) {
                            int sl = rslot - lane;
                            if (sl < 0) sl += P.w;
                            nh = ring[ring_at(sl, s_tid)];
                            prv = (s_revbits >> lane) & 1;
                        } else {
                            // real m-mer (masked) or the zero-padded top of the truncated k-mer (the mask is a no-op)
                            u64 v = (s_cur >> (2 * lane)) & P.m_mask;
                            u64 rc = rcbc(v, m);
                            u64 c = v < rc ? v : rc;
                            const u32 pc = P.literal_decycle ? dec_mem_double(coef, P.eps, c, m) : dec_class_tab(P, coef, dtab, c);
                            nh = ((u64)pc << 62) + bfc_mix(c, P.m_mask);
                            prv = c != v;
                        }
                    }
                    u64 hm = warp_min64(nh);
                    if (has_hi && P.polyA_hash < hm) hm = P.polyA_hash;
                    u64 tmask = __ballot_sync(0xffffffffu, nh == hm);
                    if (has_hi && P.polyA_hash == hm) tmask |= ((((u64)1) << (nmax - 31)) - 1) << 32;
                    const unsigned rvlo = __ballot_sync(0xffffffffu, prv);
                    int mp = __ffsll((long long)tmask) - 1;
                    bool rv0 = mp < 32 ? (rvlo >> mp) & 1u : false;
                    u64 rest = tmask & (tmask - 1);
                    if (rest) {  // equal hashes in the window: the tie rules of Kmers.cpp:389-404, in scan order
                        int cs = 0;
                        if (lane == src) {
                            int nlo = kk < 32 ? kk : 32;
                            u64 lo = get_bases(pk, (u64)t + 1 - nlo, nlo);
                            u64 hi = kk > 32 ? get_bases(pk, (u64)t + 1 - kk, kk - 32) : 0;
                            cs = canonized128(lo, hi, kk) ? 1 : 0;
                        }
                        cs = __shfl_sync(0xffffffffu, cs, src);
                        while (rest) {
                            const int i2 = __ffsll((long long)rest) - 1;
                            rest &= rest - 1;
                            const bool rvi = i2 < 32 ? (rvlo >> i2) & 1u : false;
                            if (nmax - i2 < mp) { mp = nmax - i2; rv0 = rvi; }
                            else if (nmax - i2 == mp) { if (!cs) rv0 = false; }
                        }
                    }
                    if (lane == src) { H = hm; pos = mp; rev = rv0; }
                }
            }
            if (active && t + 1 >= (u32)k) {
                const u32 i = t + 1 - k;
                bool boundary = need_kk != 0;
                if (!boundary && h < H) {
                    boundary = true;
                    H = h; pos = 0; rev = rv;
                }
                if (i == 0) {
                    pos0 = pos;
                } else if (boundary) {
                    append(i0, (int)(i - i0), pos0, old_rev);
                    i0 = i; pos0 = pos;
                }
                if (t + 1 == L) append(i0, (int)(i + 1 - i0), pos0, rev);
            }
        }
        __syncthreads();
        const u32 cnt_raw = pool_n;
        __syncthreads();  // nobody appends again before everyone has read the fill level
        const bool last = t0 + ENUM_SYNC_STEPS >= maxL;
        if (cnt_raw >= ENUM_POOL / 2 || (last && cnt_raw > 0)) {
            const u32 cnt = cnt_raw < ENUM_POOL ? cnt_raw : ENUM_POOL;
            if (tid == 0) s_base = atomicAdd(out.rec_count, (u64)cnt);
            __syncthreads();
            const u64 base = s_base;
            if (out.recs) {
                for (u32 e = tid; e < cnt; e += ENUM_THREADS) {
                    u64 g = base + e;
                    if (g >= out.rec_cap) break;
                    u64 d = pool[e];
                    int owner = (int)(d >> 49) & 127;
                    bool rvd = (d >> 48) & 1;
                    u64 rec[RW];
                    build_record<RW>(P, packed + rd_start[owner], (u32)d, (int)(d >> 32) & 255, (int)(d >> 40) & 255, rvd, rec);
                    store_record<RW>(out.recs, g, rec);
                    if (out.rec_occ) out.rec_occ[g] = ((out.kmer_base[rd_id[owner]] + (u32)d) << 1) | (u64)rvd;
                }
            }
            __syncthreads();
            if (tid == 0) pool_n = 0;
            __syncthreads();
        }
    }
}
