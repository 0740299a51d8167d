// store.cuh — the GPU-resident bucket store (north-star item 6).
//
// Replaces DenseMenuYo + Bucket + SKL (brisk/DenseMenuYo.hpp:248-396, brisk/buckets.hpp:146-225,
// brisk/SuperKmerLight.hpp:209-217).  In the reference a k-mer is found by (bucket id, compacted
// k-b nucleotides, position idx+s); that triple is a bijection of (kmer_s, minimizer_idx), so the store
// is one open-addressing table keyed on (kmer_s, minimizer_idx):
//
//   KW = 1 (2k <= 64):  16-byte slot  { u64 key;      u32 meta; u32 aux;          }   two per 32-B sector
//   KW = 2           :  32-byte slot  { u64 lo, hi;   u32 meta; u32 aux; u64 aux2 }   one per sector
//   meta = (minimizer_idx + 1) in bits 0..7 (0 = claimed but not yet published).
//   cnt[slot] (a dense u32 array beside the table): the counter's DATA.  The reference's uint8_t wraps at 256
//          (apps/counter.cpp:262-269); a u32 that is only ever incremented and read modulo 256 is exactly that,
//          and the ++ is one fire-and-forget red.add on a word that shares its 32-byte sector with the counts
//          of the neighbouring k-mers of the same superkmer: a found k-mer no longer dirties its slot's sector.
//
// insert = CAS-claim of the key word(s) (atomicCAS / atom.cas.b128), then one atomicAdd publishes the
// idx tag.  A slot never moves until the table is rehashed, so the slot index is the stable DATA handle
// that replaces the reference's realloc + protect_data.
#pragma once
#include "common.cuh"
#include "synth.cuh"

#define STORE_EMPTY (~0ull)
#define STORE_MAX_PROBE 8192u
#define META_TAG_MASK 0xFFu

struct Slot1 { u64 key; u32 meta; u32 aux; };
struct Slot2 { u64 lo, hi; u32 meta; u32 aux; u64 aux2; };

struct StoreView {
    void *slots;        // Slot1[cap] or Slot2[cap]
    u64 mask;           // cap - 1
    u32 *special;       // [64] meta words of the one key that equals the empty pattern (poly-G at k = 32 / 64)
    u32 *cnt;           // [cap + 64] counts (DATA of the counter, modulo 256); [cap + idx] belongs to special[idx]
    u64 *counters;      // [0] stored k-mers, [1] failed inserts, [2] records whose first k-mer was new, [3] dump cursor
    u32 *bucket_bits;   // 4^b bits: bucket ids in use (null = not tracked)
    unsigned char *data;  // DATA arena for data_size > 1 (null otherwise)
    u32 data_size;
    // failed inserts (probe limit hit) are parked here and replayed after a rehash
    u64 *overflow;      // [ovf_cap][3]  lo, hi, idx | add<<8
    u64 ovf_cap;
};

__device__ __forceinline__ u64 ld_cg64(const u64 *p) { return __ldcg(p); }
__device__ __forceinline__ u32 ld_cg32(const u32 *p) { return __ldcg(p); }
__device__ __forceinline__ void ld_relaxed128(const void *addr, u64 &lo, u64 &hi) {
    asm volatile("{\n .reg .b128 d;\n ld.relaxed.gpu.global.b128 d, [%2];\n mov.b128 {%0,%1}, d;\n}\n"
                 : "=l"(lo), "=l"(hi) : "l"(addr) : "memory");
}
__device__ __forceinline__ u32 ld_relaxed32(const u32 *addr) {
    u32 v;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(addr) : "memory");
    return v;
}
// the whole 32-byte slot in one request (sm_100: LDG.256): key words + meta word
__device__ __forceinline__ void ld_relaxed256(const void *addr, u64 &a, u64 &b, u64 &c, u64 &d) {
    asm volatile("ld.relaxed.gpu.global.v4.u64 {%0,%1,%2,%3}, [%4];" : "=l"(a), "=l"(b), "=l"(c), "=l"(d) : "l"(addr) : "memory");
}
__device__ __forceinline__ void cas128(void *addr, u64 clo, u64 chi, u64 slo, u64 shi, u64 &olo, u64 &ohi) {
    asm volatile("{\n .reg .b128 c, s, d;\n mov.b128 c, {%3,%4};\n mov.b128 s, {%5,%6};\n"
                 " atom.global.relaxed.gpu.cas.b128 d, [%2], c, s;\n mov.b128 {%0,%1}, d;\n}\n"
                 : "=l"(olo), "=l"(ohi) : "l"(addr), "l"(clo), "l"(chi), "l"(slo), "l"(shi) : "memory");
}

template <int KW> struct SlotOf;
template <> struct SlotOf<1> { typedef Slot1 type; };
template <> struct SlotOf<2> { typedef Slot2 type; };

template <int KW>
__device__ __forceinline__ u32 *slot_meta_ptr(const StoreView &T, u64 slot) {
    if (slot > T.mask) return T.special + (slot - T.mask - 1);
    typedef typename SlotOf<KW>::type S;
    return &reinterpret_cast<S *>(T.slots)[slot].meta;
}

// find-or-claim (kmer_s, idx).  add = increment applied to the count (1 for the counter, 0 for a plain insert).
// Returns 0 = found, 1 = newly inserted, -1 = probe limit reached (nothing changed).
template <int KW>
__device__ __forceinline__ int store_upsert(const StoreView &T, u64 lo, u64 hi, u32 idx, u32 add, u64 home, u64 &slot_out) {
    typedef typename SlotOf<KW>::type S;
    const u32 tag = idx + 1;
    if (lo == STORE_EMPTY && (KW == 1 || hi == STORE_EMPTY)) {
        const u32 old = atomicOr(T.special + idx, tag);
        slot_out = T.mask + 1 + idx;
        if (add) atomicAdd(T.cnt + slot_out, add);
        return (old & META_TAG_MASK) == 0 ? 1 : 0;
    }
    S *tab = reinterpret_cast<S *>(T.slots);
    u64 s = home;
    u32 probes = 0;
    while (probes < STORE_MAX_PROBE) {
        S *p = tab + s;
        bool match;
        // key and meta are fetched together (one 32-byte sector, one round trip); `mt` is only trusted when the key
        // it came with matched, and a zero tag sends the probe back to look again
        u32 mt;
        if (KW == 1) {
            u64 cur, w1;
            ld_relaxed128(p, cur, w1);
            mt = (u32)w1;
            if (cur == STORE_EMPTY) {
                u64 old = atomicCAS(reinterpret_cast<u64 *>(p), STORE_EMPTY, lo);
                if (old == STORE_EMPTY) {
                    atomicAdd(&p->meta, tag);
                    if (add) atomicAdd(T.cnt + s, add);
                    slot_out = s;
                    return 1;
                }
                cur = old;
                mt = 0;  // somebody else claimed the slot meanwhile: re-read its tag
            }
            match = cur == lo;
        } else {
            u64 clo, chi;
            ld_relaxed128(p, clo, chi);
            mt = ld_relaxed32(&p->meta);
            if (clo == STORE_EMPTY && chi == STORE_EMPTY) {
                u64 olo, ohi;
                cas128(p, STORE_EMPTY, STORE_EMPTY, lo, hi, olo, ohi);
                if (olo == STORE_EMPTY && ohi == STORE_EMPTY) {
                    atomicAdd(&p->meta, tag);
                    if (add) atomicAdd(T.cnt + s, add);
                    slot_out = s;
                    return 1;
                }
                clo = olo; chi = ohi;
                mt = 0;
            }
            match = clo == lo && chi == hi;
        }
        if (match) {
            u32 t = mt & META_TAG_MASK;
            if (t == 0) continue;  // claimed by another thread, tag not published yet: look again
            if (t == tag) {
                if (add) atomicAdd(T.cnt + s, add);
                slot_out = s;
                return 0;
            }
        }
        probes++;
        s = probe_slot(home, lo, KW == 1 ? 0 : hi, idx, probes, T.mask);
    }
    return -1;
}

// lookup only; returns true, the slot and its count word (meta_out) when present
template <int KW>
__device__ __forceinline__ bool store_find(const StoreView &T, u64 lo, u64 hi, u32 idx, u64 home, u64 &slot_out, u32 &meta_out) {
    typedef typename SlotOf<KW>::type S;
    const u32 tag = idx + 1;
    if (lo == STORE_EMPTY && (KW == 1 || hi == STORE_EMPTY)) {
        u32 mt = ld_cg32(T.special + idx);
        slot_out = T.mask + 1 + idx;
        meta_out = ld_cg32(T.cnt + slot_out);
        return (mt & META_TAG_MASK) == tag;
    }
    const S *tab = reinterpret_cast<const S *>(T.slots);
    u64 s = home;
    for (u32 probes = 0; probes < STORE_MAX_PROBE;) {
        const S *p = tab + s;
        bool match;
        u32 mt;
        if (KW == 1) {
            u64 cur, w1;
            ld_relaxed128(p, cur, w1);
            mt = (u32)w1;
            if (cur == STORE_EMPTY) return false;
            match = cur == lo;
        } else {
            u64 clo, chi;
            ld_relaxed128(p, clo, chi);
            mt = ld_relaxed32(&p->meta);
            if (clo == STORE_EMPTY && chi == STORE_EMPTY) return false;
            match = clo == lo && chi == hi;
        }
        if (match) {
            u32 t = mt & META_TAG_MASK;
            if (t == 0) continue;
            if (t == tag) { slot_out = s; meta_out = ld_cg32(T.cnt + s); return true; }
        }
        probes++;
        s = probe_slot(home, lo, KW == 1 ? 0 : hi, idx, probes, T.mask);
    }
    return false;
}

__device__ __forceinline__ void store_note_new(const StoreView &T, u64 bucket) {
    if (T.bucket_bits) atomicOr(T.bucket_bits + (bucket >> 5), 1u << (bucket & 31));
}
__device__ __forceinline__ void store_park(const StoreView &T, u64 lo, u64 hi, u32 idx, u32 add) {
    u64 i = atomicAdd(T.counters + 1, 1ull);
    if (T.overflow && i < T.ovf_cap) {
        T.overflow[3 * i] = lo; T.overflow[3 * i + 1] = hi; T.overflow[3 * i + 2] = (u64)idx | ((u64)add << 8);
    }
}

// ---------------------------------------------------------------------------------------------
// k-mer j (emission order) of a record whose payload words are pw[] (header already masked off)
template <int RW, int KW>
__device__ __forceinline__ void record_kmer(const BriskParams &P, const u64 (&pw)[RW], int n, int j, u64 &lo, u64 &hi) {
    u64 t[RW];
#pragma unroll
    for (int i = 0; i < RW; i++) t[i] = pw[i];
    mw_shr<RW, 1>(t, 2 * (n - 1 - j));  // n <= w <= 64: at most 126 bits
    lo = t[0] & P.kmask_lo;
    hi = KW == 2 ? (t[1] & P.kmask_hi) : 0ull;
}

enum { MODE_COUNT = 0, MODE_INSERT = 1, MODE_GET = 2 };

struct OccOut {
    const u64 *rec_occ;  // per record: (first read-order occurrence index << 1) | rev  (null: no per-occurrence output)
    u64 *slots;          // per occurrence (nullable)
    unsigned char *flags;  // per occurrence: newly (insert) / found (get) (nullable)
    unsigned char *data;   // per occurrence data_size bytes (get, nullable)
    u64 *agg;              // get: [0] += hits, [1] += sum of the DATA bytes of the hits (nullable)
};

// ---------------------------------------------------------------------------------------------
// records_kernel: one warp takes 32 records at a time, stages them in shared memory and spreads their k-mers over
// its lanes, so that every lane runs one table operation per step whatever the superkmer sizes are.
//
// A table operation is ONE probe (probe_once).  A lane whose probe did not settle its k-mer (slot taken by another
// key, or claimed but not yet published) appends the k-mer to the warp's retry queue in shared memory and the warp
// moves on; whenever the queue holds 32 entries the warp runs them as one full-width probe round.  The number of
// probe rounds a warp executes is then the AVERAGE number of probes per k-mer, not the maximum over its 32 lanes,
// and a lane never idles through another lane's collision chain.
#define REC_QUEUE 64  // retry queue entries per warp (<= 31 left over + 32 from one round)

struct KmerOp { u64 lo, hi, home, oi; u32 misc; bool active; };

enum { PR_FOUND = 0, PR_NEW = 1, PR_ABSENT = 2, PR_NEXT = 3, PR_AGAIN = 4 };

// one probe of slot s for (lo, hi, idx).  MODE_GET never writes.  PR_NEXT: the slot holds another key; PR_AGAIN: the
// slot holds this k-mer's key words but its tag is not published yet — look at the same slot again.
template <int KW, int MODE>
__device__ __forceinline__ int probe_once(const StoreView &T, u64 lo, u64 hi, u32 idx, u64 s) {
    typedef typename SlotOf<KW>::type S;
    const u32 tag = idx + 1;
    S *p = reinterpret_cast<S *>(T.slots) + s;
    u64 clo, chi;
    u32 mt;
    if (KW == 1) { ld_relaxed128(p, clo, chi); mt = (u32)chi; }
    else { u64 w2, w3; ld_relaxed256(p, clo, chi, w2, w3); mt = (u32)w2; }
    const bool empty = clo == STORE_EMPTY && (KW == 1 || chi == STORE_EMPTY);
    if (empty) {
        if (MODE == MODE_GET) return PR_ABSENT;
        u64 olo, ohi = STORE_EMPTY;
        if (KW == 1) olo = atomicCAS(reinterpret_cast<u64 *>(p), STORE_EMPTY, lo);
        else cas128(p, STORE_EMPTY, STORE_EMPTY, lo, hi, olo, ohi);
        if (olo == STORE_EMPTY && ohi == STORE_EMPTY) {
            atomicAdd(&p->meta, tag);
            if (MODE == MODE_COUNT) atomicAdd(T.cnt + s, 1u);
            return PR_NEW;
        }
        // somebody else claimed the slot meanwhile: same key words -> wait for its tag, else move on
        return (olo == lo && (KW == 1 || ohi == hi)) ? PR_AGAIN : PR_NEXT;
    }
    if (clo == lo && (KW == 1 || chi == hi)) {
        const u32 t = mt & META_TAG_MASK;
        if (t == 0) return PR_AGAIN;
        if (t == tag) {
            if (MODE == MODE_COUNT) atomicAdd(T.cnt + s, 1u);
            return PR_FOUND;
        }
    }
    return PR_NEXT;
}

template <int RW, int KW>
struct RecWarpSmem {
    static constexpr int PSTRIDE = 2 * RW + 3;  // odd: records of neighbouring lanes start in different banks; 3 zero words of slack
    u32 payload[32 * PSTRIDE];
    u64 gbase[32];
    u64 occ[32];
    u32 pack[32];  // exclusive prefix of n | n << 12 | idx0 << 20
    // retry queue
    u64 q_lo[REC_QUEUE];
    u64 q_hi[KW == 2 ? REC_QUEUE : 1];
    u64 q_home[REC_QUEUE];
    u64 q_oi[REC_QUEUE];
    u32 q_misc[REC_QUEUE];  // idx | first-of-record flag << 8 | probes << 16
};

template <int RW, int KW, int MODE>
__global__ void __launch_bounds__(256) records_kernel(const BriskParams P, const u64 *__restrict__ recs, u64 n_recs,
                                                      StoreView T, OccOut O) {
    __shared__ RecWarpSmem<RW, KW> smem[8];
    RecWarpSmem<RW, KW> &S = smem[threadIdx.x >> 5];
    constexpr int PSTRIDE = RecWarpSmem<RW, KW>::PSTRIDE;
    const int lane = threadIdx.x & 31;
    const u32 lane_le = 0xffffffffu >> (31 - lane), lane_lt = lane_le >> 1;
    const u64 warp = ((u64)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const u64 n_warps = ((u64)gridDim.x * blockDim.x) >> 5;
    const u64 top_mask = rec_payload_mask_top(P);
    u32 new_keys = 0, new_recs = 0;
    u64 agg_hits = 0, agg_sum = 0;
    int qn = 0;  // queue length (warp-uniform)

    // MODE_GET: the count word of a hit is requested when the hit is known and consumed one probe round later, after
    // that round's slot load has come back, so that it never adds a second dependent round trip to a step
    bool pend_hit = false;
    u32 pend_cv = 0;
    u64 pend_oi = 0;
    auto consume_count = [&]() {
        if (MODE == MODE_GET && pend_hit) {
            agg_sum += pend_cv & 0xffu;
            if (O.rec_occ && O.data && T.data_size == 1) O.data[pend_oi] = (unsigned char)pend_cv;
        }
        pend_hit = false;
    };
    // one probe round over the k-mers the lanes hold; unsettled ones go (back) to the queue
    auto round = [&](bool active, u64 lo, u64 hi, u64 home, u64 oi, u32 misc) {
        int res = PR_NEXT;
        u64 slot = ~0ull;
        if (active) {
            const u32 idx = misc & 0xffu, probes = misc >> 16;
            if (lo == STORE_EMPTY && (KW == 1 || hi == STORE_EMPTY)) {  // the key that equals the empty pattern
                slot = T.mask + 1 + idx;
                if (MODE == MODE_GET) {
                    res = (ld_cg32(T.special + idx) & META_TAG_MASK) == idx + 1 ? PR_FOUND : PR_ABSENT;
                } else {
                    res = (atomicOr(T.special + idx, idx + 1) & META_TAG_MASK) == 0 ? PR_NEW : PR_FOUND;
                    if (MODE == MODE_COUNT) atomicAdd(T.cnt + slot, 1u);
                }
            } else if (probes >= STORE_MAX_PROBE) {
                res = PR_ABSENT;  // probe limit: parked (insert) / absent (get)
                if (MODE != MODE_GET) store_park(T, lo, hi, idx, MODE == MODE_COUNT ? 1u : 0u);
            } else {
                slot = probe_slot(home, lo, KW == 1 ? 0 : hi, idx, probes, T.mask);
                res = probe_once<KW, MODE>(T, lo, hi, idx, slot);
            }
        }
        const bool again = active && res >= PR_NEXT;
        const u32 am = __ballot_sync(0xffffffffu, again);
        if (again) {
            const int q = qn + __popc(am & lane_lt);
            S.q_lo[q] = lo;
            if (KW == 2) S.q_hi[q] = hi;
            S.q_home[q] = home;
            S.q_oi[q] = oi;
            S.q_misc[q] = res == PR_NEXT ? misc + (1u << 16) : misc;
        }
        qn += __popc(am);
        if (MODE == MODE_GET) {
            consume_count();
            const bool hit = active && !again && res == PR_FOUND;
            if (hit) {
                agg_hits++;
                pend_cv = ld_relaxed32(T.cnt + slot);
                pend_hit = true;
                pend_oi = oi;
            }
            if (active && !again && O.rec_occ) {
                if (O.flags) O.flags[oi] = hit;
                if (O.slots) O.slots[oi] = hit ? slot : ~0ull;
                if (O.data) {
                    if (T.data_size == 1) { if (!hit) O.data[oi] = 0; }
                    else for (u32 q = 0; q < T.data_size; q++)
                        O.data[oi * T.data_size + q] = hit ? T.data[slot * T.data_size + q] : 0;
                }
            }
        } else if (active && !again) {
            if (res == PR_NEW) {
                new_keys++;
                if (misc & 0x100u) new_recs++;
            }
            if (O.rec_occ) {
                if (O.flags) O.flags[oi] = res == PR_NEW;
                if (O.slots) O.slots[oi] = res == PR_ABSENT ? ~0ull : slot;
            }
        }
    };
    // run the newest 32 (or all, if fewer) queued k-mers as one round
    auto retry = [&]() {
        const int take = qn < 32 ? qn : 32;
        qn -= take;
        __syncwarp();
        const bool active = lane < take;
        u64 lo = 0, hi = 0, home = 0, oi = 0;
        u32 misc = 0;
        if (active) {
            const int q = qn + lane;
            lo = S.q_lo[q];
            if (KW == 2) hi = S.q_hi[q];
            home = S.q_home[q];
            oi = S.q_oi[q];
            misc = S.q_misc[q];
        }
        __syncwarp();
        round(active, lo, hi, home, oi, misc);
    };

    for (u64 r0 = warp * 32; r0 < n_recs; r0 += n_warps * 32) {
        const u64 ri = r0 + lane;
        int n = 0;
        __syncwarp();  // the lanes are done with the previous group's staging
        {
            u64 w[RW];
            int idx0 = 0;
            u64 bucket = 0, gbase = 0;
            if (ri < n_recs) {
                const ulonglong2 *src = reinterpret_cast<const ulonglong2 *>(recs + ri * RW);
#pragma unroll
                for (int i = 0; i < RW; i += 2) {
                    ulonglong2 v = __ldcs(src + (i >> 1));
                    w[i] = v.x; w[i + 1] = v.y;
                }
                rec_unpack_header(P, w[RW - 1], bucket, n, idx0);
                w[RW - 1] &= top_mask;
                // every k-mer of an inserted record ends up stored, so its bucket is in use: one bit per record (checked
                // before the atomic), not one atomic per new k-mer
                if (MODE != MODE_GET && T.bucket_bits) {
                    const u32 bit = 1u << (bucket & 31);
                    if (!(__ldg(T.bucket_bits + (bucket >> 5)) & bit)) atomicOr(T.bucket_bits + (bucket >> 5), bit);
                }
                // the minimizer m-mer of the record (nucleotides [idx0, idx0+m) of its first k-mer) places the whole superkmer
                gbase = group_home(mw_extract64<RW, 3>(w, 2 * (n - 1 + idx0)) & P.m_mask, bucket, T.mask);
                S.occ[lane] = O.rec_occ ? O.rec_occ[ri] : 0ull;
            } else {
#pragma unroll
                for (int i = 0; i < RW; i++) w[i] = 0;
            }
            u32 *pp = S.payload + lane * PSTRIDE;
#pragma unroll
            for (int i = 0; i < RW; i++) { pp[2 * i] = (u32)w[i]; pp[2 * i + 1] = (u32)(w[i] >> 32); }
            pp[2 * RW] = 0; pp[2 * RW + 1] = 0; pp[2 * RW + 2] = 0;
            S.gbase[lane] = gbase;
            // inclusive prefix sum of n over the warp
            int incl = n;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                int v = __shfl_up_sync(0xffffffffu, incl, d);
                if (lane >= d) incl += v;
            }
            n = incl - n;  // from here on: the exclusive prefix = first output slot of this lane's record
            S.pack[lane] = (u32)n | ((u32)(incl - n) << 12) | ((u32)idx0 << 20);
        }
        const int excl = n;
        const int total = __shfl_sync(0xffffffffu, excl + (int)((S.pack[lane] >> 12) & 0xffu), 31);
        __syncwarp();
        // k-mer of output slot c + lane; its home sector is requested (L2 prefetch) one step ahead of its probe
        auto prepare = [&](int c) -> KmerOp {
            KmerOp K;
            K.lo = 0; K.hi = 0; K.home = 0; K.oi = 0; K.misc = 0;
            const int o = c + lane;
            // source record of output slot o = the last record that starts at or before o: records are in lane order,
            // so it is (records starting before c) + (records starting in [c, o]) - 1
            const u32 starts = __reduce_or_sync(0xffffffffu, (excl >= c && excl < c + 32) ? 1u << (excl - c) : 0u);
            const int before = __popc(__ballot_sync(0xffffffffu, excl < c));
            K.active = o < total;
            if (K.active) {
                const int src = before + __popc(starts & lane_le) - 1;
                const u32 pk = S.pack[src];
                const int s_n = (int)((pk >> 12) & 0xffu), j = o - (int)(pk & 0xfffu);
                const u32 idx = (pk >> 20) + (u32)j;
                // k-mer j (emission order) of the record: bits [2(n-1-j), 2(n-1-j) + 2k) of its payload
                const int sh = 2 * (s_n - 1 - j);
                const u32 *pp = S.payload + src * PSTRIDE + (sh >> 5);
                u32 x[2 * KW + 1];
#pragma unroll
                for (int i = 0; i <= 2 * KW; i++) x[i] = pp[i];
                u32 y[2 * KW];
#pragma unroll
                for (int i = 0; i < 2 * KW; i++) y[i] = __funnelshift_r(x[i], x[i + 1], sh & 31);
                K.lo = ((u64)y[0] | ((u64)y[1] << 32)) & P.kmask_lo;
                if (KW == 2) K.hi = ((u64)y[2] | ((u64)y[3] << 32)) & P.kmask_hi;
                K.home = (S.gbase[src] + idx) & T.mask;
                if (O.rec_occ) {
                    const u64 so = S.occ[src];
                    K.oi = (so >> 1) + ((so & 1) ? (u64)(s_n - 1 - j) : (u64)j);
                }
                K.misc = idx | (j == 0 ? 0x100u : 0u);
                typedef typename SlotOf<KW>::type SL;
                asm volatile("prefetch.global.L2 [%0];" ::"l"(reinterpret_cast<const SL *>(T.slots) + K.home));
            }
            return K;
        };
        KmerOp nx = prepare(0);
        for (int c = 0; c < total; c += 32) {
            const KmerOp cur = nx;
            if (c + 32 < total) nx = prepare(c + 32);
            round(cur.active, cur.lo, cur.hi, cur.home, cur.oi, cur.misc);
            while (qn >= 32) retry();
        }
    }
    while (qn > 0) retry();
    consume_count();
    if (MODE == MODE_GET && O.agg) {
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) {
            agg_hits += __shfl_xor_sync(0xffffffffu, agg_hits, d);
            agg_sum += __shfl_xor_sync(0xffffffffu, agg_sum, d);
        }
        if (lane == 0 && agg_hits) { atomicAdd(O.agg, agg_hits); atomicAdd(O.agg + 1, agg_sum); }
    }
    if (MODE != MODE_GET) {
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) {
            new_keys += __shfl_xor_sync(0xffffffffu, new_keys, d);
            new_recs += __shfl_xor_sync(0xffffffffu, new_recs, d);
        }
        if (lane == 0) {
            if (new_keys) atomicAdd(T.counters + 0, (u64)new_keys);
            if (new_recs) atomicAdd(T.counters + 2, (u64)new_recs);
        }
    }
}

// explicit k-mers: Brisk::insert / Brisk::get.  kmers = {lo, hi, idx, pad} as brisk_kmer (24 bytes).
struct KmerIn { u64 lo, hi; unsigned char idx; unsigned char pad[7]; };

template <int KW, int MODE>
__global__ void __launch_bounds__(256) kmers_kernel(const BriskParams P, const KmerIn *__restrict__ kmers, u64 n, StoreView T,
                                                    u64 *slots, unsigned char *flags, unsigned char *data) {
    u32 new_keys = 0;
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x) {
        {
            u64 lo = kmers[i].lo & P.kmask_lo, hi = kmers[i].hi & P.kmask_hi;
            u32 idx = kmers[i].idx;
            const u64 home = (group_home(minimizer_word(P, lo, hi, (int)idx), bucket_of_kmer(P, lo, hi, (int)idx), T.mask) + idx) & T.mask;
            u64 slot = ~0ull;
            if (MODE == MODE_GET) {
                u32 mt = 0;
                bool found = store_find<KW>(T, lo, hi, idx, home, slot, mt);
                if (slots) slots[i] = found ? slot : ~0ull;
                if (flags) flags[i] = found;
                if (data) {
                    if (T.data_size == 1) data[i] = found ? (unsigned char)mt : 0;
                    else for (u32 q = 0; q < T.data_size; q++) data[i * T.data_size + q] = found ? T.data[slot * T.data_size + q] : 0;
                }
            } else {
                const u32 add = MODE == MODE_COUNT ? 1u : 0u;
                int res = store_upsert<KW>(T, lo, hi, idx, add, home, slot);
                if (res == 1) {
                    new_keys++;
                    store_note_new(T, bucket_of_kmer(P, lo, hi, (int)idx));
                } else if (res < 0) {
                    store_park(T, lo, hi, idx, add);
                }
                if (slots) slots[i] = res < 0 ? ~0ull : slot;
                if (flags) flags[i] = res == 1;
            }
        }
    }
    if (MODE != MODE_GET) {
        // lanes may have left the loop at different times; plain per-thread atomics keep this simple
        if (new_keys) atomicAdd(T.counters + 0, (u64)new_keys);
    }
}

// replay of parked inserts after a rehash
template <int KW>
__global__ void __launch_bounds__(256) replay_kernel(const BriskParams P, const u64 *__restrict__ parked, u64 n, StoreView T) {
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x) {
        u64 lo = parked[3 * i], hi = parked[3 * i + 1], x = parked[3 * i + 2];
        u32 idx = (u32)(x & 0xFF), add = (u32)(x >> 8);
        u64 slot;
        const u64 home = (group_home(minimizer_word(P, lo, hi, (int)idx), bucket_of_kmer(P, lo, hi, (int)idx), T.mask) + idx) & T.mask;
        int res = store_upsert<KW>(T, lo, hi, idx, add, home, slot);
        if (res == 1) {
            atomicAdd(T.counters + 0, 1ull);
            store_note_new(T, bucket_of_kmer(P, lo, hi, (int)idx));
        } else if (res < 0) {
            store_park(T, lo, hi, idx, add);
        }
    }
}

// rehash: every occupied slot of `from` is re-seated in `to` with its meta / aux / DATA
template <int KW>
__global__ void __launch_bounds__(256) rehash_kernel(const BriskParams P, StoreView from, StoreView to) {
    typedef typename SlotOf<KW>::type S;
    const S *src = reinterpret_cast<const S *>(from.slots);
    S *dst = reinterpret_cast<S *>(to.slots);
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i <= from.mask; i += (u64)gridDim.x * blockDim.x) {
        S v = src[i];
        u64 lo = *reinterpret_cast<const u64 *>(&v);
        u64 hi = KW == 2 ? reinterpret_cast<const u64 *>(&v)[1] : 0;
        if (lo == STORE_EMPTY && (KW == 1 || hi == STORE_EMPTY)) continue;
        u32 idx = (v.meta & META_TAG_MASK) - 1;
        const u64 home = (group_home(minimizer_word(P, lo, hi, (int)idx), bucket_of_kmer(P, lo, hi, (int)idx), to.mask) + idx) & to.mask;
        u64 s = home;
        for (u32 probes = 0;;) {
            S *p = dst + s;
            bool won;
            if (KW == 1) {
                won = atomicCAS(reinterpret_cast<u64 *>(p), STORE_EMPTY, lo) == STORE_EMPTY;
            } else {
                u64 olo, ohi;
                cas128(p, STORE_EMPTY, STORE_EMPTY, lo, hi, olo, ohi);
                won = olo == STORE_EMPTY && ohi == STORE_EMPTY;
            }
            if (won) {
                p->meta = v.meta;
                p->aux = v.aux;
                to.cnt[s] = from.cnt[i];
                if (KW == 2) reinterpret_cast<u64 *>(p)[3] = reinterpret_cast<const u64 *>(&v)[3];
                if (from.data) for (u32 q = 0; q < from.data_size; q++) to.data[s * to.data_size + q] = from.data[i * from.data_size + q];
                break;
            }
            probes++;
            s = probe_slot(home, lo, hi, idx, probes, to.mask);
        }
    }
}

// dump: compacts occupied slots into (key, data) arrays (unordered)
template <int KW>
__global__ void __launch_bounds__(256) dump_kernel(StoreView T, KmerIn *keys, unsigned char *data, u64 cap_out, int n_special) {
    typedef typename SlotOf<KW>::type S;
    const S *tab = reinterpret_cast<const S *>(T.slots);
    const u64 total = T.mask + 1 + (u64)n_special;
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (u64)gridDim.x * blockDim.x) {
        u64 lo, hi = 0;
        u32 meta;
        if (i <= T.mask) {
            lo = *reinterpret_cast<const u64 *>(tab + i);
            if (KW == 2) hi = reinterpret_cast<const u64 *>(tab + i)[1];
            if (lo == STORE_EMPTY && (KW == 1 || hi == STORE_EMPTY)) continue;
            meta = tab[i].meta;
        } else {
            meta = T.special[i - T.mask - 1];
            if ((meta & META_TAG_MASK) == 0) continue;
            lo = STORE_EMPTY;
            hi = KW == 2 ? STORE_EMPTY : 0;
        }
        u64 o = atomicAdd(T.counters + 3, 1ull);
        if (o < cap_out) {
            if (keys) {
                KmerIn kk;
                kk.lo = lo; kk.hi = hi; kk.idx = (unsigned char)((meta & META_TAG_MASK) - 1);
                for (int q = 0; q < 7; q++) kk.pad[q] = 0;
                keys[o] = kk;
            }
            if (data) {
                if (T.data_size == 1) data[o] = (unsigned char)T.cnt[i];
                else for (u32 q = 0; q < T.data_size; q++) data[o * T.data_size + q] = T.data[i * T.data_size + q];
            }
        }
    }
}

// out[0] += sum of checksum_mix over the stored k-mers, out[1] += sum of their DATA bytes, out[2] += their number
template <int KW>
__global__ void __launch_bounds__(256) checksum_kernel(StoreView T, int n_special, u64 *out) {
    typedef typename SlotOf<KW>::type S;
    const S *tab = reinterpret_cast<const S *>(T.slots);
    const u64 total = T.mask + 1 + (u64)n_special;
    u64 a = 0, b = 0, c = 0;
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (u64)gridDim.x * blockDim.x) {
        u64 lo, hi = 0;
        u32 meta;
        if (i <= T.mask) {
            lo = *reinterpret_cast<const u64 *>(tab + i);
            if (KW == 2) hi = reinterpret_cast<const u64 *>(tab + i)[1];
            if (lo == STORE_EMPTY && (KW == 1 || hi == STORE_EMPTY)) continue;
            meta = tab[i].meta;
        } else {
            meta = T.special[i - T.mask - 1];
            if ((meta & META_TAG_MASK) == 0) continue;
            lo = STORE_EMPTY;
            hi = KW == 2 ? STORE_EMPTY : 0;
        }
        const u32 d = T.cnt[i] & 0xffu;
        a += checksum_mix(lo, hi, (meta & META_TAG_MASK) - 1, d);
        b += d;
        c += 1;
    }
    for (int d = 16; d > 0; d >>= 1) {
        a += __shfl_xor_sync(0xffffffffu, a, d);
        b += __shfl_xor_sync(0xffffffffu, b, d);
        c += __shfl_xor_sync(0xffffffffu, c, d);
    }
    if ((threadIdx.x & 31) == 0 && c) { atomicAdd(out, a); atomicAdd(out + 1, b); atomicAdd(out + 2, c); }
}

template <int KW>
__global__ void __launch_bounds__(256) init_table_kernel(void *slots, u64 cap) {
    // every 16 bytes: either an empty key (+ for KW = 2 the zero meta/aux half) or {EMPTY, 0}
    ulonglong2 *p = reinterpret_cast<ulonglong2 *>(slots);
    const u64 n16 = cap * (KW == 1 ? 1 : 2);
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += (u64)gridDim.x * blockDim.x) {
        if (KW == 1) p[i] = make_ulonglong2(STORE_EMPTY, 0ull);
        else p[i] = (i & 1) ? make_ulonglong2(0ull, 0ull) : make_ulonglong2(STORE_EMPTY, STORE_EMPTY);
    }
}

__global__ void __launch_bounds__(256) popcount_kernel(const u32 *bits, u64 n_words, u64 *out) {
    u64 acc = 0;
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n_words; i += (u64)gridDim.x * blockDim.x) acc += __popc(bits[i]);
    for (int d = 16; d > 0; d >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, d);
    if ((threadIdx.x & 31) == 0 && acc) atomicAdd(out, acc);
}
