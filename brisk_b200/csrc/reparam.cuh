// reparam.cuh — Brisk::reallocate (brisk/Brisk.hpp:202-224): the index rebuilt under (k, m + 2, b + 2).  Every stored
// (kmer_s, DATA) keeps its kmer_s as it is stored; its new minimizer_idx comes from update_kmer (brisk/Kmers.cpp:88-97),
// i.e. the literal get_minimizer (brisk/Kmers.cpp:367-408: 64-bit truncation, tie rules) over the k-mer with the new
// m, mirrored (k - m - pos) when get_minimizer reports the reverse strand.  Two old entries with the same kmer_s (it was
// stored under two minimizer positions) become ONE new entry: the reference keeps the DATA of
// whichever it enumerates last (bucket order, then its arenas' insertion order); here the entry with the largest old
// minimizer_idx wins, which is deterministic.
#pragma once
#include "store.cuh"

// get_minimizer (brisk/Kmers.cpp:367-408) by one thread; P describes the NEW (k, m)
__device__ __forceinline__ void get_minimizer_literal(const BriskParams &P, u64 lo, u64 hi, int &min_position, bool &reversed) {
    const int k = P.k, m = P.m;
    u64 cur = lo;  // uint64_t cur_seq = seq: the truncation of Q3
    u64 fwd = cur & P.m_mask;
    u64 rc = rcbc(fwd, m);
    u64 mini = fwd < rc ? fwd : rc;
    u64 hmin = bfc_hash_literal(P.coef, P.eps, mini, P.m_mask, m);
    reversed = mini != fwd;
    min_position = 0;
    for (int i = 1; i <= k - m; i++) {
        cur >>= 2;
        fwd = cur & P.m_mask;
        rc = rcbc(fwd, m);
        const u64 mmer = fwd < rc ? fwd : rc;
        const u64 nh = bfc_hash_literal(P.coef, P.eps, mmer, P.m_mask, m);
        if (nh < hmin) {
            min_position = i; mini = mmer; reversed = mini != fwd; hmin = nh;
        } else if (nh == hmin) {
            if (k - m - i < min_position) {
                min_position = k - m - i; mini = mmer; reversed = mini != fwd; hmin = nh;
            } else if (k - m - i == min_position) {
                if (!canonized128(lo, hi, k)) { min_position = k - m - i; mini = mmer; reversed = false; }
            }
        }
    }
}

// pass 1: every old entry is seated in the new table under its new minimizer_idx; prio[new slot] = max over the old
// entries that landed there of (old minimizer_idx + 1) << 40 | old slot
template <int KW>
__global__ void __launch_bounds__(256) reparam_insert_kernel(const BriskParams P2, StoreView from, StoreView to, int n_special, u64 *prio) {
    typedef typename SlotOf<KW>::type S;
    const S *src = reinterpret_cast<const S *>(from.slots);
    const u64 total = from.mask + 1 + (u64)n_special;
    u32 new_keys = 0;
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (u64)gridDim.x * blockDim.x) {
        u64 lo, hi = 0;
        u32 old_idx;
        if (i <= from.mask) {
            lo = *reinterpret_cast<const u64 *>(src + i);
            if (KW == 2) hi = reinterpret_cast<const u64 *>(src + i)[1];
            if (lo == STORE_EMPTY && (KW == 1 || hi == STORE_EMPTY)) continue;
            old_idx = (src[i].meta & META_TAG_MASK) - 1;
        } else {
            const u32 mt = from.special[i - from.mask - 1];
            if ((mt & META_TAG_MASK) == 0) continue;
            old_idx = (mt & META_TAG_MASK) - 1;
            lo = STORE_EMPTY; hi = KW == 2 ? STORE_EMPTY : 0;
        }
        // update_kmer works on the k-mer VALUE (masked), also for the all-ones key
        const u64 vlo = lo & P2.kmask_lo, vhi = hi & P2.kmask_hi;
        int pos; bool rev;
        get_minimizer_literal(P2, vlo, vhi, pos, rev);
        const u32 idx = (u32)(rev ? P2.k - P2.m - pos : pos);
        const u64 bucket = bucket_of_kmer(P2, vlo, vhi, (int)idx);
        const u64 home = (group_home(minimizer_word(P2, vlo, vhi, (int)idx), bucket, to.mask) + idx) & to.mask;
        u64 slot;
        const int res = store_upsert<KW>(to, lo, hi, idx, 0u, home, slot);
        if (res < 0) { store_park(to, lo, hi, idx, 0u); continue; }
        if (res == 1) { new_keys++; store_note_new(to, bucket); }
        atomicMax((unsigned long long *)prio + slot, ((u64)(old_idx + 1) << 40) | i);
    }
    if (new_keys) atomicAdd((unsigned long long *)to.counters + 0, (unsigned long long)new_keys);
}
// pass 2: DATA of the winning old entry
__global__ void __launch_bounds__(256) reparam_data_kernel(StoreView from, StoreView to, int n_special, const u64 *__restrict__ prio) {
    const u64 total = to.mask + 1 + (u64)n_special;
    for (u64 s = (u64)blockIdx.x * blockDim.x + threadIdx.x; s < total; s += (u64)gridDim.x * blockDim.x) {
        const u64 p = prio[s];
        if (!p) continue;
        const u64 old = p & ((1ull << 40) - 1);
        to.cnt[s] = from.cnt[old];
        if (from.data) for (u32 q = 0; q < from.data_size; q++) to.data[s * to.data_size + q] = from.data[old * from.data_size + q];
    }
}
