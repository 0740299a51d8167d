// fasta.cuh — FASTA ingest on the device: getLineFasta + clean_dna of the reference's demo app
// (apps/counter.cpp:128-190) for a whole text buffer at once.
//
// What the reference's reader loop hands to count_sequence (one reader thread): for every record — the lines between
// two header lines, a header being the first line of the file or any line starting with '>' — the maximal runs of
// [ACGTacgt] of its concatenated lines ('\n' is transparent, every other byte cuts, '\r' included), upper-cased, in
// file order.  That is a 5-state automaton over the bytes:
//     H   inside a header line               Lb / Lr  at a line start, last sequence byte was a cut (b) / a base (r)
//                                            Sb / Sr  inside a sequence line, same b / r
// and automaton transitions compose associatively, so the text is processed as a two-level scan:
//   fasta_summarize_kernel : every thread classifies its 128-byte chunk into three 128-bit masks (newline, '>', base;
//                            one shared-memory table lookup per byte) and derives, with carry-propagating mask
//                            arithmetic, the exit state, bases emitted and runs started for all 5 entry states
//                            (fa_masks); the CTA composes its 256 chunks into one tile summary
//   fasta_tiles_kernel     : one CTA composes the tile summaries from the start state H -> entry state, first base
//                            index and first run index of every tile; totals; the closing offset
//   fasta_emit_kernel      : every CTA reloads its chunk summaries, scans them, and every thread recomputes the masks of its
//                            chunk for its now known entry state, writing the selected bytes upper-cased and densely,
//                            and one offset per run
// The text is read twice, the bases written once, a 16-byte summary per 128-byte chunk written and read back:
// about 3.3 bytes of HBM traffic per text byte.
#pragma once
#include "common.cuh"

#define FA_THREADS 256
#define FA_CHUNK 128  // bytes per thread: one 128-byte line; counts of a chunk still fit a byte
#define FA_TILE (FA_THREADS * FA_CHUNK)

enum { FA_H = 0, FA_LB = 1, FA_LR = 2, FA_SB = 3, FA_SR = 4 };

__device__ __forceinline__ bool fasta_is_base(unsigned char c) {
    c &= 0xDF;  // upper-case
    return c == 'A' || c == 'C' || c == 'G' || c == 'T';
}

// byte classes and their transition maps (3 bits per entry state, H..Sr from the low bits)
enum { FA_NL = 0, FA_GT = 1, FA_BASE = 2, FA_OTHER = 3 };
#define FA_MAP(h, lb, lr, sb, sr) ((u32)(h) | ((u32)(lb) << 3) | ((u32)(lr) << 6) | ((u32)(sb) << 9) | ((u32)(sr) << 12))
__device__ __forceinline__ u32 fasta_class(unsigned char c) { return c == '\n' ? FA_NL : c == '>' ? FA_GT : fasta_is_base(c) ? FA_BASE : FA_OTHER; }
__device__ __forceinline__ u32 fasta_map(u32 cls) {
    // newline: transparent inside a record (the b / r memory survives), ends a header.  '>': header at a line start,
    // a cut elsewhere.  base: opens or extends a run.  other: a cut.
    return cls == FA_NL ? FA_MAP(FA_LB, FA_LB, FA_LR, FA_LB, FA_LR)
         : cls == FA_GT ? FA_MAP(FA_H, FA_H, FA_H, FA_SB, FA_SB)
         : cls == FA_BASE ? FA_MAP(FA_H, FA_SR, FA_SR, FA_SR, FA_SR) : FA_MAP(FA_H, FA_SB, FA_SB, FA_SB, FA_SB);
}
// one automaton step; returns the new state, sets emit (a base is output) and start (it opens a run: the odd states
// Lb / Sb remember a cut)
__device__ __forceinline__ u32 fasta_step(u32 s, u32 cls, u32 map, bool &emit, bool &start) {
    emit = cls == FA_BASE && s != FA_H;
    start = emit && (s & 1u);
    return (map >> (3 * s)) & 7u;
}

// transition function of a stretch of text: per entry state the exit state and what was emitted on the way
struct FaSummary {
    u32 exits;   // 5 x 3 bits
    u32 nb[5];   // bases emitted
    u32 nr[5];   // runs started
    __device__ __forceinline__ u32 exit_of(u32 s) const { return (exits >> (3 * s)) & 7u; }
};
__device__ __forceinline__ FaSummary fa_identity() {
    FaSummary f;
    f.exits = 0u | (1u << 3) | (2u << 6) | (3u << 9) | (4u << 12);
#pragma unroll
    for (int s = 0; s < 5; s++) { f.nb[s] = 0; f.nr[s] = 0; }
    return f;
}
__device__ __forceinline__ u32 fa_pick(const u32 (&a)[5], u32 i) {  // a[i] with static indexing (registers)
    return i == 0 ? a[0] : i == 1 ? a[1] : i == 2 ? a[2] : i == 3 ? a[3] : a[4];
}
// a then b
struct FaCompose {
    __device__ __forceinline__ FaSummary operator()(const FaSummary &a, const FaSummary &b) const {
        FaSummary f;
        f.exits = 0;
#pragma unroll
        for (int s = 0; s < 5; s++) {
            const u32 mid = a.exit_of(s);
            f.exits |= b.exit_of(mid) << (3 * s);
            f.nb[s] = a.nb[s] + fa_pick(b.nb, mid);
            f.nr[s] = a.nr[s] + fa_pick(b.nr, mid);
        }
        return f;
    }
};

struct FaChunkBytes {  // the chunk's bytes in registers (eight 16-byte loads when the text is 16-byte aligned)
    uint4 q[FA_CHUNK / 16];
};
__device__ __forceinline__ void fa_load_chunk(const unsigned char *__restrict__ text, u64 beg, u64 n, bool aligned, FaChunkBytes &B) {
    if (aligned && beg + FA_CHUNK <= n) {
        const uint4 *src = reinterpret_cast<const uint4 *>(text + beg);
#pragma unroll
        for (int i = 0; i < FA_CHUNK / 16; i++) B.q[i] = __ldg(src + i);
    } else {
#pragma unroll
        for (int i = 0; i < FA_CHUNK / 16; i++) {
            u32 w[4] = {0, 0, 0, 0};
            for (int j = 0; j < 16; j++) {
                const u64 a = beg + 16 * i + j;
                if (a < n) w[j >> 2] |= (u32)__ldg(text + a) << (8 * (j & 3));
            }
            B.q[i] = make_uint4(w[0], w[1], w[2], w[3]);
        }
    }
}

// ---- the automaton over a chunk as mask arithmetic ---------------------------------------------------------------
// bit i of a 128-bit mask = byte i of the chunk.  Classification: one lookup per byte in a 256-entry table whose
// entries carry the class bit at bit 0 (newline), 8 ('>') or 16 (base): eight lookups, each shifted by its byte's
// position, add up to three 8-bit masks at once.
struct M128 { u64 lo, hi; };
__device__ __forceinline__ M128 m_and(M128 a, M128 b) { return {a.lo & b.lo, a.hi & b.hi}; }
__device__ __forceinline__ M128 m_or(M128 a, M128 b) { return {a.lo | b.lo, a.hi | b.hi}; }
__device__ __forceinline__ M128 m_andn(M128 a, M128 b) { return {a.lo & ~b.lo, a.hi & ~b.hi}; }  // a & ~b
__device__ __forceinline__ M128 m_shl1(M128 a, u64 in) { return {(a.lo << 1) | in, (a.hi << 1) | (a.lo >> 63)}; }
__device__ __forceinline__ M128 m_add(M128 a, M128 b) {
    M128 r;
    r.lo = a.lo + b.lo;
    r.hi = a.hi + b.hi + (r.lo < a.lo ? 1ull : 0ull);
    return r;
}
__device__ __forceinline__ u32 m_bit(M128 a, int i) { return (u32)((i < 64 ? a.lo >> i : a.hi >> (i - 64)) & 1ull); }
__device__ __forceinline__ u32 m_popc(M128 a) { return __popcll(a.lo) + __popcll(a.hi); }
__device__ __forceinline__ M128 m_first(int n) {  // bits [0, n)
    M128 r;
    r.lo = n >= 64 ? ~0ull : ((1ull << n) - 1);
    r.hi = n <= 64 ? 0ull : (n >= 128 ? ~0ull : ((1ull << (n - 64)) - 1));
    return r;
}

__device__ __forceinline__ void fa_init_lut(u32 *lut) {  // [256], filled by the CTA
    for (int c = threadIdx.x; c < 256; c += blockDim.x)
        lut[c] = (c == '\n' ? 1u : 0u) | (c == '>' ? 1u << 8 : 0u) | (fasta_is_base((unsigned char)c) ? 1u << 16 : 0u);
}
struct FaChunkMasks { M128 N, G, B; };
__device__ __forceinline__ FaChunkMasks fa_classify(const FaChunkBytes &Bt, const u32 *lut) {
    FaChunkMasks m;
    u64 n[2] = {0, 0}, g[2] = {0, 0}, b[2] = {0, 0};
#pragma unroll
    for (int i = 0; i < FA_CHUNK / 16; i++) {
        const u32 w[4] = {Bt.q[i].x, Bt.q[i].y, Bt.q[i].z, Bt.q[i].w};
#pragma unroll
        for (int h = 0; h < 2; h++) {  // 8 bytes: bits 0..7 / 8..15 / 16..23 of acc = their newline / '>' / base masks
            u32 acc = 0;
#pragma unroll
            for (int j = 0; j < 8; j++) acc += lut[(w[2 * h + (j >> 2)] >> (8 * (j & 3))) & 255u] << j;
            const int grp = 2 * i + h;  // 8-byte group 0..15
            n[grp >> 3] |= (u64)(acc & 255u) << (8 * (grp & 7));
            g[grp >> 3] |= (u64)((acc >> 8) & 255u) << (8 * (grp & 7));
            b[grp >> 3] |= (u64)((acc >> 16) & 255u) << (8 * (grp & 7));
        }
    }
    m.N = {n[0], n[1]}; m.G = {g[0], g[1]}; m.B = {b[0], b[1]};
    return m;
}
// The chunk of n bytes entered in state `entry`: K = the bases that are emitted, St = the ones that open a run, and the
// exit state.  (Checked against the byte-by-byte automaton above for every entry state on random chunks; the derivation:
//   header lines  HL : runs of non-newline bytes that start at a '>' on a line start, or at byte 0 when the chunk is
//                      entered inside a header — adding the start bits to the non-newline mask ripples a carry through
//                      exactly those runs and clears them;
//   emitted       K  : bases outside header lines;
//   run starts    St : emitted bases whose previous byte is neither an emitted base nor a run of transparent newlines
//                      that follows one (the newline that ends a header line is a cut, not transparent) — the same
//                      carry trick over the transparent newlines, with the entry state's b / r memory as carry-in.)
__device__ __forceinline__ u32 fa_masks(const FaChunkMasks &C, int n, u32 entry, M128 &K, M128 &St) {
    const M128 V = m_first(n);
    const u64 inH = entry == FA_H, atL = (entry == FA_LB || entry == FA_LR), r = (entry == FA_LR || entry == FA_SR);
    const M128 N = m_and(C.N, V);
    const M128 M = m_andn(V, N);
    const M128 LS = m_and(m_shl1(N, atL), V);
    M128 Sx = m_and(C.G, LS);
    Sx.lo |= inH & M.lo & 1ull;
    const M128 HL = m_andn(M, m_add(M, Sx));
    M128 hn = m_shl1(HL, 0);
    hn.lo |= inH;
    const M128 HN = m_and(N, hn);
    K = m_andn(m_and(C.B, V), HL);
    const M128 T = m_andn(N, HN);
    const u64 reff = r & (inH ^ 1ull);
    const M128 S2 = m_and(m_shl1(K, reff), T);
    const M128 F = m_andn(T, m_add(T, S2));
    const M128 KF = m_or(K, F);
    St = m_andn(K, m_shl1(KF, reff));
    if (n == 0) return entry;
    const int last = n - 1;
    if (m_bit(HL, last)) return FA_H;
    const u32 l = m_bit(N, last), rb = m_bit(KF, last);
    return l ? (rb ? FA_LR : FA_LB) : (rb ? FA_SR : FA_SB);
}

__device__ __forceinline__ FaSummary fa_chunk_summary(const FaChunkMasks &C, u64 beg, u64 n) {
    const int len = beg >= n ? 0 : (n - beg < FA_CHUNK ? (int)(n - beg) : FA_CHUNK);
    FaSummary f;
    f.exits = 0;
#pragma unroll
    for (u32 s = 0; s < 5; s++) {
        M128 K, St;
        const u32 ex = fa_masks(C, len, s, K, St);
        f.exits |= ex << (3 * s);
        f.nb[s] = m_popc(K);
        f.nr[s] = m_popc(St);
    }
    return f;
}

// a chunk summary in 16 bytes (the counts of a 128-byte chunk fit a byte): x = exits, y/z = nb[0..4] nr[0..2], w = nr[3..4]
__device__ __forceinline__ uint4 fa_pack(const FaSummary &f) {
    uint4 v;
    v.x = f.exits;
    v.y = f.nb[0] | (f.nb[1] << 8) | (f.nb[2] << 16) | (f.nb[3] << 24);
    v.z = f.nb[4] | (f.nr[0] << 8) | (f.nr[1] << 16) | (f.nr[2] << 24);
    v.w = f.nr[3] | (f.nr[4] << 8);
    return v;
}
__device__ __forceinline__ FaSummary fa_unpack(const uint4 v) {
    FaSummary f;
    f.exits = v.x;
    f.nb[0] = v.y & 255; f.nb[1] = (v.y >> 8) & 255; f.nb[2] = (v.y >> 16) & 255; f.nb[3] = v.y >> 24;
    f.nb[4] = v.z & 255; f.nr[0] = (v.z >> 8) & 255; f.nr[1] = (v.z >> 16) & 255; f.nr[2] = v.z >> 24;
    f.nr[3] = v.w & 255; f.nr[4] = (v.w >> 8) & 255;
    return f;
}

__device__ __forceinline__ FaSummary fa_shfl_up(const FaSummary &f, int d) {
    FaSummary r;
    r.exits = __shfl_up_sync(0xffffffffu, f.exits, d);
#pragma unroll
    for (int s = 0; s < 5; s++) {
        r.nb[s] = __shfl_up_sync(0xffffffffu, f.nb[s], d);
        r.nr[s] = __shfl_up_sync(0xffffffffu, f.nr[s], d);
    }
    return r;
}
// exclusive scan of one summary per thread over a CTA of NW warps (warp shuffles, then the NW warp totals); *total
// (optional) receives the composition of the whole CTA
template <int NW>
__device__ __forceinline__ FaSummary fa_block_exclusive_scan(const FaSummary &mine, FaSummary *warp_tot /* [32] shared */, FaSummary *total = nullptr) {
    FaCompose op;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    FaSummary incl = mine;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const FaSummary o = fa_shfl_up(incl, d);
        if (lane >= d) incl = op(o, incl);
    }
    if (lane == 31) warp_tot[w] = incl;
    __syncthreads();
    if (w == 0) {
        FaSummary t = lane < NW ? warp_tot[lane] : fa_identity();
#pragma unroll
        for (int d = 1; d < NW; d <<= 1) {
            const FaSummary o = fa_shfl_up(t, d);
            if (lane >= d) t = op(o, t);
        }
        const FaSummary ex = fa_shfl_up(t, 1);
        if (lane < NW) warp_tot[lane] = lane == 0 ? fa_identity() : ex;  // exclusive prefix of the warps
        if (lane == NW - 1) warp_tot[NW] = t;                             // everything
    }
    __syncthreads();
    if (total) *total = warp_tot[NW];
    const FaSummary prev = fa_shfl_up(incl, 1);
    const FaSummary within = lane == 0 ? fa_identity() : prev;
    return op(warp_tot[w], within);
}

__global__ void __launch_bounds__(FA_THREADS, 2) fasta_summarize_kernel(const unsigned char *__restrict__ text, u64 n, FaSummary *__restrict__ tiles,
                                                                     uint4 *__restrict__ chunks) {
    __shared__ FaSummary warp_tot[FA_THREADS / 32 + 1];
    __shared__ u32 lut[256];
    fa_init_lut(lut);
    const u64 beg = (u64)blockIdx.x * FA_TILE + (u64)threadIdx.x * FA_CHUNK;
    FaChunkBytes B;
    fa_load_chunk(text, beg, n, (reinterpret_cast<uintptr_t>(text) & 15) == 0, B);
    __syncthreads();
    const FaChunkMasks C = fa_classify(B, lut);
    const FaSummary mine = fa_chunk_summary(C, beg, n);
    chunks[(u64)blockIdx.x * FA_THREADS + threadIdx.x] = fa_pack(mine);
    // the tile's summary = the composition of its 256 chunks: warp totals only (no per-thread prefix is needed here)
    FaCompose op;
    FaSummary incl = mine;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const FaSummary o = fa_shfl_up(incl, d);
        if (lane >= d) incl = op(o, incl);
    }
    if (lane == 31) warp_tot[w] = incl;
    __syncthreads();
    if (threadIdx.x == 0) {
        FaSummary t = warp_tot[0];
        for (int i = 1; i < FA_THREADS / 32; i++) t = op(t, warp_tot[i]);
        tiles[blockIdx.x] = t;
    }
}

// per tile: entry state | first base index | first run index, from the start state H (the first line is a header)
struct FaTileEntry {
    u32 state, base0, run0, pad;
};
__global__ void __launch_bounds__(1024) fasta_tiles_kernel(const FaSummary *__restrict__ tiles, u32 n_tiles, FaTileEntry *__restrict__ entry,
                                                           u64 *__restrict__ totals, u64 *__restrict__ offsets) {
    __shared__ FaSummary warp_tot[33];
    const u32 per = (n_tiles + 1023) / 1024;
    const u32 t0 = min(n_tiles, threadIdx.x * per), t1 = min(n_tiles, t0 + per);
    FaSummary seg = fa_identity();
    FaCompose op;
    for (u32 t = t0; t < t1; t++) seg = op(seg, tiles[t]);
    const FaSummary before = fa_block_exclusive_scan<32>(seg, warp_tot);
    // walk the segment again from the state this thread's segment is entered in
    u32 state = before.exit_of(FA_H), base0 = before.nb[FA_H], run0 = before.nr[FA_H];
    for (u32 t = t0; t < t1; t++) {
        FaTileEntry e;
        e.state = state; e.base0 = base0; e.run0 = run0; e.pad = 0;
        entry[t] = e;
        const FaSummary f = tiles[t];
        base0 += fa_pick(f.nb, state);
        run0 += fa_pick(f.nr, state);
        state = f.exit_of(state);
    }
    if (t1 == n_tiles && t0 < t1) {  // the thread that owns the last tile
        totals[0] = base0;
        totals[1] = run0;
        offsets[run0] = base0;  // closing offset
    }
}

// shared staging of a tile's output bytes: 4 bytes of padding per 128 keep the lanes of a warp (whose outputs start
// about 128 bytes apart) on different banks
__device__ __forceinline__ u32 fa_stage_at(u32 off) { return off + ((off >> 7) << 2); }

__global__ void __launch_bounds__(FA_THREADS, 2) fasta_emit_kernel(const unsigned char *__restrict__ text, u64 n, const FaTileEntry *__restrict__ entry,
                                                                const uint4 *__restrict__ chunks, unsigned char *__restrict__ bases,
                                                                u64 *__restrict__ offsets) {
    __shared__ FaSummary warp_tot[FA_THREADS / 32 + 1];
    __shared__ unsigned char stage[FA_TILE + (FA_TILE >> 7) * 4 + 16];
    __shared__ u32 s_total;
    __shared__ u32 lut[256];
    fa_init_lut(lut);
    const u64 beg = (u64)blockIdx.x * FA_TILE + (u64)threadIdx.x * FA_CHUNK;
    const FaSummary mine = fa_unpack(chunks[(u64)blockIdx.x * FA_THREADS + threadIdx.x]);
    const FaSummary before = fa_block_exclusive_scan<FA_THREADS / 32>(mine, warp_tot);
    const FaTileEntry e = entry[blockIdx.x];
    const u32 state = before.exit_of(e.state);
    u32 lb = fa_pick(before.nb, e.state);  // index of this thread's first base inside the tile's output
    u64 r = (u64)e.run0 + fa_pick(before.nr, e.state);
    if (threadIdx.x == FA_THREADS - 1) s_total = lb + fa_pick(mine.nb, state);
    FaChunkBytes B;
    fa_load_chunk(text, beg, n, (reinterpret_cast<uintptr_t>(text) & 15) == 0, B);
    __syncthreads();  // lut
    const int len = beg >= n ? 0 : (n - beg < FA_CHUNK ? (int)(n - beg) : FA_CHUNK);
    const FaChunkMasks C = fa_classify(B, lut);
    M128 K, St;
    fa_masks(C, len, state, K, St);
    // one offset per run that starts in the chunk: the number of emitted bases before it
    for (int half = 0; half < 2; half++) {
        u64 sm = half ? St.hi : St.lo;
        while (sm) {
            const int p = __ffsll((long long)sm) - 1;
            sm &= sm - 1;
            const u32 before_p = half ? __popcll(K.lo) + __popcll(K.hi & ((1ull << p) - 1)) : __popcll(K.lo & ((1ull << p) - 1));
            offsets[r++] = (u64)e.base0 + lb + before_p;
        }
    }
    // the emitted bytes, upper-cased, into the tile's staging area
#pragma unroll
    for (int i = 0; i < FA_CHUNK / 16; i++) {
        const u32 w[4] = {B.q[i].x, B.q[i].y, B.q[i].z, B.q[i].w};
        const u32 k16 = (u32)((i < 4 ? K.lo >> (16 * i) : K.hi >> (16 * (i - 4))) & 0xffffu);
#pragma unroll
        for (int j = 0; j < 16; j++) {
            if ((k16 >> j) & 1u) stage[fa_stage_at(lb++)] = (unsigned char)((w[j >> 2] >> (8 * (j & 3))) & 0xDFu);
        }
    }
    __syncthreads();
    // the tile's bases are one contiguous piece of the output: copy it out in aligned 4-byte words
    const u32 total = s_total;
    const u64 g0 = e.base0, g1 = g0 + total;
    for (u64 w = (g0 >> 2) + threadIdx.x; (w << 2) < g1; w += FA_THREADS) {
        const u64 a = w << 2;
        if (a >= g0 && a + 4 <= g1) {
            const u32 o = (u32)(a - g0);
            const u32 v = (u32)stage[fa_stage_at(o)] | ((u32)stage[fa_stage_at(o + 1)] << 8) | ((u32)stage[fa_stage_at(o + 2)] << 16) |
                          ((u32)stage[fa_stage_at(o + 3)] << 24);
            *reinterpret_cast<u32 *>(bases + a) = v;
        } else {
            for (u64 x = a < g0 ? g0 : a; x < a + 4 && x < g1; x++) bases[x] = stage[fa_stage_at((u32)(x - g0))];
        }
    }
}
