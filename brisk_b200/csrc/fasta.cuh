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
//   fasta_summarize_kernel : every thread runs its 128-byte chunk from all 5 entry states (exit state, bases emitted,
//                            runs started per entry state); the CTA composes its 256 chunks into one tile summary
//   fasta_tiles_kernel     : one CTA composes the tile summaries from the start state H -> entry state, first base
//                            index and first run index of every tile; totals; the closing offset
//   fasta_emit_kernel      : every CTA reloads its chunk summaries, scans them, and every thread replays its chunk from
//                            its now known entry state, writing upper-cased bases densely and one offset per run
// The text is read twice, the bases written once, a 16-byte summary per 128-byte chunk written and read back:
// about 3.3 bytes of HBM traffic per text byte.
#pragma once
#include <cub/block/block_scan.cuh>

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

// 5 bits (one per entry state) -> 5 bytes of 0 / 1
__device__ __forceinline__ u64 fa_spread(u32 m) {
    return (((((u64)m * 0x0101010101ull) & 0x1008040201ull) + 0x7f7f7f7f7full) >> 7) & 0x0101010101ull;
}

// Summary of the chunk [beg, beg + FA_CHUNK) cut at n.  The five entry states are run TOGETHER as three 5-bit masks
// (bit s = the run that entered in state s): Hm = inside a header line, Lm = at a line start, Rm = the last sequence byte
// was a base.  Every byte class is then a handful of mask operations:
//   '\n'  : headers end (their b / r memory restarts at b), everybody is at a line start
//   '>'   : the runs at a line start enter a header; for the others it is a cut
//   base  : every run outside a header emits it; it opens a read where Rm is clear
//   other : a cut
// Bases are counted in one scalar between changes of Hm (all runs outside a header emit the same bases) and flushed
// into the five byte-wide counters when Hm changes; read starts are rare and added directly.
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
__device__ __forceinline__ unsigned char fa_byte(const FaChunkBytes &B, int i, int j) {  // byte j of quad i (static i)
    const u32 w = j < 4 ? B.q[i].x : j < 8 ? B.q[i].y : j < 12 ? B.q[i].z : B.q[i].w;
    return (unsigned char)(w >> (8 * (j & 3)));
}

__device__ __forceinline__ FaSummary fa_chunk_summary(const FaChunkBytes &B, u64 beg, u64 n) {
    u32 Hm = 1u, Lm = 6u, Rm = 20u;  // entry states H, Lb, Lr, Sb, Sr = bits 0..4
    u32 run = 0;                     // bases seen since Hm last changed
    u64 nbp = 0, nrp = 0;            // 5 byte-wide counters each
    const int len = beg >= n ? 0 : (n - beg < FA_CHUNK ? (int)(n - beg) : FA_CHUNK);
#pragma unroll
    for (int i = 0; i < FA_CHUNK / 16; i++) {
#pragma unroll
        for (int j = 0; j < 16; j++) {
            if (16 * i + j < len) {
                const u32 cls = fasta_class(fa_byte(B, i, j));
                if (cls == FA_BASE) {
                    const u32 st = ~Hm & ~Rm & 31u;
                    if (st) nrp += fa_spread(st);
                    run++;
                    Rm = 31u;
                    Lm = 0u;
                } else {
                    const u32 newH = cls == FA_NL ? 0u : cls == FA_GT ? (Hm | Lm) : Hm;
                    if (newH != Hm) {
                        nbp += fa_spread(~Hm & 31u) * run;
                        run = 0;
                    }
                    Rm = cls == FA_NL ? (Rm & ~Hm) : 0u;
                    Lm = cls == FA_NL ? 31u : 0u;
                    Hm = newH;
                }
            }
        }
    }
    nbp += fa_spread(~Hm & 31u) * run;
    FaSummary f;
    f.exits = 0;
#pragma unroll
    for (int s = 0; s < 5; s++) {
        const u32 h = (Hm >> s) & 1u, l = (Lm >> s) & 1u, r = (Rm >> s) & 1u;
        const u32 st = h ? FA_H : (l ? FA_LB + r : FA_SB + r);
        f.exits |= st << (3 * s);
        f.nb[s] = (u32)(nbp >> (8 * s)) & 255u;
        f.nr[s] = (u32)(nrp >> (8 * s)) & 255u;
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

__global__ void __launch_bounds__(FA_THREADS) fasta_summarize_kernel(const unsigned char *__restrict__ text, u64 n, FaSummary *__restrict__ tiles,
                                                                     uint4 *__restrict__ chunks) {
    typedef cub::BlockScan<FaSummary, FA_THREADS> Scan;
    __shared__ typename Scan::TempStorage tmp;
    const u64 beg = (u64)blockIdx.x * FA_TILE + (u64)threadIdx.x * FA_CHUNK;
    FaChunkBytes B;
    fa_load_chunk(text, beg, n, (reinterpret_cast<uintptr_t>(text) & 15) == 0, B);
    FaSummary mine = fa_chunk_summary(B, beg, n), incl;
    chunks[(u64)blockIdx.x * FA_THREADS + threadIdx.x] = fa_pack(mine);
    Scan(tmp).InclusiveScan(mine, incl, FaCompose());
    if (threadIdx.x == FA_THREADS - 1) tiles[blockIdx.x] = incl;
}

// per tile: entry state | first base index | first run index, from the start state H (the first line is a header)
struct FaTileEntry {
    u32 state, base0, run0, pad;
};
__global__ void __launch_bounds__(1024) fasta_tiles_kernel(const FaSummary *__restrict__ tiles, u32 n_tiles, FaTileEntry *__restrict__ entry,
                                                           u64 *__restrict__ totals, u64 *__restrict__ offsets) {
    typedef cub::BlockScan<FaSummary, 1024> Scan;
    __shared__ typename Scan::TempStorage tmp;
    const u32 per = (n_tiles + 1023) / 1024;
    const u32 t0 = threadIdx.x * per, t1 = min(n_tiles, t0 + per);
    FaSummary seg = fa_identity();
    FaCompose op;
    for (u32 t = t0; t < t1; t++) seg = op(seg, tiles[t]);
    FaSummary before;
    Scan(tmp).ExclusiveScan(seg, before, fa_identity(), op);
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

__global__ void __launch_bounds__(FA_THREADS) fasta_emit_kernel(const unsigned char *__restrict__ text, u64 n, const FaTileEntry *__restrict__ entry,
                                                                const uint4 *__restrict__ chunks, unsigned char *__restrict__ bases,
                                                                u64 *__restrict__ offsets) {
    typedef cub::BlockScan<FaSummary, FA_THREADS> Scan;
    __shared__ typename Scan::TempStorage tmp;
    __shared__ unsigned char stage[FA_TILE + (FA_TILE >> 7) * 4 + 16];
    __shared__ u32 s_total;
    const u64 beg = (u64)blockIdx.x * FA_TILE + (u64)threadIdx.x * FA_CHUNK;
    FaSummary mine = fa_unpack(chunks[(u64)blockIdx.x * FA_THREADS + threadIdx.x]), before;
    Scan(tmp).ExclusiveScan(mine, before, fa_identity(), FaCompose());
    const FaTileEntry e = entry[blockIdx.x];
    u32 state = before.exit_of(e.state);
    u32 lb = fa_pick(before.nb, e.state);  // index of this thread's first base inside the tile's output
    u64 r = (u64)e.run0 + fa_pick(before.nr, e.state);
    if (threadIdx.x == FA_THREADS - 1) s_total = lb + fa_pick(mine.nb, state);
    FaChunkBytes B;
    fa_load_chunk(text, beg, n, (reinterpret_cast<uintptr_t>(text) & 15) == 0, B);
    const int len = beg >= n ? 0 : (n - beg < FA_CHUNK ? (int)(n - beg) : FA_CHUNK);
#pragma unroll
    for (int i = 0; i < FA_CHUNK / 16; i++) {
#pragma unroll
        for (int j = 0; j < 16; j++) {
            if (16 * i + j < len) {
                const unsigned char c = fa_byte(B, i, j);
                const u32 cls = fasta_class(c);
                bool em, st;
                state = fasta_step(state, cls, fasta_map(cls), em, st);
                if (em) {
                    if (st) offsets[r++] = (u64)e.base0 + lb;
                    stage[fa_stage_at(lb++)] = c & 0xDF;
                }
            }
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
