// brisk_capi.cu — host side of libbrisk_b200.so: the C ABI of include/brisk_b200.h over the kernels in
// enumerate.cuh / partition.cuh / store.cuh.  There is no CPU implementation behind these entry points:
// without a CUDA device brisk_create fails.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include <cub/device/device_scan.cuh>

#include "../../include/brisk_b200.h"
#include "common.cuh"
#include "enumerate.cuh"
#include "store.cuh"

static thread_local std::string g_err;
static int fail(int code, const std::string &msg) {
    g_err = msg;
    return code;
}
#define CK(call)                                                                                           \
    do {                                                                                                   \
        cudaError_t e_ = (call);                                                                           \
        if (e_ != cudaSuccess)                                                                             \
            return fail(BRISK_E_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_) + " (" + __FILE__ + ":" + \
                                          std::to_string(__LINE__) + ")");                                 \
    } while (0)

struct DevBuf {
    void *p = nullptr;
    size_t cap = 0;
    cudaError_t ensure(size_t need) {
        if (need <= cap) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
        size_t want = need + need / 8 + 256;
        cudaError_t e = cudaMalloc(&p, want);
        if (e == cudaSuccess) cap = want;
        return e;
    }
    void release() {
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
    }
};

struct brisk_handle {
    BriskParams P;
    int device = 0;
    uint32_t flags = 0;
    uint32_t data_size = 1;
    uint64_t hint = 0;
    cudaStream_t stream = nullptr;
    cudaStream_t own_stream = nullptr;
    cudaEvent_t ev[6] = {};
    float last_ms[5] = {0, 0, 0, 0, 0};
    uint32_t last_launches = 0;
    // store
    void *slots = nullptr;
    uint64_t cap = 0;
    uint32_t *special = nullptr;
    u64 *counters = nullptr;  // 8 u64
    uint32_t *bucket_bits = nullptr;
    uint64_t bitmap_words = 0;
    unsigned char *data = nullptr;
    u64 *overflow = nullptr;
    uint64_t ovf_cap = 0;
    uint64_t n_keys = 0, n_new_recs = 0;
    // scratch
    DevBuf d_bases, d_offsets, d_packed, d_recs, d_rec_occ, d_kmer_base, d_cnt, d_out_slots, d_out_flags, d_out_data,
        d_kmers, d_cub;
    uint64_t *d_rec_count = nullptr;  // device counter
    std::mutex mu;
};

static bool is_device_ptr(const void *p) {
    if (!p) return false;
    cudaPointerAttributes a;
    cudaError_t e = cudaPointerGetAttributes(&a, p);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}

static uint64_t pow2ceil(uint64_t x) {
    uint64_t p = 1;
    while (p < x) p <<= 1;
    return p;
}

static int fill_params(BriskParams &P, int k, int m, int b, uint32_t flags) {
    // Brisk.hpp:50-51 asserts m odd and m < k; README.md:9 k <= 64; minimizer lives in a u64 with the class in
    // bits 62-63 (hashing.cpp:18) so m <= 31; the bucket directory of the reference is 4 * 4^b bytes.
    if (!(m % 2 == 1) || !(m < k) || k > 64 || k < 2 || m > BRISK_MAX_M || b < 1 || b > m || b > 16) return BRISK_E_PARAM;
    memset(&P, 0, sizeof(P));
    P.k = k; P.m = m; P.b = b;
    P.w = k - m + 1;
    P.s = (m - b + 1) / 2;
    int wb = 1;
    while ((1 << wb) < P.w) wb++;
    P.wb = wb;
    P.payload_bits = 2 * (2 * k - m);
    int bits = P.payload_bits + 2 * b + 2 * wb;
    int words = (bits + 63) / 64;
    words += words & 1;
    if (words > BRISK_MAX_RW) return BRISK_E_PARAM;
    P.rec_words = words;
    P.key_words = k <= 32 ? 1 : 2;
    P.literal_decycle = (flags & BRISK_F_LITERAL_DECYCLE) ? 1 : 0;
    P.m_mask = (m == 32) ? ~0ull : (((u64)1 << (2 * m)) - 1);
    if (k >= 64) { P.kmask_lo = ~0ull; P.kmask_hi = ~0ull; }      // the k = 64 mask fix
    else if (k > 32) { P.kmask_lo = ~0ull; P.kmask_hi = ((u64)1 << (2 * k - 64)) - 1; }
    else if (k == 32) { P.kmask_lo = ~0ull; P.kmask_hi = 0; }
    else { P.kmask_lo = ((u64)1 << (2 * k)) - 1; P.kmask_hi = 0; }
    P.bucket_mask = ((u64)1 << (2 * b)) - 1;
    P.eps = 0.000001;  // Decycling.h:19
    // DecyclingSet::DecyclingSet (Decycling.cpp:7-13), same libm as the reference build
    double unit = 2 * M_PI / m;
    for (size_t i = 4; i < (size_t)4 * m; i += 4) {
        P.coef[i + 1] = sin(unit * (i / 4));
        P.coef[i + 2] = 2 * P.coef[i + 1];
        P.coef[i + 3] = 3 * P.coef[i + 1];
    }
    for (int q = 0; q < m; q++) {
        P.tw_c[q] = cos(unit * q);
        P.tw_s[q] = sin(unit * q);
    }
    P.rot_c = cos(unit); P.rot_s = sin(unit);
    P.in_c = cos(unit * (m - 1)); P.in_s = sin(unit * (m - 1));
    P.guard = 1e-9;
    P.polyA_hash = bfc_hash_literal(P.coef, P.eps, 0, P.m_mask, m);
    return BRISK_OK;
}

static StoreView view_of(brisk_handle *h) {
    StoreView T;
    T.slots = h->slots;
    T.mask = h->cap - 1;
    T.special = h->special;
    T.counters = h->counters;
    T.bucket_bits = h->bucket_bits;
    T.data = h->data;
    T.data_size = h->data_size;
    T.overflow = h->overflow;
    T.ovf_cap = h->ovf_cap;
    return T;
}

static size_t slot_bytes(const brisk_handle *h) { return h->P.key_words == 1 ? sizeof(Slot1) : sizeof(Slot2); }

static int alloc_table(brisk_handle *h, uint64_t cap, void **slots, unsigned char **data) {
    size_t bytes = cap * slot_bytes(h);
    if (h->flags & BRISK_F_MANAGED) CK(cudaMallocManaged(slots, bytes));
    else CK(cudaMalloc(slots, bytes));
    int blocks = (int)std::min<uint64_t>((cap + 255) / 256, 148 * 16);
    if (h->P.key_words == 1) init_table_kernel<1><<<blocks, 256, 0, h->stream>>>(*slots, cap);
    else init_table_kernel<2><<<blocks, 256, 0, h->stream>>>(*slots, cap);
    CK(cudaGetLastError());
    *data = nullptr;
    if (h->data_size > 1) {
        size_t db = (cap + 64) * (size_t)h->data_size;
        if (h->flags & BRISK_F_MANAGED) CK(cudaMallocManaged((void **)data, db));
        else CK(cudaMalloc((void **)data, db));
        CK(cudaMemsetAsync(*data, 0, db, h->stream));
    }
    return BRISK_OK;
}

static int grow_to(brisk_handle *h, uint64_t new_cap) {
    if (new_cap <= h->cap) return BRISK_OK;
    void *ns = nullptr;
    unsigned char *nd = nullptr;
    int rc = alloc_table(h, new_cap, &ns, &nd);
    if (rc) return rc;
    if (h->slots) {
        StoreView from = view_of(h);
        StoreView to = from;
        to.slots = ns; to.mask = new_cap - 1; to.data = nd;
        int blocks = (int)std::min<uint64_t>((h->cap + 255) / 256, 148 * 16);
        if (h->P.key_words == 1) rehash_kernel<1><<<blocks, 256, 0, h->stream>>>(h->P, from, to);
        else rehash_kernel<2><<<blocks, 256, 0, h->stream>>>(h->P, from, to);
        CK(cudaGetLastError());
        if (h->data) {  // DATA of the special key follows the arena's end
            CK(cudaMemcpyAsync(nd + new_cap * h->data_size, h->data + h->cap * h->data_size, 64 * (size_t)h->data_size,
                               cudaMemcpyDeviceToDevice, h->stream));
        }
        CK(cudaStreamSynchronize(h->stream));
        cudaFree(h->slots);
        if (h->data) cudaFree(h->data);
    }
    h->slots = ns;
    h->data = nd;
    h->cap = new_cap;
    return BRISK_OK;
}

extern "C" const char *brisk_last_error(void) { return g_err.c_str(); }
extern "C" const char *brisk_version(void) { return "brisk_b200 0.1 (sm_100a)"; }

extern "C" int brisk_create(uint8_t k, uint8_t m, uint8_t b, uint32_t data_size, int device, uint64_t capacity_hint,
                            uint32_t flags, brisk_t **out) {
    if (!out) return fail(BRISK_E_PARAM, "out is null");
    *out = nullptr;
    BriskParams P;
    if (fill_params(P, k, m, b, flags) != BRISK_OK)
        return fail(BRISK_E_PARAM, "bad parameters: need m odd, m < k <= 64, m <= 31, 1 <= b <= min(m, 16)");
    if (data_size == 0 || data_size > 64) return fail(BRISK_E_PARAM, "data_size must be 1..64");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return fail(BRISK_E_CUDA, "no CUDA device: libbrisk_b200 has no CPU path");
    }
    if (device < 0) CK(cudaGetDevice(&device));
    if (device >= ndev) return fail(BRISK_E_PARAM, "device index out of range");
    CK(cudaSetDevice(device));
    // the store is probed at random: fetch single 32-byte sectors from HBM instead of 64-byte pairs
    cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, 32);
    cudaGetLastError();
    brisk_handle *h = new brisk_handle();
    h->P = P;
    h->device = device;
    h->flags = flags;
    h->data_size = data_size;
    h->hint = capacity_hint;
    CK(cudaStreamCreateWithFlags(&h->own_stream, cudaStreamNonBlocking));
    h->stream = h->own_stream;
    for (int i = 0; i < 6; i++) CK(cudaEventCreate(&h->ev[i]));
    CK(cudaMalloc((void **)&h->counters, 8 * sizeof(uint64_t)));
    CK(cudaMemsetAsync(h->counters, 0, 8 * sizeof(uint64_t), h->stream));
    CK(cudaMalloc((void **)&h->d_rec_count, sizeof(uint64_t)));
    if (flags & BRISK_F_MANAGED) CK(cudaMallocManaged((void **)&h->special, 64 * sizeof(uint32_t)));
    else CK(cudaMalloc((void **)&h->special, 64 * sizeof(uint32_t)));
    CK(cudaMemsetAsync(h->special, 0, 64 * sizeof(uint32_t), h->stream));
    h->bitmap_words = (((uint64_t)1 << (2 * b)) + 31) / 32;
    CK(cudaMalloc((void **)&h->bucket_bits, h->bitmap_words * sizeof(uint32_t)));
    CK(cudaMemsetAsync(h->bucket_bits, 0, h->bitmap_words * sizeof(uint32_t), h->stream));
    h->ovf_cap = 1u << 20;
    CK(cudaMalloc((void **)&h->overflow, h->ovf_cap * 3 * sizeof(uint64_t)));
    uint64_t cap = pow2ceil(std::max<uint64_t>(capacity_hint * 2, capacity_hint ? 1u << 10 : 1u << 16));
    int rc = grow_to(h, cap);
    if (rc) return rc;
    CK(cudaStreamSynchronize(h->stream));
    *out = h;
    return BRISK_OK;
}

extern "C" int brisk_destroy(brisk_t *h) {
    if (!h) return BRISK_OK;
    cudaSetDevice(h->device);
    cudaStreamSynchronize(h->stream);
    DevBuf *bufs[] = {&h->d_bases, &h->d_offsets, &h->d_packed, &h->d_recs, &h->d_rec_occ, &h->d_kmer_base, &h->d_cnt,
                      &h->d_out_slots, &h->d_out_flags, &h->d_out_data, &h->d_kmers, &h->d_cub};
    for (DevBuf *b : bufs) b->release();
    cudaFree(h->slots);
    cudaFree(h->data);
    cudaFree(h->special);
    cudaFree(h->counters);
    cudaFree(h->bucket_bits);
    cudaFree(h->overflow);
    cudaFree(h->d_rec_count);
    for (int i = 0; i < 6; i++) cudaEventDestroy(h->ev[i]);
    cudaStreamDestroy(h->own_stream);
    delete h;
    return BRISK_OK;
}

// ------------------------------------------------------------------------------------------------
struct ReadBatch {
    const unsigned char *d_bases = nullptr;
    const u64 *d_offsets = nullptr;
    u64 *d_packed = nullptr;
    uint64_t n_reads = 0, total_bytes = 0;
};

// copies (if needed) and packs a batch of reads
static int stage_reads(brisk_handle *h, const char *bases, const uint64_t *offsets, uint64_t n_reads, ReadBatch &B) {
    B.n_reads = n_reads;
    uint64_t total = 0;
    if (is_device_ptr(offsets)) {
        CK(cudaMemcpyAsync(&total, offsets + n_reads, sizeof(uint64_t), cudaMemcpyDeviceToHost, h->stream));
        CK(cudaStreamSynchronize(h->stream));
        B.d_offsets = (const u64 *)offsets;
    } else {
        total = offsets[n_reads];
        CK(h->d_offsets.ensure((n_reads + 1) * sizeof(uint64_t)));
        CK(cudaMemcpyAsync(h->d_offsets.p, offsets, (n_reads + 1) * sizeof(uint64_t), cudaMemcpyHostToDevice, h->stream));
        B.d_offsets = (const u64 *)h->d_offsets.p;
    }
    B.total_bytes = total;
    if (is_device_ptr(bases)) {
        B.d_bases = (const unsigned char *)bases;
    } else {
        CK(h->d_bases.ensure(total + 16));
        CK(cudaMemcpyAsync(h->d_bases.p, bases, total, cudaMemcpyHostToDevice, h->stream));
        B.d_bases = (const unsigned char *)h->d_bases.p;
    }
    uint64_t n_words = total / 32 + n_reads + 8;
    CK(h->d_packed.ensure(n_words * sizeof(uint64_t)));
    B.d_packed = (u64 *)h->d_packed.p;
    CK(cudaEventRecord(h->ev[0], h->stream));
    if (n_reads) {
        // tail words that get_bits64 may touch past the last read must be readable; zero the pad
        CK(cudaMemsetAsync(B.d_packed + (n_words - 8), 0, 8 * sizeof(uint64_t), h->stream));
        unsigned blocks = (unsigned)((n_reads + 7) / 8);
        pack_reads_kernel<<<blocks, 256, 0, h->stream>>>(B.d_bases, B.d_offsets, n_reads, B.d_packed);
        CK(cudaGetLastError());
        h->last_launches++;
    }
    CK(cudaEventRecord(h->ev[1], h->stream));
    return BRISK_OK;
}

template <int RW>
static cudaError_t launch_enum(brisk_handle *h, const ReadBatch &B, EnumOut out) {
    size_t smem = ((size_t)h->P.w * ENUM_THREADS + ENUM_POOL) * sizeof(u64) + 4 * h->P.m * sizeof(double);
    cudaError_t e = cudaFuncSetAttribute(enumerate_kernel<RW>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    unsigned blocks = (unsigned)((B.n_reads + ENUM_THREADS - 1) / ENUM_THREADS);
    enumerate_kernel<RW><<<blocks, ENUM_THREADS, smem, h->stream>>>(h->P, B.d_packed, B.d_offsets, B.n_reads, out);
    return cudaGetLastError();
}

__global__ void kmer_counts_kernel(const u64 *offsets, u64 n_reads, int k, u64 *counts) {
    u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n_reads) {
        u64 len = offsets[i + 1] - offsets[i];
        counts[i] = len >= (u64)k ? len - k + 1 : 0;
    }
}

// enumerates a staged batch into h->d_recs; with_occ also fills d_rec_occ / d_kmer_base
static int enumerate_batch(brisk_handle *h, const ReadBatch &B, bool with_occ, uint64_t &n_recs, uint64_t &n_occ) {
    n_recs = 0;
    n_occ = 0;
    if (B.n_reads == 0) return BRISK_OK;
    const int RW = h->P.rec_words;
    if (with_occ) {
        CK(h->d_cnt.ensure((B.n_reads + 1) * sizeof(uint64_t)));
        CK(h->d_kmer_base.ensure((B.n_reads + 1) * sizeof(uint64_t)));
        unsigned blocks = (unsigned)((B.n_reads + 1 + 255) / 256);
        CK(cudaMemsetAsync((u64 *)h->d_cnt.p + B.n_reads, 0, sizeof(uint64_t), h->stream));
        kmer_counts_kernel<<<blocks, 256, 0, h->stream>>>(B.d_offsets, B.n_reads, h->P.k, (u64 *)h->d_cnt.p);
        size_t tmp = 0;
        CK(cub::DeviceScan::ExclusiveSum(nullptr, tmp, (u64 *)h->d_cnt.p, (u64 *)h->d_kmer_base.p, (int)(B.n_reads + 1), h->stream));
        CK(h->d_cub.ensure(tmp));
        CK(cub::DeviceScan::ExclusiveSum(h->d_cub.p, tmp, (u64 *)h->d_cnt.p, (u64 *)h->d_kmer_base.p, (int)(B.n_reads + 1), h->stream));
        CK(cudaMemcpyAsync(&n_occ, (u64 *)h->d_kmer_base.p + B.n_reads, sizeof(uint64_t), cudaMemcpyDeviceToHost, h->stream));
        h->last_launches += 3;
    }
    // expected density of superkmers is 2/(w+1) per k-mer; leave slack, retry on overflow
    uint64_t upper = B.total_bytes;  // >= number of k-mer occurrences
    uint64_t guess = (uint64_t)(3.0 * (double)upper / (h->P.w + 1)) + B.n_reads + 4096;
    uint64_t cap = std::min(upper + 1, guess);
    for (int attempt = 0; attempt < 2; attempt++) {
        CK(h->d_recs.ensure(cap * RW * sizeof(uint64_t)));
        if (with_occ) CK(h->d_rec_occ.ensure(cap * sizeof(uint64_t)));
        CK(cudaMemsetAsync(h->d_rec_count, 0, sizeof(uint64_t), h->stream));
        EnumOut out;
        out.recs = (u64 *)h->d_recs.p;
        out.rec_cap = cap;
        out.rec_count = (u64 *)h->d_rec_count;
        out.rec_occ = with_occ ? (u64 *)h->d_rec_occ.p : nullptr;
        out.kmer_base = with_occ ? (const u64 *)h->d_kmer_base.p : nullptr;
        cudaError_t e = RW == 2 ? launch_enum<2>(h, B, out) : RW == 4 ? launch_enum<4>(h, B, out) : launch_enum<6>(h, B, out);
        CK(e);
        h->last_launches++;
        CK(cudaMemcpyAsync(&n_recs, h->d_rec_count, sizeof(uint64_t), cudaMemcpyDeviceToHost, h->stream));
        CK(cudaStreamSynchronize(h->stream));
        if (n_recs <= cap) break;
        cap = n_recs;
        if (attempt == 1) return fail(BRISK_E_CUDA, "superkmer count changed between two enumerations");
    }
    return BRISK_OK;
}

template <int RW, int KW, int MODE>
static cudaError_t launch_records_t(brisk_handle *h, uint64_t n_recs, OccOut O) {
    if (n_recs == 0) return cudaSuccess;
    uint64_t want = (n_recs + 255) / 256;  // one warp per 32 records
    unsigned blocks = (unsigned)std::min<uint64_t>(want, 148 * 8 * 4);
    records_kernel<RW, KW, MODE><<<blocks, 256, 0, h->stream>>>(h->P, (const u64 *)h->d_recs.p, n_recs, view_of(h), O);
    return cudaGetLastError();
}
template <int RW, int KW>
static cudaError_t launch_records_m(brisk_handle *h, int mode, uint64_t n_recs, OccOut O) {
    switch (mode) {
        case MODE_COUNT: return launch_records_t<RW, KW, MODE_COUNT>(h, n_recs, O);
        case MODE_INSERT: return launch_records_t<RW, KW, MODE_INSERT>(h, n_recs, O);
        default: return launch_records_t<RW, KW, MODE_GET>(h, n_recs, O);
    }
}
static cudaError_t launch_records(brisk_handle *h, int mode, uint64_t n_recs, OccOut O) {
    const int RW = h->P.rec_words, KW = h->P.key_words;
    if (KW == 1) {
        if (RW == 2) return launch_records_m<2, 1>(h, mode, n_recs, O);
        if (RW == 4) return launch_records_m<4, 1>(h, mode, n_recs, O);
        return launch_records_m<6, 1>(h, mode, n_recs, O);
    }
    if (RW == 2) return launch_records_m<2, 2>(h, mode, n_recs, O);
    if (RW == 4) return launch_records_m<4, 2>(h, mode, n_recs, O);
    return launch_records_m<6, 2>(h, mode, n_recs, O);
}

// makes room for `incoming` more k-mer occurrences before an insert
static int pregrow(brisk_handle *h, uint64_t incoming) {
    uint64_t est_new = incoming;
    if (h->hint) est_new = std::min<uint64_t>(incoming, h->hint > h->n_keys ? h->hint - h->n_keys : 0);
    uint64_t need = h->n_keys + est_new;
    if (need * 10 > h->cap * 7) return grow_to(h, pow2ceil(need * 2));
    return BRISK_OK;
}

// reads the counters back; replays parked inserts after growing until none is left
static int settle(brisk_handle *h) {
    for (int round = 0; round < 40; round++) {
        uint64_t c[4];
        CK(cudaMemcpyAsync(c, h->counters, sizeof(c), cudaMemcpyDeviceToHost, h->stream));
        CK(cudaStreamSynchronize(h->stream));
        h->n_keys = c[0];
        h->n_new_recs = c[2];
        uint64_t parked = c[1];
        if (parked == 0) return BRISK_OK;
        if (parked > h->ovf_cap) return fail(BRISK_E_FULL, "bucket store overflow: more than 2^20 inserts hit the probe limit; "
                                                          "give brisk_create a capacity_hint or call brisk_reserve");
        // move the parked entries aside, grow, replay
        DevBuf tmp;
        CK(tmp.ensure(parked * 3 * sizeof(uint64_t)));
        CK(cudaMemcpyAsync(tmp.p, h->overflow, parked * 3 * sizeof(uint64_t), cudaMemcpyDeviceToDevice, h->stream));
        CK(cudaMemsetAsync(h->counters + 1, 0, sizeof(uint64_t), h->stream));
        int rc = grow_to(h, std::max<uint64_t>(h->cap * 2, pow2ceil(2 * (h->n_keys + parked))));
        if (rc) { tmp.release(); return rc; }
        unsigned blocks = (unsigned)std::min<uint64_t>((parked + 255) / 256, 148 * 8);
        if (h->P.key_words == 1) replay_kernel<1><<<blocks, 256, 0, h->stream>>>(h->P, (const u64 *)tmp.p, parked, view_of(h));
        else replay_kernel<2><<<blocks, 256, 0, h->stream>>>(h->P, (const u64 *)tmp.p, parked, view_of(h));
        CK(cudaGetLastError());
        CK(cudaStreamSynchronize(h->stream));
        tmp.release();
    }
    return fail(BRISK_E_FULL, "bucket store could not absorb the parked inserts");
}

static void finish_timing(brisk_handle *h) {
    // ev: 0 start/pack begin, 1 pack end, 2 enumerate end, 3 partition end, 4 insert end
    cudaEventSynchronize(h->ev[4]);
    for (int i = 0; i < 4; i++) cudaEventElapsedTime(&h->last_ms[i], h->ev[i], h->ev[i + 1]);
    cudaEventElapsedTime(&h->last_ms[4], h->ev[0], h->ev[4]);
}

// shared driver of count / insert / get over reads
static int run_reads(brisk_handle *h, const char *bases, const uint64_t *offsets, uint64_t n_reads, int mode,
                     uint64_t *slots, uint8_t *flags, void *data_out, uint64_t *agg_out = nullptr) {
    if (!h) return fail(BRISK_E_PARAM, "null handle");
    if (n_reads && (!bases || !offsets)) return fail(BRISK_E_PARAM, "null input");
    std::lock_guard<std::mutex> lock(h->mu);
    CK(cudaSetDevice(h->device));
    h->last_launches = 0;
    ReadBatch B;
    int rc = stage_reads(h, bases, offsets, n_reads, B);
    if (rc) return rc;
    const bool with_occ = slots || flags || data_out;
    uint64_t n_recs = 0, n_occ = 0;
    rc = enumerate_batch(h, B, with_occ, n_recs, n_occ);
    if (rc) return rc;
    CK(cudaEventRecord(h->ev[2], h->stream));
    CK(cudaEventRecord(h->ev[3], h->stream));
    if (mode != MODE_GET) {
        rc = pregrow(h, B.total_bytes);
        if (rc) return rc;
    }
    OccOut O;
    O.rec_occ = with_occ ? (const u64 *)h->d_rec_occ.p : nullptr;
    O.slots = nullptr; O.flags = nullptr; O.data = nullptr; O.agg = nullptr;
    if (agg_out) {
        CK(cudaMemsetAsync(h->counters + 5, 0, 2 * sizeof(uint64_t), h->stream));
        O.agg = h->counters + 5;
    }
    // per-occurrence outputs go through device staging when the caller's buffers are on the host
    bool s_dev = is_device_ptr(slots), f_dev = is_device_ptr(flags), d_dev = is_device_ptr(data_out);
    if (slots) {
        if (s_dev) O.slots = (u64 *)slots;
        else { CK(h->d_out_slots.ensure((n_occ + 1) * sizeof(uint64_t))); O.slots = (u64 *)h->d_out_slots.p; }
    }
    if (flags) {
        if (f_dev) O.flags = flags;
        else { CK(h->d_out_flags.ensure(n_occ + 1)); O.flags = (unsigned char *)h->d_out_flags.p; }
    }
    if (data_out) {
        if (d_dev) O.data = (unsigned char *)data_out;
        else { CK(h->d_out_data.ensure((n_occ + 1) * h->data_size)); O.data = (unsigned char *)h->d_out_data.p; }
    }
    CK(launch_records(h, mode, n_recs, O));
    if (n_recs) h->last_launches++;
    CK(cudaEventRecord(h->ev[4], h->stream));
    if (slots && !s_dev) CK(cudaMemcpyAsync(slots, O.slots, n_occ * sizeof(uint64_t), cudaMemcpyDeviceToHost, h->stream));
    if (flags && !f_dev) CK(cudaMemcpyAsync(flags, O.flags, n_occ, cudaMemcpyDeviceToHost, h->stream));
    if (data_out && !d_dev) CK(cudaMemcpyAsync(data_out, O.data, n_occ * h->data_size, cudaMemcpyDeviceToHost, h->stream));
    if (agg_out) CK(cudaMemcpyAsync(agg_out, h->counters + 5, 2 * sizeof(uint64_t), cudaMemcpyDeviceToHost, h->stream));
    if (mode != MODE_GET) {
        rc = settle(h);
        if (rc) return rc;
    } else {
        CK(cudaStreamSynchronize(h->stream));
    }
    finish_timing(h);
    return BRISK_OK;
}

extern "C" int brisk_query_reads(brisk_t *h, const char *bases, const uint64_t *offsets, uint64_t n_reads, uint64_t *hits,
                                 uint64_t *total) {
    uint64_t agg[2] = {0, 0};
    int rc = run_reads(h, bases, offsets, n_reads, MODE_GET, nullptr, nullptr, nullptr, agg);
    if (hits) *hits = agg[0];
    if (total) *total = agg[1];
    return rc;
}

extern "C" int brisk_count_reads(brisk_t *h, const char *bases, const uint64_t *offsets, uint64_t n_reads) {
    return run_reads(h, bases, offsets, n_reads, MODE_COUNT, nullptr, nullptr, nullptr);
}
extern "C" int brisk_insert_reads(brisk_t *h, const char *bases, const uint64_t *offsets, uint64_t n_reads, uint64_t *slots,
                                  uint8_t *newly) {
    return run_reads(h, bases, offsets, n_reads, MODE_INSERT, slots, newly, nullptr);
}
extern "C" int brisk_get_reads(brisk_t *h, const char *bases, const uint64_t *offsets, uint64_t n_reads, uint8_t *found,
                               void *data_out) {
    return run_reads(h, bases, offsets, n_reads, MODE_GET, nullptr, found, data_out);
}

extern "C" int brisk_enumerate_reads(brisk_t *h, const char *bases, const uint64_t *offsets, uint64_t n_reads, uint64_t *recs,
                                     uint64_t cap_recs, uint64_t *n_recs_out, uint32_t *rec_words) {
    if (!h) return fail(BRISK_E_PARAM, "null handle");
    std::lock_guard<std::mutex> lock(h->mu);
    CK(cudaSetDevice(h->device));
    h->last_launches = 0;
    if (rec_words) *rec_words = (uint32_t)h->P.rec_words;
    ReadBatch B;
    int rc = stage_reads(h, bases, offsets, n_reads, B);
    if (rc) return rc;
    uint64_t n_recs = 0, n_occ = 0;
    rc = enumerate_batch(h, B, false, n_recs, n_occ);
    if (rc) return rc;
    CK(cudaEventRecord(h->ev[2], h->stream));
    CK(cudaEventRecord(h->ev[3], h->stream));
    CK(cudaEventRecord(h->ev[4], h->stream));
    if (n_recs_out) *n_recs_out = n_recs;
    if (recs) {
        if (n_recs > cap_recs) return fail(BRISK_E_CAPACITY, "record buffer too small");
        CK(cudaMemcpyAsync(recs, h->d_recs.p, n_recs * h->P.rec_words * sizeof(uint64_t),
                           is_device_ptr(recs) ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, h->stream));
    }
    CK(cudaStreamSynchronize(h->stream));
    finish_timing(h);
    return BRISK_OK;
}

// ------------------------------------------------------------------------------------------------
static int run_kmers(brisk_handle *h, const brisk_kmer *kmers, uint64_t n, int mode, uint64_t *slots, uint8_t *flags,
                     void *data_out) {
    if (!h) return fail(BRISK_E_PARAM, "null handle");
    if (n && !kmers) return fail(BRISK_E_PARAM, "null input");
    std::lock_guard<std::mutex> lock(h->mu);
    CK(cudaSetDevice(h->device));
    h->last_launches = 0;
    if (n == 0) return BRISK_OK;
    const KmerIn *d_k;
    if (is_device_ptr(kmers)) d_k = (const KmerIn *)kmers;
    else {
        CK(h->d_kmers.ensure(n * sizeof(KmerIn)));
        CK(cudaMemcpyAsync(h->d_kmers.p, kmers, n * sizeof(KmerIn), cudaMemcpyHostToDevice, h->stream));
        d_k = (const KmerIn *)h->d_kmers.p;
    }
    if (mode != MODE_GET) {
        int rc = pregrow(h, n);
        if (rc) return rc;
    }
    bool s_dev = is_device_ptr(slots), f_dev = is_device_ptr(flags), d_dev = is_device_ptr(data_out);
    u64 *o_slots = nullptr;
    unsigned char *o_flags = nullptr, *o_data = nullptr;
    if (slots) { if (s_dev) o_slots = (u64 *)slots; else { CK(h->d_out_slots.ensure(n * 8)); o_slots = (u64 *)h->d_out_slots.p; } }
    if (flags) { if (f_dev) o_flags = flags; else { CK(h->d_out_flags.ensure(n)); o_flags = (unsigned char *)h->d_out_flags.p; } }
    if (data_out) { if (d_dev) o_data = (unsigned char *)data_out; else { CK(h->d_out_data.ensure(n * h->data_size)); o_data = (unsigned char *)h->d_out_data.p; } }
    for (int i = 0; i < 5; i++) CK(cudaEventRecord(h->ev[i], h->stream));
    unsigned blocks = (unsigned)std::min<uint64_t>((n + 255) / 256, 148 * 8 * 4);
    StoreView T = view_of(h);
    const int KW = h->P.key_words;
#define LAUNCH_K(KWV, MODEV) kmers_kernel<KWV, MODEV><<<blocks, 256, 0, h->stream>>>(h->P, d_k, n, T, o_slots, o_flags, o_data)
    if (KW == 1) { if (mode == MODE_COUNT) LAUNCH_K(1, MODE_COUNT); else if (mode == MODE_INSERT) LAUNCH_K(1, MODE_INSERT); else LAUNCH_K(1, MODE_GET); }
    else { if (mode == MODE_COUNT) LAUNCH_K(2, MODE_COUNT); else if (mode == MODE_INSERT) LAUNCH_K(2, MODE_INSERT); else LAUNCH_K(2, MODE_GET); }
#undef LAUNCH_K
    CK(cudaGetLastError());
    h->last_launches++;
    CK(cudaEventRecord(h->ev[4], h->stream));
    if (slots && !s_dev) CK(cudaMemcpyAsync(slots, o_slots, n * 8, cudaMemcpyDeviceToHost, h->stream));
    if (flags && !f_dev) CK(cudaMemcpyAsync(flags, o_flags, n, cudaMemcpyDeviceToHost, h->stream));
    if (data_out && !d_dev) CK(cudaMemcpyAsync(data_out, o_data, n * h->data_size, cudaMemcpyDeviceToHost, h->stream));
    if (mode != MODE_GET) {
        int rc = settle(h);
        if (rc) return rc;
    } else {
        CK(cudaStreamSynchronize(h->stream));
    }
    finish_timing(h);
    return BRISK_OK;
}

extern "C" int brisk_insert_kmers(brisk_t *h, const brisk_kmer *kmers, uint64_t n, uint64_t *slots, uint8_t *newly, int count) {
    return run_kmers(h, kmers, n, count ? MODE_COUNT : MODE_INSERT, slots, newly, nullptr);
}
extern "C" int brisk_get_kmers(brisk_t *h, const brisk_kmer *kmers, uint64_t n, uint64_t *slots, void *data_out) {
    return run_kmers(h, kmers, n, MODE_GET, slots, nullptr, data_out);
}

extern "C" void *brisk_data_ptr(brisk_t *h, uint64_t slot) {
    if (!h || slot == BRISK_NO_SLOT || slot >= h->cap + 64) return nullptr;
    if (h->data_size == 1) {
        if (slot >= h->cap) return (unsigned char *)(h->special + (slot - h->cap)) + 3;
        size_t off = slot * slot_bytes(h) + (h->P.key_words == 1 ? offsetof(Slot1, meta) : offsetof(Slot2, meta)) + 3;
        return (unsigned char *)h->slots + off;  // the DATA byte is the top byte of the little-endian meta word
    }
    return h->data + slot * h->data_size;
}

extern "C" int brisk_dump(brisk_t *h, brisk_kmer *keys, void *data, uint64_t cap, uint64_t *n) {
    if (!h) return fail(BRISK_E_PARAM, "null handle");
    std::lock_guard<std::mutex> lock(h->mu);
    CK(cudaSetDevice(h->device));
    CK(cudaStreamSynchronize(h->stream));
    if (n) *n = h->n_keys;
    if (!keys && !data) return BRISK_OK;
    if (cap < h->n_keys) return fail(BRISK_E_CAPACITY, "dump buffers too small");
    uint64_t nk = h->n_keys;
    bool k_dev = is_device_ptr(keys), d_dev = is_device_ptr(data);
    KmerIn *o_keys = nullptr;
    unsigned char *o_data = nullptr;
    if (keys) { if (k_dev) o_keys = (KmerIn *)keys; else { CK(h->d_kmers.ensure((nk + 1) * sizeof(KmerIn))); o_keys = (KmerIn *)h->d_kmers.p; } }
    if (data) { if (d_dev) o_data = (unsigned char *)data; else { CK(h->d_out_data.ensure((nk + 1) * h->data_size)); o_data = (unsigned char *)h->d_out_data.p; } }
    CK(cudaMemsetAsync(h->counters + 3, 0, sizeof(uint64_t), h->stream));
    unsigned blocks = (unsigned)std::min<uint64_t>((h->cap + 255) / 256, 148 * 16);
    if (h->P.key_words == 1) dump_kernel<1><<<blocks, 256, 0, h->stream>>>(view_of(h), o_keys, o_data, nk, 64);
    else dump_kernel<2><<<blocks, 256, 0, h->stream>>>(view_of(h), o_keys, o_data, nk, 64);
    CK(cudaGetLastError());
    uint64_t got = 0;
    CK(cudaMemcpyAsync(&got, h->counters + 3, sizeof(uint64_t), cudaMemcpyDeviceToHost, h->stream));
    if (keys && !k_dev) CK(cudaMemcpyAsync(keys, o_keys, nk * sizeof(KmerIn), cudaMemcpyDeviceToHost, h->stream));
    if (data && !d_dev) CK(cudaMemcpyAsync(data, o_data, nk * h->data_size, cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    if (got != nk) return fail(BRISK_E_CUDA, "dump found " + std::to_string(got) + " keys, expected " + std::to_string(nk));
    return BRISK_OK;
}

extern "C" int brisk_stats(brisk_t *h, uint64_t *nb_buckets, uint64_t *nb_skmers, uint64_t *nb_kmers, uint64_t *device_bytes,
                           uint64_t *largest_bucket) {
    if (!h) return fail(BRISK_E_PARAM, "null handle");
    std::lock_guard<std::mutex> lock(h->mu);
    CK(cudaSetDevice(h->device));
    uint64_t nb = 0;
    CK(cudaMemsetAsync(h->counters + 4, 0, sizeof(uint64_t), h->stream));
    unsigned blocks = (unsigned)std::min<uint64_t>((h->bitmap_words + 255) / 256, 148 * 16);
    popcount_kernel<<<blocks, 256, 0, h->stream>>>(h->bucket_bits, h->bitmap_words, h->counters + 4);
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(&nb, h->counters + 4, sizeof(uint64_t), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    if (nb_buckets) *nb_buckets = nb;
    if (nb_skmers) *nb_skmers = h->n_new_recs;
    if (nb_kmers) *nb_kmers = h->n_keys;
    if (largest_bucket) *largest_bucket = 0;
    if (device_bytes) {
        *device_bytes = h->cap * slot_bytes(h) + h->bitmap_words * 4 + h->ovf_cap * 24 +
                        (h->data ? (h->cap + 64) * h->data_size : 0);
    }
    return BRISK_OK;
}

extern "C" int brisk_reserve(brisk_t *h, uint64_t n_keys) {
    if (!h) return fail(BRISK_E_PARAM, "null handle");
    std::lock_guard<std::mutex> lock(h->mu);
    CK(cudaSetDevice(h->device));
    h->hint = std::max(h->hint, n_keys);
    int rc = grow_to(h, pow2ceil(std::max<uint64_t>(n_keys * 2, 1u << 16)));
    if (rc) return rc;
    CK(cudaStreamSynchronize(h->stream));
    return BRISK_OK;
}
extern "C" int brisk_capacity(brisk_t *h, uint64_t *slots) {
    if (!h || !slots) return fail(BRISK_E_PARAM, "null argument");
    *slots = h->cap;
    return BRISK_OK;
}
extern "C" int brisk_sync(brisk_t *h) {
    if (!h) return fail(BRISK_E_PARAM, "null handle");
    CK(cudaSetDevice(h->device));
    CK(cudaStreamSynchronize(h->stream));
    return BRISK_OK;
}
extern "C" int brisk_set_stream(brisk_t *h, void *stream) {
    if (!h) return fail(BRISK_E_PARAM, "null handle");
    std::lock_guard<std::mutex> lock(h->mu);
    CK(cudaSetDevice(h->device));
    CK(cudaStreamSynchronize(h->stream));
    h->stream = stream ? (cudaStream_t)stream : h->own_stream;
    return BRISK_OK;
}
extern "C" int brisk_last_timing(brisk_t *h, float ms[5], uint32_t *n_launches) {
    if (!h) return fail(BRISK_E_PARAM, "null handle");
    if (ms) memcpy(ms, h->last_ms, sizeof(h->last_ms));
    if (n_launches) *n_launches = h->last_launches;
    return BRISK_OK;
}

// ---- multi-GPU entry points (filled in by comm.cuh once the exchange step lands)
extern "C" uint32_t brisk_owner_of_bucket(uint64_t bucket_id, uint8_t b, uint32_t n_ranks) {
    // bucket ids are the middle bits of the hashed minimizer; the minimum of a window is biased towards small
    // hashes, so the id is scrambled before the range split to keep the shards even
    (void)b;
    uint64_t x = bucket_id * 0x9E3779B97F4A7C15ull;
    x ^= x >> 32;
    return (uint32_t)(((x & 0xffffffffull) * n_ranks) >> 32);
}
extern "C" int brisk_comm_unique_id(uint8_t id[128]) {
    (void)id;
    return fail(BRISK_E_COMM, "NCCL exchange not built yet");
}
extern "C" int brisk_comm_init(brisk_t *h, const uint8_t id[128], int rank, int n_ranks) {
    (void)h; (void)id; (void)rank; (void)n_ranks;
    return fail(BRISK_E_COMM, "NCCL exchange not built yet");
}
