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
#include "partition.cuh"
#include "fasta.cuh"
#include "store.cuh"
#include "sort.cuh"
#include "kff.cuh"
#include "reparam.cuh"
#include "comm.cuh"

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

#define BRISK_MAX_CHUNKS 8
#define BRISK_MIN_CHUNK_BYTES (8u << 20)

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
    bool use_ring = true;
    uint32_t long_limit = 2048;   // reads longer than this many bases go through the segmented launch (0: never)
    uint64_t seg_warm = 256;      // k-mers a segment warms up over before its first k-mer
    uint64_t seg_retries = 0;     // segmented launches repeated because neighbouring segments disagreed
    bool region_sort = true;
    cudaStream_t stream = nullptr;
    cudaStream_t own_stream = nullptr;
    cudaEvent_t ev[6] = {};
    cudaStream_t copy_stream = nullptr;
    cudaEvent_t chunk_ev[BRISK_MAX_CHUNKS] = {};
    cudaEvent_t offsets_ev = nullptr;
    cudaStream_t store_stream = nullptr;  // run_reads_overlapped: the store kernel of a batch's first chunks (low priority)
    cudaEvent_t ov_ev[2] = {};            // [0] the handle's stream is ready for that kernel, [1] the kernel is done
    bool pending_on_side = false;         // the pending store kernel runs on store_stream (its records are in d_send)
    bool pending_settle = false;  // a count / insert call left its store kernel running: settle() before the table is used again
    bool timing_dirty = false;    // stage times of the last call not read back yet
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
    uint32_t *cnt = nullptr;  // [cap + 64] counts beside the table (StoreView::cnt)
    u64 *overflow = nullptr;
    uint64_t ovf_cap = 0;
    uint64_t n_keys = 0, n_new_recs = 0;
    uint64_t pending_incoming = 0, keys_before = 0;  // the insert whose new-key rate settle() will measure
    double new_rate = 1.0;                           // new keys per incoming occurrence of the last insert
    // scratch
    DevBuf d_bases, d_offsets, d_packed, d_recs, d_rec_occ, d_kmer_base, d_cnt, d_out_slots, d_out_flags, d_out_data,
        d_kmers, d_cub, d_long;
    // partition / exchange (sharded index and the partition entry point)
    DevBuf d_fa_text, d_fa_bases, d_fa_offsets, d_fa_a, d_fa_b, d_fa_chunks;  // FASTA ingest
    DevBuf d_send, d_send_occ, d_recv, d_recv_occ, d_occ_off, d_ret_slots, d_ret_flags, d_ret_data;
    int2 *dec_tab = nullptr;  // grouped decycling weights (BriskParams::dec_tab)
    u64 *d_part2 = nullptr;  // [64] totals, [65] starts, [64] cursors of the table-region partition
    u64 *d_part = nullptr;  // [64] totals, [65] starts, [64] cursors, [64*64] all-gathered totals
    Comm *comm = nullptr;
    uint64_t *d_rec_count = nullptr;  // device counter
    std::recursive_mutex mu;  // recursive: the FASTA entry points hold it across parse + run_reads
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
    // fixed-point tables of the rolled decycling class (enumerate.cuh) and the band inside which the literal sum
    // decides.  Error of R, R(rot) in units of 2^-22: table rounding 0.5 per term and component in 2^-24 units, bit
    // planes of at most m terms, X = A +- 2B -> 1.5 m / 4; the rotation by (cos, sin) multiplies by at most sqrt(2);
    // rotation constants rounded at 2^-30 (|X| <= 3 m); two truncating high-multiplies: 2.
    for (int q = 0; q < 32; q++) {
        const double c = cos(unit * q), sn = sin(unit * q);
        P.dec_tw[q] = make_int2((int)llround(c * 16777216.0), (int)llround(sn * 16777216.0));
        P.dec_rot[q] = make_int2((int)llround(c * 1073741824.0), (int)llround(sn * 1073741824.0));
    }
    P.rot1_c30 = P.dec_rot[1].x; P.rot1_s30 = P.dec_rot[1].y;
    {
        const double err = 1.5 * m / 4.0 * 1.41422 + 3.0 * m / 256.0 + 2.0;
        P.dec_th = (int)ceil(P.eps * 4194304.0 + err) + 2;
    }
    P.polyA_hash = bfc_hash_literal(P.coef, P.eps, 0, P.m_mask, m);
    return BRISK_OK;
}

// grouped decycling tables (2^24 fixed point): entry (g, v) = sum_j code(nucleotide j of v) * w^(m-1-(3g+j)), positions >= m weigh 0
static std::vector<int2> decycle_group_table(int m) {
    const int groups = (m + 2) / 3;
    const double unit = 2 * M_PI / m;
    std::vector<int2> tab((size_t)groups * 64);
    for (int g = 0; g < groups; g++)
        for (int v = 0; v < 64; v++) {
            double sr = 0, si = 0;
            for (int j = 0; j < 3; j++) {
                const int pos = 3 * g + j, code = (v >> (2 * j)) & 3;
                if (pos >= m) break;
                sr += code * cos(unit * (m - 1 - pos));
                si += code * sin(unit * (m - 1 - pos));
            }
            tab[(size_t)g * 64 + v] = make_int2((int)llround(sr * 16777216.0), (int)llround(si * 16777216.0));
        }
    return tab;
}

static StoreView view_of(brisk_handle *h) {
    StoreView T;
    T.slots = h->slots;
    T.mask = h->cap - 1;
    T.special = h->special;
    T.cnt = h->cnt;
    T.counters = h->counters;
    T.bucket_bits = h->bucket_bits;
    T.data = h->data;
    T.data_size = h->data_size;
    T.overflow = h->overflow;
    T.ovf_cap = h->ovf_cap;
    return T;
}

static size_t slot_bytes(const brisk_handle *h) { return h->P.key_words == 1 ? sizeof(Slot1) : sizeof(Slot2); }

static int alloc_table(brisk_handle *h, uint64_t cap, void **slots, unsigned char **data, uint32_t **cnt) {
    size_t bytes = cap * slot_bytes(h);
    if (h->flags & BRISK_F_MANAGED) CK(cudaMallocManaged(slots, bytes));
    else CK(cudaMalloc(slots, bytes));
    int blocks = (int)std::min<uint64_t>((cap + 255) / 256, 148 * 16);
    if (h->P.key_words == 1) init_table_kernel<1><<<blocks, 256, 0, h->stream>>>(*slots, cap);
    else init_table_kernel<2><<<blocks, 256, 0, h->stream>>>(*slots, cap);
    CK(cudaGetLastError());
    {
        size_t cb = (cap + 64) * sizeof(uint32_t);
        if (h->flags & BRISK_F_MANAGED) CK(cudaMallocManaged((void **)cnt, cb));
        else CK(cudaMalloc((void **)cnt, cb));
        CK(cudaMemsetAsync(*cnt, 0, cb, h->stream));
    }
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
    uint32_t *nc = nullptr;
    int rc = alloc_table(h, new_cap, &ns, &nd, &nc);
    if (rc) return rc;
    if (h->slots) {
        StoreView from = view_of(h);
        StoreView to = from;
        to.slots = ns; to.mask = new_cap - 1; to.data = nd; to.cnt = nc;
        int blocks = (int)std::min<uint64_t>((h->cap + 255) / 256, 148 * 16);
        if (h->P.key_words == 1) rehash_kernel<1><<<blocks, 256, 0, h->stream>>>(h->P, from, to);
        else rehash_kernel<2><<<blocks, 256, 0, h->stream>>>(h->P, from, to);
        CK(cudaGetLastError());
        // counts (and DATA) of the special key follow the arena's end
        CK(cudaMemcpyAsync(nc + new_cap, h->cnt + h->cap, 64 * sizeof(uint32_t), cudaMemcpyDeviceToDevice, h->stream));
        if (h->data) {
            CK(cudaMemcpyAsync(nd + new_cap * h->data_size, h->data + h->cap * h->data_size, 64 * (size_t)h->data_size,
                               cudaMemcpyDeviceToDevice, h->stream));
        }
        CK(cudaStreamSynchronize(h->stream));
        cudaFree(h->slots);
        cudaFree(h->cnt);
        if (h->data) cudaFree(h->data);
    }
    h->slots = ns;
    h->data = nd;
    h->cnt = nc;
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
    h->use_ring = k <= 32;
    if (const char *e = getenv("BRISK_ENUM_RING")) h->use_ring = atoi(e) != 0;  // development switch
    if (const char *e = getenv("BRISK_REGION_SORT")) h->region_sort = atoi(e) != 0;  // development switch
    if (const char *e = getenv("BRISK_LONG_READ")) h->long_limit = (uint32_t)std::max(0, atoi(e));  // tests: segment shorter reads too
    if (h->long_limit && h->long_limit < (uint32_t)k) h->long_limit = k;
    if (const char *e = getenv("BRISK_SEG_WARM")) h->seg_warm = (uint64_t)std::max(1, atoi(e));
    CK(cudaStreamCreateWithFlags(&h->own_stream, cudaStreamNonBlocking));
    h->stream = h->own_stream;
    for (int i = 0; i < 6; i++) CK(cudaEventCreate(&h->ev[i]));
    CK(cudaStreamCreateWithFlags(&h->copy_stream, cudaStreamNonBlocking));
    for (int i = 0; i < BRISK_MAX_CHUNKS; i++) CK(cudaEventCreateWithFlags(&h->chunk_ev[i], cudaEventDisableTiming));
    CK(cudaEventCreateWithFlags(&h->offsets_ev, cudaEventDisableTiming));
    CK(cudaMalloc((void **)&h->counters, 8 * sizeof(uint64_t)));
    CK(cudaMemsetAsync(h->counters, 0, 8 * sizeof(uint64_t), h->stream));
    CK(cudaMalloc((void **)&h->d_rec_count, 32 * sizeof(uint64_t)));  // [0]: the batch's record counter; [16 + c]: chunk c of a pipelined exchange
    {
        std::vector<int2> tab = decycle_group_table(m);
        CK(cudaMalloc((void **)&h->dec_tab, tab.size() * sizeof(int2)));
        CK(cudaMemcpy(h->dec_tab, tab.data(), tab.size() * sizeof(int2), cudaMemcpyHostToDevice));
        h->P.dec_tab = h->dec_tab;
        h->P.dec_groups = (m + 2) / 3;
    }
    CK(cudaMalloc((void **)&h->d_part2, (64 + 65 + 64 + 7) * sizeof(uint64_t)));
    CK(cudaMalloc((void **)&h->d_part, (64 + 65 + 64 + 64 * 64 + 7) * sizeof(uint64_t)));
    if (flags & BRISK_F_MANAGED) CK(cudaMallocManaged((void **)&h->special, 64 * sizeof(uint32_t)));
    else CK(cudaMalloc((void **)&h->special, 64 * sizeof(uint32_t)));
    CK(cudaMemsetAsync(h->special, 0, 64 * sizeof(uint32_t), h->stream));
    h->bitmap_words = (((uint64_t)1 << (2 * b)) + 31) / 32;
    CK(cudaMalloc((void **)&h->bucket_bits, h->bitmap_words * sizeof(uint32_t)));
    CK(cudaMemsetAsync(h->bucket_bits, 0, h->bitmap_words * sizeof(uint32_t), h->stream));
    h->ovf_cap = 1u << 20;
    CK(cudaMalloc((void **)&h->overflow, h->ovf_cap * 3 * sizeof(uint64_t)));
    // Table size: a load factor of about 0.25 at capacity_hint keys when the device has the memory for it (the store
    // kernels are bound by 32-byte sector traffic, i.e. by the number of probes per k-mer: 1.33 at load 0.5, 1.1 at
    // 0.25; on cfg 3 the count kernel goes from 2.7 to 2.1 ms and the get kernel from 2.5 to 2.1 ms), else 0.5.
    uint64_t cap = pow2ceil(std::max<uint64_t>(capacity_hint * 2, capacity_hint ? 1u << 10 : 1u << 16));
    if (capacity_hint) {
        size_t free_b = 0, total_b = 0;
        if (cudaMemGetInfo(&free_b, &total_b) == cudaSuccess) {
            const uint64_t per_slot = (P.key_words == 1 ? sizeof(Slot1) : sizeof(Slot2)) + sizeof(uint32_t) + (data_size > 1 ? data_size : 0);
            const uint64_t wide = pow2ceil(capacity_hint * 4);
            if (wide > cap && wide * per_slot <= free_b / 4) cap = wide;
        } else cudaGetLastError();
    }
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
    if (h->store_stream) cudaStreamSynchronize(h->store_stream);
    DevBuf *bufs[] = {&h->d_bases, &h->d_offsets, &h->d_packed, &h->d_recs, &h->d_rec_occ, &h->d_kmer_base, &h->d_cnt,
                      &h->d_out_slots, &h->d_out_flags, &h->d_out_data, &h->d_kmers, &h->d_cub, &h->d_long, &h->d_send, &h->d_send_occ,
                      &h->d_recv, &h->d_recv_occ, &h->d_occ_off, &h->d_ret_slots, &h->d_ret_flags, &h->d_ret_data,
                      &h->d_fa_text, &h->d_fa_bases, &h->d_fa_offsets, &h->d_fa_a, &h->d_fa_b, &h->d_fa_chunks};
    for (DevBuf *b : bufs) b->release();
    if (h->comm) {
        if (h->comm->comm) g_nccl.CommDestroy(h->comm->comm);
        if (h->comm->stream) cudaStreamDestroy(h->comm->stream);
        for (int i = 0; i < 2; i++) if (h->comm->aux[i]) cudaStreamDestroy(h->comm->aux[i]);
        for (int i = 0; i < COMM_MAX_CHUNKS; i++) if (h->comm->ev_enum[i]) cudaEventDestroy(h->comm->ev_enum[i]);
        if (h->comm->ev_counts) cudaEventDestroy(h->comm->ev_counts);
        if (h->comm->ev_done) cudaEventDestroy(h->comm->ev_done);
        if (h->comm->h_pinned) cudaFreeHost(h->comm->h_pinned);
        cudaFree(h->comm->d_gather);
        delete h->comm;
    }
    cudaFree(h->d_part);
    cudaFree(h->d_part2);
    cudaFree(h->dec_tab);
    cudaFree(h->slots);
    cudaFree(h->cnt);
    cudaFree(h->data);
    cudaFree(h->special);
    cudaFree(h->counters);
    cudaFree(h->bucket_bits);
    cudaFree(h->overflow);
    cudaFree(h->d_rec_count);
    for (int i = 0; i < 6; i++) cudaEventDestroy(h->ev[i]);
    for (int i = 0; i < BRISK_MAX_CHUNKS; i++) cudaEventDestroy(h->chunk_ev[i]);
    if (h->offsets_ev) cudaEventDestroy(h->offsets_ev);
    cudaStreamDestroy(h->copy_stream);
    if (h->store_stream) cudaStreamDestroy(h->store_stream);
    for (int i = 0; i < 2; i++) if (h->ov_ev[i]) cudaEventDestroy(h->ov_ev[i]);
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
    // host input: the bases arrive in chunks on the copy stream, chunk c = reads [chunk_first[c], chunk_first[c + 1]),
    // ready when h->chunk_ev[c] has fired; pack + enumerate of chunk c overlap the copy of chunk c + 1
    std::vector<uint64_t> chunk_first, chunk_byte;  // chunk_byte[c] = byte offset of the chunk's first read
    bool chunk_events = false;
};


// copies (if needed) and packs a batch of reads
static int stage_reads(brisk_handle *h, const char *bases, const uint64_t *offsets, uint64_t n_reads, ReadBatch &B) {
    B.n_reads = n_reads;
    uint64_t total = 0;
    static const uint64_t zero_offsets[1] = {0};
    if (n_reads == 0) { offsets = zero_offsets; bases = nullptr; }  // a rank of a sharded index may have no reads: (NULL, NULL, 0) is allowed
    std::vector<uint64_t> h_off_copy;
    if (n_reads && is_device_ptr(offsets) && !is_device_ptr(bases)) {
        // bases on the host, offsets on the device: the chunk planner reads the offsets on the host
        h_off_copy.resize(n_reads + 1);
        CK(cudaMemcpyAsync(h_off_copy.data(), offsets, (n_reads + 1) * sizeof(uint64_t), cudaMemcpyDeviceToHost, h->stream));
        CK(cudaStreamSynchronize(h->stream));
        offsets = h_off_copy.data();
    }
    if (is_device_ptr(offsets)) {
        CK(cudaMemcpyAsync(&total, offsets + n_reads, sizeof(uint64_t), cudaMemcpyDeviceToHost, h->stream));
        CK(cudaStreamSynchronize(h->stream));
        B.d_offsets = (const u64 *)offsets;
    } else {
        total = offsets[n_reads];
        CK(h->d_offsets.ensure((n_reads + 1) * sizeof(uint64_t)));
        // on the copy stream, ahead of the bases: the handle's stream may still be busy with the previous call's store
        // kernel (which does not read the offsets), and a copy queued behind it would reach the copy engine after all
        // the chunk copies of this call
        CK(cudaMemcpyAsync(h->d_offsets.p, offsets, (n_reads + 1) * sizeof(uint64_t), cudaMemcpyHostToDevice, h->copy_stream));
        CK(cudaEventRecord(h->offsets_ev, h->copy_stream));
        CK(cudaStreamWaitEvent(h->stream, h->offsets_ev, 0));
        B.d_offsets = (const u64 *)h->d_offsets.p;
    }
    B.total_bytes = total;
    B.chunk_first.assign(1, 0);
    if (is_device_ptr(bases)) {
        B.d_bases = (const unsigned char *)bases;
        B.chunk_first.push_back(n_reads);
        B.chunk_byte = {0, total};
    } else {
        CK(h->d_bases.ensure(total + 16));
        B.d_bases = (const unsigned char *)h->d_bases.p;
        // chunks of about total / 8 bytes on read boundaries (the host offsets are at hand), copied on the copy stream
        // (the last of 8 chunks is half the size of the others: what follows the last byte's arrival is its pack + enumerate)
        const uint64_t want = std::max<uint64_t>(BRISK_MIN_CHUNK_BYTES, (2 * total + 2 * BRISK_MAX_CHUNKS - 2) / (2 * BRISK_MAX_CHUNKS - 1));
        uint64_t first = 0;
        while (first < n_reads && B.chunk_first.size() < BRISK_MAX_CHUNKS + 1) {  // n chunks = n + 1 boundaries
            // last read whose end stays within `want` bytes of the chunk's start
            const uint64_t limit = offsets[first] + want;
            uint64_t lo = first + 1, hi = n_reads;
            while (lo < hi) {
                uint64_t mid = (lo + hi + 1) / 2;
                if (offsets[mid] <= limit) lo = mid; else hi = mid - 1;
            }
            first = lo;
            B.chunk_first.push_back(first);
        }
        if (B.chunk_first.back() != n_reads) B.chunk_first.back() = n_reads;
        if (B.chunk_first.size() == 1) B.chunk_first.push_back(n_reads);
        B.chunk_events = true;
        for (uint64_t f : B.chunk_first) B.chunk_byte.push_back(offsets[f]);
        for (size_t c = 0; c + 1 < B.chunk_first.size(); c++) {
            const uint64_t b0 = offsets[B.chunk_first[c]], b1 = offsets[B.chunk_first[c + 1]];
            if (b1 > b0) CK(cudaMemcpyAsync((char *)h->d_bases.p + b0, bases + b0, b1 - b0, cudaMemcpyHostToDevice, h->copy_stream));
            CK(cudaEventRecord(h->chunk_ev[c], h->copy_stream));
        }
    }
    uint64_t n_words = total / 32 + n_reads + 8;
    CK(h->d_packed.ensure(n_words * sizeof(uint64_t)));
    B.d_packed = (u64 *)h->d_packed.p;
    CK(cudaEventRecord(h->ev[0], h->stream));
    // tail words that get_bits64 may touch past the last read must be readable; zero the pad
    if (n_reads) CK(cudaMemsetAsync(B.d_packed + (n_words - 8), 0, 8 * sizeof(uint64_t), h->stream));
    return BRISK_OK;
}

// 2-bit packing of chunk c of a staged batch (after its bases have arrived)
static int pack_chunk(brisk_handle *h, const ReadBatch &B, size_t c) {
    const uint64_t r0 = B.chunk_first[c], r1 = B.chunk_first[c + 1];
    if (B.chunk_events) CK(cudaStreamWaitEvent(h->stream, h->chunk_ev[c], 0));
    if (r1 > r0) {
        unsigned blocks = (unsigned)((r1 - r0 + PACK_READS_PER_CTA - 1) / PACK_READS_PER_CTA);
        pack_reads_kernel<<<blocks, 256, 0, h->stream>>>(B.d_bases, B.d_offsets, r0, r1, B.d_packed);
        CK(cudaGetLastError());
        h->last_launches++;
    }
    return BRISK_OK;
}

static EnumLong no_long() { EnumLong l; l.list = nullptr; l.cap = 0; l.count = nullptr; l.limit = 0; return l; }
static EnumSegIn no_segs() { EnumSegIn g; g.segs = nullptr; g.n_segs = 0; g.at0 = nullptr; g.at1 = nullptr; return g; }

template <int KT, int MT, int RW, bool RING, bool LIT, bool SEG>
static cudaError_t launch_enum_t(brisk_handle *h, const ReadBatch &B, uint64_t r0, uint64_t r1, EnumOut out, cudaStream_t st, EnumLong lg,
                                 EnumSegIn sg) {
    typedef typename KM<KT, MT>::hash_t hash_t;
    size_t smem = (RING ? ((size_t)h->P.w * ENUM_THREADS * sizeof(hash_t) + 7) / 8 * 8 : 0) + (ENUM_THREADS / 32) * ENUM_WPOOL * sizeof(u64) +
                  4 * h->P.m * sizeof(double) + 64 * (size_t)h->P.dec_groups * sizeof(int2);
    cudaError_t e = cudaFuncSetAttribute(enumerate_kernel<KT, MT, RW, RING, LIT, SEG>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    // one thread per read of [r0, r1), or (SEG) per segment of [0, n_segs)
    const uint64_t first = SEG ? 0 : r0, end = SEG ? sg.n_segs : r1;
    unsigned blocks = (unsigned)((end - first + ENUM_THREADS - 1) / ENUM_THREADS);
    enumerate_kernel<KT, MT, RW, RING, LIT, SEG><<<blocks, ENUM_THREADS, smem, st>>>(h->P, B.d_packed, B.d_offsets, first, end, out, lg, sg);
    return cudaGetLastError();
}
// generic (k, m): run-time parameters; the literal decycling mode (a debug / verification switch) and the segmented
// launch for long reads only exist here
template <int RW, bool SEG>
static cudaError_t launch_enum_generic(brisk_handle *h, const ReadBatch &B, uint64_t r0, uint64_t r1, EnumOut out, cudaStream_t st, EnumLong lg,
                                       EnumSegIn sg) {
    // k > 32: most window positions of a rescan are recomputed anyway, so the ring only costs shared memory
    if (h->P.literal_decycle)
        return h->use_ring ? launch_enum_t<0, 0, RW, true, true, SEG>(h, B, r0, r1, out, st, lg, sg)
                           : launch_enum_t<0, 0, RW, false, true, SEG>(h, B, r0, r1, out, st, lg, sg);
    return h->use_ring ? launch_enum_t<0, 0, RW, true, false, SEG>(h, B, r0, r1, out, st, lg, sg)
                       : launch_enum_t<0, 0, RW, false, false, SEG>(h, B, r0, r1, out, st, lg, sg);
}
static cudaError_t launch_enum(brisk_handle *h, const ReadBatch &B, uint64_t r0, uint64_t r1, EnumOut out, cudaStream_t st = nullptr,
                               EnumLong lg = no_long()) {
    if (!st) st = h->stream;
    const int k = h->P.k, m = h->P.m, RW = h->P.rec_words;
    const EnumSegIn sg = no_segs();
    if (!h->P.literal_decycle && !(h->flags & BRISK_F_GENERIC_ENUM)) {
        // the parameter sets of BASELINE.json's configurations: k and m compiled in
        if (k == 63 && m == 21 && RW == 4 && !h->use_ring) return launch_enum_t<63, 21, 4, false, false, false>(h, B, r0, r1, out, st, lg, sg);
        if (k == 64 && m == 23 && RW == 4 && !h->use_ring) return launch_enum_t<64, 23, 4, false, false, false>(h, B, r0, r1, out, st, lg, sg);
        if (k == 31 && m == 15 && RW == 2 && h->use_ring) return launch_enum_t<31, 15, 2, true, false, false>(h, B, r0, r1, out, st, lg, sg);
        if (k == 31 && m == 19 && RW == 2 && h->use_ring) return launch_enum_t<31, 19, 2, true, false, false>(h, B, r0, r1, out, st, lg, sg);
    }
    return RW == 2 ? launch_enum_generic<2, false>(h, B, r0, r1, out, st, lg, sg)
                   : RW == 4 ? launch_enum_generic<4, false>(h, B, r0, r1, out, st, lg, sg) : launch_enum_generic<6, false>(h, B, r0, r1, out, st, lg, sg);
}
static cudaError_t launch_enum_segments(brisk_handle *h, const ReadBatch &B, EnumOut out, EnumSegIn sg) {
    const int RW = h->P.rec_words;
    return RW == 2 ? launch_enum_generic<2, true>(h, B, 0, 0, out, h->stream, no_long(), sg)
                   : RW == 4 ? launch_enum_generic<4, true>(h, B, 0, 0, out, h->stream, no_long(), sg)
                             : launch_enum_generic<6, true>(h, B, 0, 0, out, h->stream, no_long(), sg);
}

__global__ void kmer_counts_kernel(const u64 *offsets, u64 n_reads, int k, u64 *counts) {
    u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n_reads) {
        u64 len = offsets[i + 1] - offsets[i];
        counts[i] = len >= (u64)k ? len - k + 1 : 0;
    }
}

// The long reads of a batch (listed in h->d_long by the plain launch), cut into segments of ENUM_SEG_KMERS k-mers
// (enumerate.cuh: EnumSeg).  A segment warms up from a rescan `warm` k-mers before its first k-mer; the neighbours'
// states at the segment boundaries are compared, and when any pair differs the long reads are enumerated again with an
// 8 times longer warm-up, finally with one thread per read (no warm-up at all: exact by construction).  The records go
// behind the batch's record counter; n_recs is updated.
static int enumerate_long_reads(brisk_handle *h, const ReadBatch &B, EnumOut out, uint64_t n_long, uint64_t &n_recs) {
    const uint64_t short_recs = n_recs;
    u64 *list = (u64 *)h->d_long.p;
    u64 *mism = (u64 *)h->d_rec_count + 9;
    DevBuf nseg, sstart, segs, at0, at1, scan_tmp;
    auto cleanup = [&]() { for (DevBuf *b : {&nseg, &sstart, &segs, &at0, &at1, &scan_tmp}) b->release(); };
    auto body = [&]() -> int {
        CK(nseg.ensure((n_long + 1) * sizeof(u64)));
        CK(sstart.ensure((n_long + 1) * sizeof(u64)));
        const unsigned lb = (unsigned)((n_long + 1 + 255) / 256);
        enumerate_seg_count_kernel<<<lb, 256, 0, h->stream>>>(B.d_offsets, list, n_long, h->P.k, (u64 *)nseg.p);
        size_t tmp = 0;
        CK(cub::DeviceScan::ExclusiveSum(nullptr, tmp, (u64 *)nseg.p, (u64 *)sstart.p, (int)(n_long + 1), h->stream));
        CK(scan_tmp.ensure(tmp));
        CK(cub::DeviceScan::ExclusiveSum(scan_tmp.p, tmp, (u64 *)nseg.p, (u64 *)sstart.p, (int)(n_long + 1), h->stream));
        uint64_t n_segs = 0;
        CK(cudaMemcpyAsync(&n_segs, (u64 *)sstart.p + n_long, sizeof(u64), cudaMemcpyDeviceToHost, h->stream));
        CK(cudaStreamSynchronize(h->stream));
        h->last_launches += 3;
        if (n_segs == 0) return BRISK_OK;
        CK(segs.ensure(n_segs * sizeof(EnumSeg)));
        CK(at0.ensure(n_segs * sizeof(EnumSegState)));
        CK(at1.ensure(n_segs * sizeof(EnumSegState)));
        uint64_t warm = h->seg_warm;
        for (int attempt = 0;; attempt++) {
            const bool serial = attempt == 3;  // last resort: every long read is one segment
            EnumSegIn sg;
            sg.segs = (const EnumSeg *)segs.p; sg.at0 = (EnumSegState *)at0.p; sg.at1 = (EnumSegState *)at1.p;
            if (!serial) {
                enumerate_seg_fill_kernel<<<(unsigned)((n_long + 255) / 256), 256, 0, h->stream>>>(B.d_offsets, list, n_long, h->P.k, warm, (const u64 *)sstart.p,
                                                                                                  (EnumSeg *)segs.p);
                sg.n_segs = n_segs;
            } else {
                // segment size "infinite": reuse the fill kernel's layout with one entry per read
                std::vector<uint64_t> hl(n_long), ho(B.n_reads + 1);
                CK(cudaMemcpy(hl.data(), list, n_long * sizeof(u64), cudaMemcpyDeviceToHost));
                CK(cudaMemcpy(ho.data(), B.d_offsets, (B.n_reads + 1) * sizeof(u64), cudaMemcpyDeviceToHost));
                std::vector<EnumSeg> hs;
                for (uint64_t r : hl) {
                    const uint64_t len = ho[r + 1] - ho[r];
                    if (len < (uint64_t)h->P.k) continue;
                    EnumSeg g; g.read = r; g.kmer0 = 0; g.kmer1 = len - h->P.k + 1; g.start = 0;
                    hs.push_back(g);
                }
                CK(cudaMemcpy(segs.p, hs.data(), hs.size() * sizeof(EnumSeg), cudaMemcpyHostToDevice));
                sg.n_segs = hs.size();
            }
            CK(cudaMemsetAsync(mism, 0, sizeof(u64), h->stream));
            CK(cudaMemcpyAsync(h->d_rec_count, &short_recs, sizeof(u64), cudaMemcpyHostToDevice, h->stream));
            CK(launch_enum_segments(h, B, out, sg));
            if (!serial) enumerate_seg_verify_kernel<<<(unsigned)((sg.n_segs + 255) / 256), 256, 0, h->stream>>>(sg.segs, sg.at0, sg.at1, sg.n_segs, mism);
            CK(cudaGetLastError());
            h->last_launches += 3;
            uint64_t bad = 0;
            CK(cudaMemcpyAsync(&n_recs, h->d_rec_count, sizeof(u64), cudaMemcpyDeviceToHost, h->stream));
            CK(cudaMemcpyAsync(&bad, mism, sizeof(u64), cudaMemcpyDeviceToHost, h->stream));
            CK(cudaStreamSynchronize(h->stream));
            h->seg_retries += bad ? 1 : 0;
            if (!bad || serial) return BRISK_OK;
            warm *= 8;
        }
    };
    int rc = body();
    cleanup();
    return rc;
}

// enumerates a staged batch into h->d_recs; with_occ also fills d_rec_occ / d_kmer_base
static int enumerate_batch(brisk_handle *h, const ReadBatch &B, bool with_occ, uint64_t &n_recs, uint64_t &n_occ,
                           size_t c0 = 0, size_t c1 = ~(size_t)0) {
    n_recs = 0;
    if (c0 == 0) n_occ = 0;  // the occurrence index of the whole batch is built with its first chunk
    c1 = std::min(c1, B.chunk_first.size() - 1);  // chunks [c0, c1) of the batch (default: all of it)
    if (B.n_reads == 0) {
        CK(cudaEventRecord(h->ev[1], h->stream));
        return BRISK_OK;
    }
    const int RW = h->P.rec_words;
    if (with_occ && c0 == 0) {
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
    uint64_t upper = B.chunk_byte[c1] - B.chunk_byte[c0];  // >= number of k-mer occurrences
    uint64_t guess = (uint64_t)(3.0 * (double)upper / (h->P.w + 1)) + (B.chunk_first[c1] - B.chunk_first[c0]) + 4096;
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
        // reads longer than long_limit bases are left out by the plain launch and listed (enumerate.cuh: EnumLong)
        EnumLong lg = no_long();
        if (h->long_limit) {
            CK(h->d_long.ensure((B.n_reads + 1) * sizeof(uint64_t)));
            lg.list = (u64 *)h->d_long.p; lg.cap = B.n_reads; lg.count = (u64 *)h->d_rec_count + 8; lg.limit = h->long_limit;
            CK(cudaMemsetAsync(lg.count, 0, sizeof(uint64_t), h->stream));
        }
        // chunk by chunk: pack + enumerate of a chunk run while the next chunk's bases are still on their way; the
        // records of all chunks accumulate behind the same device counter
        for (size_t c = c0; c < c1; c++) {
            int prc = pack_chunk(h, B, c);
            if (prc) return prc;
            if (c == 0) CK(cudaEventRecord(h->ev[1], h->stream));
            const uint64_t r0 = B.chunk_first[c], r1 = B.chunk_first[c + 1];
            if (r1 == r0) continue;
            cudaError_t e = launch_enum(h, B, r0, r1, out, nullptr, lg);
            CK(e);
            h->last_launches++;
        }
        uint64_t n_long = 0;
        CK(cudaMemcpyAsync(&n_recs, h->d_rec_count, sizeof(uint64_t), cudaMemcpyDeviceToHost, h->stream));
        if (lg.list) CK(cudaMemcpyAsync(&n_long, lg.count, sizeof(uint64_t), cudaMemcpyDeviceToHost, h->stream));
        CK(cudaStreamSynchronize(h->stream));
        if (n_long) {
            int lrc = enumerate_long_reads(h, B, out, n_long, n_recs);
            if (lrc) return lrc;
        }
        if (n_recs <= cap) break;
        cap = n_recs;
        if (attempt == 1) return fail(BRISK_E_CUDA, "superkmer count changed between two enumerations");
    }
    return BRISK_OK;
}

template <int RW, int KW, int MODE>
static cudaError_t launch_records_t(brisk_handle *h, const u64 *recs, uint64_t n_recs, OccOut O, cudaStream_t st, int ctas_per_sm) {
    if (n_recs == 0) return cudaSuccess;
    uint64_t want = (n_recs + 255) / 256;  // one warp per 32 records
    // persistent grid: exactly the CTAs that are resident at once (B200: 148 SMs), every warp strides over the records
    static int per_sm = 0;
    if (per_sm == 0) {
        int v = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&v, records_kernel<RW, KW, MODE>, 256, 0) != cudaSuccess || v < 1) v = 4;
        if (const char *e = getenv("BRISK_STORE_CTAS")) v = std::max(1, atoi(e));
        per_sm = v;
    }
    unsigned blocks = (unsigned)std::min<uint64_t>(want, (uint64_t)148 * (ctas_per_sm > 0 ? std::min(ctas_per_sm, per_sm) : per_sm));
    records_kernel<RW, KW, MODE><<<blocks, 256, 0, st ? st : h->stream>>>(h->P, recs, n_recs, view_of(h), O);
    return cudaGetLastError();
}
template <int RW, int KW>
static cudaError_t launch_records_m(brisk_handle *h, int mode, const u64 *recs, uint64_t n_recs, OccOut O, cudaStream_t st, int ctas_per_sm) {
    switch (mode) {
        case MODE_COUNT: return launch_records_t<RW, KW, MODE_COUNT>(h, recs, n_recs, O, st, ctas_per_sm);
        case MODE_INSERT: return launch_records_t<RW, KW, MODE_INSERT>(h, recs, n_recs, O, st, ctas_per_sm);
        default: return launch_records_t<RW, KW, MODE_GET>(h, recs, n_recs, O, st, ctas_per_sm);
    }
}
// st = nullptr: the handle's stream; ctas_per_sm > 0 caps the persistent grid's residency
static cudaError_t launch_records(brisk_handle *h, int mode, const u64 *recs, uint64_t n_recs, OccOut O, cudaStream_t st = nullptr, int ctas_per_sm = 0) {
    const int RW = h->P.rec_words, KW = h->P.key_words;
    if (KW == 1) {
        if (RW == 2) return launch_records_m<2, 1>(h, mode, recs, n_recs, O, st, ctas_per_sm);
        if (RW == 4) return launch_records_m<4, 1>(h, mode, recs, n_recs, O, st, ctas_per_sm);
        return launch_records_m<6, 1>(h, mode, recs, n_recs, O, st, ctas_per_sm);
    }
    if (RW == 2) return launch_records_m<2, 2>(h, mode, recs, n_recs, O, st, ctas_per_sm);
    if (RW == 4) return launch_records_m<4, 2>(h, mode, recs, n_recs, O, st, ctas_per_sm);
    return launch_records_m<6, 2>(h, mode, recs, n_recs, O, st, ctas_per_sm);
}

// ------------------------------------------------------------------------------------------------ partition / exchange
#define NCK(call)                                                                                                     \
    do {                                                                                                              \
        ncclResult_t r_ = (call);                                                                                     \
        if (r_ != ncclSuccess)                                                                                        \
            return fail(BRISK_E_COMM, std::string(#call) + ": " + g_nccl.GetErrorString(r_) + " (" + __FILE__ + ":" + \
                                          std::to_string(__LINE__) + ")");                                            \
    } while (0)

static u64 *part_totals(brisk_handle *h) { return h->d_part; }
static u64 *part_starts(brisk_handle *h) { return h->d_part + 64; }
static u64 *part_cursors(brisk_handle *h) { return h->d_part + 64 + 65; }
static u64 *part_matrix(brisk_handle *h) { return h->d_part + 64 + 65 + 64; }

template <int RW>
static cudaError_t launch_partition(brisk_handle *h, u64 *scratch, const u64 *recs, const u64 *occ, uint64_t n_recs, uint32_t n_dest, int mode,
                                    u64 *dst, u64 *dst_occ, cudaStream_t stream = nullptr) {
    if (!stream) stream = h->stream;
    const int shift = 64 - 2 * h->P.b;
    u64 *totals = scratch, *starts = scratch + 64, *cursors = scratch + 64 + 65;
    unsigned b1 = (unsigned)std::min<uint64_t>((n_recs + PART_THREADS - 1) / PART_THREADS, 148 * 8);
    partition_count_kernel<RW><<<b1, PART_THREADS, 0, stream>>>(recs, n_recs, shift, n_dest, mode, totals);
    partition_starts_kernel<<<1, 32, 0, stream>>>(totals, n_dest, starts, cursors);
    unsigned b2 = (unsigned)std::min<uint64_t>((n_recs + PART_TILE - 1) / PART_TILE, 148 * 8);
    partition_scatter_kernel<RW><<<b2, PART_THREADS, 0, stream>>>(recs, n_recs, shift, n_dest, mode, cursors, dst, occ, occ ? dst_occ : nullptr);
    return cudaGetLastError();
}

// Radix partition of n_recs records (and, when occ is given, their occurrence tags) into one contiguous region per
// destination.  mode 0: destination = owner rank (sharded index, before the exchange); mode 1: destination = table
// region (STORE_REGIONS), right before the store kernel.  The per-destination totals and region starts stay on the
// device (part_totals / part_starts).
static int partition_records(brisk_handle *h, const u64 *recs, const u64 *occ, uint64_t n_recs, uint32_t n_dest, int mode, DevBuf &dst,
                             DevBuf &dst_occ) {
    const int RW = h->P.rec_words;
    u64 *scratch = mode == 0 ? h->d_part : h->d_part2;  // the rank partition's starts are still needed on the way back
    CK(cudaMemsetAsync(scratch, 0, 64 * sizeof(uint64_t), h->stream));
    CK(dst.ensure((n_recs + 1) * RW * sizeof(uint64_t)));
    if (occ) CK(dst_occ.ensure((n_recs + 1) * sizeof(uint64_t)));
    if (n_recs == 0) {
        partition_starts_kernel<<<1, 32, 0, h->stream>>>(scratch, n_dest, scratch + 64, scratch + 64 + 65);
        CK(cudaGetLastError());
        h->last_launches += 1;
        return BRISK_OK;
    }
    u64 *d = (u64 *)dst.p, *d_o = occ ? (u64 *)dst_occ.p : nullptr;
    CK(RW == 2 ? launch_partition<2>(h, scratch, recs, occ, n_recs, n_dest, mode, d, d_o)
               : RW == 4 ? launch_partition<4>(h, scratch, recs, occ, n_recs, n_dest, mode, d, d_o) : launch_partition<6>(h, scratch, recs, occ, n_recs, n_dest, mode, d, d_o));
    h->last_launches += 3;
    return BRISK_OK;
}
// before the exchange: d_recs (+ d_rec_occ) -> d_send (+ d_send_occ), one region per destination rank
static int partition_batch(brisk_handle *h, uint64_t n_recs, uint32_t n_ranks, bool with_occ) {
    return partition_records(h, (const u64 *)h->d_recs.p, with_occ ? (const u64 *)h->d_rec_occ.p : nullptr, n_recs, n_ranks, 0, h->d_send,
                             h->d_send_occ);
}

// number of k-mers of every record -> exclusive prefix (n_recs + 1 entries) in `off`
template <int RW>
__global__ void __launch_bounds__(256) rec_n_kernel(const BriskParams P, const u64 *__restrict__ recs, u64 n_recs, u64 *__restrict__ n_out) {
    u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i > n_recs) return;
    u64 n = 0;
    if (i < n_recs) {
        u64 bucket;
        int nn, idx0;
        rec_unpack_header(P, recs[i * RW + (RW - 1)], bucket, nn, idx0);
        n = (u64)nn;
    }
    n_out[i] = n;
}
static int occurrence_offsets(brisk_handle *h, const u64 *recs, uint64_t n_recs, DevBuf &off) {
    const int RW = h->P.rec_words;
    CK(h->d_cnt.ensure((n_recs + 1) * sizeof(uint64_t)));
    CK(off.ensure((n_recs + 1) * sizeof(uint64_t)));
    unsigned blocks = (unsigned)((n_recs + 1 + 255) / 256);
    if (RW == 2) rec_n_kernel<2><<<blocks, 256, 0, h->stream>>>(h->P, recs, n_recs, (u64 *)h->d_cnt.p);
    else if (RW == 4) rec_n_kernel<4><<<blocks, 256, 0, h->stream>>>(h->P, recs, n_recs, (u64 *)h->d_cnt.p);
    else rec_n_kernel<6><<<blocks, 256, 0, h->stream>>>(h->P, recs, n_recs, (u64 *)h->d_cnt.p);
    CK(cudaGetLastError());
    size_t tmp = 0;
    CK(cub::DeviceScan::ExclusiveSum(nullptr, tmp, (u64 *)h->d_cnt.p, (u64 *)off.p, (int)(n_recs + 1), h->stream));
    CK(h->d_cub.ensure(tmp));
    CK(cub::DeviceScan::ExclusiveSum(h->d_cub.p, tmp, (u64 *)h->d_cnt.p, (u64 *)off.p, (int)(n_recs + 1), h->stream));
    h->last_launches += 3;
    return BRISK_OK;
}

// owner side: results of record i go to [off[i], off[i] + n) in emission order
__global__ void __launch_bounds__(256) occ_tag_kernel(const u64 *off, u64 n_recs, u64 *tag) {
    u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n_recs) tag[i] = off[i] << 1;
}

// source side: the results that came back (send order, emission order inside a record) -> read order
template <int RW>
__global__ void __launch_bounds__(256) scatter_results_kernel(const BriskParams P, const u64 *__restrict__ send_recs,
                                                              const u64 *__restrict__ send_occ, const u64 *__restrict__ off,
                                                              u64 n_recs, const u64 *__restrict__ starts, u32 n_ranks,
                                                              const u64 *__restrict__ back_slots, const unsigned char *__restrict__ back_flags,
                                                              const unsigned char *__restrict__ back_data, u32 data_size, u64 *slots,
                                                              unsigned char *flags, unsigned char *data) {
    u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_recs) return;
    u64 bucket;
    int n, idx0;
    rec_unpack_header(P, send_recs[i * RW + (RW - 1)], bucket, n, idx0);
    u32 d = 0;
    while (d + 1 < n_ranks && starts[d + 1] <= i) d++;
    const u64 tag = send_occ[i], o = off[i];
    for (int j = 0; j < n; j++) {
        const u64 oi = (tag >> 1) + ((tag & 1) ? (u64)(n - 1 - j) : (u64)j);
        if (flags) flags[oi] = back_flags[o + j];
        if (slots) {
            u64 sl = back_slots[o + j];
            slots[oi] = sl == ~0ull ? sl : (sl | ((u64)d << 56));  // slot ids of a sharded index carry the owner rank
        }
        if (data) for (u32 q = 0; q < data_size; q++) data[oi * data_size + q] = back_data[(o + j) * data_size + q];
    }
}

// sharded index: partitions the enumerated records of this rank, exchanges them all-to-all and returns the records
// this rank owns (d_recv).  With per-occurrence outputs the plan of the way back is prepared too.
struct Exchange {
    const u64 *recs = nullptr;   // records to insert / look up here
    uint64_t n = 0;
    std::vector<uint64_t> send_occ_start, recv_occ_start;  // occurrence offsets at the region boundaries (with_occ)
};

static int exchange_records(brisk_handle *h, uint64_t n_recs, bool with_occ, Exchange &X) {
    Comm &c = *h->comm;
    const int RW = h->P.rec_words, n = c.n;
    int rc = partition_batch(h, n_recs, (uint32_t)n, with_occ);
    if (rc) return rc;
    NCK(g_nccl.AllGather(part_totals(h), part_matrix(h), (size_t)n, ncclUint64, c.comm, h->stream));
    c.h_matrix.resize((size_t)n * n);
    CK(cudaMemcpyAsync(c.h_matrix.data(), part_matrix(h), (size_t)n * n * sizeof(uint64_t), cudaMemcpyDeviceToHost, h->stream));
    if (with_occ) {
        rc = occurrence_offsets(h, (const u64 *)h->d_send.p, n_recs, h->d_occ_off);
        if (rc) return rc;
    }
    CK(cudaStreamSynchronize(h->stream));
    c.send_start.assign(n + 1, 0);
    c.recv_start.assign(n + 1, 0);
    exchange_plan(c.h_matrix.data(), n, c.rank, c.send_start.data(), c.recv_start.data());
    if (c.send_start[n] != n_recs) return fail(BRISK_E_COMM, "partition lost records");
    const uint64_t n_recv = c.recv_start[n];
    CK(h->d_recv.ensure((n_recv + 1) * RW * sizeof(uint64_t)));
    NCK(g_nccl.GroupStart());
    for (int p = 0; p < n; p++) {
        uint64_t sc = c.send_start[p + 1] - c.send_start[p], rcnt = c.recv_start[p + 1] - c.recv_start[p];
        if (sc) NCK(g_nccl.Send((const u64 *)h->d_send.p + c.send_start[p] * RW, sc * RW, ncclUint64, p, c.comm, h->stream));
        if (rcnt) NCK(g_nccl.Recv((u64 *)h->d_recv.p + c.recv_start[p] * RW, rcnt * RW, ncclUint64, p, c.comm, h->stream));
    }
    NCK(g_nccl.GroupEnd());
    X.recs = (const u64 *)h->d_recv.p;
    X.n = n_recv;
    if (with_occ) {
        rc = occurrence_offsets(h, X.recs, n_recv, h->d_recv_occ);  // d_recv_occ first holds the offsets ...
        if (rc) return rc;
        X.send_occ_start.assign(n + 1, 0);
        X.recv_occ_start.assign(n + 1, 0);
        for (int p = 0; p <= n; p++) {
            CK(cudaMemcpyAsync(&X.send_occ_start[p], (u64 *)h->d_occ_off.p + c.send_start[p], sizeof(uint64_t), cudaMemcpyDeviceToHost, h->stream));
            CK(cudaMemcpyAsync(&X.recv_occ_start[p], (u64 *)h->d_recv_occ.p + c.recv_start[p], sizeof(uint64_t), cudaMemcpyDeviceToHost, h->stream));
        }
        if (n_recv) {  // ... then the tags the store kernel reads (offset << 1: emission order, contiguous)
            occ_tag_kernel<<<(unsigned)((n_recv + 255) / 256), 256, 0, h->stream>>>((const u64 *)h->d_recv_occ.p, n_recv, (u64 *)h->d_recv_occ.p);
            CK(cudaGetLastError());
            h->last_launches++;
        }
        CK(cudaStreamSynchronize(h->stream));
    }
    return BRISK_OK;
}

// Pipelined form of enumerate + exchange for calls without per-occurrence outputs (the hot path: brisk_count_reads,
// brisk_query_reads).  The batch is enumerated in chunks on the handle's stream; as soon as a chunk's records exist the
// exchange stream partitions them by owner, all-gathers the per-destination totals (plus one status word per rank, so
// that a rank-local failure is seen by everybody and all ranks leave the call together), and sends / receives them,
// while the handle's stream is already enumerating the next chunk.  The host only waits on the exchange stream's small
// count copies (pinned memory); the chunks shrink towards the end (40/30/20/10 %), so what stays exposed after the last
// enumerate launch is the exchange of a tenth of the batch.
// Returns BRISK_OK with X filled (records in d_recv), or 1 when the ranks agreed to fall back to the unpipelined path
// (record buffer overflow somewhere), or an error every rank returns alike.
static int exchange_pipelined(brisk_handle *h, const ReadBatch &B, int local_status, Exchange &X, uint64_t &n_recs_out) {
    Comm &c = *h->comm;
    const int RW = h->P.rec_words, n = c.n;
    // chunk boundaries (reads): the host-input chunks as staged, or a geometric split of a device-resident batch
    std::vector<uint64_t> cf;
    if (B.chunk_events || B.n_reads < 65536) cf = B.chunk_first;
    else {
        // three chunks shrinking towards the end (50 / 35 / 15 % of the reads): every exchange costs a few hundred
        // microseconds of latency whatever its size, so few chunks; what stays exposed is the exchange of the last one
        const double frac[2] = {0.5, 0.85};
        cf.push_back(0);
        for (int i = 0; i < 2; i++) cf.push_back(((uint64_t)(frac[i] * (double)B.n_reads) / ENUM_THREADS) * ENUM_THREADS);
        cf.push_back(B.n_reads);
    }
    const size_t nc = cf.size() - 1;
    if (nc > COMM_MAX_CHUNKS || n >= 64) return 1;  // (word n of the partition scratch carries the status: n < 64)
    // every chunk enumerates into its own region of d_recs behind its own counter: the chunks of a device-resident batch
    // are launched alternately on the handle's stream and on an auxiliary stream, so that the CTAs of chunk c + 1 fill the
    // SMs while the last CTAs of chunk c drain (separate launches in ONE stream would leave the GPU partly idle at every
    // chunk boundary), and the records of different chunks never interleave
    std::vector<uint64_t> base(nc + 1, 0), ccap(nc, 0);
    for (size_t ci = 0; ci < nc; ci++) {
        const uint64_t bytes = B.chunk_events ? B.chunk_byte[ci + 1] - B.chunk_byte[ci]
                                              : (uint64_t)((double)B.total_bytes * (double)(cf[ci + 1] - cf[ci]) / (double)std::max<uint64_t>(1, B.n_reads)) + 1;
        ccap[ci] = std::min(bytes + 1, (uint64_t)(3.0 * (double)bytes / (h->P.w + 1)) + (cf[ci + 1] - cf[ci]) + 4096);
        base[ci + 1] = base[ci] + ccap[ci];
    }
    const uint64_t cap = base[nc];
    CK(h->d_recs.ensure((cap + 1) * RW * sizeof(uint64_t)));
    CK(h->d_send.ensure((cap + 1) * RW * sizeof(uint64_t)));
    uint64_t recv_cap = cap + cap / 4 + 4096;
    CK(h->d_recv.ensure(recv_cap * RW * sizeof(uint64_t)));
    recv_cap = h->d_recv.cap / (RW * sizeof(uint64_t));
    u64 *counters = (u64 *)h->d_rec_count + 16;
    CK(cudaMemsetAsync(counters, 0, 16 * sizeof(uint64_t), h->stream));
    // development switch BRISK_EXCHANGE_MODE: 1 = chunks one after the other on the handle's stream, 2 = alternating between
    // the handle's stream and one auxiliary stream, 3 (default) = priority-ordered auxiliary streams
    static int xmode = -1;
    if (xmode < 0) { const char *e = getenv("BRISK_EXCHANGE_MODE"); xmode = e ? atoi(e) : 3; }
    const bool two_streams = !B.chunk_events && nc > 1 && nc <= 3 && xmode >= 2 && (xmode == 2 || c.aux_ordered);
    if (two_streams) {  // the whole batch is packed first (its bases are resident), then the chunks only enumerate
        if (B.n_reads) {
            pack_reads_kernel<<<(unsigned)((B.n_reads + PACK_READS_PER_CTA - 1) / PACK_READS_PER_CTA), 256, 0, h->stream>>>(B.d_bases, B.d_offsets, 0, B.n_reads, B.d_packed);
            CK(cudaGetLastError());
            h->last_launches++;
        }
        CK(cudaEventRecord(h->ev[1], h->stream));
        for (int i = 0; i < 2; i++) CK(cudaStreamWaitEvent(c.aux[i], h->ev[1], 0));
    }
    for (size_t ci = 0; ci < nc; ci++) {
        const uint64_t r0 = cf[ci], r1 = cf[ci + 1];
        cudaStream_t st = !two_streams ? h->stream : xmode == 2 ? ((ci & 1) ? c.aux[1] : h->stream) : (ci + 1 < nc ? c.aux[ci & 1] : h->stream);
        EnumOut out;
        out.recs = (u64 *)h->d_recs.p + base[ci] * RW;
        out.rec_cap = ccap[ci];
        out.rec_count = counters + ci;
        out.rec_occ = nullptr;
        out.kmer_base = nullptr;
        if (B.chunk_events) CK(cudaStreamWaitEvent(h->stream, h->chunk_ev[ci], 0));
        if (r1 > r0) {
            if (!two_streams) {
                pack_reads_kernel<<<(unsigned)((r1 - r0 + PACK_READS_PER_CTA - 1) / PACK_READS_PER_CTA), 256, 0, h->stream>>>(B.d_bases, B.d_offsets, r0, r1, B.d_packed);
                CK(cudaGetLastError());
                h->last_launches++;
            }
            if (ci == 0 && !two_streams) CK(cudaEventRecord(h->ev[1], h->stream));
            CK(launch_enum(h, B, r0, r1, out, st));
            h->last_launches++;
        } else if (ci == 0 && !two_streams) CK(cudaEventRecord(h->ev[1], h->stream));
        CK(cudaMemcpyAsync(c.h_pinned + ci, counters + ci, sizeof(uint64_t), cudaMemcpyDeviceToHost, st));
        CK(cudaEventRecord(c.ev_enum[ci], st));
    }
    if (two_streams) for (size_t ci = 0; ci < nc; ci++) CK(cudaStreamWaitEvent(h->stream, c.ev_enum[ci], 0));
    CK(cudaEventRecord(h->ev[2], h->stream));
    uint64_t done = 0, recv_total = 0;
    u64 *g_mine = c.d_gather, *g_all = c.d_gather + (64 + 65 + 64 + 8);
    uint64_t *h_all = c.h_pinned + 16;
    int verdict = BRISK_OK;
    for (size_t ci = 0; ci < nc; ci++) {
        CK(cudaEventSynchronize(c.ev_enum[ci]));
        const uint64_t after = c.h_pinned[ci];
        int status = local_status;
        if (after > ccap[ci]) status = 1;  // record buffer overflow: everybody falls back to the unpipelined path
        const uint64_t r0 = base[ci], cnt = std::min(after, ccap[ci]);
        done += cnt;
        CK(cudaStreamWaitEvent(c.stream, c.ev_enum[ci], 0));
        // this rank's row of the gather = the partition's own totals (n words) followed by the status word
        CK(cudaMemsetAsync(g_mine, 0, 65 * sizeof(uint64_t), c.stream));
        if (status) {
            const uint64_t st64 = (uint64_t)(int64_t)status;
            CK(cudaMemcpyAsync(g_mine + n, &st64, sizeof(uint64_t), cudaMemcpyHostToDevice, c.stream));
        }
        if (cnt) {
            const u64 *src = (const u64 *)h->d_recs.p + r0 * RW;
            u64 *dst = (u64 *)h->d_send.p + r0 * RW;
            CK(RW == 2 ? launch_partition<2>(h, g_mine, src, nullptr, cnt, (uint32_t)n, 0, dst, nullptr, c.stream)
                       : RW == 4 ? launch_partition<4>(h, g_mine, src, nullptr, cnt, (uint32_t)n, 0, dst, nullptr, c.stream)
                                 : launch_partition<6>(h, g_mine, src, nullptr, cnt, (uint32_t)n, 0, dst, nullptr, c.stream));
            h->last_launches += 3;
        }
        NCK(g_nccl.AllGather(g_mine, g_all, (size_t)n + 1, ncclUint64, c.comm, c.stream));
        CK(cudaMemcpyAsync(h_all, g_all, (size_t)n * (n + 1) * sizeof(uint64_t), cudaMemcpyDeviceToHost, c.stream));
        CK(cudaEventRecord(c.ev_counts, c.stream));
        CK(cudaEventSynchronize(c.ev_counts));
        c.h_matrix.assign((size_t)n * n, 0);
        for (int s = 0; s < n; s++) {
            for (int d = 0; d < n; d++) c.h_matrix[(size_t)s * n + d] = h_all[(size_t)s * (n + 1) + d];
            const int64_t st = (int64_t)h_all[(size_t)s * (n + 1) + n];
            if (st != 0 && verdict == BRISK_OK) verdict = (int)st;
        }
        if (verdict != BRISK_OK) break;  // every rank sees the same matrix: all leave here
        c.send_start.assign(n + 1, 0);
        c.recv_start.assign(n + 1, 0);
        exchange_plan(c.h_matrix.data(), n, c.rank, c.send_start.data(), c.recv_start.data());
        if (c.send_start[n] != cnt) return fail(BRISK_E_COMM, "partition lost records");
        const uint64_t n_recv = c.recv_start[n];
        if (recv_total + n_recv > recv_cap) {  // rare: the receive arena grows, keeping what has arrived
            CK(cudaStreamSynchronize(c.stream));
            DevBuf bigger;
            CK(bigger.ensure((recv_total + n_recv) * 2 * RW * sizeof(uint64_t)));
            CK(cudaMemcpy(bigger.p, h->d_recv.p, recv_total * RW * sizeof(uint64_t), cudaMemcpyDeviceToDevice));
            h->d_recv.release();
            h->d_recv = bigger;
            recv_cap = h->d_recv.cap / (RW * sizeof(uint64_t));
        }
        NCK(g_nccl.GroupStart());
        for (int p = 0; p < n; p++) {
            const uint64_t sc = c.send_start[p + 1] - c.send_start[p], rc_ = c.recv_start[p + 1] - c.recv_start[p];
            if (sc) NCK(g_nccl.Send((const u64 *)h->d_send.p + (r0 + c.send_start[p]) * RW, sc * RW, ncclUint64, p, c.comm, c.stream));
            if (rc_) NCK(g_nccl.Recv((u64 *)h->d_recv.p + (recv_total + c.recv_start[p]) * RW, rc_ * RW, ncclUint64, p, c.comm, c.stream));
        }
        NCK(g_nccl.GroupEnd());
        recv_total += n_recv;
    }
    CK(cudaEventRecord(c.ev_done, c.stream));
    CK(cudaStreamWaitEvent(h->stream, c.ev_done, 0));
    if (verdict == 1) return 1;
    if (verdict != BRISK_OK) return fail(verdict, verdict == local_status ? g_err : std::string("another rank of the sharded index failed in this call (status ") +
                                                                                        std::to_string(verdict) + ")");
    n_recs_out = done;
    X.recs = (const u64 *)h->d_recv.p;
    X.n = recv_total;
    return BRISK_OK;
}

// the way back of per-occurrence results: owner-side arrays (d_ret_*) -> source-side arrays in send order
// (d_out_* are the caller-order staging buffers, so the send-order copies live in d_cnt-independent buffers)
static int return_results(brisk_handle *h, const Exchange &X, uint64_t n_recs, bool want_slots, bool want_flags, bool want_data,
                          DevBuf &back_slots, DevBuf &back_flags, DevBuf &back_data, u64 *o_slots, unsigned char *o_flags,
                          unsigned char *o_data) {
    Comm &c = *h->comm;
    const int n = c.n;
    const uint64_t n_back = X.send_occ_start[n];
    const uint32_t ds = h->data_size;
    if (want_slots) CK(back_slots.ensure((n_back + 1) * 8));
    if (want_flags) CK(back_flags.ensure(n_back + 1));
    if (want_data) CK(back_data.ensure((n_back + 1) * ds));
    NCK(g_nccl.GroupStart());
    for (int p = 0; p < n; p++) {
        uint64_t sc = X.recv_occ_start[p + 1] - X.recv_occ_start[p], so = X.recv_occ_start[p];
        uint64_t rcnt = X.send_occ_start[p + 1] - X.send_occ_start[p], ro = X.send_occ_start[p];
        if (want_slots) {
            if (sc) NCK(g_nccl.Send((const u64 *)h->d_ret_slots.p + so, sc, ncclUint64, p, c.comm, h->stream));
            if (rcnt) NCK(g_nccl.Recv((u64 *)back_slots.p + ro, rcnt, ncclUint64, p, c.comm, h->stream));
        }
        if (want_flags) {
            if (sc) NCK(g_nccl.Send((const char *)h->d_ret_flags.p + so, sc, ncclUint8, p, c.comm, h->stream));
            if (rcnt) NCK(g_nccl.Recv((char *)back_flags.p + ro, rcnt, ncclUint8, p, c.comm, h->stream));
        }
        if (want_data) {
            if (sc) NCK(g_nccl.Send((const char *)h->d_ret_data.p + so * ds, sc * ds, ncclUint8, p, c.comm, h->stream));
            if (rcnt) NCK(g_nccl.Recv((char *)back_data.p + ro * ds, rcnt * ds, ncclUint8, p, c.comm, h->stream));
        }
    }
    NCK(g_nccl.GroupEnd());
    if (n_recs) {
        const int RW = h->P.rec_words;
        unsigned blocks = (unsigned)((n_recs + 255) / 256);
#define LAUNCH_SC(RWV)                                                                                                              \
    scatter_results_kernel<RWV><<<blocks, 256, 0, h->stream>>>(h->P, (const u64 *)h->d_send.p, (const u64 *)h->d_send_occ.p,        \
                                                                (const u64 *)h->d_occ_off.p, n_recs, part_starts(h), (u32)n,        \
                                                                (const u64 *)back_slots.p, (const unsigned char *)back_flags.p,     \
                                                                (const unsigned char *)back_data.p, ds, o_slots, o_flags, o_data)
        if (RW == 2) LAUNCH_SC(2); else if (RW == 4) LAUNCH_SC(4); else LAUNCH_SC(6);
#undef LAUNCH_SC
        CK(cudaGetLastError());
        h->last_launches++;
    }
    return BRISK_OK;
}

// makes room for `incoming` more k-mer occurrences before an insert.  The number of NEW keys among them is estimated
// from the capacity hint while it lasts and from the new-key rate of the previous insert once the stored keys approach
// or pass the hint (a hint that is too small must not stop the table from growing ahead of the batch: with the table
// nearly full every insert walks to the probe limit and is parked).  The overflow arena is sized with the batch, so
// that even a badly underestimated batch is parked and replayed, not dropped.
// worst_case: the call hands slot ids / newly flags back, which a rehash during settle() would invalidate: besides the
// estimate, make sure that even with every incoming k-mer new the load stays below 0.9, where nothing reaches the probe
// limit, so that nothing is parked (a table that is merely re-queried with known keys does not grow for this).
static int pregrow(brisk_handle *h, uint64_t incoming, bool worst_case = false) {
    uint64_t est_new = incoming;
    if (h->hint) {
        const uint64_t by_hint = std::min<uint64_t>(incoming, h->hint > h->n_keys ? h->hint - h->n_keys : 0);
        const uint64_t by_rate = std::min<uint64_t>(incoming, (uint64_t)((1.25 * h->new_rate + 0.01) * (double)incoming));
        est_new = std::max(by_hint, by_rate);
    }
    if (worst_case && (h->n_keys + incoming) * 10 > h->cap * 9) {
        int rc = grow_to(h, pow2ceil((h->n_keys + incoming) * 2));
        if (rc) return rc;
    }
    h->pending_incoming = incoming;
    h->keys_before = h->n_keys;
    const uint64_t want_ovf = std::max<uint64_t>(1u << 20, incoming / 16);
    if (want_ovf > h->ovf_cap) {
        if (h->overflow) CK(cudaFree(h->overflow));
        h->overflow = nullptr;
        h->ovf_cap = 0;
        CK(cudaMalloc((void **)&h->overflow, want_ovf * 3 * sizeof(uint64_t)));
        h->ovf_cap = want_ovf;
    }
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
        if (round == 0 && h->pending_incoming) {  // new-key rate of this insert (parked ones count as new), for the next pregrow
            h->new_rate = std::min(1.0, (double)(h->n_keys - h->keys_before + parked) / (double)h->pending_incoming);
            h->pending_incoming = 0;
        }
        if (parked == 0) return BRISK_OK;
        if (parked > h->ovf_cap) {
            // the arena holds 1/16 of the batch's occurrences: this takes a table that was nearly full before a batch of
            // almost only new keys.  The inserts beyond the arena are lost; the handle stays usable.
            CK(cudaMemsetAsync(h->counters + 1, 0, sizeof(uint64_t), h->stream));
            CK(cudaStreamSynchronize(h->stream));
            return fail(BRISK_E_FULL, "bucket store overflow: " + std::to_string(parked - h->ovf_cap) + " inserts of the last batch hit the "
                                      "probe limit beyond the overflow arena and were dropped; call brisk_reserve ahead of such batches");
        }
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

// count / insert calls without per-occurrence outputs return once their input has been consumed; their store kernel
// is settled (counters read back, parked inserts replayed) when the handle is used next
static int settle_pending(brisk_handle *h) {
    if (!h->pending_settle) return BRISK_OK;
    h->pending_settle = false;
    if (h->pending_on_side) {  // the store kernel was left running on the side stream: the handle's stream joins it
        h->pending_on_side = false;
        CK(cudaStreamWaitEvent(h->stream, h->ov_ev[1], 0));
    }
    return settle(h);
}
#define SETTLE_PENDING(h) do { int prc_ = settle_pending(h); if (prc_) return prc_; } while (0)

static void finish_timing(brisk_handle *h) {
    h->timing_dirty = false;
    // ev: 0 start/pack begin, 1 pack end, 2 enumerate end, 3 partition end, 4 insert end
    cudaEventSynchronize(h->ev[4]);
    for (int i = 0; i < 4; i++) cudaEventElapsedTime(&h->last_ms[i], h->ev[i], h->ev[i + 1]);
    cudaEventElapsedTime(&h->last_ms[4], h->ev[0], h->ev[4]);
}

// Host input with nothing to hand back per occurrence (brisk_count_reads, brisk_query_reads, brisk_insert_reads without
// outputs) on an unsharded index: the batch arrives in chunks over PCIe for about as long as all its kernels take, so
// the chunks are cut into GROUPS, and the records of every group but the last go to the store kernel WHILE the later
// chunks are still on their way — on a second, low-priority stream with a capped residency, so that pack + enumerate
// of the arriving chunks are not queued behind it.  What is left after the last byte has arrived is the enumeration
// of the last chunk and the store kernel over the last group only.  Each group enumerates into its own region of
// d_recs and is put in table-region order on its own (d_send, same regions).  done = false: nothing has been applied
// (the first group's region was too small) and the caller runs the plain path.
static int ensure_side_stream(brisk_handle *h) {
    if (h->store_stream) return BRISK_OK;
    int lo_p = 0, hi_p = 0;
    CK(cudaDeviceGetStreamPriorityRange(&lo_p, &hi_p));
    CK(cudaStreamCreateWithPriority(&h->store_stream, cudaStreamNonBlocking, lo_p));
    for (int i = 0; i < 2; i++) CK(cudaEventCreateWithFlags(&h->ov_ev[i], cudaEventDisableTiming));
    return BRISK_OK;
}
static std::vector<size_t> overlap_cuts(const char *env_name, const char *dflt) {
    const char *e = getenv(env_name);
    std::vector<size_t> v;
    for (const char *p = e ? e : dflt; *p;) {
        char *end = nullptr;
        const long x = strtol(p, &end, 10);
        if (end == p) break;
        if (x > 0 && (v.empty() || (size_t)x > v.back())) v.push_back((size_t)x);
        p = *end ? end + 1 : end;
    }
    return v;
}
static int run_reads_overlapped(brisk_handle *h, const ReadBatch &B, int mode, OccOut O, bool &done) {
    done = false;
    // development switches: the chunk indices (of 8) at which a new group starts, "0" = off; the residency cap of the
    // store kernels that run beside the enumeration; 0 = everything on the handle's stream
    static const std::vector<size_t> cuts_put = overlap_cuts("BRISK_OVERLAP_COUNT", "3"), cuts_get = overlap_cuts("BRISK_OVERLAP_GET", "3");
    static const int ctas_env = getenv("BRISK_OVERLAP_CTAS") ? atoi(getenv("BRISK_OVERLAP_CTAS")) : 2;
    static const int stream_env = getenv("BRISK_OVERLAP_STREAM") ? atoi(getenv("BRISK_OVERLAP_STREAM")) : 1;
    const size_t nc = B.chunk_first.size() - 1;
    std::vector<size_t> cut(1, 0);
    for (size_t c : (mode == MODE_GET ? cuts_get : cuts_put)) if (c < nc) cut.push_back(c);
    cut.push_back(nc);
    const int ng = (int)cut.size() - 1;
    if (ng < 2 || nc < 4) return BRISK_OK;
    const int RW = h->P.rec_words;
    {
        int src = ensure_side_stream(h);
        if (src) return src;
    }
    cudaStream_t side = stream_env ? h->store_stream : h->stream;
    if (mode != MODE_GET) {  // room for the whole batch before its first store kernel (a rehash runs on the handle's stream)
        const uint64_t lost = B.n_reads * (uint64_t)(h->P.k - 1);
        int rc = pregrow(h, B.total_bytes > lost ? B.total_bytes - lost : B.total_bytes, false);
        if (rc) return rc;
    }
    // record regions of the groups
    std::vector<uint64_t> base(ng + 1, 0), ccap(ng, 0);
    for (int g = 0; g < ng; g++) {
        const uint64_t bytes = B.chunk_byte[cut[g + 1]] - B.chunk_byte[cut[g]], reads = B.chunk_first[cut[g + 1]] - B.chunk_first[cut[g]];
        ccap[g] = std::min(bytes + 1, (uint64_t)(3.0 * (double)bytes / (h->P.w + 1)) + reads + 4096);
    }
    if (const char *tc = getenv("BRISK_OVERLAP_TEST_CAP")) {  // tests: "g:cap" makes group g's region too small
        const int g = atoi(tc);
        const char *colon = strchr(tc, ':');
        if (colon && g >= 0 && g < ng) ccap[g] = std::min<uint64_t>(ccap[g], strtoull(colon + 1, nullptr, 10));
    }
    for (int g = 0; g < ng; g++) base[g + 1] = base[g] + ccap[g];
    CK(h->d_recs.ensure((base[ng] + 1) * RW * sizeof(uint64_t)));
    if (h->region_sort) CK(h->d_send.ensure((base[ng] + 1) * RW * sizeof(uint64_t)));
    // every group counts behind d_rec_count[0] (where the long-read launch appends): the host waits for a group's
    // enumeration before the next one is queued
    EnumLong lg = no_long();
    if (h->long_limit) {
        CK(h->d_long.ensure((B.n_reads + 1) * sizeof(uint64_t)));
        lg.list = (u64 *)h->d_long.p; lg.cap = B.n_reads; lg.count = (u64 *)h->d_rec_count + 8; lg.limit = h->long_limit;
    }
    // region order + store kernel of one group on stream st
    auto store_group = [&](const u64 *recs, uint64_t n, u64 *sorted, cudaStream_t st, int ctas) -> int {
        if (n == 0) return BRISK_OK;
        if (!st) st = h->stream;
        if (h->region_sort && n >= 4096) {
            CK(cudaMemsetAsync(h->d_part2, 0, 64 * sizeof(uint64_t), st));
            CK(RW == 2 ? launch_partition<2>(h, h->d_part2, recs, nullptr, n, STORE_REGIONS, 1, sorted, nullptr, st)
                       : RW == 4 ? launch_partition<4>(h, h->d_part2, recs, nullptr, n, STORE_REGIONS, 1, sorted, nullptr, st)
                                 : launch_partition<6>(h, h->d_part2, recs, nullptr, n, STORE_REGIONS, 1, sorted, nullptr, st));
            h->last_launches += 3;
            recs = sorted;
        }
        CK(launch_records(h, mode, recs, n, O, st, ctas));
        h->last_launches++;
        return BRISK_OK;
    };
    bool side_busy = false;  // store kernels of earlier groups may still be running on the side stream
    for (int g = 0; g < ng; g++) {
        const bool last = g + 1 == ng;
        EnumOut out;
        out.recs = (u64 *)h->d_recs.p + base[g] * RW;
        out.rec_cap = ccap[g];
        out.rec_count = (u64 *)h->d_rec_count;
        out.rec_occ = nullptr;
        out.kmer_base = nullptr;
        CK(cudaMemsetAsync(h->d_rec_count, 0, 9 * sizeof(uint64_t), h->stream));  // record counter [0], long-read counter [8]
        for (size_t c = cut[g]; c < cut[g + 1]; c++) {
            int prc = pack_chunk(h, B, c);
            if (prc) return prc;
            if (c == 0) CK(cudaEventRecord(h->ev[1], h->stream));
            const uint64_t r0 = B.chunk_first[c], r1 = B.chunk_first[c + 1];
            if (r1 == r0) continue;
            CK(launch_enum(h, B, r0, r1, out, nullptr, lg));
            h->last_launches++;
        }
        uint64_t n = 0, n_long = 0;
        CK(cudaMemcpyAsync(&n, h->d_rec_count, sizeof(uint64_t), cudaMemcpyDeviceToHost, h->stream));
        if (lg.list) CK(cudaMemcpyAsync(&n_long, lg.count, sizeof(uint64_t), cudaMemcpyDeviceToHost, h->stream));
        CK(cudaStreamSynchronize(h->stream));
        if (n_long && n <= ccap[g]) {  // reads beyond the thread-per-read limit: their records follow the group's
            int rc = enumerate_long_reads(h, B, out, n_long, n);
            if (rc) return rc;
        }
        const u64 *recs = out.recs;
        u64 *sorted = h->region_sort ? (u64 *)h->d_send.p + base[g] * RW : nullptr;
        bool fell_back = false;
        if (n > ccap[g]) {
            if (g == 0) return BRISK_OK;  // nothing applied yet: the plain path takes the whole batch
            // the group's region was too small: enumerate its chunks again the plain way, into d_recs from its start,
            // once the earlier groups' kernels are done with it; the group is then stored on the handle's stream, which
            // orders it before everything the later groups do to these buffers
            CK(cudaStreamSynchronize(side));
            side_busy = false;
            uint64_t n_occ = 0;
            int rc = enumerate_batch(h, B, false, n, n_occ, cut[g], cut[g + 1]);
            if (rc) return rc;
            if (h->region_sort) CK(h->d_send.ensure((std::max(n, base[ng]) + 1) * RW * sizeof(uint64_t)));
            recs = (const u64 *)h->d_recs.p;
            sorted = h->region_sort ? (u64 *)h->d_send.p : nullptr;
            fell_back = true;
        }
        if (last || fell_back) {
            if (last) {
                CK(cudaEventRecord(h->ev[2], h->stream));
                CK(cudaEventRecord(h->ev[3], h->stream));
            }
            if (side_busy) CK(cudaStreamWaitEvent(h->stream, h->ov_ev[1], 0));
            side_busy = false;
            int rc = store_group(recs, n, sorted, nullptr, 0);
            if (rc) return rc;
        } else {
            CK(cudaEventRecord(h->ov_ev[0], h->stream));
            CK(cudaStreamWaitEvent(side, h->ov_ev[0], 0));
            int rc = store_group(recs, n, sorted, side, ctas_env);
            if (rc) return rc;
            CK(cudaEventRecord(h->ov_ev[1], side));
            side_busy = true;
        }
        done = true;  // from here on part of the batch is in the store
    }
    return BRISK_OK;
}

// shared driver of count / insert / get over reads
static int run_reads(brisk_handle *h, const char *bases, const uint64_t *offsets, uint64_t n_reads, int mode,
                     uint64_t *slots, uint8_t *flags, void *data_out, uint64_t *agg_out = nullptr) {
    if (!h) return fail(BRISK_E_PARAM, "null handle");
    if (n_reads && (!bases || !offsets)) return fail(BRISK_E_PARAM, "null input");
    std::lock_guard<std::recursive_mutex> lock(h->mu);
    CK(cudaSetDevice(h->device));
    h->last_launches = 0;
    ReadBatch B;
    int rc = stage_reads(h, bases, offsets, n_reads, B);
    if (rc) return rc;
    const bool with_occ = slots || flags || data_out;
    const bool sharded = h->comm && h->comm->n > 1;
    // after staging: this call's host->device copies run while the previous store kernel finishes.  On a sharded index a
    // deferred failure (store full) must not make this rank leave before the collectives of the call: it is carried
    // into the exchange, where every rank learns of it
    // Device-resident input on an unsharded index: the pending store kernel of the previous call (memory-bound, on the
    // side stream, its records in d_send) and this call's pack + enumerate (ALU-bound, on the handle's stream, into
    // d_recs) share no buffer, so the previous call is only settled once this call's records exist
    static const int async_store = getenv("BRISK_ASYNC_STORE") ? atoi(getenv("BRISK_ASYNC_STORE")) : 0;
    static const int async_ctas = getenv("BRISK_ASYNC_CTAS") ? atoi(getenv("BRISK_ASYNC_CTAS")) : 2;
    const bool late_settle = !sharded && h->pending_settle && h->pending_on_side && !B.chunk_events;
    const int settle_rc = late_settle ? BRISK_OK : settle_pending(h);
    if (settle_rc && !sharded) return settle_rc;
    uint64_t n_recs = 0, n_occ = 0;
    OccOut O;
    O.rec_occ = nullptr; O.slots = nullptr; O.flags = nullptr; O.data = nullptr; O.agg = nullptr;
    if (agg_out) {
        CK(cudaMemsetAsync(h->counters + 5, 0, 2 * sizeof(uint64_t), h->stream));
        O.agg = h->counters + 5;
    }
    // per-occurrence outputs go through device staging when the caller's buffers are on the host
    const bool s_dev = is_device_ptr(slots), f_dev = is_device_ptr(flags), d_dev = is_device_ptr(data_out);
    u64 *o_slots = nullptr;
    unsigned char *o_flags = nullptr, *o_data = nullptr;
    auto stage_outputs = [&]() -> int {  // needs n_occ (known once the first chunk is enumerated)
        if (slots) {
            if (s_dev) o_slots = (u64 *)slots;
            else { CK(h->d_out_slots.ensure((n_occ + 1) * sizeof(uint64_t))); o_slots = (u64 *)h->d_out_slots.p; }
        }
        if (flags) {
            if (f_dev) o_flags = flags;
            else { CK(h->d_out_flags.ensure(n_occ + 1)); o_flags = (unsigned char *)h->d_out_flags.p; }
        }
        if (data_out) {
            if (d_dev) o_data = (unsigned char *)data_out;
            else { CK(h->d_out_data.ensure((n_occ + 1) * h->data_size)); o_data = (unsigned char *)h->d_out_data.p; }
        }
        return BRISK_OK;
    };
    // Records in table-region order: the store kernel then walks the table region by region (L2 keeps the slots a
    // batch revisits, the TLB sees one region at a time).  Occurrence tags travel with their records.
    auto by_region = [&](const u64 *&recs, const u64 *&tags, uint64_t n, DevBuf &dst, DevBuf &dst_occ) -> int {
        if (!h->region_sort || n < 4096) return BRISK_OK;
        int prc = partition_records(h, recs, tags, n, STORE_REGIONS, 1, dst, dst_occ);
        if (prc) return prc;
        recs = (const u64 *)dst.p;
        if (tags) tags = (const u64 *)dst_occ.p;
        return BRISK_OK;
    };
    DevBuf back_slots, back_flags, back_data;
    // Host input: count / insert calls return once their input is consumed, and the NEXT call's host->device copy
    // overlaps this call's last store kernel; within a call the records of the first chunks are stored while the last
    // chunks are still arriving (run_reads_overlapped: two groups; finer cuts were measured and lose to the per-group
    // costs and to the contention with the enumeration, profiles/r2_e2e_overlap.md).
    bool overlapped = false;
    if (!sharded && !with_occ && B.chunk_events) {
        O.agg = agg_out ? h->counters + 5 : nullptr;
        rc = run_reads_overlapped(h, B, mode, O, overlapped);
        if (rc) return rc;
    }
    if (!overlapped) {
        Exchange X;
        bool exchanged = false;
        if (sharded && !with_occ && !(h->flags & BRISK_F_SERIAL_EXCHANGE)) {
            rc = exchange_pipelined(h, B, settle_rc, X, n_recs);
            if (rc < 0) return rc;
            exchanged = rc == BRISK_OK;
        } else if (settle_rc) {
            return settle_rc;
        }
        if (!exchanged) {
            rc = enumerate_batch(h, B, with_occ, n_recs, n_occ);
            if (rc) return rc;
            if (late_settle) {
                rc = settle_pending(h);
                if (rc) return rc;
            }
            CK(cudaEventRecord(h->ev[2], h->stream));
            X.recs = (const u64 *)h->d_recs.p;
            X.n = n_recs;
            if (sharded) {
                rc = exchange_records(h, n_recs, with_occ, X);
                if (rc) return rc;
            }
        }
        const u64 *tags = !with_occ ? nullptr : sharded ? (const u64 *)h->d_recv_occ.p : (const u64 *)h->d_rec_occ.p;
        rc = by_region(X.recs, tags, X.n, sharded ? h->d_recs : h->d_send, sharded ? h->d_rec_occ : h->d_send_occ);
        if (rc) return rc;
        CK(cudaEventRecord(h->ev[3], h->stream));
        if (mode != MODE_GET) {
            // occurrences about to be inserted: this rank's own (bytes minus the k - 1 bases every read loses), or (sharded)
            // the received records times the k-mers per record of the batch this rank enumerated itself
            const uint64_t lost = B.n_reads * (uint64_t)(h->P.k - 1);
            const uint64_t own_occ = B.total_bytes > lost ? B.total_bytes - lost : B.total_bytes;
            const uint64_t per_rec = n_recs ? (own_occ + n_recs - 1) / n_recs : (uint64_t)h->P.w;
            // (calls that hand slot ids back need an upper bound, not an estimate)
            if (with_occ) rc = pregrow(h, sharded ? X.n * (uint64_t)h->P.w : B.total_bytes, true);
            else rc = pregrow(h, sharded ? X.n * std::min<uint64_t>(per_rec, (uint64_t)h->P.w) : own_occ, false);
            if (rc) return rc;
        }
        rc = stage_outputs();
        if (rc) return rc;
        if (!sharded) {
            O.rec_occ = tags;
            O.slots = o_slots; O.flags = o_flags; O.data = o_data;
        } else if (with_occ) {
            // the owner writes its answers in arrival order; they travel back and are put in read order below
            const uint64_t n_ret = X.recv_occ_start[h->comm->n];
            O.rec_occ = tags;
            if (slots) { CK(h->d_ret_slots.ensure((n_ret + 1) * sizeof(uint64_t))); O.slots = (u64 *)h->d_ret_slots.p; }
            if (flags) { CK(h->d_ret_flags.ensure(n_ret + 1)); O.flags = (unsigned char *)h->d_ret_flags.p; }
            if (data_out) { CK(h->d_ret_data.ensure((n_ret + 1) * h->data_size)); O.data = (unsigned char *)h->d_ret_data.p; }
        }
        const bool leave_pending = mode != MODE_GET && !with_occ && !(h->flags & BRISK_F_SYNC_INSERT);
        const bool on_side = async_store && leave_pending && !sharded && !B.chunk_events && X.n &&
                             X.recs == (const u64 *)h->d_send.p;  // (records in d_recs would be overwritten by the next enumeration)
        if (on_side) {
            rc = ensure_side_stream(h);
            if (rc) return rc;
            CK(cudaEventRecord(h->ov_ev[0], h->stream));
            CK(cudaStreamWaitEvent(h->store_stream, h->ov_ev[0], 0));
            CK(launch_records(h, mode, X.recs, X.n, O, h->store_stream, async_ctas));
            CK(cudaEventRecord(h->ov_ev[1], h->store_stream));
            CK(cudaEventRecord(h->ev[4], h->store_stream));
            h->pending_on_side = true;
        } else {
            CK(launch_records(h, mode, X.recs, X.n, O));
        }
        if (X.n) h->last_launches++;
        if (sharded && with_occ) {
            rc = return_results(h, X, n_recs, slots != nullptr, flags != nullptr, data_out != nullptr, back_slots, back_flags, back_data,
                                o_slots, o_flags, o_data);
            if (rc) return rc;
        }
        if (sharded && agg_out)  // every rank gets the totals of the whole query batch
            NCK(g_nccl.AllReduce(h->counters + 5, h->counters + 5, 2, ncclUint64, ncclSum, h->comm->comm, h->stream));
    }
    if (!h->pending_on_side) CK(cudaEventRecord(h->ev[4], h->stream));
    if (slots && !s_dev) CK(cudaMemcpyAsync(slots, o_slots, n_occ * sizeof(uint64_t), cudaMemcpyDeviceToHost, h->stream));
    if (flags && !f_dev) CK(cudaMemcpyAsync(flags, o_flags, n_occ, cudaMemcpyDeviceToHost, h->stream));
    if (data_out && !d_dev) CK(cudaMemcpyAsync(data_out, o_data, n_occ * h->data_size, cudaMemcpyDeviceToHost, h->stream));
    if (agg_out) CK(cudaMemcpyAsync(agg_out, h->counters + 5, 2 * sizeof(uint64_t), cudaMemcpyDeviceToHost, h->stream));
    if (mode != MODE_GET && !with_occ && !(h->flags & BRISK_F_SYNC_INSERT)) {
        // nothing to hand back and the input is consumed (enumeration was waited for): leave the store kernel running
        h->pending_settle = true;
        h->timing_dirty = true;
        return BRISK_OK;
    }
    if (mode != MODE_GET) {
        rc = settle(h);
    } else {
        rc = BRISK_OK;
        cudaError_t e = cudaStreamSynchronize(h->stream);
        if (e != cudaSuccess) rc = fail(BRISK_E_CUDA, std::string("cudaStreamSynchronize: ") + cudaGetErrorString(e));
    }
    back_slots.release(); back_flags.release(); back_data.release();
    if (rc) return rc;
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
    if (n_reads && (!bases || !offsets)) return fail(BRISK_E_PARAM, "null input");
    std::lock_guard<std::recursive_mutex> lock(h->mu);
    CK(cudaSetDevice(h->device));
    SETTLE_PENDING(h);
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

// brisk_enumerate_reads + the read-order tag of every record: occ[i] = (index of the record's first read-order k-mer
// occurrence << 1) | reversed.  Sorting by occ >> 1 gives SuperKmerEnumerator::next's emission order.
extern "C" int brisk_enumerate_reads_tagged(brisk_t *h, const char *bases, const uint64_t *offsets, uint64_t n_reads, uint64_t *recs,
                                            uint64_t cap_recs, uint64_t *n_recs_out, uint32_t *rec_words, uint64_t *occ) {
    if (!h) return fail(BRISK_E_PARAM, "null handle");
    if (n_reads && (!bases || !offsets)) return fail(BRISK_E_PARAM, "null input");
    std::lock_guard<std::recursive_mutex> lock(h->mu);
    CK(cudaSetDevice(h->device));
    SETTLE_PENDING(h);
    h->last_launches = 0;
    if (rec_words) *rec_words = (uint32_t)h->P.rec_words;
    ReadBatch B;
    int rc = stage_reads(h, bases, offsets, n_reads, B);
    if (rc) return rc;
    uint64_t n_recs = 0, n_occ = 0;
    rc = enumerate_batch(h, B, true, n_recs, n_occ);
    if (rc) return rc;
    for (int i = 2; i <= 4; i++) CK(cudaEventRecord(h->ev[i], h->stream));
    if (n_recs_out) *n_recs_out = n_recs;
    if (recs || occ) {
        if (n_recs > cap_recs) return fail(BRISK_E_CAPACITY, "record buffer too small");
        if (recs) CK(cudaMemcpyAsync(recs, h->d_recs.p, n_recs * h->P.rec_words * sizeof(uint64_t),
                                     is_device_ptr(recs) ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, h->stream));
        if (occ) CK(cudaMemcpyAsync(occ, h->d_rec_occ.p, n_recs * sizeof(uint64_t),
                                    is_device_ptr(occ) ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, h->stream));
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
    std::lock_guard<std::recursive_mutex> lock(h->mu);
    CK(cudaSetDevice(h->device));
    SETTLE_PENDING(h);
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
        int rc = pregrow(h, n, slots != nullptr || flags != nullptr);
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
    if (!h) return nullptr;
    if (h->pending_settle) {  // the caller is about to dereference the pointer: the store has to be at rest
        std::lock_guard<std::recursive_mutex> lock(h->mu);
        if (cudaSetDevice(h->device) != cudaSuccess || settle_pending(h) != BRISK_OK) return nullptr;
    }
    if (slot == BRISK_NO_SLOT || slot >= h->cap + 64) return nullptr;
    if (h->data_size == 1) return (unsigned char *)(h->cnt + slot);  // the DATA byte is the low byte of the little-endian count word
    return h->data + slot * h->data_size;
}

extern "C" int brisk_dump(brisk_t *h, brisk_kmer *keys, void *data, uint64_t cap, uint64_t *n) {
    if (!h) return fail(BRISK_E_PARAM, "null handle");
    std::lock_guard<std::recursive_mutex> lock(h->mu);
    CK(cudaSetDevice(h->device));
    SETTLE_PENDING(h);
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

// ---- ordered dump: the stored set in the order DenseMenuYo::next walks it at bucket level
// (increasing bucket id, brisk/DenseMenuYo.hpp:476-521); inside a bucket the reference's order is its arena's insertion
// order, here (minimizer_idx, kmer_s) ascending so that the result is deterministic.
// The sort permutes SLOT INDICES (u32) with our own LSD radix sort (sort.cuh); the key words are read from the table
// through the permutation, so the temporaries are two u32 permutations (8 bytes per stored k-mer) + the histograms.
template <int KW>
__global__ void __launch_bounds__(256) dump_slots_kernel(StoreView T, int n_special, u32 *slots_out, u64 cap_out) {
    typedef typename SlotOf<KW>::type S;
    const S *tab = reinterpret_cast<const S *>(T.slots);
    const u64 total = T.mask + 1 + (u64)n_special;
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (u64)gridDim.x * blockDim.x) {
        bool used;
        if (i <= T.mask) {
            const u64 lo = *reinterpret_cast<const u64 *>(tab + i);
            const u64 hi = KW == 2 ? reinterpret_cast<const u64 *>(tab + i)[1] : 0;
            used = !(lo == STORE_EMPTY && (KW == 1 || hi == STORE_EMPTY));
        } else {
            used = (T.special[i - T.mask - 1] & META_TAG_MASK) != 0;
        }
        if (used) {
            const u64 o = atomicAdd(T.counters + 3, 1ull);
            if (o < cap_out) slots_out[o] = (u32)i;
        }
    }
}
// key words of the k-mer stored in a slot (a special slot holds the all-ones key)
template <int KW>
struct SlotKey {
    StoreView T;
    BriskParams P;
    __device__ __forceinline__ void load(u32 slot, u64 &lo, u64 &hi, u32 &idx) const {
        typedef typename SlotOf<KW>::type S;
        if (slot <= T.mask) {
            const S *p = reinterpret_cast<const S *>(T.slots) + slot;
            lo = *reinterpret_cast<const u64 *>(p);
            hi = KW == 2 ? reinterpret_cast<const u64 *>(p)[1] : 0;
            idx = (p->meta & META_TAG_MASK) - 1;
        } else {
            lo = STORE_EMPTY;
            hi = KW == 2 ? STORE_EMPTY : 0;
            idx = (T.special[slot - T.mask - 1] & META_TAG_MASK) - 1;
        }
    }
};
template <int KW> struct KeyLo : SlotKey<KW> {
    __device__ __forceinline__ u64 operator()(u32 slot) const { u64 lo, hi; u32 idx; this->load(slot, lo, hi, idx); return lo; }
};
template <int KW> struct KeyHi : SlotKey<KW> {
    __device__ __forceinline__ u64 operator()(u32 slot) const { u64 lo, hi; u32 idx; this->load(slot, lo, hi, idx); return hi; }
};
template <int KW> struct KeyBucketIdx : SlotKey<KW> {  // (bucket id << 8) | minimizer_idx
    __device__ __forceinline__ u64 operator()(u32 slot) const {
        u64 lo, hi; u32 idx;
        this->load(slot, lo, hi, idx);
        return (bucket_of_kmer(this->P, lo, hi, (int)idx) << 8) | idx;
    }
};
template <int KW>
__global__ void __launch_bounds__(256) gather_slots_kernel(SlotKey<KW> K, const u32 *__restrict__ perm, u64 n, KmerIn *okeys, unsigned char *odata) {
    u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const u32 slot = perm[i];
    u64 lo, hi; u32 idx;
    K.load(slot, lo, hi, idx);
    if (okeys) {
        KmerIn kk;
        kk.lo = lo; kk.hi = hi; kk.idx = (unsigned char)idx;
        for (int q = 0; q < 7; q++) kk.pad[q] = 0;
        okeys[i] = kk;
    }
    if (odata) {
        if (K.T.data_size == 1) odata[i] = (unsigned char)K.T.cnt[slot];
        else for (u32 q = 0; q < K.T.data_size; q++) odata[i * K.T.data_size + q] = K.T.data[(u64)slot * K.T.data_size + q];
    }
}
// the largest number of k-mers one bucket holds, from the slots sorted by bucket id: every run start walks its run
template <int KW>
__global__ void __launch_bounds__(256) bucket_runs_kernel(KeyBucketIdx<KW> K, const u32 *__restrict__ perm, u64 n, u64 *out_max) {
    u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const u64 b = K(perm[i]) >> 8;
    if (i > 0 && (K(perm[i - 1]) >> 8) == b) return;
    u64 len = 1;
    while (i + len < n && (K(perm[i + len]) >> 8) == b) len++;
    atomicMax(out_max, len);
}

// slots of the stored k-mers sorted by (bucket id[, minimizer_idx, kmer_s]); *result points into perm_a / perm_b
template <int KW>
static int sorted_slots_t(brisk_handle *h, uint64_t nk, bool full_order, DevBuf &perm_a, DevBuf &perm_b, DevBuf &hist, u32 **result) {
    CK(perm_a.ensure((nk + 1) * sizeof(u32)));
    CK(perm_b.ensure((nk + 1) * sizeof(u32)));
    CK(hist.ensure(radix_sort_hist_bytes(nk) + 1024));
    CK(cudaMemsetAsync(h->counters + 3, 0, sizeof(uint64_t), h->stream));
    unsigned blocks = (unsigned)std::min<uint64_t>((h->cap + 255) / 256, 148 * 16);
    dump_slots_kernel<KW><<<blocks, 256, 0, h->stream>>>(view_of(h), 64, (u32 *)perm_a.p, nk);
    CK(cudaGetLastError());
    uint64_t got = 0;
    CK(cudaMemcpyAsync(&got, h->counters + 3, sizeof(uint64_t), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    if (got != nk) return fail(BRISK_E_CUDA, "dump found " + std::to_string(got) + " keys, expected " + std::to_string(nk));
    u32 *cur = (u32 *)perm_a.p, *other = (u32 *)perm_b.p;
    auto swap_to = [&](u32 *res) { if (res != cur) { other = cur; cur = res; } };
    u32 *res = cur;
    if (full_order) {
        // least-significant first: kmer_s low word, high word, then (bucket id, minimizer_idx); every pass is stable
        KeyLo<KW> klo; klo.T = view_of(h); klo.P = h->P;
        CK(radix_sort_perm(klo, cur, other, nk, 0, std::min(64, 2 * h->P.k), (u32 *)hist.p, h->stream, &res));
        swap_to(res);
        if (KW == 2) {
            KeyHi<KW> khi; khi.T = view_of(h); khi.P = h->P;
            CK(radix_sort_perm(khi, cur, other, nk, 0, 2 * h->P.k - 64, (u32 *)hist.p, h->stream, &res));
            swap_to(res);
        }
    }
    KeyBucketIdx<KW> kb; kb.T = view_of(h); kb.P = h->P;
    CK(radix_sort_perm(kb, cur, other, nk, full_order ? 0 : 8, 2 * h->P.b + 8, (u32 *)hist.p, h->stream, &res));
    swap_to(res);
    *result = cur;
    return BRISK_OK;
}

extern "C" int brisk_dump_ordered(brisk_t *h, brisk_kmer *keys, void *data, uint64_t cap, uint64_t *n) {
    if (!h) return fail(BRISK_E_PARAM, "null handle");
    uint64_t nk = 0;
    int rc = brisk_dump(h, nullptr, nullptr, 0, &nk);
    if (rc) return rc;
    if (n) *n = nk;
    if (!keys && !data) return BRISK_OK;
    if (cap < nk) return fail(BRISK_E_CAPACITY, "dump buffers too small");
    if (h->cap + 64 > (1ull << 32)) return fail(BRISK_E_CAPACITY, "ordered dump handles tables of fewer than 2^32 slots per shard");
    if (nk == 0) return BRISK_OK;
    std::lock_guard<std::recursive_mutex> lock(h->mu);
    CK(cudaSetDevice(h->device));
    SETTLE_PENDING(h);
    DevBuf perm_a, perm_b, hist, okeys, odata;
    auto cleanup = [&]() { for (DevBuf *b : {&perm_a, &perm_b, &hist, &okeys, &odata}) b->release(); };
    auto body = [&]() -> int {
        u32 *perm = nullptr;
        int r = h->P.key_words == 1 ? sorted_slots_t<1>(h, nk, true, perm_a, perm_b, hist, &perm) : sorted_slots_t<2>(h, nk, true, perm_a, perm_b, hist, &perm);
        if (r) return r;
        const bool k_dev = is_device_ptr(keys), d_dev = is_device_ptr(data);
        KmerIn *ok = nullptr;
        unsigned char *od = nullptr;
        if (keys) { if (k_dev) ok = (KmerIn *)keys; else { CK(okeys.ensure(nk * sizeof(KmerIn))); ok = (KmerIn *)okeys.p; } }
        if (data) { if (d_dev) od = (unsigned char *)data; else { CK(odata.ensure(nk * h->data_size)); od = (unsigned char *)odata.p; } }
        const unsigned blocks = (unsigned)((nk + 255) / 256);
        if (h->P.key_words == 1) { SlotKey<1> K; K.T = view_of(h); K.P = h->P; gather_slots_kernel<1><<<blocks, 256, 0, h->stream>>>(K, perm, nk, ok, od); }
        else { SlotKey<2> K; K.T = view_of(h); K.P = h->P; gather_slots_kernel<2><<<blocks, 256, 0, h->stream>>>(K, perm, nk, ok, od); }
        CK(cudaGetLastError());
        if (keys && !k_dev) CK(cudaMemcpyAsync(keys, ok, nk * sizeof(KmerIn), cudaMemcpyDeviceToHost, h->stream));
        if (data && !d_dev) CK(cudaMemcpyAsync(data, od, nk * h->data_size, cudaMemcpyDeviceToHost, h->stream));
        CK(cudaStreamSynchronize(h->stream));
        return BRISK_OK;
    };
    rc = body();
    cleanup();
    return rc;
}

// the largest number of k-mers stored under one bucket id (the reference's `largest_bucket` counts the SKL records of a
// bucket, which depends on its insertion order, brisk/DenseMenuYo.hpp:262-264; this is the order-independent analogue)
static int largest_bucket_kmers(brisk_handle *h, uint64_t &out) {
    out = 0;
    const uint64_t nk = h->n_keys;
    if (nk == 0 || h->cap + 64 > (1ull << 32)) return BRISK_OK;
    DevBuf perm_a, perm_b, hist;
    auto body = [&]() -> int {
        u32 *perm = nullptr;
        int r = h->P.key_words == 1 ? sorted_slots_t<1>(h, nk, false, perm_a, perm_b, hist, &perm) : sorted_slots_t<2>(h, nk, false, perm_a, perm_b, hist, &perm);
        if (r) return r;
        CK(cudaMemsetAsync(h->counters + 4, 0, sizeof(uint64_t), h->stream));
        const unsigned blocks = (unsigned)((nk + 255) / 256);
        if (h->P.key_words == 1) { KeyBucketIdx<1> K; K.T = view_of(h); K.P = h->P; bucket_runs_kernel<1><<<blocks, 256, 0, h->stream>>>(K, perm, nk, h->counters + 4); }
        else { KeyBucketIdx<2> K; K.T = view_of(h); K.P = h->P; bucket_runs_kernel<2><<<blocks, 256, 0, h->stream>>>(K, perm, nk, h->counters + 4); }
        CK(cudaGetLastError());
        CK(cudaMemcpyAsync(&out, h->counters + 4, sizeof(uint64_t), cudaMemcpyDeviceToHost, h->stream));
        CK(cudaStreamSynchronize(h->stream));
        return BRISK_OK;
    };
    int rc = body();
    for (DevBuf *b : {&perm_a, &perm_b, &hist}) b->release();
    return rc;
}

// ---- KFF writer: BriskWriter::write (brisk/writer.hpp:75-179) for this handle's shard
static void kff_be(std::vector<unsigned char> &v, int bytes, uint64_t x) {
    for (int i = bytes - 1; i >= 0; i--) v.push_back((unsigned char)(x >> (8 * i)));
}
static void kff_gv_section(std::vector<unsigned char> &v, const std::vector<std::pair<std::string, uint64_t>> &vars) {
    // Section_GV::close (brisk/lib/kff/kff_io.cpp:786-803): 'v', number of variables, then name\0 + value in the order of
    // the std::map it keeps them in (callers pass them sorted by name)
    v.push_back('v');
    kff_be(v, 8, vars.size());
    for (auto &kv : vars) {
        for (char ch : kv.first) v.push_back((unsigned char)ch);
        v.push_back(0);
        kff_be(v, 8, kv.second);
    }
}
static int ceil_log2_u64(uint64_t x) {  // ceil(log2(x)) as the KFF library computes field widths (kff_io.cpp:1063-1069)
    int b = 0;
    while (((uint64_t)1 << b) < x) b++;
    return b;
}

template <int KW>
static int write_kff_t(brisk_handle *h, FILE *f, uint64_t &n_blocks, uint64_t &n_sections) {
    const BriskParams &P = h->P;
    KffGeom G;
    G.k = P.k; G.m = P.m; G.data_size = (int)h->data_size;
    const uint64_t kff_max = 2 * (uint64_t)(P.k - P.m);  // writer.hpp:92
    G.nb_kmers_bytes = (ceil_log2_u64(kff_max) + 7) / 8;
    G.mini_pos_bytes = (ceil_log2_u64((uint64_t)P.k + kff_max - 1) + 7) / 8;
    G.mini_bytes = (P.m + 3) / 4;
    // file header (Kff_file::open 'w', write_encoding(0, 1, 3, 2), write_metadata: writer.hpp:23-31)
    std::vector<unsigned char> head = {'K', 'F', 'F', 1, 0, (unsigned char)((0 << 6) | (1 << 4) | (3 << 2) | 2), 0, 0};
    const std::string meta = "File generated with Brisk v1. See https://github.com/Malfoy/Brisk";
    kff_be(head, 4, meta.size());
    head.insert(head.end(), meta.begin(), meta.end());
    std::vector<std::pair<uint64_t, char>> sections;  // absolute position, type
    sections.push_back({head.size(), 'v'});
    kff_gv_section(head, {{"data_size", h->data_size}, {"k", (uint64_t)P.k}, {"max", 1}});            // writer.hpp:78-82
    sections.push_back({head.size(), 'v'});
    kff_gv_section(head, {{"data_size", h->data_size}, {"k", (uint64_t)P.k}, {"m", (uint64_t)P.m}, {"max", kff_max}});  // :86-91
    const uint64_t body0 = head.size();

    const uint64_t nk = h->n_keys;
    DevBuf link, chains, perm_a, perm_b, hist, sizes, first, offs, ords, scan_tmp, body, secpos;
    auto cleanup = [&]() { for (DevBuf *b : {&link, &chains, &perm_a, &perm_b, &hist, &sizes, &first, &offs, &ords, &scan_tmp, &body, &secpos}) b->release(); };
    std::vector<unsigned char> h_body;
    std::vector<uint64_t> h_secpos;
    uint64_t n_chains = 0, body_bytes = 0;
    auto device_part = [&]() -> int {
        if (nk == 0) return BRISK_OK;
        if (h->cap >= (1ull << 40)) return fail(BRISK_E_CAPACITY, "KFF writer handles tables of fewer than 2^40 slots");
        StoreView T = view_of(h);
        CK(link.ensure((h->cap + 1) * sizeof(u64)));
        CK(chains.ensure((nk + 1) * sizeof(KffChain)));
        const unsigned blocks = (unsigned)std::min<uint64_t>((h->cap + 255) / 256, 148 * 16);
        kff_links_kernel<KW><<<blocks, 256, 0, h->stream>>>(P, T, (u64 *)link.p);
        kff_edges_kernel<<<blocks, 256, 0, h->stream>>>((u64 *)link.p, h->cap);
        CK(cudaMemsetAsync(h->counters + 3, 0, sizeof(uint64_t), h->stream));
        kff_chains_kernel<KW><<<blocks, 256, 0, h->stream>>>(P, T, (const u64 *)link.p, 64, (KffChain *)chains.p, nk, h->counters + 3);
        CK(cudaGetLastError());
        CK(cudaMemcpyAsync(&n_chains, h->counters + 3, sizeof(uint64_t), cudaMemcpyDeviceToHost, h->stream));
        CK(cudaStreamSynchronize(h->stream));
        if (n_chains == 0 || n_chains > nk) return fail(BRISK_E_CUDA, "KFF writer: inconsistent chain count");
        if (n_chains >= (1ull << 32)) return fail(BRISK_E_CAPACITY, "KFF writer handles fewer than 2^32 superkmers per shard");
        // order the chains by bucket id, then minimizer
        CK(perm_a.ensure(n_chains * sizeof(u32)));
        CK(perm_b.ensure(n_chains * sizeof(u32)));
        CK(hist.ensure(radix_sort_hist_bytes(n_chains) + 1024));
        const unsigned cb = (unsigned)((n_chains + 255) / 256);
        sort_iota_kernel<<<cb, 256, 0, h->stream>>>((u32 *)perm_a.p, n_chains);
        KffChainKey ck; ck.chains = (const KffChain *)chains.p;
        u32 *perm = nullptr;
        CK(radix_sort_perm(ck, (u32 *)perm_a.p, (u32 *)perm_b.p, n_chains, 0, 2 * P.m, (u32 *)hist.p, h->stream, &perm));
        // layout: bytes of every chain (+ section header where a section opens), section ordinals
        for (DevBuf *b : {&sizes, &first, &offs, &ords}) CK(b->ensure((n_chains + 1) * sizeof(u64)));
        kff_sizes_kernel<<<cb, 256, 0, h->stream>>>(G, (const KffChain *)chains.p, perm, n_chains, (u64 *)sizes.p, (u64 *)first.p);
        CK(cudaMemsetAsync((u64 *)sizes.p + n_chains, 0, sizeof(u64), h->stream));
        CK(cudaMemsetAsync((u64 *)first.p + n_chains, 0, sizeof(u64), h->stream));
        size_t tmp = 0;
        CK(cub::DeviceScan::ExclusiveSum(nullptr, tmp, (u64 *)sizes.p, (u64 *)offs.p, (int)(n_chains + 1), h->stream));
        CK(scan_tmp.ensure(tmp));
        CK(cub::DeviceScan::ExclusiveSum(scan_tmp.p, tmp, (u64 *)sizes.p, (u64 *)offs.p, (int)(n_chains + 1), h->stream));
        CK(cub::DeviceScan::ExclusiveSum(scan_tmp.p, tmp, (u64 *)first.p, (u64 *)ords.p, (int)(n_chains + 1), h->stream));
        uint64_t tot[2] = {0, 0};
        CK(cudaMemcpyAsync(&tot[0], (u64 *)offs.p + n_chains, sizeof(u64), cudaMemcpyDeviceToHost, h->stream));
        CK(cudaMemcpyAsync(&tot[1], (u64 *)ords.p + n_chains, sizeof(u64), cudaMemcpyDeviceToHost, h->stream));
        CK(cudaStreamSynchronize(h->stream));
        body_bytes = tot[0];
        n_sections = tot[1];
        CK(body.ensure(body_bytes + 16));
        CK(secpos.ensure((n_sections + 1) * sizeof(u64)));
        kff_emit_kernel<KW><<<(unsigned)((n_chains + 127) / 128), 128, 0, h->stream>>>(P, G, T, (const u64 *)link.p, (const KffChain *)chains.p, perm,
                                                                                       n_chains, (const u64 *)offs.p, (const u64 *)ords.p,
                                                                                       (unsigned char *)body.p, (u64 *)secpos.p);
        CK(cudaGetLastError());
        h_body.resize(body_bytes);
        h_secpos.resize(n_sections);
        CK(cudaMemcpyAsync(h_body.data(), body.p, body_bytes, cudaMemcpyDeviceToHost, h->stream));
        CK(cudaMemcpyAsync(h_secpos.data(), secpos.p, n_sections * sizeof(u64), cudaMemcpyDeviceToHost, h->stream));
        CK(cudaStreamSynchronize(h->stream));
        return BRISK_OK;
    };
    int rc = device_part();
    cleanup();
    if (rc) return rc;
    n_blocks = n_chains;
    for (uint64_t p : h_secpos) sections.push_back({body0 + p, 'm'});
    // footer (Kff_file::write_footer, kff_io.cpp:550-568): the index of every section, positions relative to the end of
    // the index section, then a value section naming it, then the closing signature
    std::vector<unsigned char> tail;
    const uint64_t index_begin = body0 + body_bytes;
    const int64_t index_end = (int64_t)(index_begin + 17 + 9 * sections.size());
    tail.push_back('i');
    kff_be(tail, 8, sections.size());
    for (auto &sp : sections) {
        tail.push_back((unsigned char)sp.second);
        kff_be(tail, 8, (uint64_t)((int64_t)sp.first - index_end));
    }
    kff_be(tail, 8, 0);  // next_index
    kff_gv_section(tail, {{"first_index", index_begin}, {"footer_size", 9 + 2 * (12 + 8)}});
    tail.push_back('K'); tail.push_back('F'); tail.push_back('F');
    if (fwrite(head.data(), 1, head.size(), f) != head.size() || (body_bytes && fwrite(h_body.data(), 1, body_bytes, f) != body_bytes) ||
        fwrite(tail.data(), 1, tail.size(), f) != tail.size())
        return fail(BRISK_E_PARAM, "KFF writer: short write");
    return BRISK_OK;
}

extern "C" int brisk_write_kff(brisk_t *h, const char *path, uint64_t *n_blocks_out, uint64_t *n_sections_out) {
    if (!h || !path) return fail(BRISK_E_PARAM, "null argument");
    std::lock_guard<std::recursive_mutex> lock(h->mu);
    CK(cudaSetDevice(h->device));
    SETTLE_PENDING(h);
    CK(cudaStreamSynchronize(h->stream));
    FILE *f = fopen(path, "wb");
    if (!f) return fail(BRISK_E_PARAM, std::string("cannot open ") + path);
    uint64_t nb = 0, ns = 0;
    int rc = h->P.key_words == 1 ? write_kff_t<1>(h, f, nb, ns) : write_kff_t<2>(h, f, nb, ns);
    if (fclose(f) != 0 && rc == BRISK_OK) rc = fail(BRISK_E_PARAM, "KFF writer: close failed");
    if (n_blocks_out) *n_blocks_out = nb;
    if (n_sections_out) *n_sections_out = ns;
    return rc;
}

extern "C" int brisk_stats(brisk_t *h, uint64_t *nb_buckets, uint64_t *nb_skmers, uint64_t *nb_kmers, uint64_t *device_bytes,
                           uint64_t *largest_bucket) {
    if (!h) return fail(BRISK_E_PARAM, "null handle");
    std::lock_guard<std::recursive_mutex> lock(h->mu);
    CK(cudaSetDevice(h->device));
    SETTLE_PENDING(h);
    uint64_t nb = 0;
    CK(cudaMemsetAsync(h->counters + 4, 0, sizeof(uint64_t), h->stream));
    unsigned blocks = (unsigned)std::min<uint64_t>((h->bitmap_words + 255) / 256, 148 * 16);
    popcount_kernel<<<blocks, 256, 0, h->stream>>>(h->bucket_bits, h->bitmap_words, h->counters + 4);
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(&nb, h->counters + 4, sizeof(uint64_t), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    uint64_t v[4] = {nb, h->n_new_recs, h->n_keys,
                     h->cap * slot_bytes(h) + (h->cap + 64) * 4 + h->bitmap_words * 4 + h->ovf_cap * 24 + (h->data ? (h->cap + 64) * h->data_size : 0)};
    if (h->comm && h->comm->n > 1) {  // buckets are disjoint over the ranks: the totals are sums
        u64 *tmp = part_matrix(h);
        CK(cudaMemcpyAsync(tmp, v, sizeof(v), cudaMemcpyHostToDevice, h->stream));
        NCK(g_nccl.AllReduce(tmp, tmp, 4, ncclUint64, ncclSum, h->comm->comm, h->stream));
        CK(cudaMemcpyAsync(v, tmp, sizeof(v), cudaMemcpyDeviceToHost, h->stream));
        CK(cudaStreamSynchronize(h->stream));
    }
    if (nb_buckets) *nb_buckets = v[0];
    if (nb_skmers) *nb_skmers = v[1];
    if (nb_kmers) *nb_kmers = v[2];
    if (largest_bucket) {
        uint64_t lb = 0;
        int rc = largest_bucket_kmers(h, lb);
        if (rc) return rc;
        if (h->comm && h->comm->n > 1) {  // buckets are disjoint over the ranks: the maximum of the shards' maxima
            u64 *tmp = part_matrix(h);
            CK(cudaMemcpyAsync(tmp, &lb, sizeof(lb), cudaMemcpyHostToDevice, h->stream));
            NCK(g_nccl.AllReduce(tmp, tmp, 1, ncclUint64, ncclMax, h->comm->comm, h->stream));
            CK(cudaMemcpyAsync(&lb, tmp, sizeof(lb), cudaMemcpyDeviceToHost, h->stream));
            CK(cudaStreamSynchronize(h->stream));
        }
        *largest_bucket = lb;
    }
    if (device_bytes) *device_bytes = v[3];
    return BRISK_OK;
}

// order-independent checksum of the stored set of this handle (this rank's shard): out[0] = sum of a 64-bit mix of every
// (kmer_s, minimizer_idx, DATA byte), out[1] = sum of the DATA bytes, out[2] = number of stored k-mers
extern "C" int brisk_checksum(brisk_t *h, uint64_t out[3]) {
    if (!h || !out) return fail(BRISK_E_PARAM, "null argument");
    std::lock_guard<std::recursive_mutex> lock(h->mu);
    CK(cudaSetDevice(h->device));
    SETTLE_PENDING(h);
    u64 *acc = part_matrix(h);  // scratch words
    CK(cudaMemsetAsync(acc, 0, 3 * sizeof(uint64_t), h->stream));
    unsigned blocks = (unsigned)std::min<uint64_t>((h->cap + 255) / 256, 148 * 16);
    if (h->P.key_words == 1) checksum_kernel<1><<<blocks, 256, 0, h->stream>>>(view_of(h), 64, acc);
    else checksum_kernel<2><<<blocks, 256, 0, h->stream>>>(view_of(h), 64, acc);
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(out, acc, 3 * sizeof(uint64_t), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    return BRISK_OK;
}

// ---- synthetic reads on the device (synth.cuh): benchmark / test support.  `genome` and `out` are device pointers on
// the current device; the kernels run on `stream` (cudaStream_t, NULL = the default stream) and are not waited for.
extern "C" int brisk_synth_genome(void *genome, uint64_t G, uint64_t seed, void *stream) {
    if (!genome || !is_device_ptr(genome)) return fail(BRISK_E_PARAM, "genome must be a device pointer");
    if (G == 0) return BRISK_OK;
    const uint64_t threads = (G + 15) / 16;
    synth_genome_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, (cudaStream_t)stream>>>((unsigned char *)genome, G, seed);
    CK(cudaGetLastError());
    return BRISK_OK;
}
extern "C" int brisk_synth_reads(const void *genome, uint64_t G, uint32_t L, uint64_t seed, uint64_t first, uint64_t count, void *out,
                                 void *stream) {
    if (!genome || !out || !is_device_ptr(genome) || !is_device_ptr(out)) return fail(BRISK_E_PARAM, "genome and out must be device pointers");
    if (L == 0 || L > G) return fail(BRISK_E_PARAM, "need 0 < L <= G");
    if (count == 0) return BRISK_OK;
    synth_reads_kernel<<<(unsigned)((count * 32 + 255) / 256), 256, 0, (cudaStream_t)stream>>>((const unsigned char *)genome, G, L, seed, first, count,
                                                                                                (unsigned char *)out);
    CK(cudaGetLastError());
    return BRISK_OK;
}

// Brisk::reallocate (brisk/Brisk.hpp:202-224): the handle rebuilt under (k, new_m, new_b) — the reference always goes to
// (m + 2, b + 2).  See reparam.cuh for the semantics.
extern "C" int brisk_reparameterise(brisk_t *h, uint8_t new_m, uint8_t new_b) {
    if (!h) return fail(BRISK_E_PARAM, "null handle");
    std::lock_guard<std::recursive_mutex> lock(h->mu);
    CK(cudaSetDevice(h->device));
    SETTLE_PENDING(h);
    if (h->comm && h->comm->n > 1) return fail(BRISK_E_PARAM, "re-parameterisation of a sharded index is not supported (the owner of a k-mer changes with its bucket)");
    if (h->cap >= (1ull << 40)) return fail(BRISK_E_CAPACITY, "table too large");
    BriskParams P2;
    if (fill_params(P2, h->P.k, new_m, new_b, h->flags) != BRISK_OK) return fail(BRISK_E_PARAM, "bad parameters: need m odd, m < k, m <= 31, 1 <= b <= min(m, 16)");
    if (P2.rec_words > BRISK_MAX_RW) return fail(BRISK_E_PARAM, "bad parameters");
    CK(cudaStreamSynchronize(h->stream));
    // new grouped decycling tables, bucket bitmap, table of the same capacity, priorities
    std::vector<int2> tab = decycle_group_table(new_m);
    int2 *new_dec = nullptr;
    CK(cudaMalloc((void **)&new_dec, tab.size() * sizeof(int2)));
    CK(cudaMemcpy(new_dec, tab.data(), tab.size() * sizeof(int2), cudaMemcpyHostToDevice));
    P2.dec_tab = new_dec;
    P2.dec_groups = (new_m + 2) / 3;
    const uint64_t new_words = (((uint64_t)1 << (2 * new_b)) + 31) / 32;
    uint32_t *new_bits = nullptr, *new_special = nullptr;
    CK(cudaMalloc((void **)&new_bits, new_words * sizeof(uint32_t)));
    CK(cudaMemsetAsync(new_bits, 0, new_words * sizeof(uint32_t), h->stream));
    if (h->flags & BRISK_F_MANAGED) CK(cudaMallocManaged((void **)&new_special, 64 * sizeof(uint32_t)));
    else CK(cudaMalloc((void **)&new_special, 64 * sizeof(uint32_t)));
    CK(cudaMemsetAsync(new_special, 0, 64 * sizeof(uint32_t), h->stream));
    void *ns = nullptr;
    unsigned char *nd = nullptr;
    uint32_t *ncnt = nullptr;
    int rc = alloc_table(h, h->cap, &ns, &nd, &ncnt);
    if (rc) return rc;
    DevBuf prio;
    CK(prio.ensure((h->cap + 64) * sizeof(u64)));
    CK(cudaMemsetAsync(prio.p, 0, (h->cap + 64) * sizeof(u64), h->stream));
    CK(cudaMemsetAsync(h->counters, 0, 3 * sizeof(uint64_t), h->stream));
    StoreView from = view_of(h), to = from;
    to.slots = ns; to.data = nd; to.cnt = ncnt; to.special = new_special; to.bucket_bits = new_bits;
    const unsigned blocks = (unsigned)std::min<uint64_t>((h->cap + 255) / 256, 148 * 16);
    if (h->P.key_words == 1) reparam_insert_kernel<1><<<blocks, 256, 0, h->stream>>>(P2, from, to, 64, (u64 *)prio.p);
    else reparam_insert_kernel<2><<<blocks, 256, 0, h->stream>>>(P2, from, to, 64, (u64 *)prio.p);
    reparam_data_kernel<<<blocks, 256, 0, h->stream>>>(from, to, 64, (const u64 *)prio.p);
    CK(cudaGetLastError());
    uint64_t c[3];
    CK(cudaMemcpyAsync(c, h->counters, sizeof(c), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    prio.release();
    if (c[1]) return fail(BRISK_E_FULL, "re-parameterisation could not seat every k-mer");  // (same capacity, fewer or as many keys: not expected)
    cudaFree(h->slots); cudaFree(h->cnt); if (h->data) cudaFree(h->data);
    cudaFree(h->special); cudaFree(h->bucket_bits); cudaFree(h->dec_tab);
    h->slots = ns; h->data = nd; h->cnt = ncnt; h->special = new_special; h->bucket_bits = new_bits; h->bitmap_words = new_words;
    h->dec_tab = new_dec;
    h->P = P2;
    h->n_keys = c[0];
    h->n_new_recs = 0;
    h->use_ring = h->P.k <= 32;
    return BRISK_OK;
}

// long-read path (enumerate.cuh: EnumSeg): out[0] = reads longer than this many bases are enumerated in segments,
// out[1] = k-mers of warm-up, out[2] = segmented launches that had to be repeated with a longer warm-up so far
extern "C" int brisk_long_read_info(brisk_t *h, uint64_t out[3]) {
    if (!h || !out) return fail(BRISK_E_PARAM, "null argument");
    out[0] = h->long_limit; out[1] = h->seg_warm; out[2] = h->seg_retries;
    return BRISK_OK;
}

extern "C" int brisk_reserve(brisk_t *h, uint64_t n_keys) {
    if (!h) return fail(BRISK_E_PARAM, "null handle");
    std::lock_guard<std::recursive_mutex> lock(h->mu);
    CK(cudaSetDevice(h->device));
    SETTLE_PENDING(h);
    h->hint = std::max(h->hint, n_keys);
    int rc = grow_to(h, pow2ceil(std::max<uint64_t>(n_keys * 2, 1u << 16)));
    if (rc) return rc;
    CK(cudaStreamSynchronize(h->stream));
    return BRISK_OK;
}
extern "C" int brisk_capacity(brisk_t *h, uint64_t *slots) {
    if (!h || !slots) return fail(BRISK_E_PARAM, "null argument");
    std::lock_guard<std::recursive_mutex> lock(h->mu);
    CK(cudaSetDevice(h->device));
    SETTLE_PENDING(h);
    *slots = h->cap;
    return BRISK_OK;
}
extern "C" int brisk_sync(brisk_t *h) {
    if (!h) return fail(BRISK_E_PARAM, "null handle");
    std::lock_guard<std::recursive_mutex> lock(h->mu);
    CK(cudaSetDevice(h->device));
    SETTLE_PENDING(h);
    CK(cudaStreamSynchronize(h->stream));
    return BRISK_OK;
}
extern "C" int brisk_set_stream(brisk_t *h, void *stream) {
    if (!h) return fail(BRISK_E_PARAM, "null handle");
    std::lock_guard<std::recursive_mutex> lock(h->mu);
    CK(cudaSetDevice(h->device));
    SETTLE_PENDING(h);
    CK(cudaStreamSynchronize(h->stream));
    h->stream = stream ? (cudaStream_t)stream : h->own_stream;
    return BRISK_OK;
}
extern "C" int brisk_last_timing(brisk_t *h, float ms[5], uint32_t *n_launches) {
    if (!h) return fail(BRISK_E_PARAM, "null handle");
    std::lock_guard<std::recursive_mutex> lock(h->mu);
    if (h->timing_dirty) {
        CK(cudaSetDevice(h->device));
        finish_timing(h);
    }
    if (ms) memcpy(ms, h->last_ms, sizeof(h->last_ms));
    if (n_launches) *n_launches = h->last_launches;
    return BRISK_OK;
}

// ---- FASTA ingest: getLineFasta + clean_dna of apps/counter.cpp:128-190 on the device
// parses `text` into h->d_fa_bases (upper-cased ACGT runs, dense) and h->d_fa_offsets (n_reads + 1 run starts)
static int parse_fasta(brisk_handle *h, const char *text, uint64_t n, uint64_t &n_bases, uint64_t &n_reads) {
    n_bases = n_reads = 0;
    if (n >= 0xfffffffeull) return fail(BRISK_E_PARAM, "FASTA text must be shorter than 2^32 - 2 bytes per call (split at a record boundary)");
    CK(h->d_fa_offsets.ensure((n / 2 + 2) * sizeof(uint64_t)));
    if (n == 0) {
        CK(cudaMemsetAsync(h->d_fa_offsets.p, 0, sizeof(uint64_t), h->stream));
        CK(h->d_fa_bases.ensure(16));
        return BRISK_OK;
    }
    const unsigned char *d_text = (const unsigned char *)text;
    if (!is_device_ptr(text)) {
        CK(h->d_fa_text.ensure(n));
        CK(cudaMemcpyAsync(h->d_fa_text.p, text, n, cudaMemcpyHostToDevice, h->stream));
        d_text = (const unsigned char *)h->d_fa_text.p;
    }
    CK(h->d_fa_bases.ensure(n + 16));
    const uint64_t n_tiles = (n + FA_TILE - 1) / FA_TILE;
    CK(h->d_fa_a.ensure(n_tiles * sizeof(FaSummary)));
    CK(h->d_fa_b.ensure(n_tiles * sizeof(FaTileEntry)));
    u64 *totals = part_matrix(h);  // two scratch words
    CK(h->d_fa_chunks.ensure(n_tiles * FA_THREADS * sizeof(uint4)));
    fasta_summarize_kernel<<<(unsigned)n_tiles, FA_THREADS, 0, h->stream>>>(d_text, n, (FaSummary *)h->d_fa_a.p, (uint4 *)h->d_fa_chunks.p);
    fasta_tiles_kernel<<<1, 1024, 0, h->stream>>>((const FaSummary *)h->d_fa_a.p, (u32)n_tiles, (FaTileEntry *)h->d_fa_b.p, totals,
                                                  (u64 *)h->d_fa_offsets.p);
    fasta_emit_kernel<<<(unsigned)n_tiles, FA_THREADS, 0, h->stream>>>(d_text, n, (const FaTileEntry *)h->d_fa_b.p, (const uint4 *)h->d_fa_chunks.p,
                                                                       (unsigned char *)h->d_fa_bases.p, (u64 *)h->d_fa_offsets.p);
    CK(cudaGetLastError());
    h->last_launches += 3;
    uint64_t tot[2];
    CK(cudaMemcpyAsync(tot, totals, sizeof(tot), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    n_bases = tot[0];
    n_reads = tot[1];
    return BRISK_OK;
}

extern "C" int brisk_parse_fasta(brisk_t *h, const char *text, uint64_t n_bytes, char *bases, uint64_t cap_bases, uint64_t *offsets,
                                 uint64_t cap_reads, uint64_t *n_bases_out, uint64_t *n_reads_out) {
    if (!h) return fail(BRISK_E_PARAM, "null handle");
    if (n_bytes && !text) return fail(BRISK_E_PARAM, "null input");
    std::lock_guard<std::recursive_mutex> lock(h->mu);
    CK(cudaSetDevice(h->device));
    SETTLE_PENDING(h);
    uint64_t nb = 0, nr = 0;
    int rc = parse_fasta(h, text, n_bytes, nb, nr);
    if (rc) return rc;
    if (n_bases_out) *n_bases_out = nb;
    if (n_reads_out) *n_reads_out = nr;
    if (bases) {
        if (cap_bases < nb) return fail(BRISK_E_CAPACITY, "bases buffer too small");
        CK(cudaMemcpyAsync(bases, h->d_fa_bases.p, nb, is_device_ptr(bases) ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, h->stream));
    }
    if (offsets) {
        if (cap_reads < nr) return fail(BRISK_E_CAPACITY, "offsets buffer too small (needs n_reads + 1 entries)");
        CK(cudaMemcpyAsync(offsets, h->d_fa_offsets.p, (nr + 1) * sizeof(uint64_t),
                           is_device_ptr(offsets) ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, h->stream));
    }
    CK(cudaStreamSynchronize(h->stream));
    return BRISK_OK;
}

// A text of any size is handed to the parser in pieces of about 1 GB that end right before a header line ("\n>"): every
// piece starts with a header like a file does, so the pieces parse independently and no read spans two of them (the
// 32-bit positions of the tile scan cap one parse at 4 GB; the pieces also bound the scratch memory of a call).
static int fasta_pieces(brisk_handle *h, const char *text, uint64_t n, std::vector<std::pair<uint64_t, uint64_t>> &pieces) {
    const uint64_t target = 1ull << 30, window = 4ull << 20;
    const bool dev = is_device_ptr(text);
    std::vector<char> buf;
    uint64_t at = 0;
    while (n - at > target + target / 2) {
        // the last "\n>" at or before at + target, looked for in growing windows
        uint64_t cut = 0;
        for (uint64_t back = window; cut == 0; back *= 4) {
            const uint64_t hi = at + target, lo = hi > at + back ? hi - back : at + 1;
            const char *w = text + lo;
            if (dev) {
                buf.resize(hi - lo + 1);
                CK(cudaMemcpy(buf.data(), text + lo, hi - lo + 1, cudaMemcpyDeviceToHost));
                w = buf.data();
            }
            for (uint64_t i = hi - lo; i >= 1; i--)
                if (w[i] == '>' && w[i - 1] == '\n') { cut = lo + i; break; }
            if (lo == at + 1) break;
        }
        if (cut == 0) return fail(BRISK_E_PARAM, "FASTA text: no record boundary within 1 GB (one record of more than 1 GB is not supported)");
        pieces.push_back({at, cut - at});
        at = cut;
    }
    pieces.push_back({at, n - at});
    (void)h;
    return BRISK_OK;
}

// count_fasta / query_fasta of the demo app (apps/counter.cpp:197-226, :314-345) for one text buffer
extern "C" int brisk_count_fasta(brisk_t *h, const char *text, uint64_t n_bytes, uint64_t *n_reads_out) {
    if (!h) return fail(BRISK_E_PARAM, "null handle");
    if (n_bytes && !text) return fail(BRISK_E_PARAM, "null input");
    std::lock_guard<std::recursive_mutex> lock(h->mu);
    CK(cudaSetDevice(h->device));
    SETTLE_PENDING(h);
    std::vector<std::pair<uint64_t, uint64_t>> pieces;
    int rc = fasta_pieces(h, text, n_bytes, pieces);
    if (rc) return rc;
    uint64_t reads = 0;
    for (auto &pc : pieces) {
        uint64_t nb = 0, nr = 0;
        rc = parse_fasta(h, text + pc.first, pc.second, nb, nr);
        if (rc) return rc;
        reads += nr;
        rc = brisk_count_reads(h, (const char *)h->d_fa_bases.p, (const uint64_t *)h->d_fa_offsets.p, nr);
        if (rc) return rc;
    }
    if (n_reads_out) *n_reads_out = reads;
    return BRISK_OK;
}
extern "C" int brisk_query_fasta(brisk_t *h, const char *text, uint64_t n_bytes, uint64_t *hits, uint64_t *total) {
    if (!h) return fail(BRISK_E_PARAM, "null handle");
    if (n_bytes && !text) return fail(BRISK_E_PARAM, "null input");
    std::lock_guard<std::recursive_mutex> lock(h->mu);
    CK(cudaSetDevice(h->device));
    SETTLE_PENDING(h);
    std::vector<std::pair<uint64_t, uint64_t>> pieces;
    int rc = fasta_pieces(h, text, n_bytes, pieces);
    if (rc) return rc;
    uint64_t H = 0, T = 0;
    for (auto &pc : pieces) {
        uint64_t nb = 0, nr = 0, h1 = 0, t1 = 0;
        rc = parse_fasta(h, text + pc.first, pc.second, nb, nr);
        if (rc) return rc;
        rc = brisk_query_reads(h, (const char *)h->d_fa_bases.p, (const uint64_t *)h->d_fa_offsets.p, nr, &h1, &t1);
        if (rc) return rc;
        H += h1; T += t1;
    }
    if (hits) *hits = H;
    if (total) *total = T;
    return BRISK_OK;
}

// ---- partition and record-level entry points (what a rank does before and after the exchange)
extern "C" int brisk_partition_reads(brisk_t *h, const char *bases, const uint64_t *offsets, uint64_t n_reads, uint32_t n_ranks,
                                     uint64_t *recs, uint64_t cap_recs, uint64_t *starts, uint32_t *rec_words) {
    if (!h) return fail(BRISK_E_PARAM, "null handle");
    if (n_reads && (!bases || !offsets)) return fail(BRISK_E_PARAM, "null input");
    if (n_ranks == 0 || n_ranks > PART_MAX_RANKS) return fail(BRISK_E_PARAM, "n_ranks must be 1..64");
    if (!starts) return fail(BRISK_E_PARAM, "starts is null");
    std::lock_guard<std::recursive_mutex> lock(h->mu);
    CK(cudaSetDevice(h->device));
    SETTLE_PENDING(h);
    h->last_launches = 0;
    if (rec_words) *rec_words = (uint32_t)h->P.rec_words;
    ReadBatch B;
    int rc = stage_reads(h, bases, offsets, n_reads, B);
    if (rc) return rc;
    uint64_t n_recs = 0, n_occ = 0;
    rc = enumerate_batch(h, B, false, n_recs, n_occ);
    if (rc) return rc;
    CK(cudaEventRecord(h->ev[2], h->stream));
    rc = partition_batch(h, n_recs, n_ranks, false);
    if (rc) return rc;
    CK(cudaEventRecord(h->ev[3], h->stream));
    CK(cudaEventRecord(h->ev[4], h->stream));
    CK(cudaMemcpyAsync(starts, part_starts(h), (n_ranks + 1) * sizeof(uint64_t),
                       is_device_ptr(starts) ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, h->stream));
    if (recs) {
        if (n_recs > cap_recs) return fail(BRISK_E_CAPACITY, "record buffer too small");
        CK(cudaMemcpyAsync(recs, h->d_send.p, n_recs * h->P.rec_words * sizeof(uint64_t),
                           is_device_ptr(recs) ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, h->stream));
    }
    CK(cudaStreamSynchronize(h->stream));
    finish_timing(h);
    return BRISK_OK;
}

extern "C" int brisk_apply_records(brisk_t *h, const uint64_t *recs, uint64_t n_recs, int op, uint64_t *hits, uint64_t *total) {
    if (!h) return fail(BRISK_E_PARAM, "null handle");
    if (n_recs && !recs) return fail(BRISK_E_PARAM, "null input");
    if (op < 0 || op > 2) return fail(BRISK_E_PARAM, "op must be 0 (count), 1 (insert) or 2 (get)");
    std::lock_guard<std::recursive_mutex> lock(h->mu);
    CK(cudaSetDevice(h->device));
    SETTLE_PENDING(h);
    h->last_launches = 0;
    const int RW = h->P.rec_words;
    const u64 *d_r = (const u64 *)recs;
    if (n_recs && !is_device_ptr(recs)) {
        CK(h->d_recv.ensure(n_recs * RW * sizeof(uint64_t)));
        CK(cudaMemcpyAsync(h->d_recv.p, recs, n_recs * RW * sizeof(uint64_t), cudaMemcpyHostToDevice, h->stream));
        d_r = (const u64 *)h->d_recv.p;
    }
    for (int i = 0; i < 4; i++) CK(cudaEventRecord(h->ev[i], h->stream));
    if (op != MODE_GET) {
        int rc = pregrow(h, n_recs * (uint64_t)h->P.w);
        if (rc) return rc;
    }
    OccOut O;
    O.rec_occ = nullptr; O.slots = nullptr; O.flags = nullptr; O.data = nullptr; O.agg = nullptr;
    uint64_t agg[2] = {0, 0};
    if (op == MODE_GET) {
        CK(cudaMemsetAsync(h->counters + 5, 0, 2 * sizeof(uint64_t), h->stream));
        O.agg = h->counters + 5;
    }
    CK(launch_records(h, op, d_r, n_recs, O));
    if (n_recs) h->last_launches++;
    CK(cudaEventRecord(h->ev[4], h->stream));
    if (op == MODE_GET) {
        CK(cudaMemcpyAsync(agg, h->counters + 5, sizeof(agg), cudaMemcpyDeviceToHost, h->stream));
        CK(cudaStreamSynchronize(h->stream));
    } else {
        int rc = settle(h);
        if (rc) return rc;
    }
    if (hits) *hits = agg[0];
    if (total) *total = agg[1];
    finish_timing(h);
    return BRISK_OK;
}

// ---- multi-GPU entry points
extern "C" uint32_t brisk_owner_of_bucket(uint64_t bucket_id, uint8_t b, uint32_t n_ranks) {
    (void)b;
    return owner_of_bucket(bucket_id, n_ranks);
}
extern "C" int brisk_exchange_plan(const uint64_t *counts, int n_ranks, int rank, uint64_t *send_start, uint64_t *recv_start) {
    if (!counts || !send_start || !recv_start || n_ranks < 1 || rank < 0 || rank >= n_ranks) return fail(BRISK_E_PARAM, "bad argument");
    exchange_plan(counts, n_ranks, rank, send_start, recv_start);
    return BRISK_OK;
}
// the fixed-point tables of the rolled decycling class (pure host arithmetic; the CPU tests replay the kernel's formula
// against DecyclingSet::memDouble with them)
extern "C" int brisk_decycle_tables(uint8_t k, uint8_t m, uint8_t b, int32_t *tw, int32_t *rot, int32_t *th, int32_t *groups_tab) {
    BriskParams P;
    if (fill_params(P, k, m, b, 0) != BRISK_OK) return fail(BRISK_E_PARAM, "bad parameters");
    for (int q = 0; q < 32; q++) {
        if (tw) { tw[2 * q] = P.dec_tw[q].x; tw[2 * q + 1] = P.dec_tw[q].y; }
        if (rot) { rot[2 * q] = P.dec_rot[q].x; rot[2 * q + 1] = P.dec_rot[q].y; }
    }
    if (th) *th = P.dec_th;
    if (groups_tab) {
        std::vector<int2> tab = decycle_group_table(m);
        memcpy(groups_tab, tab.data(), tab.size() * sizeof(int2));
    }
    return BRISK_OK;
}
extern "C" int brisk_comm_unique_id(uint8_t id[128]) {
    if (!id) return fail(BRISK_E_PARAM, "null id");
    if (!g_nccl.load()) return fail(BRISK_E_COMM, g_nccl.error);
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
    ncclUniqueId u;
    NCK(g_nccl.GetUniqueId(&u));
    memcpy(id, &u, 128);
    return BRISK_OK;
}
extern "C" int brisk_comm_init(brisk_t *h, const uint8_t id[128], int rank, int n_ranks) {
    if (!h || !id) return fail(BRISK_E_PARAM, "null argument");
    if (n_ranks < 1 || n_ranks > PART_MAX_RANKS || rank < 0 || rank >= n_ranks) return fail(BRISK_E_PARAM, "bad rank / n_ranks");
    if (h->comm) return fail(BRISK_E_PARAM, "the handle already has a communicator");
    if (h->n_keys) return fail(BRISK_E_PARAM, "brisk_comm_init must come before the first insertion");
    if (!g_nccl.load()) return fail(BRISK_E_COMM, g_nccl.error);
    std::lock_guard<std::recursive_mutex> lock(h->mu);
    CK(cudaSetDevice(h->device));
    SETTLE_PENDING(h);
    ncclUniqueId u;
    memcpy(&u, id, 128);
    Comm *c = new Comm();
    c->rank = rank;
    c->n = n_ranks;
    ncclResult_t r = g_nccl.CommInitRank(&c->comm, n_ranks, u, rank);
    if (r != ncclSuccess) {
        delete c;
        return fail(BRISK_E_COMM, std::string("ncclCommInitRank: ") + g_nccl.GetErrorString(r));
    }
    {   // highest priority: the exchange's small kernels must get SM slots as soon as CTAs of the enumerate launch that is
        // running on the handle's stream retire, not after the whole launch has been dispatched
        int lo_p = 0, hi_p = 0;
        CK(cudaDeviceGetStreamPriorityRange(&lo_p, &hi_p));
        CK(cudaStreamCreateWithPriority(&c->stream, cudaStreamNonBlocking, hi_p));
        // the enumerate chunks before the last one run on streams of decreasing priority (the last one on the handle's
        // stream): pending CTAs are dispatched chunk after chunk, and the next chunk fills the SMs while one drains
        for (int i = 0; i < 2; i++) CK(cudaStreamCreateWithPriority(&c->aux[i], cudaStreamNonBlocking, std::min(lo_p, hi_p + 1 + i)));
        c->aux_ordered = (lo_p - hi_p) >= 3;
    }
    for (int i = 0; i < COMM_MAX_CHUNKS; i++) CK(cudaEventCreateWithFlags(&c->ev_enum[i], cudaEventDisableTiming));
    CK(cudaEventCreateWithFlags(&c->ev_counts, cudaEventDisableTiming));
    CK(cudaEventCreateWithFlags(&c->ev_done, cudaEventDisableTiming));
    CK(cudaMallocHost((void **)&c->h_pinned, (16 + 65 * 64) * sizeof(uint64_t)));
    CK(cudaMalloc((void **)&c->d_gather, (64 + 65 + 64 + 8 + 64 * 65) * sizeof(uint64_t)));  // partition scratch (totals first) + everybody's rows
    h->comm = c;
    return BRISK_OK;
}
extern "C" int brisk_comm_info(brisk_t *h, int *rank, int *n_ranks) {
    if (!h) return fail(BRISK_E_PARAM, "null handle");
    if (rank) *rank = h->comm ? h->comm->rank : 0;
    if (n_ranks) *n_ranks = h->comm ? h->comm->n : 1;
    return BRISK_OK;
}

// ------------------------------------------------------------------------------------------------
// brisk_checksum_kmers_noalloc — the checksum of brisk_checksum over an explicit list of (k-mer, DATA byte) pairs (a
// dump), for callers that must not allocate (CUDA graph capture, serving loops).  No counterpart in the reference.
// The scratch (a ticket + one partial triple per CTA) is a file-scope
// __device__ array: every GPU that loads the library holds its 384.25 KiB until exit.  Nothing here calls cudaMalloc,
// cudaFree or any allocating routine.  All calls share the array: calls that overlap on one GPU race.
#define NOALLOC_TILE 4096u                        // k-mers per CTA (256 threads x 16)
#define NOALLOC_MAX_CTAS 16384u                   // => at most 2^26 k-mers per call (BRISK_NOALLOC_MAX_KMERS)
#define NOALLOC_HEAD 32u                          // u64 words before the partials; word 0 = ticket
static_assert((uint64_t)NOALLOC_TILE * NOALLOC_MAX_CTAS == BRISK_NOALLOC_MAX_KMERS, "header and kernel disagree on the limit");
static __device__ __align__(256) unsigned long long g_noalloc_scratch[NOALLOC_HEAD + 3 * NOALLOC_MAX_CTAS];

// sums a, b, c over the CTA; the result is valid in thread 0
__device__ __forceinline__ void noalloc_block_sum(u64 &a, u64 &b, u64 &c, u64 (*sh)[3]) {
    for (int d = 16; d > 0; d >>= 1) {
        a += __shfl_xor_sync(0xffffffffu, a, d);
        b += __shfl_xor_sync(0xffffffffu, b, d);
        c += __shfl_xor_sync(0xffffffffu, c, d);
    }
    const int w = threadIdx.x >> 5;
    if ((threadIdx.x & 31) == 0) { sh[w][0] = a; sh[w][1] = b; sh[w][2] = c; }
    __syncthreads();
    if (threadIdx.x == 0) {
        a = b = c = 0;
        for (int i = 0; i < 8; i++) { a += sh[i][0]; b += sh[i][1]; c += sh[i][2]; }
    }
    __syncthreads();
}

// keys as u64 triples {lo, hi, idx | pad << 8} (brisk_kmer); every CTA leaves its partial sums in the scratch, the CTA
// that takes the last ticket adds them up and writes out[0..2]
__global__ void __launch_bounds__(256) checksum_kmers_noalloc_kernel(const u64 *__restrict__ keys, const unsigned char *__restrict__ data,
                                                                     u64 n, u64 *scratch, u64 *out) {
    __shared__ u64 sh[8][3];
    __shared__ bool last;
    u64 *part = scratch + NOALLOC_HEAD;
    u64 a = 0, b = 0, c = 0;
    const u64 base = (u64)blockIdx.x * NOALLOC_TILE;
#pragma unroll 4
    for (u32 r = 0; r < NOALLOC_TILE / 256; r++) {
        const u64 i = base + (u64)r * 256 + threadIdx.x;
        if (i < n) {
            const u64 lo = keys[3 * i], hi = keys[3 * i + 1];
            const u32 idx = (u32)(keys[3 * i + 2] & 0xffu), d = data[i];
            a += checksum_mix(lo, hi, idx, d);
            b += d;
            c += 1;
        }
    }
    noalloc_block_sum(a, b, c, sh);
    if (threadIdx.x == 0) {
        part[3 * (u64)blockIdx.x] = a;
        part[3 * (u64)blockIdx.x + 1] = b;
        part[3 * (u64)blockIdx.x + 2] = c;
        __threadfence();
        last = atomicAdd((unsigned long long *)scratch, 1ull) == (unsigned long long)gridDim.x - 1;
    }
    __syncthreads();
    if (!last) return;
    __threadfence();
    a = b = c = 0;
    for (u32 i = threadIdx.x; i < gridDim.x; i += 256) {
        a += __ldcg(part + 3 * (u64)i);
        b += __ldcg(part + 3 * (u64)i + 1);
        c += __ldcg(part + 3 * (u64)i + 2);
    }
    noalloc_block_sum(a, b, c, sh);
    if (threadIdx.x == 0) { out[0] = a; out[1] = b; out[2] = c; }
}

// once per device, with that device current: the DEVICE address of g_noalloc_scratch (its name in host code is only a
// host-side stand-in for the symbol)
static int noalloc_scratch_of_current_device(int dev, u64 **ptr) {
    static std::mutex mu;
    static u64 *addr[64] = {};
    if (dev < 0 || dev >= 64) return fail(BRISK_E_PARAM, "device ordinal out of range");
    std::lock_guard<std::mutex> lock(mu);
    if (!addr[dev]) {
        void *p = nullptr;
        CK(cudaGetSymbolAddress(&p, g_noalloc_scratch));
        if (!p) return fail(BRISK_E_CUDA, "cudaGetSymbolAddress returned a null address");
        addr[dev] = (u64 *)p;
    }
    *ptr = addr[dev];
    return BRISK_OK;
}

extern "C" int brisk_checksum_kmers_noalloc(const brisk_kmer *keys, const uint8_t *data, uint64_t n, uint64_t *out, void *stream) {
    if (!out) return fail(BRISK_E_PARAM, "null output");
    if (n && (!keys || !data)) return fail(BRISK_E_PARAM, "null input");
    if (n > BRISK_NOALLOC_MAX_KMERS)
        return fail(BRISK_E_CAPACITY, "brisk_checksum_kmers_noalloc: " + std::to_string(n) + " k-mers exceed the static scratch's limit of " +
                                          std::to_string((uint64_t)BRISK_NOALLOC_MAX_KMERS) + " per call (split the list and add the sums)");
    int dev = -1;
    CK(cudaGetDevice(&dev));
    const void *ptrs[3] = {out, n ? (const void *)keys : nullptr, n ? (const void *)data : nullptr};
    for (int i = 0; i < 3; i++) {
        if (!ptrs[i]) continue;
        cudaPointerAttributes at;
        if (cudaPointerGetAttributes(&at, ptrs[i]) != cudaSuccess || at.type != cudaMemoryTypeDevice) {
            cudaGetLastError();
            return fail(BRISK_E_PARAM, "keys, data and out must be device pointers");
        }
        if (at.device != dev) return fail(BRISK_E_PARAM, "keys, data and out must live on the current device");
    }
    if ((uintptr_t)keys & 7u || (uintptr_t)out & 7u) return fail(BRISK_E_PARAM, "keys and out must be 8-byte aligned");
    u64 *scratch = nullptr;
    int rc = noalloc_scratch_of_current_device(dev, &scratch);
    if (rc) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    // the array keeps the last call's contents: the ticket must start at zero (the partials are overwritten before use)
    CK(cudaMemsetAsync(scratch, 0, sizeof(u64), s));
    const unsigned ctas = n ? (unsigned)((n + NOALLOC_TILE - 1) / NOALLOC_TILE) : 1u;
    checksum_kmers_noalloc_kernel<<<ctas, 256, 0, s>>>((const u64 *)keys, data, n, scratch, (u64 *)out);
    CK(cudaGetLastError());
    return BRISK_OK;
}
