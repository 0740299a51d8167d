// brisk_b200/Brisk.hpp — C++ host mirror of the reference's insertion-path API over the C ABI of libbrisk_b200.so.
//
// Same names, public fields and argument meaning as the reference so that code written against
//   Parameters(k, m, b)                      brisk/parameters.hpp:9-35
//   kmer_full, kint, SuperKmerEnumerator     brisk/Kmers.hpp:26, :32-58, :62-89
//   Brisk<DATA>                              brisk/Brisk.hpp:23-42
// (apps/counter.cpp:231-310 is the model caller) keeps working, plus the batched calls the GPU path is measured on
// (count_reads / insert_reads / get_reads / query_reads / insert_kmers / get_kmers).  Self-contained: needs only the
// standard library and include/brisk_b200.h; link with -lbrisk_b200.  Nothing here computes k-mers, minimizers or
// lookups on the CPU — the enumerator and the index are both served by the CUDA library; the host code only unpacks
// records and moves pointers.  Without a CUDA device every constructor throws.
//
// Differences from the reference, all in its favour of the caller:
//   * DATA* returned by get / insert_superkmer stay valid until the store is rehashed (brisk_reserve / growth), not
//     only until the next insertion in the same bucket (README.md:109-110); with a capacity_hint they never move.
//   * the legacy pointer calls need host-dereferenceable DATA, so this wrapper creates the handle with
//     BRISK_F_MANAGED unless `device_resident` is asked for (then only the batched calls may be used).
//   * DATA must be trivially copyable, sizeof(DATA) <= 64; new DATA is zero-initialised.
#pragma once
#include <algorithm>
#include <array>
#include <cstdint>
#include <cstring>
#include <mutex>
#include <stdexcept>
#include <string>
#include <type_traits>
#include <vector>

#include "../brisk_b200.h"

typedef __uint128_t kint;
typedef __uint128_t skint;

// opaque in this implementation: the decycling table lives in the CUDA library (brisk/Decycling.h)
class DecyclingSet {
public:
    explicit DecyclingSet(unsigned m) : m(m) {}
    unsigned m;
};

class Parameters {
public:
    uint8_t k, m, b;
    uint8_t m_reduc;
    unsigned allocated_bytes;
    uint8_t compacted_size;
    uint64_t mask_large_minimizer;
    DecyclingSet *dede;
    Parameters(uint8_t k, uint8_t m, uint8_t b) : k(k), m(m), b(b) {
        m_reduc = m - b;
        mask_large_minimizer = m >= 32 ? ~(uint64_t)0 : (((uint64_t)1 << (2 * m)) - 1);
        compacted_size = k - b;
        allocated_bytes = (2u * k - m - b + 3u) / 4u;  // ceil((2k - m - b) / 4), brisk/parameters.hpp:31
        dede = new DecyclingSet(m);                    // never freed, like the reference's
    }
};

class BriskError : public std::runtime_error {
public:
    BriskError(int code, const std::string &what) : std::runtime_error(what), code(code) {}
    int code;
};
inline void brisk_check(int rc) {
    if (rc != BRISK_OK) throw BriskError(rc, std::string("brisk_b200: ") + brisk_last_error());
}

class kmer_full {
public:
    kint kmer_s = 0;            // the k-mer in the orientation in which its minimizer is canonical
    kint minimizer = 0;         // the canonical minimizer m-mer (what the enumerator leaves here, Kmers.cpp:580/:583)
    uint8_t minimizer_idx = 0;  // nucleotide offset of the minimizer from the k-mer's right end
    std::vector<int8_t> interleaved;
    DecyclingSet *dede = nullptr;
    kmer_full() {}
    kmer_full(kint value, uint8_t minimizer_idx, uint8_t minimizer_size, DecyclingSet *dede)
        : kmer_s(value), minimizer_idx(minimizer_idx), dede(dede) {
        compute_mini(minimizer_size);
    }
    void compute_mini(uint8_t m) {
        uint64_t mask = m >= 32 ? ~(uint64_t)0 : (((uint64_t)1 << (2 * m)) - 1);
        uint64_t w = (uint64_t)(kmer_s >> (2 * minimizer_idx)) & mask;
        uint64_t rc = 0, x = w ^ 0xaaaaaaaaaaaaaaaaull;  // complement is ^2 per nucleotide (A=0 C=1 T=2 G=3)
        for (int i = 0; i < m; i++) { rc = (rc << 2) | (x & 3); x >>= 2; }
        minimizer = w < rc ? w : rc;
    }
    void copy(const kmer_full &o) {
        kmer_s = o.kmer_s;
        minimizer = o.minimizer;
        minimizer_idx = o.minimizer_idx;
        dede = o.dede;
    }
    uint8_t suffix_size() const { return minimizer_idx; }
    uint8_t prefix_size(uint8_t k, uint8_t m) const { return k - m - minimizer_idx; }
    brisk_kmer abi() const {
        brisk_kmer a;
        std::memset(&a, 0, sizeof(a));
        a.lo = (uint64_t)kmer_s;
        a.hi = (uint64_t)(kmer_s >> 64);
        a.minimizer_idx = minimizer_idx;
        return a;
    }
};

namespace brisk_b200_detail {
// One process-wide enumeration handle per (k, m): SuperKmerEnumerator has no index argument in the reference, so the
// GPU enumerator it is built on lives here.  b does not influence enumeration (only the bucket id in the record).
struct EnumHandle {
    brisk_t *h = nullptr;
    uint8_t k = 0, m = 0;
    std::mutex mu;
};
inline EnumHandle &enum_handle(uint8_t k, uint8_t m) {
    static std::mutex table_mu;
    static std::vector<EnumHandle *> table;
    std::lock_guard<std::mutex> lock(table_mu);
    for (EnumHandle *e : table)
        if (e->k == k && e->m == m) return *e;
    EnumHandle *e = new EnumHandle();
    e->k = k;
    e->m = m;
    brisk_check(brisk_create(k, m, 1, 1, -1, 1, 0, &e->h));  // b = 1: a 2-bit bucket field, no directory to speak of
    table.push_back(e);
    return *e;
}
inline int index_bits(unsigned w) {
    int wb = 1;
    while ((1u << wb) < w) wb++;
    return wb;
}
// unpacks one superkmer record (layout: DESIGN.md §3) into kmer_full, emission order
inline void decode_record(const uint64_t *rec, unsigned rec_words, uint8_t k, uint8_t m, uint8_t b, DecyclingSet *dede,
                          std::vector<kmer_full> &out) {
    const int wb = index_bits(k - m + 1), sb = 64 - 2 * b, hb = 2 * b + 2 * wb;
    const uint64_t top = rec[rec_words - 1], fm = ((uint64_t)1 << wb) - 1;
    const unsigned n = (unsigned)((top >> (sb - wb)) & fm) + 1, idx0 = (unsigned)((top >> (sb - 2 * wb)) & fm);
    // payload as up to 6 words; k-mer j = payload >> 2(n-1-j), masked to 2k bits
    uint64_t w[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (unsigned i = 0; i < rec_words; i++) w[i] = rec[i];
    w[rec_words - 1] &= hb >= 64 ? 0 : (~(uint64_t)0 >> hb);
    const kint kmask = k >= 64 ? ~(kint)0 : ((((kint)1) << (2 * k)) - 1);
    for (unsigned j = 0; j < n; j++) {
        unsigned sh = 2 * (n - 1 - j), q = sh >> 6, r = sh & 63;
        uint64_t v[3];
        for (int i = 0; i < 3; i++) v[i] = r ? ((w[q + i] >> r) | (w[q + i + 1] << (64 - r))) : w[q + i];
        kint val = (((kint)v[1]) << 64) | v[0];
        out.emplace_back(val & kmask, (uint8_t)(idx0 + j), m, dede);
    }
}
}  // namespace brisk_b200_detail

// SuperKmerEnumerator(seq, k, m, dede) + next(vector<kmer_full>&): brisk/Kmers.cpp:509-603.  The whole sequence is
// enumerated on the GPU at construction (brisk_enumerate_reads_tagged); next() hands the superkmers out in the
// reference's order and returns the canonical minimizer of the superkmer (0 with an empty vector at the end).
class SuperKmerEnumerator {
public:
    SuperKmerEnumerator(std::string &s, const uint8_t k, const uint8_t m, DecyclingSet *dede) : seq(s), dede(dede), k(k), m(m) {
        if (s.size() < k) return;
        brisk_b200_detail::EnumHandle &E = brisk_b200_detail::enum_handle(k, m);
        std::lock_guard<std::mutex> lock(E.mu);
        b_ = 1;
        uint64_t offs[2] = {0, (uint64_t)s.size()}, n = 0;
        brisk_check(brisk_enumerate_reads_tagged(E.h, s.data(), offs, 1, nullptr, 0, &n, &rw_, nullptr));
        recs_.resize(n * rw_);
        std::vector<uint64_t> occ(n);
        brisk_check(brisk_enumerate_reads_tagged(E.h, s.data(), offs, 1, recs_.data(), n, &n, &rw_, occ.data()));
        order_.resize(n);
        for (uint64_t i = 0; i < n; i++) order_[i] = {occ[i] >> 1, i};
        std::sort(order_.begin(), order_.end());
    }
    kint next(std::vector<kmer_full> &kmers) {
        if (cursor_ >= order_.size()) return 0;
        const uint64_t i = order_[cursor_++][1];
        brisk_b200_detail::decode_record(recs_.data() + i * rw_, rw_, k, m, b_, dede, kmers);
        return kmers.empty() ? (kint)0 : kmers.back().minimizer;
    }
    std::string &seq;
    DecyclingSet *dede;
    uint8_t k, m;

private:
    uint8_t b_ = 0;
    uint32_t rw_ = 0;
    std::vector<uint64_t> recs_;
    std::vector<std::array<uint64_t, 2>> order_;
    size_t cursor_ = 0;
};

// Brisk<DATA>: brisk/Brisk.hpp:23-42.  `menu` (the reference's DenseMenuYo*) is replaced by the opaque C handle.
template <class DATA>
class Brisk {
    static_assert(std::is_trivially_copyable<DATA>::value && sizeof(DATA) <= 64, "DATA must be trivially copyable, <= 64 bytes");

public:
    brisk_t *menu = nullptr;
    Parameters params;

    // capacity_hint = expected number of distinct k-mers (0: grow on demand, pointers die when the store is rehashed);
    // device_resident = true keeps the store in device memory (batched calls only, the measured configuration)
    explicit Brisk(Parameters &parameters, uint64_t capacity_hint = 0, bool device_resident = false, int device = -1)
        : params(parameters), managed_(!device_resident) {
        brisk_check(brisk_create(params.k, params.m, params.b, (uint32_t)sizeof(DATA), device, capacity_hint,
                                 managed_ ? BRISK_F_MANAGED : 0u, &menu));
    }
    ~Brisk() { brisk_destroy(menu); }
    Brisk(const Brisk &) = delete;
    Brisk &operator=(const Brisk &) = delete;

    // ---- the reference's calls ------------------------------------------------------------------
    DATA *get(kmer_full &kmer) const {  // brisk/Brisk.hpp:64-69: NULL when absent
        brisk_kmer a = kmer.abi();
        uint64_t slot = BRISK_NO_SLOT;
        brisk_check(brisk_get_kmers(menu, &a, 1, &slot, nullptr));
        return slot == BRISK_NO_SLOT ? nullptr : data_at(slot);
    }
    DATA *insert(kmer_full &kmer) {
        brisk_kmer a = kmer.abi();
        uint64_t slot = BRISK_NO_SLOT;
        uint8_t newly = 0;
        brisk_check(brisk_insert_kmers(menu, &a, 1, &slot, &newly, 0));
        return data_at(slot);
    }
    // brisk/Brisk.hpp:123-147: one DATA* per k-mer, newly_inserted APPENDED to (the caller clears it, apps/counter.cpp:259)
    std::vector<DATA *> insert_superkmer(std::vector<kmer_full> &v, std::vector<bool> &newly_inserted) {
        std::vector<brisk_kmer> a(v.size());
        for (size_t i = 0; i < v.size(); i++) a[i] = v[i].abi();
        std::vector<uint64_t> slots(v.size(), BRISK_NO_SLOT);
        std::vector<uint8_t> newly(v.size(), 0);
        if (!v.empty()) brisk_check(brisk_insert_kmers(menu, a.data(), a.size(), slots.data(), newly.data(), 0));
        std::vector<DATA *> out(v.size());
        for (size_t i = 0; i < v.size(); i++) {
            out[i] = data_at(slots[i]);
            newly_inserted.push_back(newly[i] != 0);
        }
        return out;
    }
    std::vector<DATA *> get_superkmer(std::vector<kmer_full> &v) {  // brisk/Brisk.hpp:102-118
        std::vector<brisk_kmer> a(v.size());
        for (size_t i = 0; i < v.size(); i++) a[i] = v[i].abi();
        std::vector<uint64_t> slots(v.size(), BRISK_NO_SLOT);
        if (!v.empty()) brisk_check(brisk_get_kmers(menu, a.data(), a.size(), slots.data(), nullptr));
        std::vector<DATA *> out(v.size());
        for (size_t i = 0; i < v.size(); i++) out[i] = slots[i] == BRISK_NO_SLOT ? nullptr : data_at(slots[i]);
        return out;
    }
    // brisk/Brisk.hpp:152-161: the caller's *p update is guarded by a stripe of host mutexes keyed on the minimizer
    void protect_data(const kmer_full &kmer) { stripes_[stripe_of(kmer)].lock(); }
    void unprotect_data(const kmer_full &kmer) { stripes_[stripe_of(kmer)].unlock(); }

    // brisk/Brisk.hpp:166-179: walks the stored set (a snapshot taken at restart / first call) bucket by bucket in
    // increasing bucket id like DenseMenuYo::next (brisk/DenseMenuYo.hpp:476-521)
    void restart_kmer_enumeration() {
        uint64_t n = 0;
        brisk_check(brisk_dump_ordered(menu, nullptr, nullptr, 0, &n));
        snap_.assign(n, brisk_kmer());
        if (n) brisk_check(brisk_dump_ordered(menu, snap_.data(), nullptr, n, &n));
        cursor_ = 0;
        have_snap_ = true;
    }
    bool next(kmer_full &kmer) {
        if (!have_snap_) restart_kmer_enumeration();
        if (cursor_ >= snap_.size()) return false;
        const brisk_kmer &a = snap_[cursor_++];
        kmer.kmer_s = (((kint)a.hi) << 64) | a.lo;
        kmer.minimizer_idx = a.minimizer_idx;
        kmer.dede = params.dede;
        kmer.compute_mini(params.m);
        return true;
    }
    void stats(uint64_t &nb_buckets, uint64_t &nb_skmers, uint64_t &nb_kmers, uint64_t &memory_usage, uint64_t &largest_bucket) const {
        brisk_check(brisk_stats(menu, &nb_buckets, &nb_skmers, &nb_kmers, &memory_usage, &largest_bucket));
    }
    // capacity growth of the GPU table: make room for n_keys distinct k-mers
    void reserve(uint64_t n_keys) { brisk_check(brisk_reserve(menu, n_keys)); }
    // Brisk::reallocate (brisk/Brisk.hpp:202-224): the index rebuilt under (k, m + 2, b + 2)
    void reallocate() {
        brisk_check(brisk_reparameterise(menu, (uint8_t)(params.m + 2), (uint8_t)(params.b + 2)));
        params = Parameters(params.k, (uint8_t)(params.m + 2), (uint8_t)(params.b + 2));
        have_snap_ = false;
    }
    // count_reads / insert_reads return once their input is consumed; sync() waits for the store to be at rest and
    // reports a deferred store-full error (every call that reads the store does the same implicitly)
    void sync() { brisk_check(brisk_sync(menu)); }

    // ---- batched calls (flat read buffer + n_reads + 1 byte offsets; host or device pointers) ------------
    void count_reads(const char *bases, const uint64_t *offsets, uint64_t n_reads) {  // apps/counter.cpp:231-276 per read
        brisk_check(brisk_count_reads(menu, bases, offsets, n_reads));
    }
    void insert_reads(const char *bases, const uint64_t *offsets, uint64_t n_reads, uint64_t *slots, uint8_t *newly) {
        brisk_check(brisk_insert_reads(menu, bases, offsets, n_reads, slots, newly));
    }
    void get_reads(const char *bases, const uint64_t *offsets, uint64_t n_reads, uint8_t *found, DATA *data_out) {
        brisk_check(brisk_get_reads(menu, bases, offsets, n_reads, found, data_out));
    }
    void query_reads(const char *bases, const uint64_t *offsets, uint64_t n_reads, uint64_t &hits, uint64_t &total) {
        brisk_check(brisk_query_reads(menu, bases, offsets, n_reads, &hits, &total));  // apps/counter.cpp:281-310
    }
    void insert_kmers(const brisk_kmer *kmers, uint64_t n, uint64_t *slots, uint8_t *newly, bool count = false) {
        brisk_check(brisk_insert_kmers(menu, kmers, n, slots, newly, count ? 1 : 0));
    }
    void get_kmers(const brisk_kmer *kmers, uint64_t n, uint64_t *slots, DATA *data_out) {
        brisk_check(brisk_get_kmers(menu, kmers, n, slots, data_out));
    }
    // FASTA text (whole records, plain text) parsed on the GPU like getLineFasta + clean_dna (apps/counter.cpp:128-190);
    // returns the number of reads found
    uint64_t count_fasta(const char *text, uint64_t n_bytes) {
        uint64_t n_reads = 0;
        brisk_check(brisk_count_fasta(menu, text, n_bytes, &n_reads));
        return n_reads;
    }
    void query_fasta(const char *text, uint64_t n_bytes, uint64_t &hits, uint64_t &total) {
        brisk_check(brisk_query_fasta(menu, text, n_bytes, &hits, &total));
    }
    // whole-read conveniences of the reference's header (brisk/Brisk.hpp:27-28)
    void count_sequence(const std::string &seq) {
        uint64_t offs[2] = {0, (uint64_t)seq.size()};
        count_reads(seq.data(), offs, 1);
    }
    DATA *data_at(uint64_t slot) const {
        if (!managed_) throw BriskError(BRISK_E_PARAM, "brisk_b200: DATA pointers need a host-visible store (device_resident = false)");
        return reinterpret_cast<DATA *>(brisk_data_ptr(menu, slot));
    }
    // multi-GPU: one process per GPU (brisk_comm_unique_id on rank 0, broadcast by the caller)
    void comm_init(const uint8_t id[128], int rank, int n_ranks) { brisk_check(brisk_comm_init(menu, id, rank, n_ranks)); }

private:
    size_t stripe_of(const kmer_full &kmer) const { return (size_t)((uint64_t)kmer.minimizer * 0x9E3779B97F4A7C15ull >> 54); }
    bool managed_;
    std::array<std::mutex, 1024> stripes_;
    std::vector<brisk_kmer> snap_;
    size_t cursor_ = 0;
    bool have_snap_ = false;
};

// BriskWriter (brisk/writer.hpp:11-20): the index as a KFF file.  apps/counter.cpp:407-411 compiles unchanged:
//     BriskWriter writer(outfile); writer.write(counter); writer.close();
class BriskWriter {
public:
    explicit BriskWriter(std::string filename) : filename_(std::move(filename)) {}
    template <class DATA>
    void write(Brisk<DATA> &index) { brisk_check(brisk_write_kff(index.menu, filename_.c_str(), nullptr, nullptr)); }
    void close() {}

private:
    std::string filename_;
};
