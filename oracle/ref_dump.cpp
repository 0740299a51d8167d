// ref_dump.cpp — driver around the UNMODIFIED reference library (test infrastructure).
//
// This translation unit is ours; it only #includes the reference's public headers
// (brisk/Brisk.hpp and friends) from /root/reference and is linked with the
// reference's own hashing.cpp / Kmers.cpp / Decycling.cpp by oracle/Makefile.
// Nothing from the reference is copied into the repository; the binary lands in
// oracle/_ref/ (git-ignored).  It is used to
//   * generate the golden fixtures in tests/golden/ (tests/golden/make_golden.py),
//   * validate the C restatement oracle/brisk_oracle.c,
//   * time the reference's CPU path for bench.py's cpu_baseline / --impl reference.
//
// Input: a text file with one read per line (ACGT only, upper case).
//
//   ref_dump enum  reads.txt k m b out.bin     every emitted k-mer of every superkmer
//   ref_dump hash  reads.txt k m b out.bin     bfc_hash_64(canonical m-mer) per m-mer offset
//   ref_dump count reads.txt k m b out.bin     count like apps/counter.cpp, dump the index
//   ref_dump query reads.txt k m b out.bin q.txt   count reads.txt, then query q.txt
//   ref_dump realloc reads.txt k m b out.bin   count, then Brisk::reallocate() to (k, m + 2, b + 2), dump the index
//   ref_dump time  reads.txt k m b threads [max_reads]   timed count (hot path only)
//   ref_dump steps reads.txt k m b threads reads_per_step n_steps [dump.bin|- [decoy.txt]]
//                  step j = count chunk j (count_sequence) then query chunk j-1 (query_sequence), timed;
//                  then (threads = 1) the sorted dump of the final index and the aggregate answer to a decoy query
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <string>
#include <vector>

#include "Brisk.hpp"

using namespace std;

static vector<string> load_reads(const char *path, size_t max_reads = ~(size_t)0) {
    vector<string> reads;
    ifstream in(path);
    string line;
    while (reads.size() < max_reads && getline(in, line)) {
        if (!line.empty() && line[0] != '>') reads.push_back(line);
    }
    return reads;
}

#pragma pack(push, 1)
struct EnumRec {  // one emitted k-mer
    uint64_t lo, hi, mini;
    uint32_t read_id, skmer_id;
    uint8_t idx;
};
struct DumpRec {  // one stored k-mer
    uint64_t lo, hi;
    uint8_t idx, count;
};
#pragma pack(pop)

// the calls apps/counter.cpp:count_sequence makes, per superkmer
static void count_read(Brisk<uint8_t> &index, string &seq) {
    if (seq.size() < index.params.k) return;
    SuperKmerEnumerator en(seq, index.params.k, index.params.m, index.params.dede);
    vector<kmer_full> sk;
    vector<bool> newly;
    en.next(sk);
    while (sk.size() > 0) {
        kmer_full guard;
        guard.copy(sk[0]);
        newly.clear();
        index.protect_data(guard);
        vector<uint8_t *> ptrs = index.insert_superkmer(sk, newly);
        for (size_t i = 0; i < ptrs.size(); i++) {
            if (newly[i])
                *ptrs[i] = 1;
            else
                (*ptrs[i])++;
        }
        index.unprotect_data(guard);
        sk.clear();
        en.next(sk);
    }
}

static void print_stats(Brisk<uint8_t> &index) {
    uint64_t nb_buckets, nb_skmers, nb_kmers, memory, largest;
    index.stats(nb_buckets, nb_skmers, nb_kmers, memory, largest);
    printf("nb_buckets %lu\nnb_skmers %lu\nnb_kmers %lu\nlargest_bucket %lu\n", nb_buckets, nb_skmers,
           nb_kmers, largest);
}

int main(int argc, char **argv) {
    if (argc < 6) {
        fprintf(stderr, "usage: ref_dump enum|hash|count|query|time reads.txt k m b ...\n");
        return 2;
    }
    string mode = argv[1];
    uint8_t k = atoi(argv[3]), m = atoi(argv[4]), b = atoi(argv[5]);
    Parameters params(k, m, b);

    if (mode == "enum") {
        vector<string> reads = load_reads(argv[2]);
        FILE *out = fopen(argv[6], "wb");
        uint64_t n_occ = 0, n_sk = 0;
        for (size_t r = 0; r < reads.size(); r++) {
            if (reads[r].size() < k) continue;
            SuperKmerEnumerator en(reads[r], k, m, params.dede);
            vector<kmer_full> sk;
            uint32_t skid = 0;
            en.next(sk);
            while (sk.size() > 0) {
                for (kmer_full &km : sk) {
                    EnumRec rec;
                    rec.lo = (uint64_t)km.kmer_s;
                    rec.hi = (uint64_t)(km.kmer_s >> 64);
                    rec.mini = (uint64_t)km.minimizer;
                    rec.read_id = (uint32_t)r;
                    rec.skmer_id = skid;
                    rec.idx = km.minimizer_idx;
                    fwrite(&rec, sizeof(rec), 1, out);
                    n_occ++;
                }
                n_sk++;
                skid++;
                sk.clear();
                en.next(sk);
            }
        }
        fclose(out);
        printf("occurrences %lu\nsuperkmers %lu\n", n_occ, n_sk);
        return 0;
    }

    if (mode == "hash") {
        vector<string> reads = load_reads(argv[2]);
        FILE *out = fopen(argv[6], "wb");
        uint64_t m_mask = ((uint64_t)1 << (2 * m)) - 1;
        for (string &s : reads) {
            if (s.size() < m) continue;
            uint64_t f = 0;
            for (size_t i = 0; i < s.size(); i++) {
                f = ((f << 2) + ((s[i] >> 1) & 3)) & m_mask;
                if (i + 1 >= m) {
                    uint64_t c = canonize(f, (uint64_t)m);
                    uint64_t h = bfc_hash_64(c, m_mask, params.dede);
                    fwrite(&h, 8, 1, out);
                }
            }
        }
        fclose(out);
        return 0;
    }

    if (mode == "count" || mode == "query" || mode == "realloc") {
        vector<string> reads = load_reads(argv[2]);
        Brisk<uint8_t> index(params);
        for (string &s : reads) count_read(index, s);
        // realloc: Brisk::reallocate (brisk/Brisk.hpp:202-224) rebuilds the index under (k, m + 2, b + 2) before the dump
        if (mode == "realloc") { index.reallocate(); m = index.params.m; }
        // read the index back the way verif_counts does (apps/counter.cpp:90-104)
        vector<DumpRec> recs;
        kmer_full km(0, 0, m, params.dede);
        index.restart_kmer_enumeration();
        while (index.next(km)) {
            DumpRec rec;
            rec.lo = (uint64_t)km.kmer_s;
            rec.hi = (uint64_t)(km.kmer_s >> 64);
            rec.idx = km.minimizer_idx;
            uint8_t *c = index.get(km);
            rec.count = c ? *c : 0;
            if (!c) printf("MISSING_DATA\n");
            recs.push_back(rec);
            km.kmer_s = 0;
        }
        sort(recs.begin(), recs.end(), [](const DumpRec &a, const DumpRec &b) {
            if (a.hi != b.hi) return a.hi < b.hi;
            if (a.lo != b.lo) return a.lo < b.lo;
            return a.idx < b.idx;
        });
        FILE *out = fopen(argv[6], "wb");
        fwrite(recs.data(), sizeof(DumpRec), recs.size(), out);
        fclose(out);
        print_stats(index);
        printf("dumped %lu\n", recs.size());
        if (mode == "query" && argc > 7) {
            // apps/counter.cpp:281-310, per read totals; both with and without the
            // reference's early break on a zero minimizer value
            vector<string> qs = load_reads(argv[7]);
            for (string &s : qs) {
                uint64_t total = 0, hits = 0, nk = 0;
                if (s.size() >= k) {
                    SuperKmerEnumerator en(s, k, m, params.dede);
                    vector<kmer_full> sk;
                    en.next(sk);
                    while (sk.size() > 0) {
                        vector<uint8_t *> v = index.get_superkmer(sk);
                        for (uint8_t *p : v) {
                            nk++;
                            if (p) { total += *p; hits++; }
                        }
                        sk.clear();
                        en.next(sk);
                    }
                }
                printf("q %lu %lu %lu\n", nk, hits, total);
            }
        }
        return 0;
    }

    if (mode == "time") {
        int threads = argc > 6 ? atoi(argv[6]) : 1;
        size_t max_reads = argc > 7 ? strtoull(argv[7], 0, 10) : ~(size_t)0;
        vector<string> reads = load_reads(argv[2], max_reads);
        uint64_t occ = 0;
        for (string &s : reads) occ += s.size() >= k ? s.size() - k + 1 : 0;
        auto t0 = chrono::steady_clock::now();
        Brisk<uint8_t> index(params);
        auto t1 = chrono::steady_clock::now();
#pragma omp parallel for num_threads(threads) schedule(dynamic, 64)
        for (size_t r = 0; r < reads.size(); r++) count_read(index, reads[r]);
        auto t2 = chrono::steady_clock::now();
        double ctor = chrono::duration<double>(t1 - t0).count();
        double cnt = chrono::duration<double>(t2 - t1).count();
        printf("threads %d\nreads %lu\noccurrences %lu\nctor_s %.6f\ncount_s %.6f\n", threads, reads.size(),
               occ, ctor, cnt);
        print_stats(index);
        return 0;
    }
    if (mode == "steps") {
        int threads = atoi(argv[6]);
        size_t per = strtoull(argv[7], 0, 10);
        size_t n_steps = strtoull(argv[8], 0, 10);
        vector<string> reads = load_reads(argv[2], per * n_steps);
        auto t0 = chrono::steady_clock::now();
        Brisk<uint8_t> index(params);
        double ctor = chrono::duration<double>(chrono::steady_clock::now() - t0).count();
        printf("threads %d\nctor_s %.6f\n", threads, ctor);
        for (size_t j = 0; j < n_steps; j++) {
            size_t lo = j * per, hi = min(reads.size(), (j + 1) * per);
            uint64_t occ = 0, qocc = 0, hits = 0, total = 0;
            for (size_t r = lo; r < hi; r++) occ += reads[r].size() >= k ? reads[r].size() - k + 1 : 0;
            auto a = chrono::steady_clock::now();
#pragma omp parallel for num_threads(threads) schedule(dynamic, 64)
            for (size_t r = lo; r < hi; r++) count_read(index, reads[r]);
            auto b2 = chrono::steady_clock::now();
            if (j > 0) {
#pragma omp parallel for num_threads(threads) schedule(dynamic, 64) reduction(+ : qocc, hits, total)
                for (size_t r = lo - per; r < lo; r++) {
                    string &s = reads[r];
                    if (s.size() < k) continue;
                    SuperKmerEnumerator en(s, k, m, params.dede);
                    vector<kmer_full> sk;
                    en.next(sk);
                    while (sk.size() > 0) {
                        vector<uint8_t *> v = index.get_superkmer(sk);
                        for (uint8_t *p : v) {
                            qocc++;
                            if (p) { total += *p; hits++; }
                        }
                        sk.clear();
                        en.next(sk);
                    }
                }
            }
            auto c = chrono::steady_clock::now();
            printf("step %lu %.6f %.6f %lu %lu %lu %lu\n", j, chrono::duration<double>(b2 - a).count(),
                   chrono::duration<double>(c - b2).count(), occ, qocc, hits, total);
            fflush(stdout);
        }
        print_stats(index);
        // optional: sorted dump of the final index (argv[9]) and the aggregate answer to a decoy query file (argv[10]);
        // both only make sense with threads = 1
        if (argc > 9 && strcmp(argv[9], "-") != 0) {
            vector<DumpRec> recs;
            kmer_full km(0, 0, m, params.dede);
            index.restart_kmer_enumeration();
            while (index.next(km)) {
                DumpRec rec;
                rec.lo = (uint64_t)km.kmer_s;
                rec.hi = (uint64_t)(km.kmer_s >> 64);
                rec.idx = km.minimizer_idx;
                uint8_t *c = index.get(km);
                rec.count = c ? *c : 0;
                if (!c) printf("MISSING_DATA\n");
                recs.push_back(rec);
                km.kmer_s = 0;
            }
            sort(recs.begin(), recs.end(), [](const DumpRec &a, const DumpRec &b) {
                if (a.hi != b.hi) return a.hi < b.hi;
                if (a.lo != b.lo) return a.lo < b.lo;
                return a.idx < b.idx;
            });
            FILE *out = fopen(argv[9], "wb");
            fwrite(recs.data(), sizeof(DumpRec), recs.size(), out);
            fclose(out);
            printf("dumped %lu\n", recs.size());
        }
        if (argc > 10) {
            vector<string> qs = load_reads(argv[10]);
            uint64_t nk = 0, hits = 0, total = 0;
            for (string &s : qs) {
                if (s.size() < k) continue;
                SuperKmerEnumerator en(s, k, m, params.dede);
                vector<kmer_full> sk;
                en.next(sk);
                while (sk.size() > 0) {
                    vector<uint8_t *> v = index.get_superkmer(sk);
                    for (uint8_t *p : v) {
                        nk++;
                        if (p) { total += *p; hits++; }
                    }
                    sk.clear();
                    en.next(sk);
                }
            }
            printf("decoy %lu %lu %lu\n", nk, hits, total);
        }
        return 0;
    }
    fprintf(stderr, "unknown mode\n");
    return 2;
}
