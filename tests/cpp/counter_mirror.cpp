// counter_mirror.cpp — the reference's demo counter (apps/counter.cpp:231-310) written against the C++ mirror header,
// used by tests/test_gpu_cpp_mirror.py: counts every line of a text file of reads through the LEGACY per-superkmer API
// (SuperKmerEnumerator::next -> protect_data -> insert_superkmer -> *p update -> unprotect_data), then verifies like
// verif_counts (:84-126) through next()/get(), then counts the same reads again through the batched count_reads on a
// second index, and prints both dumps as "lo hi idx count" lines for the test to compare with the golden fixture.
#include <cstdio>
#include <fstream>
#include <iostream>
#include <string>
#include <vector>

#include "brisk_b200/Brisk.hpp"

using namespace std;

static void count_sequence(Brisk<uint8_t> &counter, string &sequence) {
    if (sequence.size() < counter.params.k) return;
    SuperKmerEnumerator enumerator(sequence, counter.params.k, counter.params.m, counter.params.dede);
    vector<kmer_full> superkmer;
    vector<bool> newly_inserted;
    kint minimizer = enumerator.next(superkmer);
    (void)minimizer;
    while (superkmer.size() > 0) {
        kmer_full local;
        local.copy(superkmer[0]);
        newly_inserted.clear();
        counter.protect_data(local);
        vector<uint8_t *> vec = counter.insert_superkmer(superkmer, newly_inserted);
        for (size_t i = 0; i < vec.size(); i++) {
            uint8_t *data_pointer = vec[i];
            if (newly_inserted[i]) *data_pointer = 1;
            else (*data_pointer)++;
        }
        counter.unprotect_data(local);
        superkmer.clear();
        enumerator.next(superkmer);
    }
}

static uint64_t query_sequence(Brisk<uint8_t> &counter, string &sequence, uint64_t &hits) {
    uint64_t total = 0;
    if (sequence.size() < counter.params.k) return 0;
    SuperKmerEnumerator enumerator(sequence, counter.params.k, counter.params.m, counter.params.dede);
    vector<kmer_full> superkmer;
    enumerator.next(superkmer);
    while (superkmer.size() > 0) {
        vector<uint8_t *> vec = counter.get_superkmer(superkmer);
        for (uint8_t *p : vec)
            if (p != NULL) { total += *p; hits++; }
        superkmer.clear();
        enumerator.next(superkmer);
    }
    return total;
}

static void dump(Brisk<uint8_t> &counter, const char *tag) {
    kmer_full kmer;
    counter.restart_kmer_enumeration();
    uint64_t n = 0;
    while (counter.next(kmer)) {
        uint8_t *count = counter.get(kmer);
        if (count == NULL) { printf("ERROR next() returned a k-mer that get() does not find\n"); return; }
        printf("%s %llu %llu %u %u\n", tag, (unsigned long long)(uint64_t)kmer.kmer_s, (unsigned long long)(uint64_t)(kmer.kmer_s >> 64),
               (unsigned)kmer.minimizer_idx, (unsigned)*count);
        n++;
    }
    uint64_t nb_buckets, nb_skmers, nb_kmers, memory, largest;
    counter.stats(nb_buckets, nb_skmers, nb_kmers, memory, largest);
    printf("STATS %s kmers %llu buckets %llu dumped %llu\n", tag, (unsigned long long)nb_kmers, (unsigned long long)nb_buckets, (unsigned long long)n);
}

int main(int argc, char **argv) {
    if (argc < 5) { fprintf(stderr, "usage: counter_mirror reads.txt k m b [out.kff]\n"); return 2; }
    try {
        Parameters params(atoi(argv[2]), atoi(argv[3]), atoi(argv[4]));
        vector<string> reads;
        ifstream in(argv[1]);
        for (string line; getline(in, line);)
            if (!line.empty()) reads.push_back(line);
        Brisk<uint8_t> legacy(params);
        for (string &r : reads) count_sequence(legacy, r);
        dump(legacy, "LEGACY");
        uint64_t hits = 0, total = 0;
        for (string &r : reads) total += query_sequence(legacy, r, hits);
        printf("QUERY LEGACY hits %llu total %llu\n", (unsigned long long)hits, (unsigned long long)total);
        // the same through the batched path
        Brisk<uint8_t> batched(params);
        string flat;
        vector<uint64_t> offs(1, 0);
        for (string &r : reads) { flat += r; offs.push_back(flat.size()); }
        batched.count_reads(flat.data(), offs.data(), reads.size());
        dump(batched, "BATCHED");
        uint64_t h2 = 0, t2 = 0;
        batched.query_reads(flat.data(), offs.data(), reads.size(), h2, t2);
        printf("QUERY BATCHED hits %llu total %llu\n", (unsigned long long)h2, (unsigned long long)t2);
        // Brisk<DATA> with a wider DATA (a parallel arena instead of the byte inside the slot): the same legacy loop with
        // uint32_t counts, no capacity hint, so the store grows (and moves its DATA) several times on the way
        {
            Brisk<uint32_t> wide(params);
            uint64_t occurrences = 0;
            for (string &r : reads) {
                if (r.size() < params.k) continue;
                SuperKmerEnumerator e(r, params.k, params.m, params.dede);
                vector<kmer_full> sk;
                vector<bool> fresh;
                e.next(sk);
                while (sk.size() > 0) {
                    fresh.clear();
                    vector<uint32_t *> v = wide.insert_superkmer(sk, fresh);
                    for (size_t i = 0; i < v.size(); i++) { if (fresh[i]) *v[i] = 1000000; else (*v[i])++; occurrences++; }
                    sk.clear();
                    e.next(sk);
                }
            }
            kmer_full km;
            uint64_t n = 0, sum = 0, bad = 0;
            wide.restart_kmer_enumeration();
            while (wide.next(km)) {
                uint32_t *p = wide.get(km);
                if (p == NULL || *p < 1000000) bad++;
                else sum += *p - 999999;
                n++;
            }
            printf("WIDE kmers %llu sum %llu occurrences %llu bad %llu\n", (unsigned long long)n, (unsigned long long)sum,
                   (unsigned long long)occurrences, (unsigned long long)bad);
        }
        // the KFF dump exactly as apps/counter.cpp:407-411 writes it, then Brisk::reallocate (brisk/Brisk.hpp:202-224)
        if (argc > 5) {
            BriskWriter writer(argv[5]);
            writer.write(batched);
            writer.close();
            batched.reallocate();
            uint64_t nb_buckets, nb_skmers, nb_kmers, memory, largest;
            batched.stats(nb_buckets, nb_skmers, nb_kmers, memory, largest);
            printf("REALLOC m %u b %u kmers %llu buckets %llu\n", (unsigned)batched.params.m, (unsigned)batched.params.b, (unsigned long long)nb_kmers,
                   (unsigned long long)nb_buckets);
        }
        // absent key -> NULL (brisk/DenseMenuYo.hpp:331-333)
        kmer_full absent((kint)0x123456789abcdefull, 3, params.m, params.dede);
        printf("ABSENT %s\n", legacy.get(absent) == NULL ? "null" : "found");
    } catch (const std::exception &e) {
        printf("EXCEPTION %s\n", e.what());
        return 1;
    }
    return 0;
}
