// kff_dump.cpp — reads a KFF file with the reference's OWN reader (Kff_reader of brisk/lib/kff, compiled from the
// reference checkout by oracle/Makefile into oracle/_ref/kff_dump) and prints one line per k-mer:
//     <hi hex> <lo hex> <data bytes hex>
// (kmer as a 2k-bit integer in the file's 2-bit codes, first nucleotide most significant).  Test infrastructure: it
// checks that a file written by brisk_write_kff is a KFF file the reference's library parses, and what it holds.
#include <cstdint>
#include <cstdio>
#include <string>

#include "kff_io.hpp"

int main(int argc, char **argv) {
    if (argc < 2) { fprintf(stderr, "usage: kff_dump file.kff\n"); return 2; }
    Kff_reader reader(argv[1]);
    uint64_t n = 0;
    uint8_t *kmer = nullptr, *data = nullptr;
    while (reader.has_next()) {
        if (!reader.next_kmer(kmer, data)) break;
        const uint64_t k = reader.k, ds = reader.data_size;
        // Kff_reader::next_kmer hands out a copy shifted so that the k-mer ENDS on a byte boundary: ceil(k / 4) bytes, the
        // k-mer right-aligned in them (leading bits of the first byte belong to the previous nucleotides)
        unsigned __int128 v = 0;
        const uint64_t cb = (k + 3) / 4;
        for (uint64_t i = 0; i < cb; i++) v = (v << 8) | kmer[i];
        if (k < 64) v &= (((unsigned __int128)1) << (2 * k)) - 1;
        printf("%016lx %016lx ", (unsigned long)(uint64_t)(v >> 64), (unsigned long)(uint64_t)v);
        for (uint64_t i = 0; i < ds; i++) printf("%02x", data[i]);
        printf("\n");
        n++;
    }
    fprintf(stderr, "kmers %lu k %lu data_size %lu\n", (unsigned long)n, (unsigned long)reader.k, (unsigned long)reader.data_size);
    return 0;
}
