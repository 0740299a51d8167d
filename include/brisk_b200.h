/*
 * brisk_b200.h — C ABI of libbrisk_b200.so, the B200-native k-mer insertion path of Brisk.
 *
 * The reference has no FFI: its insertion path sits behind a header-only C++ template API
 * (brisk/Brisk.hpp:23-42, brisk/Kmers.hpp:32-89, brisk/parameters.hpp:9-35) called by
 * apps/counter.cpp.  This header is the boundary a binding would target; the C++ mirror of
 * the reference API (include/brisk_b200/Brisk.hpp) and the Python host class
 * (brisk_b200/brisk.py) are thin wrappers over exactly these entry points.
 *
 * Conventions
 *  - every call returns 0 on success, <0 on error (BRISK_E_*); brisk_last_error() gives text.
 *    The reference has no error channel (absent key -> NULL, brisk/DenseMenuYo.hpp:331-333);
 *    "absent" is reported through found[] / slot == BRISK_NO_SLOT, never as an error.
 *  - `bases`/`offsets`: a flat buffer of read characters and n_reads+1 byte offsets, read r =
 *    bases[offsets[r] .. offsets[r+1]).  Any byte is accepted and coded (c>>1)&3 like
 *    brisk/Kmers.cpp:442-444 (A=0 C=1 T=2 G=3); reads shorter than k are skipped
 *    (apps/counter.cpp:233).
 *  - pointers may be host or device pointers (detected with cudaPointerGetAttributes);
 *    host buffers are copied inside the call.  No torch types, plain pointers and sizes.
 *  - a k-mer is given as the reference's kmer_full does (brisk/Kmers.hpp:32-58): the 2k-bit
 *    value kmer_s (lo, hi words) in the orientation the enumerator chose, plus
 *    minimizer_idx.  The stored set is keyed on (kmer_s, minimizer_idx) exactly as the
 *    reference's buckets are (brisk/SuperKmerLight.hpp:209-217).
 *  - there is no CPU fallback: without a CUDA device brisk_create fails with BRISK_E_CUDA.
 */
#ifndef BRISK_B200_H
#define BRISK_B200_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct brisk_handle brisk_t;

#define BRISK_OK 0
#define BRISK_E_PARAM (-1)    /* bad k/m/b (Brisk.hpp:50-51 asserts m odd, m < k) or bad argument */
#define BRISK_E_CUDA (-2)     /* CUDA runtime error, no device, allocation failure */
#define BRISK_E_CAPACITY (-3) /* an output buffer given by the caller is too small */
#define BRISK_E_FULL (-4)     /* the bucket store could not grow */
#define BRISK_E_COMM (-5)     /* NCCL error */

#define BRISK_NO_SLOT (~(uint64_t)0)

/* flags for brisk_create */
#define BRISK_F_MANAGED 1u        /* store in cudaMallocManaged memory: brisk_data_ptr() is host-dereferenceable */
#define BRISK_F_LITERAL_DECYCLE 2u /* debug: evaluate DecyclingSet::memDouble literally for every m-mer */
#define BRISK_F_SYNC_INSERT 4u     /* brisk_count_reads / brisk_insert_reads return only when the store kernel has finished
                                    * (default: they return once the input buffers may be reused; the insertion completes in
                                    * stream order before any later call on the handle reads the store, and a store-full
                                    * error of such a call is reported by the next call or by brisk_sync) */

#define BRISK_F_GENERIC_ENUM 8u     /* debug: enumerate with the run-time-(k, m) kernel even when a compiled-in
                                    * specialisation exists for the parameter set */

#define BRISK_F_SERIAL_EXCHANGE 16u  /* debug: sharded index without the chunk-pipelined exchange (enumerate everything, then
                                    * partition, exchange, store one after the other) */

/* one k-mer as the reference's kmer_full carries it */
typedef struct {
    uint64_t lo, hi;       /* kmer_s */
    uint8_t minimizer_idx; /* nucleotide offset of the minimizer from the k-mer's right end */
    uint8_t pad[7];
} brisk_kmer;

/* Parameters(k,m,b) + Brisk<DATA>(params): brisk/parameters.hpp:24-34, brisk/Brisk.hpp:47-52.
 * data_size = sizeof(DATA) (1 for the counter's uint8_t).  capacity_hint = expected number of
 * distinct stored k-mers (0 = size from the batches seen); the table is laid out for a load factor of about 0.25 at
 * that many keys when it takes less than a quarter of the device's free memory (0.5 otherwise) and grows on its own
 * beyond.  device < 0 = current device. */
int brisk_create(uint8_t k, uint8_t m, uint8_t b, uint32_t data_size, int device, uint64_t capacity_hint,
                 uint32_t flags, brisk_t **out);
int brisk_destroy(brisk_t *h);

/* SuperKmerEnumerator::next over a batch of reads (brisk/Kmers.cpp:509-603): writes one packed
 * superkmer record per emitted superkmer (rec_words u64 each, layout in DESIGN.md) and returns
 * their number.  `recs` may be NULL to only count.  Order of records is unspecified. */
int brisk_enumerate_reads(brisk_t *h, const char *bases, const uint64_t *offsets, uint64_t n_reads,
                          uint64_t *recs, uint64_t cap_recs, uint64_t *n_recs, uint32_t *rec_words);

/* The same, plus the read-order tag of every record: occ[i] = (index of the record's first k-mer occurrence in read
 * order << 1) | reversed.  Sorting by occ >> 1 restores SuperKmerEnumerator::next's emission order (the C++ mirror's
 * SuperKmerEnumerator is built on this call). */
int brisk_enumerate_reads_tagged(brisk_t *h, const char *bases, const uint64_t *offsets, uint64_t n_reads,
                                 uint64_t *recs, uint64_t cap_recs, uint64_t *n_recs, uint32_t *rec_words, uint64_t *occ);

/* count_sequence for every read (apps/counter.cpp:231-276): insert_superkmer (brisk/Brisk.hpp:123-147)
 * fused with the counter's update  new -> 1, else ++ (uint8 wrap, apps/counter.cpp:262-269).
 * Returns when `bases` / `offsets` may be reused; the insertion completes in stream order before any later call on the
 * handle touches the store (see BRISK_F_SYNC_INSERT, brisk_sync). */
int brisk_count_reads(brisk_t *h, const char *bases, const uint64_t *offsets, uint64_t n_reads);

/* insert_superkmer without touching DATA: slot of every k-mer occurrence in read order (k-mer i of
 * read r at index sum_{r'<r} max(0,len-k+1) + i) and whether it was newly inserted
 * (`newly_inserted`, brisk/Brisk.hpp:123).  slots/newly may be NULL. */
int brisk_insert_reads(brisk_t *h, const char *bases, const uint64_t *offsets, uint64_t n_reads,
                       uint64_t *slots, uint8_t *newly);

/* query_sequence / get_superkmer (apps/counter.cpp:281-310, brisk/Brisk.hpp:102-118): per k-mer
 * occurrence in read order, found[i] and the DATA bytes (data_size each; zero when absent). */
int brisk_get_reads(brisk_t *h, const char *bases, const uint64_t *offsets, uint64_t n_reads,
                    uint8_t *found, void *data_out);

/* query_sequence as the counter app uses it (apps/counter.cpp:281-310): number of k-mer occurrences found and
 * the sum of their DATA bytes (`total`), nothing per occurrence. */
int brisk_query_reads(brisk_t *h, const char *bases, const uint64_t *offsets, uint64_t n_reads, uint64_t *hits,
                      uint64_t *total);

/* Brisk::insert on explicit k-mers (size-1 superkmers through insert_superkmer).
 * count != 0 also applies the counter update. */
int brisk_insert_kmers(brisk_t *h, const brisk_kmer *kmers, uint64_t n, uint64_t *slots, uint8_t *newly,
                       int count);
/* Brisk::get (brisk/Brisk.hpp:64-69): slot or BRISK_NO_SLOT, and DATA bytes (nullable outputs). */
int brisk_get_kmers(brisk_t *h, const brisk_kmer *kmers, uint64_t n, uint64_t *slots, void *data_out);

/* address of the DATA of a slot; host-dereferenceable only with BRISK_F_MANAGED.  Slots are stable
 * until the store grows (the reference's pointers die at the next bucket realloc, README.md:109-110). */
void *brisk_data_ptr(brisk_t *h, uint64_t slot);

/* restart_kmer_enumeration + next + get (brisk/Brisk.hpp:166-179): every stored (kmer_s,
 * minimizer_idx, DATA), unordered.  keys/data may be NULL to only count. */
int brisk_dump(brisk_t *h, brisk_kmer *keys, void *data, uint64_t cap, uint64_t *n);

/* The same set in the order DenseMenuYo::next walks it at bucket level (brisk/DenseMenuYo.hpp:476-521): increasing
 * bucket id; inside a bucket by (minimizer_idx, kmer_s) (the reference's inner order is its arena's insertion order).
 * On a sharded index every rank dumps its own buckets. */
int brisk_dump_ordered(brisk_t *h, brisk_kmer *keys, void *data, uint64_t cap, uint64_t *n);

/* BriskWriter::write (brisk/writer.hpp:75-179): the stored set of this handle (its shard of a sharded index) as a KFF
 * file — header with Brisk's encoding (A 0, C 1, G 3, T 2) and metadata string, the two value sections, then per
 * bucket id in increasing order and per minimizer one Section_Minimizer whose blocks are superkmers re-assembled from
 * the stored k-mers (sequence without the minimizer, minimizer position, the DATA of every k-mer), the section index
 * and the footer.  The reference's blocks are whatever its insertion order compacted; the (k-mer, minimizer position,
 * DATA) content is the same.  Bucket arithmetic is 64-bit (the reference's `max_mini` overflows at b = 16,
 * writer.hpp:94).  n_blocks / n_sections (nullable) receive what was written. */
int brisk_write_kff(brisk_t *h, const char *path, uint64_t *n_blocks, uint64_t *n_sections);

/* Brisk::stats (brisk/Brisk.hpp:194-197, brisk/DenseMenuYo.hpp:545-568). nb_skmers = superkmer records
 * that created at least one k-mer; largest_bucket = the largest number of k-mers stored under one bucket id (the
 * reference counts the SKL records of a bucket, which depends on its insertion order). */
int brisk_stats(brisk_t *h, uint64_t *nb_buckets, uint64_t *nb_skmers, uint64_t *nb_kmers,
                uint64_t *device_bytes, uint64_t *largest_bucket);

/* order-independent checksum of the stored set held by this handle (its shard of a sharded index): out[0] = sum of a
 * 64-bit mix of every stored (kmer_s, minimizer_idx, DATA byte), out[1] = sum of the DATA bytes, out[2] = number of
 * stored k-mers.  The sums over the shards equal the sums of an unsharded index over the same reads. */
int brisk_checksum(brisk_t *h, uint64_t out[3]);

/* brisk_checksum's three sums over an explicit list of k-mers and their DATA bytes (what brisk_dump returns): equal to
 * brisk_checksum of the handle the list was dumped from; the sums of several lists add up.  NO COUNTERPART IN THE
 * REFERENCE: added for callers that cannot allocate (CUDA graph capture, tight
 * serving loops, harnesses that forbid cudaMalloc).  The call allocates nothing: its scratch is a static __device__
 * array of 384.25 KiB that every GPU loading the library holds until exit, which limits one call to
 * BRISK_NOALLOC_MAX_KMERS k-mers (more: BRISK_E_CAPACITY; split the list and add the sums).  keys, data (1 byte per
 * k-mer) and out (3 x uint64) are DEVICE pointers on the current device; the work is queued on `stream` (cudaStream_t,
 * NULL = default stream) and not waited for.  All calls on a GPU share the scratch: calls that may overlap on the GPU
 * (different streams without an ordering between them) race; calls on one stream are safe back to back.  The first
 * call on a device resolves the scratch's address (cudaGetSymbolAddress): make it before a graph capture starts. */
#define BRISK_NOALLOC_MAX_KMERS (1ull << 26)
int brisk_checksum_kmers_noalloc(const brisk_kmer *keys, const uint8_t *data, uint64_t n, uint64_t *out, void *stream);

/* benchmark / test support: the synthetic generator of brisk_b200/synth.py on the device,
 * bit-identical to it.  genome / out are device pointers on the current device, stream a cudaStream_t (NULL = default
 * stream); the kernels are queued, not waited for.  Reads first .. first+count-1 of the set, L bytes each. */
int brisk_synth_genome(void *genome, uint64_t G, uint64_t seed, void *stream);
int brisk_synth_reads(const void *genome, uint64_t G, uint32_t L, uint64_t seed, uint64_t first, uint64_t count, void *out,
                      void *stream);

/* Brisk::reallocate (brisk/Brisk.hpp:202-224, which always goes to m + 2, b + 2): the handle rebuilt under
 * (k, new_m, new_b).  Every stored kmer_s keeps its value and DATA; its minimizer_idx is recomputed by update_kmer
 * (brisk/Kmers.cpp:88-97: the literal get_minimizer with the new m, mirrored when it reports the reverse strand).  Old
 * entries that only differed in their minimizer_idx become one entry (the reference keeps the DATA of the one it
 * enumerates last; here: of the one with the largest old minimizer_idx).  Single-GPU handles only. */
int brisk_reparameterise(brisk_t *h, uint8_t new_m, uint8_t new_b);

/* Long reads (contigs, chromosomes): a read longer than out[0] bases (2048; environment BRISK_LONG_READ at brisk_create)
 * is enumerated in segments of 1024 k-mers by different threads, each warming up over out[1] k-mers (BRISK_SEG_WARM)
 * before its segment; the result is verified against the neighbouring segments' states and repeated with a longer
 * warm-up (finally with one thread per read) when they disagree — out[2] counts those repeats.  The superkmers are
 * exactly SuperKmerEnumerator::next's either way.  Reads of 2^32 bases or more are not supported. */
int brisk_long_read_info(brisk_t *h, uint64_t out[3]);

/* grow the store so that n_keys distinct k-mers fit (the GPU analogue of Brisk::reallocate) */
int brisk_reserve(brisk_t *h, uint64_t n_keys);
int brisk_capacity(brisk_t *h, uint64_t *slots);
int brisk_sync(brisk_t *h);
/* run every later call on the caller's CUDA stream (cudaStream_t; NULL = the handle's own stream) */
int brisk_set_stream(brisk_t *h, void *stream);

/* timing of the last brisk_count_reads / brisk_insert_reads / brisk_get_reads call, measured with CUDA
 * events on the handle's stream: [0] pack, [1] enumerate, [2] partition, [3] insert/get, [4] whole
 * device section (ms); n_launches = kernels launched by the call. */
int brisk_last_timing(brisk_t *h, float ms[5], uint32_t *n_launches);

/* ---- FASTA ingest on the device: what getLineFasta + clean_dna (apps/counter.cpp:128-190) hand to count_sequence
 * with one reader thread — per record (first line of the text or any line starting with '>' = header) the maximal runs
 * of [ACGTacgt] of its concatenated lines, upper-cased, in file order.  `text` (host or device) must hold whole
 * records.  brisk_parse_fasta takes texts shorter than 2^32 - 2 bytes; brisk_count_fasta / brisk_query_fasta take
 * texts of any size (they walk them in pieces of about 1 GB cut right before a header line).
 * brisk_parse_fasta: the reads as a flat buffer + n_reads + 1 offsets (outputs nullable to only count);
 * brisk_count_fasta / brisk_query_fasta: count_fasta (apps/counter.cpp:197-226) / query_fasta (:314-345). */
int brisk_parse_fasta(brisk_t *h, const char *text, uint64_t n_bytes, char *bases, uint64_t cap_bases, uint64_t *offsets,
                      uint64_t cap_reads, uint64_t *n_bases, uint64_t *n_reads);
int brisk_count_fasta(brisk_t *h, const char *text, uint64_t n_bytes, uint64_t *n_reads);
int brisk_query_fasta(brisk_t *h, const char *text, uint64_t n_bytes, uint64_t *hits, uint64_t *total);

/* ---- the two halves of the routed path, usable on their own.
 * brisk_partition_reads: enumerate + radix partition of the superkmer records by owner rank, as a rank does
 * before the exchange (replaces the lock-stripe routing of DenseMenuYo::insert_kmer_vector,
 * brisk/DenseMenuYo.hpp:248-308).  recs (nullable) receives the records grouped by destination, starts[n_ranks+1]
 * the first record of every destination's region.
 * brisk_apply_records: what the owner does with the records it received: op 0 = count (insert + counter update),
 * 1 = insert, 2 = get (hits / total as brisk_query_reads). */
int brisk_partition_reads(brisk_t *h, const char *bases, const uint64_t *offsets, uint64_t n_reads, uint32_t n_ranks,
                          uint64_t *recs, uint64_t cap_recs, uint64_t *starts, uint32_t *rec_words);
int brisk_apply_records(brisk_t *h, const uint64_t *recs, uint64_t n_recs, int op, uint64_t *hits, uint64_t *total);

/* ---- multi-GPU: one process per GPU, bucket ids sharded over ranks, superkmers exchanged with NCCL.
 * id = 128 bytes from brisk_comm_unique_id on rank 0, broadcast by the caller.  After brisk_comm_init the
 * read-level calls (count / insert / get / query _reads) and brisk_stats are COLLECTIVE: every rank calls them
 * in the same order, each with its own share of the reads (possibly none).  Per-occurrence outputs come back to
 * the rank that supplied the read; slot ids carry the owner rank in their top 8 bits.  brisk_dump,
 * brisk_insert_kmers / brisk_get_kmers and brisk_apply_records stay local to the rank's shard. */
int brisk_comm_unique_id(uint8_t id[128]);
int brisk_comm_init(brisk_t *h, const uint8_t id[128], int rank, int n_ranks);
int brisk_comm_info(brisk_t *h, int *rank, int *n_ranks);
/* owner rank of a bucket id (pure function; also used by the CPU tests of the routing) */
uint32_t brisk_owner_of_bucket(uint64_t bucket_id, uint8_t b, uint32_t n_ranks);
/* send / receive offsets of one exchange from the n x n matrix of record counts (row = sender); both outputs have
 * n_ranks + 1 entries.  Pure host arithmetic. */
int brisk_exchange_plan(const uint64_t *counts, int n_ranks, int rank, uint64_t *send_start, uint64_t *recv_start);

/* fixed-point tables of the rolled decycling class (pure host arithmetic, no device needed):
 * tw[64] / rot[64] = (cos, sin)(2 pi q / m) scaled by 2^24 / 2^30 for q < 32, *th = the band (units of 2^-22) inside
 * which the literal DecyclingSet::memDouble decides, groups_tab[ceil(m/3) * 64 * 2] = the 3-nucleotide partial sums. */
int brisk_decycle_tables(uint8_t k, uint8_t m, uint8_t b, int32_t *tw, int32_t *rot, int32_t *th, int32_t *groups_tab);

const char *brisk_last_error(void);
const char *brisk_version(void);

#ifdef __cplusplus
}
#endif
#endif
