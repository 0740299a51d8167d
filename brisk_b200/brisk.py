"""Host-side mirror of the reference interface for the insertion path.

Reference                                             here
  Parameters(k, m, b)      brisk/parameters.hpp:24      Parameters(k, m, b)
  Brisk<DATA>(params)      brisk/Brisk.hpp:47           Brisk(params, data_size=1, ...)
  count_sequence           apps/counter.cpp:231         Brisk.count_reads(bases, offsets)
  insert_superkmer         brisk/Brisk.hpp:123          Brisk.insert_reads / insert_kmers  (batched)
  get_superkmer / get      brisk/Brisk.hpp:102 / :64    Brisk.get_reads / get_kmers        (batched)
  next / restart           brisk/Brisk.hpp:166-179      Brisk.dump()
  stats                    brisk/Brisk.hpp:194          Brisk.stats()
  SuperKmerEnumerator      brisk/Kmers.cpp:509-603      Brisk.enumerate_reads + decode_records

Arrays are numpy arrays on the host or anything exposing a CUDA pointer through `.data_ptr()`
(torch tensors) — pointers are handed to the C ABI untouched.
"""
import ctypes as C

import numpy as np

from . import _capi

KMER_DT = np.dtype([("lo", "<u8"), ("hi", "<u8"), ("idx", "u1"), ("pad", "u1", (7,))])
NO_SLOT = np.uint64(0xFFFFFFFFFFFFFFFF)
F_MANAGED = 1
F_LITERAL_DECYCLE = 2
F_SYNC_INSERT = 4
F_GENERIC_ENUM = 8
F_SERIAL_EXCHANGE = 16


class BriskError(RuntimeError):
    pass


def lib_path():
    return _capi.LIB_PATH


class Parameters:
    """Parameters(k, m, b) — brisk/parameters.hpp:24-34 (derived constants included)."""

    def __init__(self, k, m, b):
        self.k, self.m, self.b = int(k), int(m), int(b)
        self.m_reduc = self.m - self.b
        self.compacted_size = self.k - self.b
        self.allocated_bytes = (2 * self.k - self.m - self.b + 3) // 4
        self.mask_large_minimizer = (1 << (2 * self.m)) - 1


def _ptr(a):
    if a is None:
        return None
    if hasattr(a, "data_ptr"):
        return C.c_void_p(a.data_ptr())
    if isinstance(a, (bytes, bytearray)):
        return C.cast(C.c_char_p(bytes(a)), C.c_void_p)
    return a.ctypes.data_as(C.c_void_p)


class Brisk:
    """Brisk<DATA> over libbrisk_b200.so.  data_size = sizeof(DATA) (1 = the counter's uint8_t)."""

    def __init__(self, params, data_size=1, device=-1, capacity_hint=0, flags=0):
        if not isinstance(params, Parameters):
            params = Parameters(*params)
        self.params = params
        self.data_size = data_size
        self._lib = _capi.load()
        if device < 0:
            # the library links its own CUDA runtime, whose "current device" is not torch's: follow torch when the
            # caller has selected a device there (torch.cuda.set_device), else device 0
            import sys
            torch = sys.modules.get("torch")
            device = torch.cuda.current_device() if torch is not None and torch.cuda.is_initialized() else 0
        self.device = device
        h = C.c_void_p()
        rc = self._lib.brisk_create(params.k, params.m, params.b, data_size, device, capacity_hint, flags, C.byref(h))
        self._h = h if rc == 0 else None
        self._check(rc)

    def _check(self, rc):
        if rc != 0:
            raise BriskError("brisk_b200 error %d: %s" % (rc, self._lib.brisk_last_error().decode()))

    def close(self):
        if getattr(self, "_h", None):
            self._lib.brisk_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- reads
    @staticmethod
    def _n_reads(offsets):
        return (offsets.numel() if hasattr(offsets, "numel") else len(offsets)) - 1

    def enumerate_reads(self, bases, offsets):
        """(records[n, rec_words] uint64, rec_words) — one packed record per emitted superkmer."""
        n = self._n_reads(offsets)
        cnt, rw = C.c_uint64(), C.c_uint32()
        self._check(self._lib.brisk_enumerate_reads(self._h, _ptr(bases), _ptr(offsets), n, None, 0, C.byref(cnt), C.byref(rw)))
        recs = np.zeros((cnt.value, rw.value), np.uint64)
        if cnt.value:
            self._check(self._lib.brisk_enumerate_reads(self._h, _ptr(bases), _ptr(offsets), n, _ptr(recs), cnt.value,
                                                        C.byref(cnt), C.byref(rw)))
        return recs, rw.value

    def count_superkmers(self, bases, offsets):
        """number of superkmer records the batch enumerates to (nothing is copied out)"""
        cnt, rw = C.c_uint64(), C.c_uint32()
        self._check(self._lib.brisk_enumerate_reads(self._h, _ptr(bases), _ptr(offsets), self._n_reads(offsets), None, 0,
                                                    C.byref(cnt), C.byref(rw)))
        return cnt.value

    def enumerate_reads_tagged(self, bases, offsets):
        """(records, occ) with occ[i] = (first read-order occurrence of record i << 1) | reversed; records sorted by
        occ >> 1 are in SuperKmerEnumerator::next's emission order."""
        n = self._n_reads(offsets)
        cnt, rw = C.c_uint64(), C.c_uint32()
        self._check(self._lib.brisk_enumerate_reads_tagged(self._h, _ptr(bases), _ptr(offsets), n, None, 0, C.byref(cnt), C.byref(rw), None))
        recs = np.zeros((cnt.value, rw.value), np.uint64)
        occ = np.zeros(cnt.value, np.uint64)
        if cnt.value:
            self._check(self._lib.brisk_enumerate_reads_tagged(self._h, _ptr(bases), _ptr(offsets), n, _ptr(recs), cnt.value,
                                                               C.byref(cnt), C.byref(rw), _ptr(occ)))
        order = np.argsort(occ >> np.uint64(1), kind="stable")
        return recs[order], occ[order]

    def write_kff(self, path):
        """BriskWriter::write: this handle's stored set as a KFF file; returns (blocks, minimizer sections) written."""
        nb, ns = C.c_uint64(), C.c_uint64()
        self._check(self._lib.brisk_write_kff(self._h, str(path).encode(), C.byref(nb), C.byref(ns)))
        return nb.value, ns.value

    def reallocate(self, new_m=None, new_b=None):
        """Brisk::reallocate (brisk/Brisk.hpp:202-224): rebuild under (k, m + 2, b + 2) unless told otherwise"""
        new_m = self.params.m + 2 if new_m is None else new_m
        new_b = self.params.b + 2 if new_b is None else new_b
        self._check(self._lib.brisk_reparameterise(self._h, new_m, new_b))
        self.params = Parameters(self.params.k, new_m, new_b)

    def long_read_info(self):
        """(length limit above which a read is enumerated in segments, warm-up k-mers, repeated segmented launches)"""
        out = np.zeros(3, np.uint64)
        self._check(self._lib.brisk_long_read_info(self._h, _ptr(out)))
        return int(out[0]), int(out[1]), int(out[2])

    def checksum(self):
        """(sum of a 64-bit mix of every stored (kmer_s, minimizer_idx, DATA byte), sum of the DATA bytes, number of stored
        k-mers) of this handle's shard; order-independent, the shards' sums add up to the unsharded index's."""
        out = np.zeros(3, np.uint64)
        self._check(self._lib.brisk_checksum(self._h, _ptr(out)))
        return int(out[0]), int(out[1]), int(out[2])

    def count_reads(self, bases, offsets):
        self._check(self._lib.brisk_count_reads(self._h, _ptr(bases), _ptr(offsets), self._n_reads(offsets)))

    def n_occurrences(self, offsets):
        if hasattr(offsets, "cpu"):
            offsets = offsets.cpu().numpy()
        lens = np.diff(np.asarray(offsets).astype(np.int64))
        return int(np.maximum(lens - self.params.k + 1, 0).sum())

    def insert_reads(self, bases, offsets):
        """(slots, newly) per k-mer occurrence in read order."""
        n_occ = self.n_occurrences(offsets)
        slots = np.zeros(n_occ, np.uint64)
        newly = np.zeros(n_occ, np.uint8)
        self._check(self._lib.brisk_insert_reads(self._h, _ptr(bases), _ptr(offsets), self._n_reads(offsets), _ptr(slots), _ptr(newly)))
        return slots, newly

    def get_reads(self, bases, offsets):
        """(found, data) per k-mer occurrence in read order."""
        n_occ = self.n_occurrences(offsets)
        found = np.zeros(n_occ, np.uint8)
        data = np.zeros((n_occ, self.data_size), np.uint8)
        self._check(self._lib.brisk_get_reads(self._h, _ptr(bases), _ptr(offsets), self._n_reads(offsets), _ptr(found), _ptr(data)))
        return found, (data[:, 0] if self.data_size == 1 else data)

    def query_reads(self, bases, offsets):
        """(hits, total) — what query_sequence accumulates (apps/counter.cpp:281-310)."""
        hits, total = C.c_uint64(), C.c_uint64()
        self._check(self._lib.brisk_query_reads(self._h, _ptr(bases), _ptr(offsets), self._n_reads(offsets), C.byref(hits), C.byref(total)))
        return hits.value, total.value

    # ---- FASTA ingest (getLineFasta + clean_dna of apps/counter.cpp on the device)
    @staticmethod
    def _text(text):
        if isinstance(text, (bytes, bytearray)):
            text = np.frombuffer(bytes(text), np.uint8)
        n = text.numel() if hasattr(text, "numel") else len(text)
        return text, n

    def parse_fasta(self, text):
        """(bases, offsets) of the reads the reference's reader loop would hand to count_sequence."""
        text, n = self._text(text)
        nb, nr = C.c_uint64(), C.c_uint64()
        self._check(self._lib.brisk_parse_fasta(self._h, _ptr(text) if n else None, n, None, 0, None, 0, C.byref(nb), C.byref(nr)))
        bases = np.zeros(nb.value, np.uint8)
        offs = np.zeros(nr.value + 1, np.uint64)
        self._check(self._lib.brisk_parse_fasta(self._h, _ptr(text) if n else None, n, _ptr(bases), nb.value, _ptr(offs), nr.value,
                                                C.byref(nb), C.byref(nr)))
        return bases, offs

    def count_fasta(self, text):
        text, n = self._text(text)
        nr = C.c_uint64()
        self._check(self._lib.brisk_count_fasta(self._h, _ptr(text) if n else None, n, C.byref(nr)))
        return nr.value

    def query_fasta(self, text):
        text, n = self._text(text)
        hits, total = C.c_uint64(), C.c_uint64()
        self._check(self._lib.brisk_query_fasta(self._h, _ptr(text) if n else None, n, C.byref(hits), C.byref(total)))
        return hits.value, total.value

    # ---- the two halves of the routed path (multi-GPU building blocks)
    def partition_reads(self, bases, offsets, n_ranks):
        """(records grouped by owner rank, starts[n_ranks + 1]) — what a rank sends in the exchange."""
        n = self._n_reads(offsets)
        starts = np.zeros(n_ranks + 1, np.uint64)
        rw = C.c_uint32()
        self._check(self._lib.brisk_partition_reads(self._h, _ptr(bases), _ptr(offsets), n, n_ranks, None, 0, _ptr(starts), C.byref(rw)))
        recs = np.zeros((int(starts[-1]), rw.value), np.uint64)
        if len(recs):
            self._check(self._lib.brisk_partition_reads(self._h, _ptr(bases), _ptr(offsets), n, n_ranks, _ptr(recs), len(recs),
                                                        _ptr(starts), C.byref(rw)))
        return recs, starts

    def apply_records(self, recs, op="count"):
        """Insert / count / look up superkmer records on this handle; returns (hits, total) for op='get'."""
        code = {"count": 0, "insert": 1, "get": 2}[op]
        hits, total = C.c_uint64(), C.c_uint64()
        n = recs.shape[0]
        self._check(self._lib.brisk_apply_records(self._h, _ptr(recs) if n else None, n, code, C.byref(hits), C.byref(total)))
        return hits.value, total.value

    # ---- sharded index: one process per GPU
    def comm_init(self, unique_id, rank, n_ranks):
        buf = (C.c_uint8 * 128).from_buffer_copy(bytes(unique_id))
        self._check(self._lib.brisk_comm_init(self._h, buf, rank, n_ranks))

    def comm_unique_id(self):
        buf = (C.c_uint8 * 128)()
        self._check(self._lib.brisk_comm_unique_id(buf))
        return bytes(buf)

    def comm_init_torch(self, dist, rank, n_ranks):
        """Bootstraps the library's NCCL communicator over an initialised torch.distributed group."""
        import torch
        dev = torch.device("cuda") if dist.get_backend() == "nccl" else torch.device("cpu")
        t = torch.zeros(128, dtype=torch.uint8)
        if rank == 0:
            t = torch.frombuffer(bytearray(self.comm_unique_id()), dtype=torch.uint8).clone()
        t = t.to(dev)
        dist.broadcast(t, src=0)
        self.comm_init(t.cpu().numpy().tobytes(), rank, n_ranks)

    def comm_info(self):
        r, n = C.c_int(), C.c_int()
        self._check(self._lib.brisk_comm_info(self._h, C.byref(r), C.byref(n)))
        return r.value, n.value

    def set_stream(self, cuda_stream):
        self._check(self._lib.brisk_set_stream(self._h, C.c_void_p(cuda_stream)))

    # ---- explicit k-mers
    @staticmethod
    def pack_kmers(lo, hi, idx):
        k = np.zeros(len(lo), KMER_DT)
        k["lo"], k["hi"], k["idx"] = lo, hi, idx
        return k

    def insert_kmers(self, kmers, count=False):
        slots = np.zeros(len(kmers), np.uint64)
        newly = np.zeros(len(kmers), np.uint8)
        self._check(self._lib.brisk_insert_kmers(self._h, _ptr(kmers), len(kmers), _ptr(slots), _ptr(newly), int(count)))
        return slots, newly

    def get_kmers(self, kmers):
        slots = np.zeros(len(kmers), np.uint64)
        data = np.zeros((len(kmers), self.data_size), np.uint8)
        self._check(self._lib.brisk_get_kmers(self._h, _ptr(kmers), len(kmers), _ptr(slots), _ptr(data)))
        return slots, (data[:, 0] if self.data_size == 1 else data)

    def data_ptr(self, slot):
        return self._lib.brisk_data_ptr(self._h, int(slot))

    # ---- whole index
    def dump(self, ordered=False):
        """(lo, hi, idx, data) of every stored k-mer; ordered=True: by bucket id like DenseMenuYo::next, then by
        (minimizer_idx, kmer_s)."""
        fn = self._lib.brisk_dump_ordered if ordered else self._lib.brisk_dump
        n = C.c_uint64()
        self._check(fn(self._h, None, None, 0, C.byref(n)))
        keys = np.zeros(n.value, KMER_DT)
        data = np.zeros((n.value, self.data_size), np.uint8)
        if n.value:
            self._check(fn(self._h, _ptr(keys), _ptr(data), n.value, C.byref(n)))
        return keys["lo"].copy(), keys["hi"].copy(), keys["idx"].copy(), (data[:, 0] if self.data_size == 1 else data)

    def stats(self):
        v = [C.c_uint64() for _ in range(5)]
        self._check(self._lib.brisk_stats(self._h, *[C.byref(x) for x in v]))
        names = ("nb_buckets", "nb_skmers", "nb_kmers", "memory_usage", "largest_bucket")
        return {n: x.value for n, x in zip(names, v)}

    def reserve(self, n_keys):
        self._check(self._lib.brisk_reserve(self._h, n_keys))

    def capacity(self):
        c = C.c_uint64()
        self._check(self._lib.brisk_capacity(self._h, C.byref(c)))
        return c.value

    def sync(self):
        self._check(self._lib.brisk_sync(self._h))

    def last_timing(self):
        ms = (C.c_float * 5)()
        n = C.c_uint32()
        self._check(self._lib.brisk_last_timing(self._h, ms, C.byref(n)))
        return dict(pack_ms=ms[0], enumerate_ms=ms[1], partition_ms=ms[2], store_ms=ms[3], total_ms=ms[4], launches=n.value)


def owner_of_bucket(bucket_id, b, n_ranks):
    return _capi.load().brisk_owner_of_bucket(int(bucket_id), int(b), int(n_ranks))


def exchange_plan(counts, rank):
    """(send_start, recv_start) of `rank` from the n x n matrix of record counts (row = sender)."""
    counts = np.ascontiguousarray(counts, dtype=np.uint64)
    n = counts.shape[0]
    ss, rs = np.zeros(n + 1, np.uint64), np.zeros(n + 1, np.uint64)
    rc = _capi.load().brisk_exchange_plan(_ptr(counts), n, rank, _ptr(ss), _ptr(rs))
    if rc:
        raise BriskError("brisk_exchange_plan: %d" % rc)
    return ss, rs


def bucket_ids(params, lo, hi, idx):
    """bucket id of every (kmer_s, minimizer_idx): the middle 2b bits of the hashed m-mer found at minimizer_idx
    (brisk/Brisk.hpp:135-137 after hash_kmer_minimizer_inplace; the integer rounds of brisk/hashing.cpp:10-16).  Host-side
    helper for merging the ordered dumps of the shards of a sharded index; numpy, no device."""
    k, m, b = params.k, params.m, params.b
    M64 = (1 << 64) - 1
    mask = (1 << (2 * m)) - 1
    out = np.zeros(len(lo), np.uint64)
    s = (m - b + 1) // 2
    for i in range(len(lo)):
        v = (int(hi[i]) << 64) | int(lo[i])
        key = (v >> (2 * int(idx[i]))) & mask
        key = ((~key & M64) + (key << 21)) & M64 & mask
        key ^= key >> 24
        key = ((key + (key << 3)) + (key << 8)) & M64 & mask
        key ^= key >> 14
        key = ((key + (key << 2)) + (key << 4)) & M64 & mask
        key ^= key >> 28
        key = (key + (key << 31)) & M64 & mask
        out[i] = (key >> (2 * s)) & ((1 << (2 * b)) - 1)
    return out


def merge_ordered_dumps(params, parts):
    """The ordered dumps (lo, hi, idx, data) of the shards of a sharded index merged into the order brisk_dump_ordered
    gives for an unsharded index: increasing bucket id (DenseMenuYo::next, brisk/DenseMenuYo.hpp:476-521), then
    (minimizer_idx, kmer_s).  Buckets are disjoint over the shards but interleaved (the owner of a bucket is a scramble
    of its id), so the concatenation in rank order is not in bucket order: this is the k-way merge."""
    lo, hi, ix, dat = (np.concatenate([p[i] for p in parts]) for i in range(4))
    order = np.lexsort((lo, hi, ix, bucket_ids(params, lo, hi, ix)))
    return lo[order], hi[order], ix[order], dat[order]


def decode_records(recs, params):
    """Unpacks superkmer records (layout in DESIGN.md / csrc/common.cuh) into per-superkmer lists of
    (kmer_s as int, minimizer_idx), in the reference's emission order, plus the bucket id.  Host-side
    helper for tests and tools; pure bit unpacking, no algorithm."""
    k, m, b = params.k, params.m, params.b
    w = k - m + 1
    wb = 1
    while (1 << wb) < w:
        wb += 1
    rw = recs.shape[1]
    sb = 64 - 2 * b
    hb = 2 * b + 2 * wb
    out = []
    kmask = (1 << (2 * k)) - 1
    for row in recs:
        top = int(row[rw - 1])
        bucket = top >> sb
        n = ((top >> (sb - wb)) & ((1 << wb) - 1)) + 1
        idx0 = (top >> (sb - 2 * wb)) & ((1 << wb) - 1)
        val = 0
        for i in range(rw - 1, -1, -1):
            word = int(row[i])
            if i == rw - 1:
                word &= (1 << (64 - hb)) - 1
            val = (val << 64) | word
        kmers = [((val >> (2 * (n - 1 - j))) & kmask, idx0 + j) for j in range(n)]
        out.append((bucket, kmers))
    return out
