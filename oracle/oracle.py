"""ctypes binding of oracle/liboracle.so (the CPU restatement, brisk_oracle.c).

TEST INFRASTRUCTURE ONLY — imported by tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs as the checker; never by the
product package brisk_b200.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def build(force=False):
    so = os.path.join(_HERE, "liboracle.so")
    src = os.path.join(_HERE, "brisk_oracle.c")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-s", "-C", _HERE, "liboracle.so"])
    return so


def _lib():
    global _LIB
    if _LIB is None:
        L = C.CDLL(build())
        L.oracle_new.restype = C.c_void_p
        L.oracle_new.argtypes = [C.c_int] * 3
        L.oracle_free.argtypes = [C.c_void_p]
        L.oracle_mem_double.restype = C.c_uint
        L.oracle_mem_double.argtypes = [C.c_void_p, C.c_uint64]
        L.oracle_bfc_hash.restype = C.c_uint64
        L.oracle_bfc_hash.argtypes = [C.c_void_p, C.c_uint64]
        L.oracle_bfc_hash_inv.restype = C.c_uint64
        L.oracle_bfc_hash_inv.argtypes = [C.c_void_p, C.c_uint64]
        L.oracle_rcbc.restype = C.c_uint64
        L.oracle_rcbc.argtypes = [C.c_uint64, C.c_uint64]
        L.oracle_coef.restype = C.c_double
        L.oracle_coef.argtypes = [C.c_void_p, C.c_int]
        L.oracle_get_minimizer.restype = C.c_uint64
        L.oracle_get_minimizer.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_int, C.c_void_p, C.c_void_p]
        L.oracle_hash_stream.restype = C.c_int64
        L.oracle_hash_stream.argtypes = [C.c_void_p, C.c_char_p, C.c_uint64, C.c_void_p, C.c_void_p]
        L.oracle_mmer_hashes.restype = C.c_int64
        L.oracle_mmer_hashes.argtypes = [C.c_void_p, C.c_char_p, C.c_uint64, C.c_void_p, C.c_void_p]
        L.oracle_enumerate.restype = C.c_int64
        L.oracle_enumerate.argtypes = [C.c_void_p, C.c_char_p, C.c_uint64, C.c_uint64] + [C.c_void_p] * 8
        L.oracle_count_read.argtypes = [C.c_void_p, C.c_char_p, C.c_uint64]
        L.oracle_count_reads.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64]
        L.oracle_query_read.restype = C.c_uint64
        L.oracle_query_read.argtypes = [C.c_void_p, C.c_char_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_int]
        L.oracle_insert_kmers.argtypes = [C.c_void_p] * 4 + [C.c_uint64, C.c_void_p, C.c_int]
        L.oracle_get_kmers.argtypes = [C.c_void_p] * 4 + [C.c_uint64, C.c_void_p, C.c_void_p]
        L.oracle_size.restype = C.c_uint64
        L.oracle_size.argtypes = [C.c_void_p]
        L.oracle_dump.restype = C.c_uint64
        L.oracle_dump.argtypes = [C.c_void_p] * 5
        L.oracle_stats.argtypes = [C.c_void_p] * 3
        _LIB = L
    return _LIB


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


class Oracle:
    """Parameters(k, m, b) + Brisk<uint8_t> of the reference, restated on the CPU."""

    def __init__(self, k, m, b):
        self.k, self.m, self.b = k, m, b
        self.L = _lib()
        self.h = self.L.oracle_new(k, m, b)
        if not self.h:
            raise ValueError("bad parameters (m must be odd, m < k <= 64, b <= m)")

    def __del__(self):
        if getattr(self, "h", None):
            self.L.oracle_free(self.h)
            self.h = None

    # --- scalar pieces
    def mem_double(self, mmer):
        return self.L.oracle_mem_double(self.h, mmer)

    def bfc_hash(self, mmer):
        return self.L.oracle_bfc_hash(self.h, mmer)

    def bfc_hash_inv(self, x):
        return self.L.oracle_bfc_hash_inv(self.h, x)

    def coef(self):
        return np.array([self.L.oracle_coef(self.h, i) for i in range(4 * self.m)])

    def get_minimizer(self, lo, hi, kk):
        pos = C.c_uint8()
        rev = C.c_uint8()
        mini = self.L.oracle_get_minimizer(self.h, lo, hi, kk, C.byref(pos), C.byref(rev))
        return mini, pos.value, rev.value

    # --- per read
    def hash_stream(self, seq: bytes):
        n = max(0, len(seq) - self.k + 1)
        h = np.zeros(n, np.uint64)
        r = np.zeros(n, np.uint8)
        got = self.L.oracle_hash_stream(self.h, seq, len(seq), _p(h), _p(r))
        assert got == n
        return h, r

    def mmer_hashes(self, seq: bytes):
        n = max(0, len(seq) - self.m + 1)
        h = np.zeros(n, np.uint64)
        r = np.zeros(n, np.uint8)
        got = self.L.oracle_mmer_hashes(self.h, seq, len(seq), _p(h), _p(r))
        assert got == n
        return h, r

    def enumerate(self, seq: bytes):
        """dict of arrays, one entry per emitted k-mer, in the reference's emission order."""
        n = max(0, len(seq) - self.k + 1)
        out = dict(
            lo=np.zeros(n, np.uint64), hi=np.zeros(n, np.uint64), idx=np.zeros(n, np.uint8),
            mini=np.zeros(n, np.uint64), skmer=np.zeros(n, np.uint32), ridx=np.zeros(n, np.uint32),
            pos=np.zeros(n, np.uint8), rev=np.zeros(n, np.uint8))
        got = self.L.oracle_enumerate(self.h, seq, len(seq), n, _p(out["lo"]), _p(out["hi"]), _p(out["idx"]),
                                      _p(out["mini"]), _p(out["skmer"]), _p(out["ridx"]), _p(out["pos"]),
                                      _p(out["rev"]))
        assert got == n, (got, n)
        return out

    def superkmers(self, seq: bytes):
        """[(start k-mer index in the read, n k-mers, reversed, idx of the first read k-mer)] per superkmer."""
        e = self.enumerate(seq)
        res = []
        if len(e["skmer"]) == 0:
            return res
        order = np.lexsort((e["ridx"], e["skmer"]))
        sk = e["skmer"][order]
        bounds = np.flatnonzero(np.diff(sk)) + 1
        for grp in np.split(order, bounds):
            first = grp[0]
            res.append((int(e["ridx"][first]), len(grp), int(e["rev"][first]), int(e["idx"][first])))
        return res

    # --- the index
    def count_read(self, seq: bytes):
        self.L.oracle_count_read(self.h, seq, len(seq))

    def count_reads(self, bases: np.ndarray, offsets: np.ndarray):
        bases = np.ascontiguousarray(bases, np.uint8)
        offsets = np.ascontiguousarray(offsets, np.uint64)
        self.L.oracle_count_reads(self.h, _p(bases), _p(offsets), len(offsets) - 1)

    def query_read(self, seq: bytes, literal_break=False):
        n = max(0, len(seq) - self.k + 1)
        found = np.zeros(n, np.uint8)
        data = np.zeros(n, np.uint8)
        total = self.L.oracle_query_read(self.h, seq, len(seq), _p(found), _p(data), int(literal_break))
        return found, data, total

    def insert_kmers(self, lo, hi, idx, count=True):
        lo = np.ascontiguousarray(lo, np.uint64)
        hi = np.ascontiguousarray(hi, np.uint64)
        idx = np.ascontiguousarray(idx, np.uint8)
        newly = np.zeros(len(lo), np.uint8)
        self.L.oracle_insert_kmers(self.h, _p(lo), _p(hi), _p(idx), len(lo), _p(newly), int(count))
        return newly

    def get_kmers(self, lo, hi, idx):
        lo = np.ascontiguousarray(lo, np.uint64)
        hi = np.ascontiguousarray(hi, np.uint64)
        idx = np.ascontiguousarray(idx, np.uint8)
        found = np.zeros(len(lo), np.uint8)
        data = np.zeros(len(lo), np.uint8)
        self.L.oracle_get_kmers(self.h, _p(lo), _p(hi), _p(idx), len(lo), _p(found), _p(data))
        return found, data

    def size(self):
        return self.L.oracle_size(self.h)

    def dump(self):
        """(lo, hi, idx, count) sorted by (kmer_s, idx) — the set Brisk::next/get exposes."""
        n = self.size()
        lo = np.zeros(n, np.uint64)
        hi = np.zeros(n, np.uint64)
        idx = np.zeros(n, np.uint8)
        cnt = np.zeros(n, np.uint8)
        got = self.L.oracle_dump(self.h, _p(lo), _p(hi), _p(idx), _p(cnt))
        assert got == n
        return lo, hi, idx, cnt

    def stats(self):
        nb = C.c_uint64()
        nk = C.c_uint64()
        self.L.oracle_stats(self.h, C.byref(nb), C.byref(nk))
        return dict(nb_buckets=nb.value, nb_kmers=nk.value)
