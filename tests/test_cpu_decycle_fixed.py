"""The rolled fixed-point decycling class of enumerate.cuh replayed on the host with the library's own
tables (brisk_decycle_tables, pure host arithmetic) against the oracle's DecyclingSet::memDouble: wherever the
fixed-point values are outside the threshold band the class must be the literal one; inside the band the kernel
recomputes literally, so the band only has to be rare."""
import ctypes as C

import numpy as np
import pytest

from brisk_b200 import _capi
from oracle.oracle import Oracle


def tables(k, m, b):
    lib = _capi.load()
    tw = np.zeros(64, np.int32)
    rot = np.zeros(64, np.int32)
    th = C.c_int32()
    groups = (m + 2) // 3
    gt = np.zeros(groups * 64 * 2, np.int32)
    rc = lib.brisk_decycle_tables(k, m, b, tw.ctypes.data_as(C.c_void_p), rot.ctypes.data_as(C.c_void_p), C.byref(th),
                                  gt.ctypes.data_as(C.c_void_p))
    assert rc == 0
    return tw.reshape(32, 2).astype(np.int64), rot.reshape(32, 2).astype(np.int64), int(th.value), gt.reshape(groups, 64, 2).astype(np.int64)


def mulhi(a, b):
    return (a * b) >> 32  # floor, like __mulhi on the signed 64-bit product


def classify(D0, D1):
    """dec_class_signs of enumerate.cuh: 0 when D0 > 0 > D1, 1 when D0 < 0 < D1, else 2 (either strand)"""
    if (D0 < 0) != (D1 < 0):
        return 1 if D1 > 0 else 0
    return 2


@pytest.mark.parametrize("k,m,b", [(63, 21, 14), (31, 15, 12), (64, 23, 14), (31, 19, 16), (40, 11, 8), (47, 31, 16), (21, 7, 4)])
def test_rolled_fixed_point_class_is_the_literal_class(k, m, b):
    tw, rot, th, _ = tables(k, m, b)
    o = Oracle(k, m, b)
    rng = np.random.default_rng(m)
    n_reads, L = 64, 400
    codes = rng.integers(0, 4, (n_reads, L)).astype(np.int64)
    codes[0, :] = 0          # poly-A: R = 0 everywhere, always inside the band
    codes[1, :] = np.arange(L) % 2  # ACAC...
    codes[2, :] = 3
    m_mask = (1 << (2 * m)) - 1
    f = np.zeros(n_reads, np.int64).astype(object)
    A = np.zeros((n_reads, 2), np.int64)   # X+ = sum of code * tw
    Bp = np.zeros((n_reads, 2), np.int64)  # X- = sum of ((code ^ 2) - 2) * tw
    checked = amb_total = 0
    fwd = [0] * n_reads
    for t in range(L):
        q = t % m
        code = codes[:, t]
        leave = codes[:, t - m] if t >= m else np.zeros(n_reads, np.int64)
        A += (code - leave)[:, None] * tw[q][None, :]
        Bp += ((code ^ 2) - (leave ^ 2))[:, None] * tw[q][None, :]
        for i in range(n_reads):
            fwd[i] = ((fwd[i] << 2) | int(code[i])) & m_mask
        if t + 1 < m:
            continue
        p = (q + 1) % m
        for i in range(n_reads):
            v = fwd[i]
            rc = 0
            x = v
            for _ in range(m):
                rc = (rc << 2) | ((x & 3) ^ 2)
                x >>= 2
            rv = rc < v
            canon = rc if rv else v
            Xr = Bp[i, 0] if rv else A[i, 0]
            Xi = Bp[i, 1] if rv else A[i, 1]
            assert abs(Xr) < 2 ** 31 and abs(Xi) < 2 ** 31
            D0 = mulhi(Xi, rot[p, 0]) - mulhi(Xr, rot[p, 1])
            D1 = mulhi(Xi, rot[q, 0]) - mulhi(Xr, rot[q, 1])
            if min(abs(D0), abs(D1)) <= th:
                amb_total += 1
                continue
            checked += 1
            assert classify(int(D0), int(D1)) == o.mem_double(canon), (t, i, D0, D1)
    # random reads: the band is rare (the three low-complexity reads are always inside it)
    # (small m: 4^m m-mers only, sums that vanish exactly are common; that costs speed, not correctness)
    assert checked > 0 and amb_total <= 3 * (L - m + 1) + (0.002 if m >= 11 else 0.1) * n_reads * L


@pytest.mark.parametrize("k,m,b", [(63, 21, 14), (64, 23, 14), (40, 11, 8), (47, 31, 16)])
def test_grouped_table_class_is_the_literal_class(k, m, b):
    """class of arbitrary (pseudo) m-mers, as the rescan of get_minimizer needs them for k > 32"""
    tw, rot, th, gt = tables(k, m, b)
    o = Oracle(k, m, b)
    rng = np.random.default_rng(k)
    groups = gt.shape[0]
    vals = [int(x) for x in rng.integers(0, 1 << 62, 3000)]
    vals = [v & ((1 << (2 * m)) - 1) for v in vals]
    vals += [v >> (2 * int(s)) for v, s in zip(vals[:1500], rng.integers(1, m, 1500))]  # zero-padded tops
    vals += [0, 1, 2, 3, (1 << (2 * m)) - 1]
    checked = 0
    for v in vals:
        Sr = sum(int(gt[g, (v >> (6 * g)) & 63, 0]) for g in range(groups))
        Si = sum(int(gt[g, (v >> (6 * g)) & 63, 1]) for g in range(groups))
        R = Si >> 2
        R2 = mulhi(Sr, int(rot[1, 1])) + mulhi(Si, int(rot[1, 0]))
        if v < 4:  # only the last nucleotide set: decided without the tables (dec_class_tab)
            assert o.mem_double(v) == (1 if v else 2)
            continue
        if min(abs(R), abs(R2)) <= th:
            continue
        checked += 1
        assert classify(int(R), int(R2)) == o.mem_double(v), v
    assert checked > 0.95 * len(vals) - 10
