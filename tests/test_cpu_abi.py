"""CPU-side checks: the C-ABI library loads and exports every symbol include/brisk_b200.h declares, fails
loudly without a GPU (no CPU fallback), and the host-side helpers (record decoding, bucket -> owner
routing) behave.  No compute call is made here."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _lib():
    from brisk_b200 import _capi
    if not os.path.exists(_capi.LIB_PATH):
        import __graft_entry__ as g
        g.build()
    return _capi.load()


def test_library_exports_every_declared_symbol():
    lib = _lib()
    hdr = open(os.path.join(ROOT, "include", "brisk_b200.h")).read()
    declared = set(re.findall(r"\b(brisk_[a-z_0-9]+)\s*\(", hdr))
    declared -= {"brisk_handle"}
    assert len(declared) >= 20
    from brisk_b200 import _capi
    assert declared == set(_capi.PROTOTYPES), declared ^ set(_capi.PROTOTYPES)
    for name in declared:
        assert hasattr(lib, name), name
    assert b"sm_100a" in lib.brisk_version()


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from brisk_b200 import Brisk, BriskError, Parameters
    with pytest.raises(BriskError, match="no CUDA device"):
        Brisk(Parameters(31, 15, 12))


def test_bad_parameters_are_rejected_before_touching_the_device():
    lib = _lib()
    h = C.c_void_p()
    for (k, m, b) in [(31, 14, 12), (15, 15, 4), (65, 21, 14), (31, 15, 17), (31, 15, 0), (63, 33, 14)]:
        assert lib.brisk_create(k, m, b, 1, -1, 0, 0, C.byref(h)) == -1
        assert b"bad parameters" in lib.brisk_last_error()


def test_owner_of_bucket_is_a_balanced_partition():
    lib = _lib()
    rng = np.random.default_rng(0)
    # bucket ids biased towards small values, like the middle bits of a window minimum
    ids = (rng.integers(0, 1 << 28, 200000) * rng.random(200000) ** 2).astype(np.uint64)
    for n in (1, 2, 4, 8):
        owners = np.array([lib.brisk_owner_of_bucket(int(x), 14, n) for x in ids[:20000]])
        assert owners.min() >= 0 and owners.max() < n
        share = np.bincount(owners, minlength=n) / len(owners)
        assert share.max() < 1.15 / n and share.min() > 0.85 / n


def test_decode_records_layout():
    from brisk_b200 import Parameters, decode_records
    p = Parameters(31, 15, 12)
    # one record, n = 3 k-mers, idx0 = 5, bucket 0xABCDEF: header in the top 24 + 5 + 5 bits of word 1
    n, idx0, bucket = 3, 5, 0xABCDEF
    seq = int.from_bytes(np.random.default_rng(1).bytes(9), "little") & ((1 << (2 * (n + 30))) - 1)
    top = (bucket << 40) | ((n - 1) << 35) | (idx0 << 30)
    val = seq | (top << 64)
    rec = np.array([[val & 0xFFFFFFFFFFFFFFFF, val >> 64]], np.uint64)
    (b, kmers), = decode_records(rec, p)
    assert b == bucket and [i for _, i in kmers] == [5, 6, 7]
    assert kmers[0][0] == seq >> 4 and kmers[2][0] == seq & ((1 << 62) - 1)


def test_create_flags_match_the_header():
    """the BRISK_F_* values of include/brisk_b200.h and the Python mirror are the same numbers"""
    import re
    from brisk_b200 import brisk as B
    text = open(os.path.join(os.path.dirname(__file__), "..", "include", "brisk_b200.h")).read()
    flags = {m.group(1): int(m.group(2)) for m in re.finditer(r"#define BRISK_F_(\w+)\s+(\d+)u", text)}
    assert flags == {"MANAGED": B.F_MANAGED, "LITERAL_DECYCLE": B.F_LITERAL_DECYCLE, "SYNC_INSERT": B.F_SYNC_INSERT,
                     "GENERIC_ENUM": B.F_GENERIC_ENUM, "SERIAL_EXCHANGE": B.F_SERIAL_EXCHANGE}


def test_noalloc_entry_point_checks_its_arguments_first():
    # brisk_checksum_kmers_noalloc (no counterpart in the reference): the size limit of the static scratch and the null
    # checks come before anything touches the device, so they can be seen without one
    import ctypes as C
    from brisk_b200 import _capi
    lib = _capi.load()
    out = (C.c_uint64 * 3)()
    assert lib.brisk_checksum_kmers_noalloc(None, None, 1, out, None) == -1
    assert lib.brisk_checksum_kmers_noalloc(None, None, 0, None, None) == -1
    buf = (C.c_uint8 * 24)()
    assert lib.brisk_checksum_kmers_noalloc(buf, buf, (1 << 26) + 1, out, None) == -3
    msg = lib.brisk_last_error().decode()
    assert "limit" in msg and str(1 << 26) in msg
