"""brisk_checksum_kmers_noalloc: the one entry point whose scratch is a static __device__ array (no counterpart in the
reference; added for callers that cannot allocate).  It is checked against the
existing entry point, brisk_checksum of the handle the k-mer list was dumped from — integer sums, so the bar is
equality — on calls in a row that share the scratch.  Collected last on purpose (pytest -x)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
MAX_KMERS = 1 << 26  # BRISK_NOALLOC_MAX_KMERS, include/brisk_b200.h


def _lib():
    from brisk_b200 import _capi
    return _capi.load()


def _err(lib):
    return lib.brisk_last_error().decode()


def _index_and_device_dump(k, m, b, genome, L, seed):
    """an index over a seeded read set, its checksum by the existing entry point, and its dump in device memory"""
    import torch
    from brisk_b200 import Brisk, Parameters, synth
    bases, offs = synth.read_set(genome, L, 3, seed)
    n_reads = len(offs) - 1
    idx = Brisk(Parameters(k, m, b))
    idx.count_reads(bases, offs)
    idx.count_reads(bases[:offs[n_reads // 3]], offs[:n_reads // 3 + 1])  # counts of 1 and 2 (and more)
    want = idx.checksum()
    n = want[2]
    keys = torch.zeros(n * 24, dtype=torch.uint8, device="cuda")
    data = torch.zeros(n, dtype=torch.uint8, device="cuda")
    got_n = C.c_uint64()
    lib = _lib()
    assert lib.brisk_dump(idx._h, C.c_void_p(keys.data_ptr()), C.c_void_p(data.data_ptr()), n, C.byref(got_n)) == 0, _err(lib)
    assert got_n.value == n
    return idx, want, keys, data


def _noalloc(keys, data, n, out, stream=None):
    lib = _lib()
    rc = lib.brisk_checksum_kmers_noalloc(C.c_void_p(keys.data_ptr()) if keys is not None else None,
                                          C.c_void_p(data.data_ptr()) if data is not None else None, n,
                                          C.c_void_p(out.data_ptr()), C.c_void_p(stream) if stream else None)
    return rc


def test_noalloc_checksum_equals_the_handle_checksum_on_calls_in_a_row():
    import torch
    torch.cuda.set_device(0)
    # a large list first, a small one second: whatever the first call left in the shared scratch must not leak into
    # the second; then the large one again
    _, want_a, keys_a, data_a = _index_and_device_dump(31, 15, 12, 400000, 150, 11)
    _, want_b, keys_b, data_b = _index_and_device_dump(63, 21, 14, 20000, 250, 12)
    assert want_a[2] > 2 * 4096 and 0 < want_b[2] < want_a[2]
    out = torch.full((3,), 0x5555555555555555, dtype=torch.int64, device="cuda")
    lib = _lib()
    for rnd in range(2):
        for want, keys, data in ((want_a, keys_a, data_a), (want_b, keys_b, data_b)):
            out.fill_(0x5555555555555555)
            torch.cuda.synchronize()
            free_before = torch.cuda.mem_get_info()[0]
            assert _noalloc(keys, data, want[2], out) == 0, _err(lib)
            torch.cuda.synchronize()
            if rnd:  # the first round loads the kernel's module; after that the device's free memory must not move
                assert torch.cuda.mem_get_info()[0] >= free_before
            got = tuple(int(x) & 0xFFFFFFFFFFFFFFFF for x in out.cpu().tolist())
            assert got == want


def test_noalloc_checksum_on_a_stream_back_to_back_and_in_pieces():
    import torch
    torch.cuda.set_device(0)
    _, want, keys, data = _index_and_device_dump(31, 19, 16, 300000, 150, 13)
    n = want[2]
    lib = _lib()
    s = torch.cuda.Stream()
    # two calls queued on one stream without a wait in between, into two outputs; and the list in two pieces whose
    # sums add up (what a caller does beyond the limit)
    outs = torch.zeros((4, 3), dtype=torch.int64, device="cuda")
    torch.cuda.synchronize()
    cut = n // 3
    kb = keys.view(-1, 24)
    assert _noalloc(keys, data, n, outs[0], s.cuda_stream) == 0, _err(lib)
    assert _noalloc(keys, data, n, outs[1], s.cuda_stream) == 0, _err(lib)
    assert _noalloc(kb[:cut], data[:cut], cut, outs[2], s.cuda_stream) == 0, _err(lib)
    assert _noalloc(kb[cut:], data[cut:], n - cut, outs[3], s.cuda_stream) == 0, _err(lib)
    s.synchronize()
    o = outs.cpu().numpy().astype(np.uint64)
    assert tuple(int(x) for x in o[0]) == want and tuple(int(x) for x in o[1]) == want
    assert tuple(int(x) for x in (o[2] + o[3])) == want  # uint64 wrap-around sum


def test_noalloc_checksum_rejects_what_it_cannot_do():
    import torch
    torch.cuda.set_device(0)
    lib = _lib()
    out = torch.zeros(3, dtype=torch.int64, device="cuda")
    tiny = torch.zeros(24, dtype=torch.uint8, device="cuda")
    # beyond the static scratch's limit: refused before anything is read
    assert _noalloc(tiny, tiny, MAX_KMERS + 1, out) == -3
    assert "limit" in _err(lib) and str(MAX_KMERS) in _err(lib)
    # host memory is refused, not silently read
    host = np.zeros(24, np.uint8)
    assert lib.brisk_checksum_kmers_noalloc(host.ctypes.data_as(C.c_void_p), C.c_void_p(tiny.data_ptr()), 1,
                                            C.c_void_p(out.data_ptr()), None) == -1
    assert "device pointers" in _err(lib)
    host_out = np.zeros(3, np.uint64)
    assert lib.brisk_checksum_kmers_noalloc(C.c_void_p(tiny.data_ptr()), C.c_void_p(tiny.data_ptr()), 1,
                                            host_out.ctypes.data_as(C.c_void_p), None) == -1
    # the empty list: three zeros
    out.fill_(7)
    assert _noalloc(None, None, 0, out) == 0, _err(lib)
    torch.cuda.synchronize()
    assert out.cpu().tolist() == [0, 0, 0]
