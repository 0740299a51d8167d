"""Host-input batches large enough to arrive in >= 4 chunks take run_reads_overlapped (brisk_capi.cu): the store kernel of
the first chunks runs on a second stream while the last chunks are still being copied.  Whatever the schedule, the index
must be the one the plain path builds from the same reads held in device memory (one chunk, one store kernel):
count_sequence / query_sequence of apps/counter.cpp:231-310 do not depend on how a batch is cut."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _reads(n_reads, L, genome, seed):
    """error-free reads of a random genome as one flat ASCII array + offsets (host)"""
    rng = np.random.default_rng(seed)
    g = np.frombuffer(b"ACGT", np.uint8)[rng.integers(0, 4, genome)]
    pos = rng.integers(0, genome - L, n_reads)
    bases = g[(pos[:, None] + np.arange(L)[None, :])].reshape(-1)
    offs = np.arange(n_reads + 1, dtype=np.uint64) * np.uint64(L)
    return np.ascontiguousarray(bases), offs


def _both_ways(k, m, b, bases, offs, expect_overlap):
    import torch
    from brisk_b200 import Brisk, Parameters
    n_chunks = 8 if len(bases) >= 7 * (8 << 20) else None
    host = Brisk(Parameters(k, m, b))
    dev = Brisk(Parameters(k, m, b))
    d_bases = torch.from_numpy(bases).cuda()
    d_offs = torch.from_numpy(offs.astype(np.int64)).cuda()
    for rnd in range(2):  # the second pass finds every k-mer again (counts 2) and settles the first pass's pending kernel
        host.count_reads(bases, offs)
        if rnd == 0:
            launches = host.last_timing()["launches"]
        dev.count_reads(d_bases, d_offs)
    assert host.checksum() == dev.checksum()
    assert host.stats()["nb_kmers"] == dev.stats()["nb_kmers"]
    q_host = host.query_reads(bases, offs)
    q_launches = host.last_timing()["launches"]
    assert q_host == dev.query_reads(d_bases, d_offs) == host.query_reads(d_bases, d_offs)
    n_occ = int(np.maximum(np.diff(offs.astype(np.int64)) - k + 1, 0).sum())
    assert q_host[0] == n_occ
    if expect_overlap and n_chunks:
        # 8 pack + 8 enumerate launches, then (3 partition + 1 store) per group instead of once
        assert launches == 2 * n_chunks + 8 and q_launches == 2 * n_chunks + 8
    return launches


@pytest.mark.parametrize("k,m,b,L", [(63, 21, 14, 250), (31, 15, 12, 150)])
def test_overlapped_store_builds_the_same_index(k, m, b, L):
    n_reads = (72 << 20) // L
    bases, offs = _reads(n_reads, L, 4_000_000, 21)
    _both_ways(k, m, b, bases, offs, True)


@pytest.mark.parametrize("where", ["first_group", "second_group"])
def test_overlapped_store_with_a_long_read(where):
    # one read of 6000 bases (> the 2048-base limit of the thread-per-read launch): the segmented launch appends its
    # records to the group whose chunks hold it
    L, n_reads = 250, (72 << 20) // 250
    bases, offs = _reads(n_reads, L, 4_000_000, 22)
    at = n_reads // 10 if where == "first_group" else n_reads - n_reads // 10
    offs = np.concatenate([offs[:at + 1], offs[at + 24:]])  # reads at .. at+23 become one
    launches = _both_ways(63, 21, 14, bases, offs, False)
    assert launches > 0


@pytest.mark.parametrize("group", [0, 1])
def test_overlapped_store_when_a_record_region_is_too_small(group, monkeypatch):
    # the first group's region too small: nothing has been stored yet and the plain path takes the batch; the second
    # group's: its chunks are enumerated again the plain way after the first group's kernels are done
    monkeypatch.setenv("BRISK_OVERLAP_TEST_CAP", "%d:1000" % group)
    L, n_reads = 250, (72 << 20) // 250
    bases, offs = _reads(n_reads, L, 4_000_000, 23)
    _both_ways(63, 21, 14, bases, offs, False)
