"""Long reads (one thread per read does not survive contigs).  Reads longer than 2048 bases are enumerated in
segments of 1024 k-mers by different threads, each warming up from a rescan before its segment and checked against its
neighbours; the superkmers must be exactly the ones SuperKmerEnumerator::next emits for the whole read,
for k <= 32 and for k > 32 (where the state machine depends on history)."""
import os

import numpy as np
import pytest

import golden_util as G

pytestmark = pytest.mark.gpu


def _oracle_superkmers(k, m, b, reads):
    from oracle.oracle import Oracle
    o = Oracle(k, m, b)
    out = []
    for seq in reads:
        e = o.enumerate(seq)
        if len(e["lo"]) == 0:
            continue
        bounds = np.flatnonzero(np.diff(e["skmer"])) + 1
        for grp in np.split(np.arange(len(e["lo"])), bounds):
            out.append(tuple(((int(e["hi"][i]) << 64) | int(e["lo"][i]), int(e["idx"][i])) for i in grp))
    return sorted(out)


def _long_reads(seed, lens):
    rng = np.random.default_rng(seed)
    reads = [bytes(rng.choice(list(b"ACGT"), int(n))) for n in lens]
    # low-complexity stretches inside a long read: ties in every window
    reads.append(reads[0][:3000] + b"ACACACAC" * 400 + reads[0][3000:9000] + b"A" * 2500 + reads[0][9000:12000])
    return reads


@pytest.mark.parametrize("kmb", [(63, 21, 14), (31, 15, 12), (64, 23, 14), (31, 19, 16), (47, 17, 12)])
def test_segmented_enumeration_of_long_reads_is_the_reference_enumeration(kmb):
    from brisk_b200 import Brisk, Parameters, decode_records
    k, m, b = kmb
    reads = _long_reads(k, [40000, 2049, 2048, 5000, 100, 12345, 250, 3071 + k, 70000])
    idx = Brisk(Parameters(k, m, b))
    assert idx.long_read_info()[0] == 2048
    bases, offs = G.flat(reads)
    recs, rw = idx.enumerate_reads(bases, offs)
    got = sorted(tuple(kmers) for _, kmers in decode_records(recs, idx.params))
    assert got == _oracle_superkmers(k, m, b, reads)


@pytest.mark.parametrize("kmb", [(63, 21, 14), (31, 15, 12)])
def test_counting_long_reads_matches_the_oracle_and_short_warm_up_is_caught(kmb, monkeypatch):
    from brisk_b200 import Brisk, Parameters
    from oracle.oracle import Oracle
    k, m, b = kmb
    reads = _long_reads(7 + k, [30000, 9000, 200, 2500])
    bases, offs = G.flat(reads)
    o = Oracle(k, m, b)
    o.count_reads(bases, offs)
    want = G.dump_sha(*o.dump())
    idx = Brisk(Parameters(k, m, b))
    idx.count_reads(bases, offs)
    assert G.dump_sha(*idx.dump()) == want
    # per-occurrence outputs in read order through the segmented path
    found, data = idx.get_reads(bases, offs)
    assert found.all() and len(found) == sum(max(0, len(r) - k + 1) for r in reads)
    # a warm-up of one k-mer cannot bring a segment into the reference's state when k > 32 (the state depends on
    # history): the disagreement between neighbours is detected and the launch repeated until the result is exact
    monkeypatch.setenv("BRISK_SEG_WARM", "1")
    short = Brisk(Parameters(k, m, b))
    short.count_reads(bases, offs)
    assert G.dump_sha(*short.dump()) == want
    if k > 32:
        assert short.long_read_info()[2] >= 1
    # segments for every read longer than 300 bases: many more segments, same answer
    monkeypatch.setenv("BRISK_LONG_READ", "300")
    monkeypatch.delenv("BRISK_SEG_WARM")
    low = Brisk(Parameters(k, m, b))
    assert low.long_read_info()[0] == 300
    low.count_reads(bases, offs)
    assert G.dump_sha(*low.dump()) == want


def test_fasta_contig_goes_through_the_segmented_path():
    """config 1 (data/test.fa: one record of 6,230 bases) and data/debug/test.fa (27,313 bases) are long reads now"""
    for name in ["test_fa_k63", "debug_fa_k31", "debug_fa_k63", "test_fa_k64"]:
        from brisk_b200 import Brisk, Parameters
        c = G.case(name)
        reads = G.reads_of(c)
        assert max(len(r) for r in reads) > 2048
        idx = Brisk(Parameters(c["k"], c["m"], c["b"]))
        idx.count_reads(*G.flat(reads))
        assert G.dump_sha(*idx.dump()) == c["dump_sha256"]
