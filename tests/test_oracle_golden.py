"""Pins oracle/brisk_oracle.c against the fixtures produced by the unmodified reference
(tests/golden/make_golden.py -> oracle/_ref/ref_dump).  CPU only."""
import hashlib

import numpy as np
import pytest

import golden_util as G
from oracle.oracle import Oracle


@pytest.mark.parametrize("name", G.case_names())
def test_enumerator_stream(name):
    c = G.case(name)
    o = Oracle(c["k"], c["m"], c["b"])
    h = hashlib.sha256()
    n_occ = n_sk = 0
    sizes = {}
    for rid, seq in enumerate(G.reads_of(c)):
        e = o.enumerate(seq)
        n = len(e["lo"])
        if n == 0:
            continue
        rec = np.zeros(n, G.ENUM_DT)
        rec["lo"], rec["hi"], rec["mini"], rec["idx"] = e["lo"], e["hi"], e["mini"], e["idx"]
        rec["read"], rec["skmer"] = rid, e["skmer"]
        h.update(rec.tobytes())
        n_occ += n
        n_sk += int(e["skmer"].max()) + 1
        for s in np.bincount(e["skmer"]):
            sizes[int(s)] = sizes.get(int(s), 0) + 1
    assert n_occ == c["occurrences"]
    assert n_sk == c["superkmers_emitted"]
    assert h.hexdigest() == c["enum_sha256"]
    hist = c["skmer_size_hist"]
    assert all(sizes.get(i, 0) == hist[i] for i in range(len(hist)))


@pytest.mark.parametrize("name", G.case_names())
def test_order_hash_stream(name):
    c = G.case(name)
    o = Oracle(c["k"], c["m"], c["b"])
    h = hashlib.sha256()
    for seq in G.reads_of(c):
        hh, _ = o.mmer_hashes(seq)
        h.update(hh.tobytes())
    assert h.hexdigest() == c["hash_sha256"]


@pytest.mark.parametrize("name", G.case_names())
def test_index_dump(name):
    c = G.case(name)
    o = Oracle(c["k"], c["m"], c["b"])
    bases, offs = G.flat(G.reads_of(c))
    o.count_reads(bases, offs)
    lo, hi, idx, cnt = o.dump()
    st = o.stats()
    assert st["nb_kmers"] == c["nb_kmers"]
    assert st["nb_buckets"] == c["nb_buckets"]
    assert int(cnt.astype(np.int64).sum()) == c["sum_counts"]
    assert G.dump_sha(lo, hi, idx, cnt) == c["dump_sha256"]
    if "query" in c:
        for seq, (nk, hits, total) in zip(G.query_reads_of(c), c["query"]):
            found, data, tot = o.query_read(seq)
            assert len(found) == nk and int(found.sum()) == hits and tot == total


def test_known_answers_from_survey():
    # data/test.fa k63 m21 b14 -> 6,105 distinct k-mers, 239 superkmers, 237 buckets;
    # data/debug/test.fa k31 m15 b12 -> 27,283 / 2,751 / 2,751; k64 (mask fix) -> 6,103 distinct.
    assert (G.case("test_fa_k63")["nb_kmers"], G.case("test_fa_k63")["nb_buckets"]) == (6105, 237)
    assert (G.case("debug_fa_k31")["nb_kmers"], G.case("debug_fa_k31")["nb_skmers"]) == (27283, 2751)
    assert G.case("test_fa_k64")["nb_kmers"] == 6103


def test_hash_inverse_and_arrays():
    o = Oracle(63, 21, 14)
    rng = np.random.default_rng(0)
    for x in rng.integers(0, 1 << 42, 200):
        hsh = o.bfc_hash(int(x))
        assert o.bfc_hash_inv(hsh & ((1 << 42) - 1)) == int(x)
        assert (hsh >> 62) == o.mem_double(int(x))
    z = np.load(G.GOLD + "/test_fa_k63.npz")
    e = o.enumerate(G.reads_of(G.case("test_fa_k63"))[0])
    first = z["enum_read"] == 0
    assert np.array_equal(e["lo"], z["enum_lo"][first]) and np.array_equal(e["idx"], z["enum_idx"][first])
