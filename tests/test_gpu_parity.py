"""Parity of the CUDA path (through the C ABI) with the oracle and with the committed golden fixtures
that the unmodified reference produced (tests/golden/make_golden.py).  Bit-exact: integer / byte work."""
import hashlib

import numpy as np
import pytest

import golden_util as G

pytestmark = pytest.mark.gpu

SMALL = ["test_fa_k63", "debug_fa_k31", "test_fa_k64", "test_fa_k31_m19", "debug_fa_k63", "debug_fa_k32",
         "debug_fa_k33", "debug_fa_k47", "debug_fa_k21_m7", "ties_k31_m15", "ties_k63_m21", "ties_k64_m23",
         "ties_k31_m19", "ties_k21_m9", "ties_k40_m11"]
SYNTH = ["synth_cfg2", "synth_cfg3", "synth_cfg4", "synth_cfg5"]


def make_index(c, **kw):
    from brisk_b200 import Brisk, Parameters
    return Brisk(Parameters(c["k"], c["m"], c["b"]), **kw)


def oracle_superkmers(c, reads):
    """sorted list of superkmers, each a tuple of (kmer_s, idx) in emission order — from the oracle"""
    from oracle.oracle import Oracle
    o = Oracle(c["k"], c["m"], c["b"])
    out = []
    for seq in reads:
        e = o.enumerate(seq)
        if len(e["lo"]) == 0:
            continue
        bounds = np.flatnonzero(np.diff(e["skmer"])) + 1
        for grp in np.split(np.arange(len(e["lo"])), bounds):
            out.append(tuple(((int(e["hi"][i]) << 64) | int(e["lo"][i]), int(e["idx"][i])) for i in grp))
    return sorted(out)


@pytest.mark.parametrize("flags", [0, 2, 8], ids=["rolled", "literal", "rolled_runtime_km"])
@pytest.mark.parametrize("name", SMALL + SYNTH[:1] + SYNTH[1:2])
def test_enumerate_matches_oracle(name, flags):
    from brisk_b200 import decode_records
    c = G.case(name)
    reads = G.reads_of(c)
    idx = make_index(c, flags=flags)
    bases, offs = G.flat(reads)
    recs, rw = idx.enumerate_reads(bases, offs)
    assert len(recs) == c["superkmers_emitted"]
    got = sorted(tuple(kmers) for _, kmers in decode_records(recs, idx.params))
    assert sum(len(s) for s in got) == c["occurrences"]
    assert got == oracle_superkmers(c, reads)
    sizes = np.bincount([len(s) for s in got], minlength=len(c["skmer_size_hist"]))
    assert list(sizes[:len(c["skmer_size_hist"])]) == c["skmer_size_hist"]


@pytest.mark.parametrize("flags", [0, 2, 8], ids=["rolled", "literal", "rolled_runtime_km"])
@pytest.mark.parametrize("name", SMALL + SYNTH)
def test_count_dump_matches_reference_golden(name, flags):
    c = G.case(name)
    idx = make_index(c, flags=flags)
    bases, offs = G.flat(G.reads_of(c))
    idx.count_reads(bases, offs)
    lo, hi, ix, cnt = idx.dump()
    st = idx.stats()
    assert st["nb_kmers"] == c["nb_kmers"] == len(lo)
    assert st["nb_buckets"] == c["nb_buckets"]
    assert int(cnt.astype(np.int64).sum()) == c["sum_counts"]
    assert G.dump_sha(lo, hi, ix, cnt) == c["dump_sha256"]


@pytest.mark.parametrize("name", ["test_fa_k63", "ties_k31_m15"])
def test_dump_equals_golden_arrays(name):
    c = G.case(name)
    z = np.load(G.GOLD + "/" + c["arrays"])
    idx = make_index(c)
    bases, offs = G.flat(G.reads_of(c))
    idx.count_reads(bases, offs)
    lo, hi, ix, cnt = idx.dump()
    order = np.lexsort((ix, lo, hi))
    assert np.array_equal(lo[order], z["dump_lo"]) and np.array_equal(hi[order], z["dump_hi"])
    assert np.array_equal(ix[order], z["dump_idx"]) and np.array_equal(cnt[order], z["dump_count"])


@pytest.mark.parametrize("name", SYNTH)
def test_batched_get_matches_reference_and_oracle(name):
    from oracle.oracle import Oracle
    c = G.case(name)
    reads = G.reads_of(c)
    idx = make_index(c)
    o = Oracle(c["k"], c["m"], c["b"])
    # inserts in two batches with a query batch in between (config 3: gets interleaved with inserts)
    half = len(reads) // 2
    b1, o1 = G.flat(reads[:half])
    b2, o2 = G.flat(reads[half:])
    idx.count_reads(b1, o1)
    o.count_reads(b1, o1)
    qb, qo = G.flat(reads[half:half + 40])
    found, data = idx.get_reads(qb, qo)
    pos = 0
    for seq in reads[half:half + 40]:
        f, d, _ = o.query_read(seq)
        assert np.array_equal(found[pos:pos + len(f)], f) and np.array_equal(data[pos:pos + len(f)], d)
        pos += len(f)
    idx.count_reads(b2, o2)
    queries = G.query_reads_of(c)
    qb, qo = G.flat(queries)
    found, data = idx.get_reads(qb, qo)
    pos = 0
    for seq, (nk, hits, total) in zip(queries, c["query"]):
        f = found[pos:pos + nk]
        d = data[pos:pos + nk]
        assert int(f.sum()) == hits and int(d.astype(np.int64).sum()) == total
        pos += nk
    assert pos == len(found)
    hits, total = idx.query_reads(qb, qo)
    assert hits == sum(q[1] for q in c["query"]) and total == sum(q[2] for q in c["query"])


def test_insert_reads_slots_and_newly():
    from oracle.oracle import Oracle
    c = G.case("synth_cfg3")
    reads = G.reads_of(c)[:300]
    idx = make_index(c)
    o = Oracle(c["k"], c["m"], c["b"])
    bases, offs = G.flat(reads)
    slots, newly = idx.insert_reads(bases, offs)
    assert int(newly.sum()) == idx.stats()["nb_kmers"]
    # every occurrence of one (kmer, idx) key gets the same slot; exactly one of them is "newly inserted"
    keys = []
    for seq in reads:
        e = o.enumerate(seq)
        order = np.argsort(e["ridx"], kind="stable")
        keys += [((int(e["hi"][i]) << 64) | int(e["lo"][i]), int(e["idx"][i])) for i in order]
    assert len(keys) == len(slots)
    by_key = {}
    for kk, s, nw in zip(keys, slots, newly):
        by_key.setdefault(kk, []).append((int(s), int(nw)))
    assert len(by_key) == int(newly.sum())
    seen_slots = set()
    for kk, lst in by_key.items():
        assert len({s for s, _ in lst}) == 1 and sum(nw for _, nw in lst) == 1
        seen_slots.add(lst[0][0])
    assert len(seen_slots) == len(by_key)
    # a second insert of the same reads finds everything
    slots2, newly2 = idx.insert_reads(bases, offs)
    assert int(newly2.sum()) == 0 and np.array_equal(slots, slots2)
    # plain inserts leave DATA untouched (zero), then counting adds up
    found, data = idx.get_reads(bases, offs)
    assert found.all() and not data.any()


@pytest.mark.parametrize("name", ["ties_k63_m21", "ties_k31_m15", "test_fa_k64"])
def test_explicit_kmers_insert_get(name):
    from brisk_b200 import Brisk
    from oracle.oracle import Oracle
    c = G.case(name)
    z = np.load(G.GOLD + "/" + c["arrays"])
    idx = make_index(c)
    o = Oracle(c["k"], c["m"], c["b"])
    lo, hi, ix = z["enum_lo"], z["enum_hi"], z["enum_idx"]
    km = Brisk.pack_kmers(lo, hi, ix)
    slots, newly = idx.insert_kmers(km, count=True)
    o.insert_kmers(lo, hi, ix, count=True)
    assert int(newly.sum()) == o.size() == idx.stats()["nb_kmers"]
    f, d = o.get_kmers(lo, hi, ix)
    s2, d2 = idx.get_kmers(km)
    assert np.array_equal(s2, slots) and np.array_equal(d2, d) and f.all()
    # absent keys: same k-mers under a different minimizer_idx
    other = Brisk.pack_kmers(lo, hi, (ix.astype(np.int64) + 1) % (c["k"] - c["m"] + 1))
    s3, _ = idx.get_kmers(other)
    fo, _ = o.get_kmers(lo, hi, ((ix.astype(np.int64) + 1) % (c["k"] - c["m"] + 1)).astype(np.uint8))
    assert np.array_equal(s3 != np.uint64(0xFFFFFFFFFFFFFFFF), fo.astype(bool))
    glo, ghi, gix, gcnt = idx.dump()
    assert G.dump_sha(glo, ghi, gix, gcnt) == c["dump_sha256"]


def test_store_grows_and_keeps_counts():
    c = G.case("synth_cfg2")
    reads = G.reads_of(c)
    idx = make_index(c, capacity_hint=1000)  # far too small: forces parking + rehash
    cap0 = idx.capacity()
    for i in range(0, len(reads), 500):
        b, o = G.flat(reads[i:i + 500])
        idx.count_reads(b, o)
    assert idx.capacity() > cap0
    lo, hi, ix, cnt = idx.dump()
    assert G.dump_sha(lo, hi, ix, cnt) == c["dump_sha256"]


def test_wrapping_uint8_counts_and_sentinel_key():
    # poly-G at k = 32 is the key whose bits equal the empty-slot pattern; 300 copies wrap the uint8 count
    from brisk_b200 import Brisk, Parameters
    from oracle.oracle import Oracle
    reads = [b"G" * 100, b"C" * 100, b"G" * 331, b"A" * 400, b"ACGT" * 90]
    for (k, m, b) in [(32, 15, 10), (64, 23, 14), (31, 15, 12)]:
        idx = Brisk(Parameters(k, m, b))
        o = Oracle(k, m, b)
        bases, offs = G.flat(reads)
        for _ in range(2):
            idx.count_reads(bases, offs)
            o.count_reads(bases, offs)
        lo, hi, ix, cnt = idx.dump()
        assert G.dump_sha(lo, hi, ix, cnt) == G.dump_sha(*o.dump())
        f, d = idx.get_reads(bases, offs)
        pos = 0
        for seq in reads:
            fo, do, _ = o.query_read(seq)
            assert np.array_equal(f[pos:pos + len(fo)], fo) and np.array_equal(d[pos:pos + len(fo)], do)
            pos += len(fo)
        # the all-ones k-mer itself (poly-G), 300 times under two different minimizer_idx
        ones = np.uint64(0xFFFFFFFFFFFFFFFF)
        klo = np.full(600, ones if k >= 32 else np.uint64((1 << (2 * k)) - 1))
        khi = np.full(600, ones if k == 64 else np.uint64(0))
        kix = np.repeat(np.array([3, 7], np.uint8), 300)
        s1, nw = idx.insert_kmers(Brisk.pack_kmers(klo, khi, kix), count=True)
        onw = o.insert_kmers(klo, khi, kix, count=True)
        assert int(nw.sum()) == int(onw.sum()) and len(set(s1.tolist())) == 2
        s2, d2 = idx.get_kmers(Brisk.pack_kmers(klo[[0, 599]], khi[[0, 599]], kix[[0, 599]]))
        _, od = o.get_kmers(klo[[0, 599]], khi[[0, 599]], kix[[0, 599]])
        assert list(d2) == list(od) and list(s2) == [s1[0], s1[599]]
        assert G.dump_sha(*idx.dump()) == G.dump_sha(*o.dump())


def test_ragged_and_empty_inputs():
    from brisk_b200 import Brisk, Parameters
    from oracle.oracle import Oracle
    rng = np.random.default_rng(11)
    reads = [b"", b"ACGT", bytes(rng.choice(list(b"ACGT"), 30)), bytes(rng.choice(list(b"ACGT"), 31)),
             bytes(rng.choice(list(b"ACGTacgtNnRY"), 500)), b"", bytes(rng.choice(list(b"ACGT"), 5000)),
             bytes(rng.choice(list(b"ACGT"), 33))] + [bytes(rng.choice(list(b"ACGT"), int(n))) for n in rng.integers(1, 400, 300)]
    idx = Brisk(Parameters(31, 15, 12))
    o = Oracle(31, 15, 12)
    bases, offs = G.flat(reads)
    idx.count_reads(bases, offs)
    o.count_reads(bases, offs)
    assert G.dump_sha(*idx.dump()) == G.dump_sha(*o.dump())
    # nothing at all
    empty = Brisk(Parameters(31, 15, 12))
    empty.count_reads(np.zeros(0, np.uint8), np.zeros(1, np.uint64))
    assert empty.stats()["nb_kmers"] == 0 and len(empty.dump()[0]) == 0


def test_device_resident_inputs():
    import torch
    c = G.case("synth_cfg5")
    idx = make_index(c)
    bases, offs = G.flat(G.reads_of(c))
    tb = torch.from_numpy(bases.copy()).cuda()
    to = torch.from_numpy(offs.astype(np.int64)).cuda()
    torch.cuda.synchronize()
    idx.count_reads(tb, to)
    assert G.dump_sha(*idx.dump()) == c["dump_sha256"]
    t = idx.last_timing()
    assert t["launches"] >= 3 and t["total_ms"] > 0


@pytest.mark.parametrize("name", ["test_fa_k63", "debug_fa_k31", "ties_k63_m21", "synth_cfg3"])
def test_tagged_enumeration_is_in_reference_order(name):
    """brisk_enumerate_reads_tagged sorted by its tag = the exact emission order of SuperKmerEnumerator::next."""
    from brisk_b200 import decode_records
    from oracle.oracle import Oracle
    c = G.case(name)
    reads = G.reads_of(c)[:300]
    idx = make_index(c)
    bases, offs = G.flat(reads)
    recs, occ = idx.enumerate_reads_tagged(bases, offs)
    got = [tuple(kmers) for _, kmers in decode_records(recs, idx.params)]
    o = Oracle(c["k"], c["m"], c["b"])
    want = []
    for seq in reads:
        e = o.enumerate(seq)
        if len(e["lo"]) == 0:
            continue
        bounds = np.flatnonzero(np.diff(e["skmer"])) + 1
        for grp in np.split(np.arange(len(e["lo"])), bounds):
            want.append(tuple(((int(e["hi"][i]) << 64) | int(e["lo"][i]), int(e["idx"][i])) for i in grp))
    assert got == want


@pytest.mark.parametrize("kmb", [(31, 15, 12), (63, 21, 14)])
def test_chunked_host_input_equals_device_input(kmb):
    """Host buffers larger than 16 MB arrive in chunks that are pipelined with the kernels; the result must be the one
    of the same reads handed over as one device-resident batch (and, on a sample, the oracle's)."""
    import torch
    from brisk_b200 import Brisk, Parameters, synth
    from oracle.oracle import Oracle
    k, m, b = kmb
    L, n = 150, 240000  # 36 MB of bases -> 5 chunks
    gen = synth.genome(2_000_000, 11)
    arr = synth.reads(gen, L, 11, 0, n).reshape(-1)
    offs = np.arange(n + 1, dtype=np.uint64) * np.uint64(L)
    host = Brisk(Parameters(k, m, b))
    host.count_reads(arr, offs)
    dev = Brisk(Parameters(k, m, b))
    dev.count_reads(torch.from_numpy(arr).cuda(), torch.from_numpy(offs.astype(np.int64)).cuda())
    assert host.stats()["nb_kmers"] == dev.stats()["nb_kmers"]
    assert G.dump_sha(*host.dump()) == G.dump_sha(*dev.dump())
    assert host.query_reads(arr, offs) == dev.query_reads(arr, offs)
    assert host.query_reads(arr, offs)[0] == n * (L - k + 1)
    # per-occurrence outputs with chunked input (only the copy / enumerate stages are pipelined there)
    f1, d1 = host.get_reads(arr, offs)
    f2, d2 = dev.get_reads(torch.from_numpy(arr).cuda(), torch.from_numpy(offs.astype(np.int64)).cuda())
    assert np.array_equal(f1, f2) and np.array_equal(d1, d2) and f1.all()
    # oracle on the first reads: counts of their k-mers in the whole set are >= their counts in the sample
    o = Oracle(k, m, b)
    o.count_reads(arr[:2000 * L], offs[:2001])
    lo, hi, ix, cnt = o.dump()
    slots, data = host.get_kmers(Brisk.pack_kmers(lo, hi, ix))
    assert (slots != np.uint64(0xFFFFFFFFFFFFFFFF)).all() and (data >= cnt).all()


@pytest.mark.parametrize("kmb,L", [((63, 21, 14), 250), ((31, 19, 16), 150), ((64, 23, 14), 250)])
def test_full_batch_size_properties(kmb, L):
    """BASELINE-size batches (2^20 reads per call) through size-independent properties: every inserted occurrence is
    found again, the counts add up to the occurrences, re-inserting creates nothing, a decoy genome hits (almost)
    nothing, and the store agrees with itself after growing."""
    import torch
    from brisk_b200 import Brisk, Parameters, synth
    k, m, b = kmb
    n = 1 << 20
    G_ = 8_000_000  # 2 batches = 65x coverage: counts stay below the uint8 wrap
    gen = synth.genome(G_, 21)
    occ = n * (L - k + 1)
    offs = torch.from_numpy(np.arange(n + 1, dtype=np.int64) * L).cuda()
    idx = Brisk(Parameters(k, m, b))  # no capacity hint: the store has to grow on its own
    batches = [torch.from_numpy(synth.reads(gen, L, 21, j * n, n).reshape(-1)).cuda() for j in range(2)]
    for bt in batches:
        idx.count_reads(bt, offs)
    st = idx.stats()
    lo, hi, ix, cnt = idx.dump()
    assert len(lo) == st["nb_kmers"]
    assert int(cnt.astype(np.int64).sum()) == 2 * occ  # coverage stays far below the uint8 wrap
    assert len(set(zip(lo[:200000].tolist(), hi[:200000].tolist(), ix[:200000].tolist()))) == 200000
    for bt in batches:
        hits, total = idx.query_reads(bt, offs)
        assert hits == occ and total >= occ
    # idempotence of insert: nothing new, counts untouched
    slots, newly = idx.insert_reads(batches[0][:4000 * L], offs[:4001])
    assert not newly.any() and (slots != np.uint64(0xFFFFFFFFFFFFFFFF)).all()
    assert idx.stats()["nb_kmers"] == st["nb_kmers"]
    # a decoy genome shares (almost) no k-mer
    decoy = synth.genome(2_000_000, 99)
    q = torch.from_numpy(synth.reads(decoy, L, 99, 0, 100000).reshape(-1)).cuda()
    hits, _ = idx.query_reads(q, offs[:100001])
    assert hits < 10
    # growing the store keeps every key and count
    idx.reserve(4 * st["nb_kmers"])
    assert G.dump_sha(*idx.dump()) == G.dump_sha(lo, hi, ix, cnt)
    assert idx.query_reads(batches[1], offs)[0] == occ


@pytest.mark.parametrize("name", ["test_fa_k63", "debug_fa_k31", "synth_cfg3", "synth_cfg5", "test_fa_k64"])
def test_ordered_dump_walks_buckets_in_id_order(name):
    """brisk_dump_ordered = the set of brisk_dump, bucket ids non-decreasing (DenseMenuYo::next's outer order), (minimizer_idx, kmer_s)
    ascending inside a bucket; bucket ids recomputed on the host from the oracle's bfc_hash."""
    from oracle.oracle import Oracle
    c = G.case(name)
    idx = make_index(c)
    bases, offs = G.flat(G.reads_of(c))
    idx.count_reads(bases, offs)
    lo, hi, ix, cnt = idx.dump(ordered=True)
    assert G.dump_sha(lo, hi, ix, cnt) == c["dump_sha256"]
    o = Oracle(c["k"], c["m"], c["b"])
    k, m, b = c["k"], c["m"], c["b"]
    s, m_mask = (m - b + 1) // 2, (1 << (2 * m)) - 1
    step = max(1, len(lo) // 3000)  # the host loop is slow: check a regular sample of consecutive pairs
    prev = None
    for i in range(0, len(lo)):
        if i % step > 1:
            prev = None
            continue
        val = (int(hi[i]) << 64) | int(lo[i])
        w = (val >> (2 * int(ix[i]))) & m_mask
        bucket = ((o.bfc_hash(w) & m_mask) >> (2 * s)) & ((1 << (2 * b)) - 1)
        key = (bucket, int(ix[i]), val)
        if prev is not None:
            assert prev < key
        prev = key


@pytest.mark.gpu
def test_count_returns_early_and_every_reader_sees_the_finished_store():
    """brisk_count_reads returns once its input is consumed (include/brisk_b200.h, BRISK_F_SYNC_INSERT); whatever is
    called next on the handle — query, stats, dump, capacity, another count over a reused host buffer — must see the
    completed insertion.  Same batches through a BRISK_F_SYNC_INSERT handle give the same dump."""
    from brisk_b200 import Brisk, Parameters, synth
    from brisk_b200.brisk import F_SYNC_INSERT
    k, m, b, L = 31, 15, 12, 150
    gen = synth.genome(200_000, 11)
    n = 20_000
    host = np.empty(n * L, np.uint8)
    offs = np.arange(n + 1, dtype=np.uint64) * np.uint64(L)
    dumps = []
    for flags in (0, F_SYNC_INSERT):
        idx = Brisk(Parameters(k, m, b), flags=flags)  # no capacity hint: the store grows while batches are in flight
        for j in range(4):
            host[:] = synth.reads(gen, L, 11, j * n, n).reshape(-1)  # the buffer of the previous call is overwritten
            idx.count_reads(host, offs)
            if j == 1:
                hits, total = idx.query_reads(host, offs)
                assert hits == n * (L - k + 1)
            if j == 2:
                assert idx.stats()["nb_kmers"] > 0
        lo, hi, ix, cnt = idx.dump()
        order = np.lexsort((ix, hi, lo))
        dumps.append((lo[order], hi[order], ix[order], cnt[order]))
        idx.close()
    for a, c in zip(dumps[0], dumps[1]):
        assert np.array_equal(a, c)
    assert int(dumps[0][3].astype(np.int64).sum()) == 4 * n * (L - k + 1)  # 60x coverage: far from the uint8 wrap
