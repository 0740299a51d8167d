"""Parity at size ("parity at scale is by property, not by oracle"): the reference-generated fixtures of
tests/golden/golden_mid.json (tests/golden/make_golden_mid.py ran the UNMODIFIED reference, one thread) on synthetic
read sets of the shapes of configs 2-5 — config 2 at its full stated size — replayed on the GPU the way config 3
prescribes: reads counted in batches, after batch j the reads of batch j-1 queried; then the sorted dump (sha256),
stats() and a decoy query.  The reads are regenerated on the box with brisk_b200.synth (nothing but hashes and integers
is committed).  Bit-exact."""
import json
import os

import numpy as np
import pytest

import golden_util as G

pytestmark = pytest.mark.gpu

MID = {c["name"]: c for c in json.load(open(os.path.join(G.GOLD, "golden_mid.json")))["cases"]}


def _batches(c):
    from brisk_b200 import synth
    s = c["synth"]
    gen = synth.genome(s["G"], s["seed"])
    n, per, L = s["n_reads"], c["batch_reads"], s["L"]
    for first in range(0, n, per):
        cnt = min(per, n - first)
        yield synth.reads(gen, L, s["seed"], first, cnt).reshape(-1), np.arange(cnt + 1, dtype=np.uint64) * np.uint64(L)


@pytest.mark.parametrize("flags", [0, 8], ids=["compiled_km", "runtime_km"])
@pytest.mark.parametrize("name", sorted(MID))
def test_batched_count_with_interleaved_queries_matches_the_reference(name, flags):
    from brisk_b200 import Brisk, Parameters, synth
    c = MID[name]
    if flags and name == "mid_cfg2_full":
        pytest.skip("the full-size case runs once")
    s = c["synth"]
    idx = Brisk(Parameters(c["k"], c["m"], c["b"]), flags=flags)  # no capacity hint: the store grows under the batches
    prev = None
    for j, (bases, offs) in enumerate(_batches(c)):
        idx.count_reads(bases, offs)
        occ, qocc, hits, total = c["steps"][j]
        assert occ == (len(offs) - 1) * (s["L"] - c["k"] + 1)
        if prev is not None:
            h, t = idx.query_reads(*prev)
            assert (h, t) == (hits, total), (name, j)
            assert qocc == (len(prev[1]) - 1) * (s["L"] - c["k"] + 1)
        prev = (bases, offs)
    st = idx.stats()
    assert st["nb_kmers"] == c["nb_kmers"] and st["nb_buckets"] == c["nb_buckets"]
    lo, hi, ix, cnt = idx.dump()
    assert int(cnt.astype(np.int64).sum()) == c["sum_counts"]
    assert [int(x) for x in np.bincount(cnt, minlength=256)] == c["count_histogram"]
    assert G.dump_sha(lo, hi, ix, cnt) == c["dump_sha256"]
    decoy = synth.genome(s["G"], s["decoy_seed"])
    q = synth.reads(decoy, s["L"], s["decoy_seed"], 0, s["n_decoy"]).reshape(-1)
    qo = np.arange(s["n_decoy"] + 1, dtype=np.uint64) * np.uint64(s["L"])
    h, t = idx.query_reads(q, qo)
    assert [s["n_decoy"] * (s["L"] - c["k"] + 1), h, t] == c["decoy"]


def test_hint_four_times_too_small_with_a_full_batch():
    """a capacity hint far below the real number of keys must not lose inserts: the table grows ahead of the
    batches from the observed new-key rate, parked inserts are replayed."""
    import torch
    from brisk_b200 import Brisk, Parameters, synth
    k, m, b, L, n = 31, 15, 12, 150, 1 << 20
    gen = synth.genome(40_000_000, 31)
    offs = torch.from_numpy(np.arange(n + 1, dtype=np.int64) * L).cuda()
    idx = Brisk(Parameters(k, m, b), capacity_hint=10_000_000)   # about 4x too small for two batches of new keys
    occ = n * (L - k + 1)
    for j in range(2):
        bt = torch.from_numpy(synth.reads(gen, L, 31, j * n, n).reshape(-1)).cuda()
        idx.count_reads(bt, offs)
        hits, total = idx.query_reads(bt, offs)
        assert hits == occ and total >= occ
    lo, hi, ix, cnt = idx.dump()
    assert int(cnt.astype(np.int64).sum()) == 2 * occ
    assert idx.stats()["nb_kmers"] == len(lo) > 30_000_000


def test_device_read_generator_is_bit_identical_to_the_host_generator():
    """reads are generated on the device for the big configurations; same bytes as brisk_b200.synth."""
    import torch
    from brisk_b200 import synth
    for (G, L, seed, first, count) in [(1_000_003, 150, 5, 0, 5000), (777_777, 250, 3, 123_456, 3001), (400, 150, 9, 7, 64)]:
        gen_h = synth.genome(G, seed)
        gen_d = synth.device_genome(G, seed)
        assert np.array_equal(gen_d.cpu().numpy(), gen_h)
        rd = synth.device_reads(gen_d, L, seed, first, count)
        torch.cuda.synchronize()
        assert np.array_equal(rd.cpu().numpy(), synth.reads(gen_h, L, seed, first, count).reshape(-1))


def test_checksum_is_order_independent_and_additive_over_shards():
    """brisk_checksum: the sums of two indexes over disjoint bucket sets add up to the sums of the index over everything
    (what bench.py's parity_check relies on at N > 1), and equal the host-side sums of the dump."""
    from brisk_b200 import Brisk, Parameters, synth
    k, m, b, L = 63, 21, 14, 250
    bases, offs = synth.read_set(300_000, L, 8, 17)
    whole = Brisk(Parameters(k, m, b))
    whole.count_reads(bases, offs)
    cs = whole.checksum()
    lo, hi, ix, cnt = whole.dump()
    assert cs[2] == len(lo) == whole.stats()["nb_kmers"] and cs[1] == int(cnt.astype(np.int64).sum())
    # two virtual ranks through the partition entry points
    src = Brisk(Parameters(k, m, b))
    recs, starts = src.partition_reads(bases, offs, 2)
    parts = []
    for r in range(2):
        shard = Brisk(Parameters(k, m, b))
        shard.apply_records(np.ascontiguousarray(recs[int(starts[r]):int(starts[r + 1])]), "count")
        parts.append(shard.checksum())
    M = (1 << 64) - 1
    assert ((parts[0][0] + parts[1][0]) & M, parts[0][1] + parts[1][1], parts[0][2] + parts[1][2]) == cs
    # a second pass over the same reads changes the counts, hence the mix sum, not the number of keys
    whole.count_reads(bases, offs)
    cs2 = whole.checksum()
    assert cs2[2] == cs[2] and cs2[1] == 2 * cs[1] and cs2[0] != cs[0]
