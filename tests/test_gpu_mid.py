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
