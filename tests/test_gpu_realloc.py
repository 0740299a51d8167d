"""brisk_reparameterise = Brisk::reallocate (brisk/Brisk.hpp:202-224), the (m + 2, b + 2) rebuild: SURVEY.md §8(f) row 4.
Fixtures: tests/golden/golden_realloc.json, from the UNMODIFIED reference (ref_dump realloc).  The key set and the bucket
count must be the reference's; DATA too wherever the result does not depend on the reference's enumeration order (old
entries that only differed in their minimizer_idx merge, and the reference keeps the one it walks last — here the one with
the largest old minimizer_idx, for which the fixture holds the expected dump, computed from the reference's own dumps)."""
import hashlib
import json
import os

import numpy as np
import pytest

import golden_util as G

pytestmark = pytest.mark.gpu
CASES = json.load(open(os.path.join(G.GOLD, "golden_realloc.json")))["cases"]


@pytest.mark.parametrize("c", CASES, ids=[c["inputs"] for c in CASES])
def test_reallocate_matches_the_reference(c):
    from brisk_b200 import Brisk, Parameters
    reads = G.reads_of(G.case(c["inputs"]))
    idx = Brisk(Parameters(c["k"], c["m"], c["b"]))
    bases, offs = G.flat(reads)
    idx.count_reads(bases, offs)
    assert idx.stats()["nb_kmers"] == c["kmers_before"]
    idx.reallocate()
    assert (idx.params.m, idx.params.b) == (c["new_m"], c["new_b"])
    st = idx.stats()
    assert st["nb_kmers"] == c["nb_kmers"] and st["nb_buckets"] == c["nb_buckets"]
    lo, hi, ix, cnt = idx.dump()
    assert G.dump_sha(lo, hi, ix, np.zeros_like(cnt)) == c["keys_sha256"]
    assert G.dump_sha(lo, hi, ix, cnt) == c["dump_sha256_largest_idx_rule"]
    if c["rules_agree"]:
        assert G.dump_sha(lo, hi, ix, cnt) == c["dump_sha256"]
    # the rebuilt index keeps working under its new parameters: every k-mer of the reads is found under the key the
    # enumerator now produces or not at all (the reference's reallocate re-keys by kmer_s alone), new reads can be counted
    h1, _ = idx.query_reads(bases, offs)
    idx.count_reads(bases, offs)
    h2, _ = idx.query_reads(bases, offs)
    assert h2 == sum(max(0, len(r) - c["k"] + 1) for r in reads) and h1 <= h2
