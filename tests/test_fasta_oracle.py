"""Pins the FASTA-ingest restatement (oracle/fasta_oracle.py: getLineFasta + clean_dna of apps/counter.cpp) to what the
unmodified reference app reported for the crafted files of tests/golden/fasta/ (tests/golden/make_fasta_golden.py)."""
import json
import os
import re

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = json.load(open(os.path.join(HERE, "golden", "fasta_golden.json")))["cases"]


def text_of(name):
    return open(os.path.join(HERE, "golden", "fasta", name), "rb").read()


@pytest.mark.parametrize("c", GOLD, ids=lambda c: "%s-k%d" % (c["file"], c["k"]))
def test_oracle_reads_give_the_reference_apps_counts(c):
    from oracle.fasta_oracle import fasta_reads
    from oracle.oracle import Oracle
    o = Oracle(c["k"], c["m"], c["b"])
    for r in fasta_reads(text_of(c["file"])):
        o.count_read(r)
    st = o.stats()
    assert st["nb_kmers"] == c["nb_kmers"] and st["nb_buckets"] == c["nb_buckets"]


def spec_reads(data):
    """the closed form of the reader loop: maximal ACGT runs per record, newlines transparent"""
    out = []
    lines = data.split(b"\n")
    recs, cur = [], None
    for i, ln in enumerate(lines):
        if i == 0 or ln.startswith(b">"):
            if cur is not None:
                recs.append(cur)
            cur = b""
        else:
            cur += ln
    if cur is not None:
        recs.append(cur)
    for r in recs:
        out += [m.group(0).upper() for m in re.finditer(rb"[ACGTacgt]+", r)]
    return out


def test_oracle_equals_closed_form_on_fuzzed_texts():
    from oracle.fasta_oracle import fasta_reads
    rng = np.random.default_rng(3)
    alphabet = list(b"ACGTacgtNn>\n\n\r -X")
    for _ in range(300):
        n = int(rng.integers(0, 200))
        data = bytes(rng.choice(alphabet, n).astype(np.uint8))
        assert fasta_reads(data) == spec_reads(data), data
