"""brisk_write_kff (the reference: BriskWriter::write, brisk/writer.hpp:75-179).  The reference's
own writer crashes in this container, so the file is checked two ways: a reader written from the format
(tests/kff_reader.py) must recover exactly the golden (kmer_s, minimizer_idx, count) multiset that the unmodified
reference's index holds (tests/golden/golden.json), and — where oracle/_ref/kff_dump exists — the reference's OWN
KFF library (Kff_reader of brisk/lib/kff) must parse the file and return the same k-mers and counts."""
import os
import subprocess

import numpy as np
import pytest

import golden_util as G
from kff_reader import kmers_of, read_kff

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KFF_DUMP = os.path.join(ROOT, "oracle", "_ref", "kff_dump")


def _index(c):
    from brisk_b200 import Brisk, Parameters
    idx = Brisk(Parameters(c["k"], c["m"], c["b"]))
    bases, offs = G.flat(G.reads_of(c))
    idx.count_reads(bases, offs)
    return idx


@pytest.mark.parametrize("name", ["test_fa_k63", "synth_cfg5", "synth_cfg2", "synth_cfg3", "test_fa_k64", "ties_k31_m15", "debug_fa_k32"])
def test_kff_file_holds_the_golden_kmers(name, tmp_path):
    from oracle.oracle import Oracle
    c = G.case(name)
    k, m, b = c["k"], c["m"], c["b"]
    idx = _index(c)
    path = str(tmp_path / (name + ".kff"))
    n_blocks, n_sections = idx.write_kff(path)
    f = read_kff(path)
    # header and value sections as BriskWriter writes them (writer.hpp:23-31, :78-91)
    assert f["version"] == (1, 0) and f["encoding"] == (0 << 6 | 1 << 4 | 3 << 2 | 2)
    assert f["metadata"] == b"File generated with Brisk v1. See https://github.com/Malfoy/Brisk"
    assert f["value_sections"][0][1] == {"k": k, "data_size": 1, "max": 1}
    assert f["value_sections"][1][1] == {"k": k, "m": m, "data_size": 1, "max": 2 * (k - m)}
    assert f["value_sections"][-1][1]["first_index"] == f["index"]["pos"] and f["value_sections"][-1][1]["footer_size"] == 49
    # the index lists every section before it, positions relative to its own end
    listed = [(f["index"]["end"] + rel, t) for (t, rel) in f["index"]["entries"]]
    assert listed == [s for s in f["sections"] if s[0] < f["index"]["pos"]]
    assert len(f["minimizer_sections"]) == n_sections and sum(len(s["blocks"]) for s in f["minimizer_sections"]) == n_blocks
    # the k-mers: exactly the golden multiset, minimizer_idx included
    km = kmers_of(f)
    assert len(km) == c["nb_kmers"]
    lo = np.array([x[0] & 0xFFFFFFFFFFFFFFFF for x in km], np.uint64)
    hi = np.array([x[0] >> 64 for x in km], np.uint64)
    ix = np.array([x[1] for x in km], np.uint8)
    cnt = np.array([x[2][0] for x in km], np.uint8)
    assert G.dump_sha(lo, hi, ix, cnt) == c["dump_sha256"]
    # superkmers were re-assembled: fewer blocks than k-mers wherever reads overlap, never more than k - m + 1 k-mers per block
    sizes = [blk[0] for s in f["minimizer_sections"] for blk in s["blocks"]]
    assert max(sizes) <= k - m + 1
    if name.startswith("synth") or name.startswith("test_fa"):
        assert len(sizes) * 3 < len(km)
    # sections walk the buckets in increasing id (DenseMenuYo order), one section per minimizer
    o = Oracle(k, m, b)
    s_, m_mask = (m - b + 1) // 2, (1 << (2 * m)) - 1
    buckets = [((o.bfc_hash(s["minimizer"]) & m_mask) >> (2 * s_)) & ((1 << (2 * b)) - 1) for s in f["minimizer_sections"]]
    assert buckets == sorted(buckets)
    assert len({s["minimizer"] for s in f["minimizer_sections"]}) == len(f["minimizer_sections"])
    # the minimizer sits where the block says: it is the m-mer at minimizer_idx of every k-mer of the block
    for (kmer, idx_, _), sec_min in zip(km[:2000], [s["minimizer"] for s in f["minimizer_sections"] for blk in s["blocks"] for _ in range(blk[0])][:2000]):
        assert (kmer >> (2 * idx_)) & m_mask == sec_min
    if os.path.exists(KFF_DUMP):
        out = subprocess.run([KFF_DUMP, path], capture_output=True, text=True, timeout=300)
        assert out.returncode == 0, out.stderr[-500:]
        theirs = sorted((int(a, 16) << 64 | int(b_, 16), int(d, 16)) for a, b_, d in (ln.split() for ln in out.stdout.splitlines()))
        ours = sorted((x[0], x[2][0]) for x in km)
        assert theirs == ours


def test_kff_all_ones_kmer_and_wider_data(tmp_path):
    """poly-G at k = 32 is the key that lives outside the table; DATA of 4 bytes goes out as stored"""
    from brisk_b200 import Brisk, Parameters
    reads = [b"G" * 100, b"ACGT" * 30, b"C" * 70]
    idx = Brisk(Parameters(32, 15, 10))
    bases, offs = G.flat(reads)
    idx.count_reads(bases, offs)
    idx.count_reads(bases, offs)
    path = str(tmp_path / "polyg.kff")
    idx.write_kff(path)
    km = kmers_of(read_kff(path))
    lo, hi, ix, cnt = idx.dump()
    assert sorted((int(a), int(i), int(c)) for a, i, c in zip(lo, ix, cnt)) == sorted((x[0], x[1], x[2][0]) for x in km)
    assert ((1 << 64) - 1) in {x[0] for x in km}
