"""The C++ host mirror of the reference API (include/brisk_b200/Brisk.hpp) driven like apps/counter.cpp drives the
reference (tests/cpp/counter_mirror.cpp): legacy per-superkmer calls (SuperKmerEnumerator::next, protect_data,
insert_superkmer, pointer updates, get_superkmer, next/get) and the batched calls must both reproduce the golden
dump that the unmodified reference produced."""
import os
import subprocess

import numpy as np
import pytest

import golden_util as G

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def binary(tmp_path_factory):
    out = str(tmp_path_factory.mktemp("cpp") / "counter_mirror")
    lib = os.path.join(ROOT, "brisk_b200")
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-Wall", "-I" + os.path.join(ROOT, "include"),
                           os.path.join(ROOT, "tests", "cpp", "counter_mirror.cpp"), "-L" + lib, "-lbrisk_b200",
                           "-Wl,-rpath," + lib, "-o", out])
    return out


def parse(out, tag):
    rows = [ln.split()[1:] for ln in out.splitlines() if ln.startswith(tag + " ")]
    a = np.array(rows, dtype=np.uint64).reshape(-1, 4)
    return a[:, 0], a[:, 1], a[:, 2].astype(np.uint8), a[:, 3].astype(np.uint8)


@pytest.mark.parametrize("name", ["ties_k31_m15", "ties_k63_m21", "ties_k64_m23", "debug_fa_k31", "test_fa_k63"])
def test_cpp_mirror_reproduces_reference_dump(binary, name, tmp_path):
    c = G.case(name)
    reads = G.reads_of(c)
    path = tmp_path / "reads.txt"
    path.write_bytes(b"\n".join(reads) + b"\n")
    kff = tmp_path / "out.kff"
    res = subprocess.run([binary, str(path), str(c["k"]), str(c["m"]), str(c["b"]), str(kff)], capture_output=True, text=True, timeout=600)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-2000:]
    assert "EXCEPTION" not in res.stdout and "ERROR" not in res.stdout
    for tag in ("LEGACY", "BATCHED"):
        lo, hi, ix, cnt = parse(res.stdout, tag)
        assert len(lo) == c["nb_kmers"], tag
        assert G.dump_sha(lo, hi, ix, cnt) == c["dump_sha256"], tag
        assert "STATS %s kmers %d buckets %d dumped %d" % (tag, c["nb_kmers"], c["nb_buckets"], c["nb_kmers"]) in res.stdout
        assert "QUERY %s hits %d total " % (tag, c["occurrences"]) in res.stdout
    ql = [ln for ln in res.stdout.splitlines() if ln.startswith("QUERY ")]
    assert ql[0].split()[2:] == ql[1].split()[2:]
    assert "ABSENT null" in res.stdout
    assert "WIDE kmers %d sum %d occurrences %d bad 0" % (c["nb_kmers"], c["occurrences"], c["occurrences"]) in res.stdout
    # BriskWriter through the mirror: the file holds the golden k-mers (tests/kff_reader.py); reallocate went to (m + 2, b + 2)
    from kff_reader import kmers_of, read_kff
    km = kmers_of(read_kff(str(kff)))
    lo = np.array([x[0] & 0xFFFFFFFFFFFFFFFF for x in km], np.uint64)
    hi = np.array([x[0] >> 64 for x in km], np.uint64)
    assert G.dump_sha(lo, hi, np.array([x[1] for x in km], np.uint8), np.array([x[2][0] for x in km], np.uint8)) == c["dump_sha256"]
    assert "REALLOC m %d b %d kmers " % (c["m"] + 2, c["b"] + 2) in res.stdout
