#!/usr/bin/env python
"""Generates tests/golden/*.json|npz by running the UNMODIFIED reference
(oracle/_ref/ref_dump, built from /root/reference by `make -C oracle ref`).

Run in the build container only (it needs /root/reference and oracle/_ref):
    python tests/golden/make_golden.py
The GPU box never runs this; it only reads the committed fixtures.

For every case the fixture records
  * sha256 of the enumerator stream (every emitted k-mer: kmer_s, minimizer value,
    read id, superkmer id, minimizer_idx — ref_dump.cpp:EnumRec, 33 B each),
  * sha256 of the per-m-mer order hash stream (bfc_hash_64 of the canonical m-mer),
  * sha256 of the sorted index dump ((kmer_s, minimizer_idx, count), 18 B each),
  * the reference's stats() output,
and, for the small cases, the arrays themselves (npz).
"""
import hashlib
import json
import os
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from brisk_b200 import synth  # noqa: E402

REF = "/root/reference"
ENUM_DT = np.dtype([("lo", "<u8"), ("hi", "<u8"), ("mini", "<u8"), ("read", "<u4"), ("skmer", "<u4"), ("idx", "u1")])
DUMP_DT = np.dtype([("lo", "<u8"), ("hi", "<u8"), ("idx", "u1"), ("count", "u1")])


def clean_dna(rec):
    """apps/counter.cpp:130-169 (clean_dna) + :173-190 (getLineFasta): a record is cut at
    every run of non-ACGT characters; each piece is counted as its own sequence."""
    import re
    return [p for p in re.split("[^ACGTacgt]+", rec) if p]


def fasta_records(path):
    return [p.upper() for r in fasta_records_raw(path) for p in clean_dna(r)]


def fasta_records_raw(path):
    recs, cur = [], []
    for line in open(path):
        line = line.strip()
        if line.startswith(">"):
            if cur:
                recs.append("".join(cur))
            cur = []
        elif line:
            cur.append(line)
    if cur:
        recs.append("".join(cur))
    return recs


def tie_reads():
    """Low-complexity reads that make equal m-mers share a window (tie rules)."""
    rng = np.random.default_rng(7)
    out = []
    for unit in ["A", "C", "T", "G", "AC", "AT", "CG", "ACG", "AAT", "ACGT", "AACCT", "ACGTTGC", "AAAAAAC",
                 "ACACACGTGTGT", "AGCTAGCTAGGATCC"]:
        for L in (150, 251):
            out.append((unit * (L // len(unit) + 1))[:L])
    # random read with an exact tandem duplication and with its own reverse complement appended
    comp = str.maketrans("ACGT", "TGCA")
    for _ in range(20):
        core = "".join("ACGT"[i] for i in rng.integers(0, 4, 40))
        left = "".join("ACGT"[i] for i in rng.integers(0, 4, 30))
        out.append(left + core + core + core[::-1].translate(comp) + left[:25])
        out.append(core + core[::-1].translate(comp) + core + left + core[:33] + core[:33])
    # palindromic-ish and near-repeats
    for _ in range(20):
        core = "".join("ACGT"[i] for i in rng.integers(0, 4, 23))
        mut = list(core)
        mut[int(rng.integers(0, 23))] = "ACGT"[int(rng.integers(0, 4))]
        out.append(core * 3 + "".join(mut) * 3 + core * 2)
    # edge lengths around k for k = 31, 63, 64
    base = "".join("ACGT"[i] for i in rng.integers(0, 4, 300))
    for L in (20, 30, 31, 32, 33, 62, 63, 64, 65, 66, 100, 127, 128, 129):
        out.append(base[7:7 + L])
    return out


def run(binary, mode, reads_path, k, m, b, out_path, *extra):
    cmd = [os.path.join(ROOT, "oracle", "_ref", binary), mode, reads_path, str(k), str(m), str(b), out_path, *extra]
    return subprocess.run(cmd, check=True, capture_output=True, text=True).stdout


def sha(path):
    return hashlib.sha256(open(path, "rb").read()).hexdigest()


def make_case(name, reads, k, m, b, keep_arrays, query=None, inputs_file=None):
    binary = "ref_dump_k64" if k == 64 else "ref_dump"
    case = dict(name=name, k=k, m=m, b=b, n_reads=len(reads), reference_build=binary)
    if inputs_file:
        case["inputs"] = inputs_file
    with tempfile.TemporaryDirectory() as t:
        rp = os.path.join(t, "reads.txt")
        open(rp, "w").write("\n".join(reads) + "\n")
        o = run(binary, "enum", rp, k, m, b, os.path.join(t, "enum.bin"))
        case["occurrences"] = int(o.split("occurrences ")[1].split()[0])
        case["superkmers_emitted"] = int(o.split("superkmers ")[1].split()[0])
        case["enum_sha256"] = sha(os.path.join(t, "enum.bin"))
        run(binary, "hash", rp, k, m, b, os.path.join(t, "hash.bin"))
        case["hash_sha256"] = sha(os.path.join(t, "hash.bin"))
        extra = []
        if query is not None:
            qp = os.path.join(t, "q.txt")
            open(qp, "w").write("\n".join(query) + "\n")
            extra = [qp]
        o = run(binary, "query" if query is not None else "count", rp, k, m, b, os.path.join(t, "dump.bin"), *extra)
        assert "MISSING_DATA" not in o
        for key in ("nb_buckets", "nb_skmers", "nb_kmers", "largest_bucket"):
            case[key] = int(o.split(key + " ")[1].split()[0])
        case["dump_sha256"] = sha(os.path.join(t, "dump.bin"))
        if query is not None:
            case["query"] = [[int(x) for x in ln.split()[1:]] for ln in o.splitlines() if ln.startswith("q ")]
        enum = np.fromfile(os.path.join(t, "enum.bin"), ENUM_DT)
        dump = np.fromfile(os.path.join(t, "dump.bin"), DUMP_DT)
        assert len(dump) == case["nb_kmers"]
        case["sum_counts"] = int(dump["count"].astype(np.int64).sum())
        sizes = np.bincount(np.unique(enum["read"].astype(np.int64) * (1 << 20) + enum["skmer"], return_counts=True)[1])
        case["skmer_size_hist"] = [int(x) for x in sizes]
        if keep_arrays:
            np.savez_compressed(os.path.join(HERE, name + ".npz"),
                                **{"enum_" + f: enum[f] for f in ENUM_DT.names},
                                **{"dump_" + f: dump[f] for f in DUMP_DT.names})
            case["arrays"] = name + ".npz"
    return case


def main():
    os.makedirs(os.path.join(HERE, "inputs"), exist_ok=True)
    cases = []

    def write_inputs(fname, reads):
        open(os.path.join(HERE, "inputs", fname), "w").write("\n".join(reads) + "\n")
        return "inputs/" + fname

    test_fa = fasta_records(os.path.join(REF, "data", "test.fa"))
    debug_fa = fasta_records(os.path.join(REF, "data", "debug", "test.fa"))
    f1 = write_inputs("test_fa.txt", test_fa)
    f2 = write_inputs("debug_fa.txt", debug_fa)
    ties = tie_reads()
    f3 = write_inputs("ties.txt", ties)

    # config 1 and the survey's pinned known answers
    cases.append(make_case("test_fa_k63", test_fa, 63, 21, 14, True, inputs_file=f1))
    cases.append(make_case("debug_fa_k31", debug_fa, 31, 15, 12, False, inputs_file=f2))
    cases.append(make_case("test_fa_k64", test_fa, 64, 23, 14, True, inputs_file=f1))
    cases.append(make_case("test_fa_k31_m19", test_fa, 31, 19, 16, False, inputs_file=f1))
    cases.append(make_case("debug_fa_k63", debug_fa, 63, 21, 14, False, inputs_file=f2))
    cases.append(make_case("debug_fa_k32", debug_fa, 32, 15, 10, False, inputs_file=f2))
    cases.append(make_case("debug_fa_k33", debug_fa, 33, 13, 8, False, inputs_file=f2))
    cases.append(make_case("debug_fa_k47", debug_fa, 47, 17, 12, False, inputs_file=f2))
    cases.append(make_case("debug_fa_k21_m7", debug_fa, 21, 7, 4, False, inputs_file=f2))
    # tie rules / edge lengths
    for (k, m, b) in [(31, 15, 12), (63, 21, 14), (64, 23, 14), (31, 19, 16), (21, 9, 6), (40, 11, 8)]:
        cases.append(make_case("ties_k%d_m%d" % (k, m), ties, k, m, b, k in (31, 63), inputs_file=f3))
    # scaled-down versions of configs 2-5 (synthetic generator)
    for name, G, L, cov, seed, k, m, b in [
        ("synth_cfg2", 40000, 150, 12, 2, 31, 15, 12),
        ("synth_cfg3", 40000, 250, 8, 3, 63, 21, 14),
        ("synth_cfg4", 40000, 250, 6, 4, 64, 23, 14),
        ("synth_cfg5", 40000, 150, 10, 5, 31, 19, 16),
    ]:
        bases, offs = synth.read_set(G, L, cov, seed)
        reads = [bytes(bases[int(offs[i]):int(offs[i + 1])]).decode() for i in range(len(offs) - 1)]
        qb, qo = synth.read_set(G, L, 1, seed + 1000)  # decoy genome: expected hits ~ 0
        query = reads[:50] + [bytes(qb[int(qo[i]):int(qo[i + 1])]).decode() for i in range(50)]
        c = make_case(name, reads, k, m, b, False, query=query)
        c["synth"] = dict(G=G, L=L, coverage=cov, seed=seed, decoy_seed=seed + 1000)
        cases.append(c)

    json.dump(dict(reference_commit="1f579a07b515c0bcd01ecf275e895b059cf41995", cases=cases),
              open(os.path.join(HERE, "golden.json"), "w"), indent=1)
    for c in cases:
        print(c["name"], c["occurrences"], c["nb_kmers"], c["superkmers_emitted"], c["nb_skmers"], c["nb_buckets"])


if __name__ == "__main__":
    main()
