"""Generates the FASTA-ingest fixtures: writes the crafted FASTA files of tests/golden/fasta/ and records what the
UNMODIFIED reference app (oracle/_ref/counter = apps/counter.cpp built by oracle/Makefile) reports for them with one
thread (`nb kmers`, `bucket used`, and that its --mode 2 self-check passes).  Run in the build container:
    python tests/golden/make_fasta_golden.py
"""
import json
import os
import re
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
COUNTER = os.path.join(ROOT, "oracle", "_ref", "counter")


def rnd(rng, n, alphabet=b"ACGT"):
    return bytes(rng.choice(list(alphabet), n).astype(np.uint8))


def files():
    rng = np.random.default_rng(5)
    f = {}
    f["simple.fa"] = b">r1 some description\n" + rnd(rng, 300) + b"\n>r2\n" + rnd(rng, 200) + b"\n"
    # multi-line records, lower case, N runs, empty records, blank lines, no trailing newline
    body = b""
    for r in range(12):
        body += b">rec%d\n" % r
        seq = bytearray(rnd(rng, int(rng.integers(50, 700))))
        for _ in range(int(rng.integers(0, 4))):
            p = int(rng.integers(0, len(seq)))
            seq[p:p + int(rng.integers(1, 9))] = b"N" * 8
        if r % 3 == 0:
            seq = bytearray(bytes(seq).lower())
        w = int(rng.integers(30, 90))
        body += b"\n".join(bytes(seq[i:i + w]) for i in range(0, len(seq), w)) + b"\n"
        if r == 4:
            body += b">empty\n>another empty\n"
        if r == 7:
            body += b"\n\n"
    f["multiline_n.fa"] = body + b">last\n" + rnd(rng, 150)
    # CRLF line ends ('\r' cuts every line), IUPAC codes, spaces, '>' inside a sequence line
    f["crlf_iupac.fa"] = (b">w\r\n" + rnd(rng, 120) + b"\r\n" + rnd(rng, 130) + b"\r\n>x\n" + rnd(rng, 90) + b"RYKM" + rnd(rng, 95) +
                          b" " + rnd(rng, 80) + b">notaheader" + rnd(rng, 75) + b"\n" + rnd(rng, 66) + b"\n")
    # first line is not a header: it is consumed as one anyway
    f["no_header.fa"] = rnd(rng, 100) + b"\n" + rnd(rng, 180) + b"\n>h\n" + rnd(rng, 140) + b"\n"
    # a record that starts and ends with N, and one with nothing but N
    f["edges.fa"] = b">a\nNNNN" + rnd(rng, 160) + b"NN\n>b\nNNNNNNNN\n>c\n" + rnd(rng, 100) + b"\nN\n" + rnd(rng, 120) + b"\n"
    return f


def main():
    os.makedirs(os.path.join(HERE, "fasta"), exist_ok=True)
    cases = []
    for name, data in files().items():
        path = os.path.join(HERE, "fasta", name)
        open(path, "wb").write(data)
        for (k, m, b) in [(31, 15, 6), (21, 9, 4), (63, 21, 8)]:
            out = subprocess.run([COUNTER, "-f", path, "-k", str(k), "-m", str(m), "-b", str(b), "--mode", "2", "-t", "1"],
                                 capture_output=True, text=True, check=True).stdout
            assert "All counts are correct" in out, out
            cases.append(dict(file=name, k=k, m=m, b=b,
                              nb_kmers=int(re.search(r"nb kmers: ([\d,]+)", out).group(1).replace(",", "")),
                              nb_buckets=int(re.search(r"([\d,]+) bucket used", out).group(1).replace(",", ""))))
    json.dump(dict(generator="tests/golden/make_fasta_golden.py", reference="oracle/_ref/counter -t 1 --mode 2", cases=cases),
              open(os.path.join(HERE, "fasta_golden.json"), "w"), indent=1)
    print(len(cases), "cases")


if __name__ == "__main__":
    main()
