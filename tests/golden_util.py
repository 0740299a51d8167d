"""Shared helpers: load the committed golden cases (tests/golden/golden.json)."""
import hashlib
import json
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = os.path.join(HERE, "golden")
ENUM_DT = np.dtype([("lo", "<u8"), ("hi", "<u8"), ("mini", "<u8"), ("read", "<u4"), ("skmer", "<u4"), ("idx", "u1")])
DUMP_DT = np.dtype([("lo", "<u8"), ("hi", "<u8"), ("idx", "u1"), ("count", "u1")])


def cases():
    return json.load(open(os.path.join(GOLD, "golden.json")))["cases"]


def case(name):
    for c in cases():
        if c["name"] == name:
            return c
    raise KeyError(name)


def case_names():
    return [c["name"] for c in cases()]


def reads_of(c):
    """list of bytes"""
    if "inputs" in c:
        return [ln.strip().encode() for ln in open(os.path.join(GOLD, c["inputs"])) if ln.strip()]
    from brisk_b200 import synth
    s = c["synth"]
    bases, offs = synth.read_set(s["G"], s["L"], s["coverage"], s["seed"])
    return [bytes(bases[int(offs[i]):int(offs[i + 1])]) for i in range(len(offs) - 1)]


def query_reads_of(c):
    from brisk_b200 import synth
    s = c["synth"]
    reads = reads_of(c)
    qb, qo = synth.read_set(s["G"], s["L"], 1, s["decoy_seed"])
    return reads[:50] + [bytes(qb[int(qo[i]):int(qo[i + 1])]) for i in range(50)]


def flat(reads):
    bases = np.frombuffer(b"".join(reads), np.uint8)
    offs = np.zeros(len(reads) + 1, np.uint64)
    offs[1:] = np.cumsum([len(r) for r in reads])
    return bases, offs


def dump_sha(lo, hi, idx, cnt):
    """sha256 of the (lo, hi, idx, count) records in the reference dumper's byte layout."""
    rec = np.zeros(len(lo), DUMP_DT)
    rec["lo"], rec["hi"], rec["idx"], rec["count"] = lo, hi, idx, cnt
    order = np.lexsort((rec["idx"], rec["lo"], rec["hi"]))
    return hashlib.sha256(rec[order].tobytes()).hexdigest()
