"""A KFF reader written from the format as brisk/lib/kff/kff_io.cpp writes it (test infrastructure): header, value
sections, minimizer sections with their blocks, index section, footer.  Returns everything it saw so that the tests
can check the structure as well as the k-mers."""
import math
import struct


def _be(b):
    return int.from_bytes(b, "big")


def _ceil_log2(x):
    b = 0
    while (1 << b) < x:
        b += 1
    return b


def read_kff(path):
    d = open(path, "rb").read()
    assert d[:3] == b"KFF" and d[-3:] == b"KFF", "missing signature"
    out = {"version": (d[3], d[4]), "encoding": d[5], "uniqueness": d[6], "canonicity": d[7]}
    msz = _be(d[8:12])
    out["metadata"] = d[12:12 + msz]
    pos = 12 + msz
    end = len(d) - 3
    gv = {}
    out["sections"] = []   # (position, type)
    out["value_sections"] = []
    out["minimizer_sections"] = []  # dict(minimizer, blocks=[(n, mini_pos, seq_int, seq_len, data bytes)])
    out["index"] = None
    while pos < end:
        t = chr(d[pos])
        out["sections"].append((pos, t))
        if t == "v":
            n = _be(d[pos + 1:pos + 9])
            p = pos + 9
            vars_ = {}
            for _ in range(n):
                e = d.index(b"\0", p)
                name = d[p:e].decode()
                vars_[name] = _be(d[e + 1:e + 9])
                p = e + 9
            gv.update(vars_)
            out["value_sections"].append((pos, vars_))
            pos = p
        elif t == "m":
            k, m, mx, ds = gv["k"], gv["m"], gv["max"], gv["data_size"]
            nkb = (_ceil_log2(mx) + 7) // 8
            mpb = (_ceil_log2(k + mx - 1) + 7) // 8
            mb = (m + 3) // 4
            mini = _be(d[pos + 1:pos + 1 + mb])
            nb = _be(d[pos + 1 + mb:pos + 9 + mb])
            p = pos + 9 + mb
            blocks = []
            for _ in range(nb):
                n = _be(d[p:p + nkb]) if nkb else 1
                p += nkb
                mp = _be(d[p:p + mpb])
                p += mpb
                sl = n + k - 1 - m
                sb = (sl + 3) // 4
                seq = _be(d[p:p + sb])
                p += sb
                data = d[p:p + n * ds]
                p += n * ds
                blocks.append((n, mp, seq, sl, data))
            out["minimizer_sections"].append({"pos": pos, "minimizer": mini, "blocks": blocks})
            pos = p
        elif t == "i":
            n = _be(d[pos + 1:pos + 9])
            p = pos + 9
            entries = []
            for _ in range(n):
                entries.append((chr(d[p]), int.from_bytes(d[p + 1:p + 9], "big", signed=True)))
                p += 9
            nxt = int.from_bytes(d[p:p + 8], "big", signed=True)
            p += 8
            out["index"] = {"pos": pos, "end": p, "entries": entries, "next": nxt}
            pos = p
        else:
            raise AssertionError("unknown section %r at %d" % (t, pos))
    assert pos == end
    out["vars"] = gv
    return out


def kmers_of(parsed):
    """every (kmer_s, minimizer_idx, data bytes) the file holds, k-mers as integers (first nucleotide most significant)"""
    k, m, ds = parsed["vars"]["k"], parsed["vars"]["m"], parsed["vars"]["data_size"]
    kmask = (1 << (2 * k)) - 1
    res = []
    for sec in parsed["minimizer_sections"]:
        mini = sec["minimizer"]
        for (n, mp, seq, sl, data) in sec["blocks"]:
            suffix_len = sl - mp
            prefix = seq >> (2 * suffix_len)
            suffix = seq & ((1 << (2 * suffix_len)) - 1)
            full = (((prefix << (2 * m)) | mini) << (2 * suffix_len)) | suffix
            L = n + k - 1
            for j in range(n):
                kmer = (full >> (2 * (L - k - j))) & kmask
                idx = (k - m - mp) + j
                res.append((kmer, idx, bytes(data[j * ds:(j + 1) * ds])))
    return res
