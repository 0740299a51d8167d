"""CPU restatement of the reference's FASTA ingest (apps/counter.cpp) — TEST INFRASTRUCTURE ONLY, like the rest of
oracle/: imported by tests/ as the checker of the GPU FASTA parser, never by the product package.

  getLineFasta  apps/counter.cpp:172-190   one record per call: the first line (the header) is dropped, the following
                                           lines are concatenated until the next line starts with '>' (or EOF); when
                                           `previous` holds a carried-over tail it is returned instead of reading
  clean_dna     apps/counter.cpp:128-168   keep the prefix before the first character outside ACGTacgt, carry the part
                                           starting at the next ACGTacgt character over to the next call, upper-case
  count_fasta   apps/counter.cpp:197-226   loop `while (in.good() or prev.size() != 0)`; empty results are skipped

Net effect (single reader thread): the reads are the maximal runs of [ACGTacgt] of every record's concatenated
sequence lines ('\\n' removed, anything else — '\\r' included — cuts), upper-cased, in file order.  The first line of the
file is always consumed as a header.  Pure Python loops: small inputs only."""

ACGT = set(b"ACGTacgt")


def _getline(data, pos):
    """std::getline: the line without its '\\n', next position, and whether EOF was hit while reading"""
    j = data.find(b"\n", pos)
    if j < 0:
        return data[pos:], len(data), True
    return data[pos:j], j + 1, False


def clean_dna(s, previous):
    """apps/counter.cpp:128-168 -> (str, previous)"""
    i, dog = 0, True
    while i < len(s) and dog:                   # :131-144
        if s[i] not in ACGT:
            dog = False
        i += 1
    if not dog:
        iforreturn = i - 1                      # :147
        while i < len(s) and not dog:           # :148-159
            if s[i] in ACGT:
                dog = True
            i += 1
        previous = s[i - 1:] if dog else b""    # :160-164
        s = s[:iforreturn]                      # :165
    return s.upper(), previous                  # :167


def fasta_reads(data: bytes):
    """the sequence of non-empty strings count_fasta hands to count_sequence with one thread"""
    out = []
    pos, good, prev = 0, True, b""
    while good or prev:                         # :213
        if not prev:                            # getLineFasta :174-182
            _, pos, eof = _getline(data, pos)
            if eof:
                good = False
            result = b""
            while pos < len(data) and data[pos:pos + 1] != b">":   # peek() != '>' and != EOF
                line, pos, eof = _getline(data, pos)
                result += line
                if eof:
                    good = False
            if pos >= len(data):                # peek() returned EOF: eofbit set
                good = False
        else:                                   # :183-186
            result, prev = prev, b""
        line, prev = clean_dna(result, prev)
        if line != b"":                         # :221
            out.append(line)
    return out


def flat(reads):
    import numpy as np
    bases = np.frombuffer(b"".join(reads), np.uint8)
    offs = np.zeros(len(reads) + 1, np.uint64)
    if reads:
        offs[1:] = np.cumsum([len(r) for r in reads])
    return bases, offs
