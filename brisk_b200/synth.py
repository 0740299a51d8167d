"""Synthetic reads for the benchmark configs.

Counter-based, so any read can be regenerated independently and the same bytes are
produced on any machine:

    genome[i] = "ACGT"[splitmix64(seed ^ (0x9E3779B97F4A7C15 * (i + 1))) & 3]
    read r    : start  = splitmix64(seed + 1 + 2r) mod (G - L + 1)
                strand = splitmix64(seed + 1 + 2r + 1) & 1
                bases  = genome[start : start + L], reverse-complemented when strand = 1

Error-free, upper-case, no N.  n_reads = floor(G * coverage / L).
"""
import numpy as np

_GOLD = np.uint64(0x9E3779B97F4A7C15)


def splitmix64(x):
    x = np.asarray(x, dtype=np.uint64)
    with np.errstate(over="ignore"):
        z = x + _GOLD
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        return z ^ (z >> np.uint64(31))


_ACGT = np.frombuffer(b"ACGT", dtype=np.uint8)
_COMP = np.zeros(256, np.uint8)
for _a, _b in zip(b"ACGT", b"TGCA"):
    _COMP[_a] = _b


def genome(G, seed, chunk=1 << 24):
    out = np.empty(G, np.uint8)
    for lo in range(0, G, chunk):
        hi = min(G, lo + chunk)
        i = np.arange(lo + 1, hi + 1, dtype=np.uint64)
        with np.errstate(over="ignore"):
            v = splitmix64(np.uint64(seed) ^ (_GOLD * i))
        out[lo:hi] = _ACGT[(v & np.uint64(3)).astype(np.intp)]
    return out


def n_reads_for(G, L, coverage):
    return (G * coverage) // L


def reads(gen, L, seed, first, count):
    """(count, L) uint8 array of reads first .. first+count-1 of the set."""
    G = len(gen)
    r = np.arange(first, first + count, dtype=np.uint64)
    with np.errstate(over="ignore"):
        base = np.uint64(seed + 1) + np.uint64(2) * r
        start = (splitmix64(base) % np.uint64(G - L + 1)).astype(np.int64)
        strand = (splitmix64(base + np.uint64(1)) & np.uint64(1)).astype(bool)
    idx = start[:, None] + np.arange(L, dtype=np.int64)[None, :]
    out = gen[idx]
    rc = _COMP[out[strand][:, ::-1]]
    out[strand] = rc
    return out


def read_set(G, L, coverage, seed, max_reads=None):
    """Flat (bases, offsets) of the whole set (or its first max_reads reads)."""
    gen = genome(G, seed)
    n = n_reads_for(G, L, coverage)
    if max_reads is not None:
        n = min(n, max_reads)
    arr = reads(gen, L, seed, 0, n)
    offsets = np.arange(n + 1, dtype=np.uint64) * np.uint64(L)
    return arr.reshape(-1), offsets


# ---- the same generator on the device (libbrisk_b200: brisk_synth_genome / brisk_synth_reads, csrc/synth.cuh)
def device_genome(G, seed, device=None):
    """torch uint8 tensor of G bases on the current CUDA device, bit-identical to genome(G, seed)"""
    import torch
    from . import _capi
    lib = _capi.load()
    out = torch.empty(G, dtype=torch.uint8, device=device or "cuda")
    rc = lib.brisk_synth_genome(out.data_ptr(), G, seed, torch.cuda.current_stream().cuda_stream)
    if rc:
        raise RuntimeError(lib.brisk_last_error().decode())
    return out


def device_reads(gen, L, seed, first, count, out=None):
    """torch uint8 tensor of count * L bases (reads first .. first+count-1), bit-identical to reads(...).reshape(-1)"""
    import torch
    from . import _capi
    lib = _capi.load()
    if out is None:
        out = torch.empty(count * L, dtype=torch.uint8, device=gen.device)
    rc = lib.brisk_synth_reads(gen.data_ptr(), gen.numel(), L, seed, first, count, out.data_ptr(), torch.cuda.current_stream().cuda_stream)
    if rc:
        raise RuntimeError(lib.brisk_last_error().decode())
    return out
