"""device-resident FASTA text (one record per read, 250 bp, header ">r<i>") -> parse only, and parse + count"""
import sys, time, numpy as np, ctypes as C
sys.path.insert(0, '.')
import torch
from brisk_b200 import Brisk, Parameters, synth
n, L = 1 << 20, 250
gen = synth.genome(10_000_000, 7)
reads = synth.reads(gen, L, 7, 0, n)
hdr = np.frombuffer(b">r0000000\n", np.uint8)
rec = np.empty((n, len(hdr) + L + 1), np.uint8)
rec[:, :len(hdr)] = hdr
rec[:, len(hdr):len(hdr) + L] = reads
rec[:, -1] = 10
text = torch.from_numpy(rec.reshape(-1)).cuda()
idx = Brisk(Parameters(63, 21, 14), capacity_hint=21_000_000)
nb, nr = C.c_uint64(), C.c_uint64()
for rep in range(4):
    torch.cuda.synchronize(); t0 = time.time()
    idx._check(idx._lib.brisk_parse_fasta(idx._h, C.c_void_p(text.data_ptr()), text.numel(), None, 0, None, 0, C.byref(nb), C.byref(nr)))
    dt = time.time() - t0
    print("parse %d bytes -> %d reads %d bases: %.3f ms = %.1f GB/s of text" % (text.numel(), nr.value, nb.value, dt * 1e3, text.numel() / dt / 1e9), flush=True)
for rep in range(2):
    torch.cuda.synchronize(); t0 = time.time()
    idx.count_fasta(text)
    dt = time.time() - t0
    print("count_fasta: %.3f ms" % (dt * 1e3), idx.last_timing())
