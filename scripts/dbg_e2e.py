"""development aid: one count + one query over a pinned host batch, with the library's own stage / copy timing"""
import sys, time, numpy as np
sys.path.insert(0, '.')
import torch
from brisk_b200 import Brisk, Parameters, synth
k, m, b, L = 63, 21, 14, 250
G = 100_000_000
gen = synth.genome(G, 3)
n = 1 << 20
hb = []
for j in range(3):
    t = torch.empty(n * L, dtype=torch.uint8, pin_memory=True)
    t.numpy()[:] = synth.reads(gen, L, 3, j * n, n).reshape(-1)
    hb.append(t)
offs = torch.from_numpy(np.arange(n + 1, dtype=np.int64) * L).pin_memory()
for flags in (0,):
    idx = Brisk(Parameters(k, m, b), capacity_hint=2 * G + (1 << 20), flags=flags)
    for rep in range(2):
        for j in range(3):
            torch.cuda.synchronize(); t0 = time.time(); idx.count_reads(hb[j], offs); t1 = time.time()
            r = idx.query_reads(hb[j], offs); t2 = time.time()
            idx.last_timing()
            print("flags", flags, "count returns after %.2f ms, query %.2f ms, both %.2f ms" % ((t1 - t0) * 1e3, (t2 - t1) * 1e3, (t2 - t0) * 1e3), flush=True)
    idx.close()
