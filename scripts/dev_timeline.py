#!/usr/bin/env python
"""Device-resident step of bench.py (cfg 3, one GPU) without reading the stage times between the calls: the count call's
store kernel may then run beside the query call's enumeration (BRISK_ASYNC_STORE).  Development tool."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from brisk_b200 import Brisk, Parameters, synth  # noqa: E402

K, M, B, L, G, SEED = 63, 21, 14, 250, 100_000_000, 3
N = 1 << 20
steps, warm = 8, 3
mode = os.environ.get("MODE", "count+query")
gen = synth.device_genome(G, SEED)
dev = [synth.device_reads(gen, L, SEED, j * N, N) for j in range(steps + warm + 1)]
offs = torch.from_numpy(np.arange(N + 1, dtype=np.int64) * L).cuda()
torch.cuda.synchronize()
idx = Brisk(Parameters(K, M, B), capacity_hint=2 * G + (1 << 20))
idx.count_reads(dev[0], offs)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
ok = True
for j in range(1, steps + warm + 1):
    if j == warm + 1:
        idx.sync()
        torch.cuda.synchronize()
        e0.record()
    idx.count_reads(dev[j], offs)
    if mode == "count+query":
        hits, _ = idx.query_reads(dev[j - 1], offs)
        ok &= hits == N * (L - K + 1)
idx.sync()
e1.record()
torch.cuda.synchronize()
print("%s %s: %.3f ms per step, queries ok %s, checksum %s" % (os.environ.get("TAG", ""), mode, e0.elapsed_time(e1) / steps, ok, "%016x" % idx.checksum()[0]))
