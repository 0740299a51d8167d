#!/usr/bin/env python
"""Host wall time of the count call and of the query call of bench.py's host-buffer (e2e) step, cfg 3, one GPU:
where the 13 ms of an e2e step go.  Development tool (run under the BRISK_OVERLAP_* switches)."""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from brisk_b200 import Brisk, Parameters, synth  # noqa: E402

K, M, B, L, G, SEED = 63, 21, 14, 250, 100_000_000, 3
N = 1 << 20
steps, warm = 8, 3
gen = synth.device_genome(G, SEED)
host = []
for j in range(steps + warm + 1):
    d = synth.device_reads(gen, L, SEED, j * N, N)
    t = torch.empty(N * L, dtype=torch.uint8, pin_memory=True)
    t.copy_(d)
    host.append(t)
offs = torch.from_numpy(np.arange(N + 1, dtype=np.int64) * L).pin_memory()
torch.cuda.synchronize()
idx = Brisk(Parameters(K, M, B), capacity_hint=2 * G + (1 << 20))
idx.count_reads(host[0], offs)
tc, tq = [], []
for j in range(1, steps + warm + 1):
    t0 = time.perf_counter()
    idx.count_reads(host[j], offs)
    t1 = time.perf_counter()
    hits, _ = idx.query_reads(host[j - 1], offs)
    t2 = time.perf_counter()
    assert hits == N * (L - K + 1)
    if j > warm:
        tc.append((t1 - t0) * 1e3)
        tq.append((t2 - t1) * 1e3)
print("%s count %.3f ms  query %.3f ms  step %.3f ms" % (os.environ.get("TAG", ""), np.mean(tc), np.mean(tq), np.mean(tc) + np.mean(tq)))
