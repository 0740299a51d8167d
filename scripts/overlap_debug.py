#!/usr/bin/env python
"""host-input (overlapped) index against a device-input (plain) index, step by step (development tool)"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from brisk_b200 import Brisk, Parameters, synth  # noqa: E402

K, M, B, L, G, SEED = 63, 21, 14, 250, 100_000_000, 3
N = 1 << 20
steps = 6
gen = synth.device_genome(G, SEED)
host, dev = [], []
for j in range(steps + 1):
    d = synth.device_reads(gen, L, SEED, j * N, N)
    t = torch.empty(N * L, dtype=torch.uint8, pin_memory=True)
    t.copy_(d)
    host.append(t)
    dev.append(d)
offs = torch.from_numpy(np.arange(N + 1, dtype=np.int64) * L).pin_memory()
offs_d = offs.cuda()
torch.cuda.synchronize()
a = Brisk(Parameters(K, M, B), capacity_hint=2 * G + (1 << 20))
b = Brisk(Parameters(K, M, B), capacity_hint=2 * G + (1 << 20))
want = N * (L - K + 1)
for j in range(steps + 1):
    a.count_reads(host[j], offs)
    la = a.last_timing()["launches"] if os.environ.get("READ_TIMING") else -1
    b.count_reads(dev[j], offs_d)
    if j:
        qa = a.query_reads(host[j - 1], offs)
        qb = b.query_reads(dev[j - 1], offs_d)
        qab = a.query_reads(dev[j - 1], offs_d)
    else:
        qa = qb = qab = None
    print(os.environ.get("TAG", ""), j, "launches", la, "checksum equal", a.checksum() == b.checksum(), "query host", qa, "dev-index", qb,
          "host-index dev-input", qab, "want", want)
