"""quick device-resident timing of count / query over one batch shape (development aid)"""
import sys, time, numpy as np
sys.path.insert(0, '.')
import torch
from brisk_b200 import Brisk, Parameters, synth
cfgs = {"cfg2": (31,15,12,150), "cfg3": (63,21,14,250), "cfg5": (31,19,16,150), "cfg4": (64,23,14,250)}
names = sys.argv[1:] or ["cfg2", "cfg3"]
G = 10_000_000
gen = synth.genome(G, 7)
for name in names:
    k,m,b,L = cfgs[name]
    n = 1 << 20
    batches = []
    for j in range(3):
        arr = synth.reads(gen, L, 7, j*n, n).reshape(-1)
        batches.append(torch.from_numpy(arr).cuda())
    offs = torch.from_numpy((np.arange(n+1, dtype=np.int64)*L)).cuda()
    occ = n*(L-k+1)
    idx = Brisk(Parameters(k,m,b), capacity_hint=(2 if k>32 else 1)*G + (1<<20))
    for j in range(3):
        torch.cuda.synchronize(); t0=time.time(); idx.count_reads(batches[j], offs); dt=time.time()-t0
        t = idx.last_timing()
        print(name, "count", j, "wall %.2f ms" % (dt*1e3), {a: round(v,3) for a,v in t.items()}, "dev Gocc/s %.2f" % (occ/t["total_ms"]/1e6), flush=True)
    torch.cuda.synchronize(); t0=time.time(); r = idx.query_reads(batches[0], offs); dt=time.time()-t0
    print(name, "query", r, "wall %.2f ms" % (dt*1e3), {a: round(v,3) for a,v in idx.last_timing().items()}, idx.stats())
