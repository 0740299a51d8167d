import sys, time, numpy as np
sys.path.insert(0, '.')
from brisk_b200 import Brisk, Parameters, synth
for (k,m,b) in [(32,15,10),(64,23,14)]:
    idx = Brisk(Parameters(k,m,b))
    ones = np.uint64(0xFFFFFFFFFFFFFFFF)
    klo = np.full(600, ones); khi = np.full(600, ones if k==64 else np.uint64(0))
    kix = np.repeat(np.array([3,7],np.uint8),300)
    s1, nw = idx.insert_kmers(Brisk.pack_kmers(klo,khi,kix), count=True)
    print(k, nw.sum(), set(s1.tolist()), idx.stats(), idx.dump())
# first timing probe: config 2 shape, 1/4 size
t0=time.time()
bases, offs = synth.read_set(5_000_000, 150, 12, 2)
print("gen", time.time()-t0, len(offs)-1)
import torch
tb = torch.from_numpy(bases).cuda(); to = torch.from_numpy(offs.astype(np.int64)).cuda()
for rep in range(3):
    idx = Brisk(Parameters(31,15,12), capacity_hint=6_000_000)
    torch.cuda.synchronize()
    t0=time.time(); idx.count_reads(tb,to); dt=time.time()-t0
    n_occ = (len(offs)-1)*(150-31+1)
    print("count", dt, idx.last_timing(), idx.stats(), "Gocc/s", n_occ/dt/1e9)
    t0=time.time(); idx.count_reads(bases,offs); dt=time.time()-t0
    print("count host-in", dt, idx.last_timing(), "Gocc/s", n_occ/dt/1e9)
