"""partition kernels on one GPU (virtual 8 ranks) for an ncu capture: cfg3 batch shape"""
import sys, numpy as np, torch
sys.path.insert(0, '.')
from brisk_b200 import Brisk, Parameters, synth
gen = synth.genome(10_000_000, 7)
n = 1 << 20
arr = torch.from_numpy(synth.reads(gen, 250, 7, 0, n).reshape(-1)).cuda()
offs = torch.from_numpy(np.arange(n + 1, dtype=np.int64) * 250).cuda()
idx = Brisk(Parameters(63, 21, 14))
import ctypes as C
starts = np.zeros(9, np.uint64); rw = C.c_uint32()
for _ in range(3):
    idx._check(idx._lib.brisk_partition_reads(idx._h, C.c_void_p(arr.data_ptr()), C.c_void_p(offs.data_ptr()), n, 8, None, 0, starts.ctypes.data_as(C.c_void_p), C.byref(rw)))
    print(starts, idx.last_timing())
