"""host->device copy rate of this box from pinned memory (what bounds the e2e figure of bench.py)"""
import torch
N = 256 << 20
h = torch.empty(N, dtype=torch.uint8).pin_memory()
d = torch.empty(N, dtype=torch.uint8, device="cuda")
def run(piece_mb, reps=6):
    p = piece_mb << 20
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for _ in range(2):
        for o in range(0, N, p): d[o:o + p].copy_(h[o:o + p], non_blocking=True)
    torch.cuda.synchronize()
    a.record()
    for _ in range(reps):
        for o in range(0, N, p): d[o:o + p].copy_(h[o:o + p], non_blocking=True)
    b.record(); torch.cuda.synchronize()
    return reps * N / a.elapsed_time(b) / 1e6
for mb in (256, 64, 32, 16, 8, 4, 8, 32, 256):
    print("H2D 256 MB in pieces of %3d MB: %.1f GB/s" % (mb, run(mb)), flush=True)
