"""host->device copy rate of this box from pinned memory (what bounds the e2e figure of bench.py).
Alone:        python scripts/h2d_bw.py
All ranks:    python -m torch.distributed.run --nproc-per-node N ... scripts/h2d_bw.py   (every rank copies to its own GPU
              at the same time; rank 0 prints the per-rank and aggregate rates: the host's H2D ceiling for N concurrent feeds)"""
import os
import torch
rank, world, local = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
if world > 1:
    import torch.distributed as dist
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
N = 256 << 20
h = torch.empty(N, dtype=torch.uint8).pin_memory()
d = torch.empty(N, dtype=torch.uint8, device="cuda")
def run(piece_mb, reps=6):
    p = piece_mb << 20
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for _ in range(2):
        for o in range(0, N, p): d[o:o + p].copy_(h[o:o + p], non_blocking=True)
    torch.cuda.synchronize()
    if world > 1: dist.barrier(); torch.cuda.synchronize()
    a.record()
    for _ in range(reps):
        for o in range(0, N, p): d[o:o + p].copy_(h[o:o + p], non_blocking=True)
    b.record(); torch.cuda.synchronize()
    return reps * N / a.elapsed_time(b) / 1e6
for mb in (256, 32, 8, 32, 256):
    r = run(mb)
    if world > 1:
        t = torch.tensor([r], device="cuda")
        out = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(out, t)
        if rank == 0:
            rates = [float(x.item()) for x in out]
            print("H2D x%d concurrent, 256 MB in pieces of %3d MB: per rank %s GB/s, aggregate %.1f GB/s" % (world, mb, ["%.1f" % x for x in rates], sum(rates)), flush=True)
    else:
        print("H2D 256 MB in pieces of %3d MB: %.1f GB/s" % (mb, r), flush=True)
if world > 1: dist.destroy_process_group()
