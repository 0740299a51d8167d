"""One rank of the multi-GPU parity check (run under torchrun by tests/test_gpu_sharding.py or by hand):
every rank supplies its share of the reads of a golden case to the COLLECTIVE count / query / get / insert calls;
the union of the shard dumps must be the reference's golden dump and the per-occurrence answers must equal the
single-GPU ones."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import golden_util as G  # noqa: E402


def main():
    import torch
    import torch.distributed as dist
    from brisk_b200 import Brisk, Parameters
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    for name in ["synth_cfg2", "synth_cfg3", "synth_cfg4", "synth_cfg5", "debug_fa_k31", "ties_k63_m21"]:
        c = G.case(name)
        reads = G.reads_of(c)
        mine = reads[rank::world]
        bases, offs = G.flat(mine)
        idx = Brisk(Parameters(c["k"], c["m"], c["b"]), device=local)
        idx.comm_init_torch(dist, rank, world)
        assert idx.comm_info() == (rank, world) and idx.device == local
        idx.count_reads(bases, offs)
        # routed query: all ranks get the totals of the whole batch
        hits, total = idx.query_reads(bases, offs)
        assert hits == c["occurrences"], (name, hits, c["occurrences"])
        st = idx.stats()
        assert st["nb_kmers"] == c["nb_kmers"] and st["nb_buckets"] == c["nb_buckets"], (name, st)
        # per-occurrence answers come back to the rank that asked, in read order
        found, data = idx.get_reads(bases, offs)
        ref = Brisk(Parameters(c["k"], c["m"], c["b"]), device=local)  # single-GPU index over ALL reads as the comparison
        ab, ao = G.flat(reads)
        ref.count_reads(ab, ao)
        rfound, rdata = ref.get_reads(bases, offs)
        assert np.array_equal(found, rfound) and np.array_equal(data, rdata), name
        assert found.all() and int(data.astype(np.int64).sum()) == total_of(ref, bases, offs)
        assert total == sum_over_ranks(dist, torch, int(data.astype(np.int64).sum()))
        # decoy queries: mostly absent
        if "synth" in c:
            q = G.query_reads_of(c)[50:]
            qb, qo = G.flat(q[rank::world])
            f2, d2 = idx.get_reads(qb, qo)
            rf2, rd2 = ref.get_reads(qb, qo)
            assert np.array_equal(f2, rf2) and np.array_equal(d2, rd2), name
        # insert_reads on a fresh sharded index: newly flags = first occurrence in the global order is not defined
        # across ranks, but their number is: one per distinct k-mer
        # (room for everything up front: the slot ids of the second pass below are compared with the first pass's, and
        # slot ids are only stable while the store does not grow)
        idx2 = Brisk(Parameters(c["k"], c["m"], c["b"]), device=local, capacity_hint=4 * c["nb_kmers"])
        idx2.comm_init_torch(dist, rank, world)
        slots, newly = idx2.insert_reads(bases, offs)
        t = torch.tensor([int(newly.sum())], device="cuda")
        dist.all_reduce(t)
        assert int(t.item()) == c["nb_kmers"], (name, int(t.item()))
        if len(slots):  # a rank may have no reads at all (single-record inputs): it still takes part in every call
            owners = (slots >> np.uint64(56)).astype(np.int64)
            assert owners.min() >= 0 and owners.max() < world and (slots != np.uint64(0xFFFFFFFFFFFFFFFF)).all()
        s2, n2 = idx2.insert_reads(bases, offs)
        assert np.array_equal(s2, slots) and not n2.any(), name
        # union of the shard dumps = the golden dump
        parts = [None] * world
        dist.all_gather_object(parts, idx.dump())
        if rank == 0:
            lo, hi, ix, cnt = (np.concatenate([p[i] for p in parts]) for i in range(4))
            assert len(lo) == c["nb_kmers"]
            assert G.dump_sha(lo, hi, ix, cnt) == c["dump_sha256"], name
            assert min(len(p[0]) for p in parts) > 0.5 * len(lo) / world
            print("case %s ok: %d k-mers over %d ranks %s" % (name, len(lo), world, [len(p[0]) for p in parts]), flush=True)
        idx.close(); idx2.close(); ref.close()
    dist.barrier()
    if rank == 0:
        print("MGPU OK", flush=True)
    dist.destroy_process_group()


def sum_over_ranks(dist, torch, v):
    t = torch.tensor([v], device="cuda")
    dist.all_reduce(t)
    return int(t.item())


def total_of(ref, bases, offs):
    return ref.query_reads(bases, offs)[1]


if __name__ == "__main__":
    main()
