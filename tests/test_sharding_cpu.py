"""World-size-2 `gloo` test of the host side of the sharded path on the CPU: bucket -> owner routing
(brisk_owner_of_bucket), the send / receive layout of the exchange (brisk_exchange_plan) and the all-to-all itself.
The superkmers come from the oracle (test infrastructure); every rank enumerates its half of the reads, routes whole
superkmers to the owner of their bucket, and the owners count them.  The union of the shards must be the reference's
golden dump, and no key may live on two ranks.  No CUDA call is made."""
import os
import socket

import numpy as np
import pytest

import golden_util as G

CASES = ["debug_fa_k31", "ties_k63_m21", "test_fa_k64"]


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _superkmers_with_bucket(o, c, reads):
    """[(bucket, [(lo, hi, idx), ...])] from the oracle; bucket = middle 2b bits of the hashed minimizer
    (brisk/Brisk.hpp:135-137)."""
    k, m, b = c["k"], c["m"], c["b"]
    s = (m - b + 1) // 2
    m_mask = (1 << (2 * m)) - 1
    out = []
    for seq in reads:
        e = o.enumerate(seq)
        if len(e["lo"]) == 0:
            continue
        bounds = np.flatnonzero(np.diff(e["skmer"])) + 1
        for grp in np.split(np.arange(len(e["lo"])), bounds):
            i = grp[0]
            val = (int(e["hi"][i]) << 64) | int(e["lo"][i])
            w = (val >> (2 * int(e["idx"][i]))) & m_mask
            bucket = ((o.bfc_hash(w) & m_mask) >> (2 * s)) & ((1 << (2 * b)) - 1)
            out.append((bucket, [(int(e["lo"][j]), int(e["hi"][j]), int(e["idx"][j])) for j in grp]))
    return out


def _worker(rank, world, port, name, q):
    import torch
    import torch.distributed as dist
    from brisk_b200 import exchange_plan, owner_of_bucket
    from oracle.oracle import Oracle
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    try:
        c = G.case(name)
        reads = G.reads_of(c)
        if len(reads) == 1:  # one long record: cut it into overlapping pieces so that both ranks have work
            seq, L, k = reads[0], 400, c["k"]
            reads = [seq[i:i + L + k - 1] for i in range(0, len(seq) - k + 1, L)]
        mine = reads[rank::world]
        o = Oracle(c["k"], c["m"], c["b"])
        sks = _superkmers_with_bucket(o, c, mine)
        # partition by owner: rows of (lo, hi, idx) k-mers, whole superkmers stay together
        per_dest = [[] for _ in range(world)]
        for bucket, kmers in sks:
            per_dest[owner_of_bucket(bucket, c["b"], world)].extend(kmers)
        counts_row = torch.tensor([len(x) for x in per_dest], dtype=torch.int64)
        rows = [torch.zeros(world, dtype=torch.int64) for _ in range(world)]
        dist.all_gather(rows, counts_row)
        matrix = torch.stack(rows).numpy().astype(np.uint64)
        ss, rs = exchange_plan(matrix, rank)
        assert int(ss[-1]) == sum(len(x) for x in per_dest)
        send = np.array([km for d in range(world) for km in per_dest[d]], dtype=np.uint64).reshape(-1, 3)
        recv = np.zeros((int(rs[-1]), 3), np.uint64)
        ops = []
        for p in range(world):
            a, b_ = int(ss[p]), int(ss[p + 1])
            ra, rb = int(rs[p]), int(rs[p + 1])
            if p == rank:
                recv[ra:rb] = send[a:b_]
                continue
            if b_ > a:
                ops.append(dist.isend(torch.from_numpy(send[a:b_].view(np.int64).copy()), p))
            if rb > ra:
                t = torch.zeros((rb - ra, 3), dtype=torch.int64)
                ops.append((dist.irecv(t, p), t, ra, rb))
        for op in ops:
            if isinstance(op, tuple):
                op[0].wait()
                recv[op[2]:op[3]] = op[1].numpy().view(np.uint64)
            else:
                op.wait()
        shard = Oracle(c["k"], c["m"], c["b"])
        if len(recv):
            shard.insert_kmers(recv[:, 0], recv[:, 1], recv[:, 2].astype(np.uint8), count=True)
        lo, hi, ix, cnt = shard.dump()
        parts = [None] * world
        dist.all_gather_object(parts, (lo, hi, ix, cnt))
        if rank == 0:
            q.put(("ok", parts, matrix.tolist()))
    except Exception as e:  # pragma: no cover
        if rank == 0:
            q.put(("err", repr(e), None))
        raise
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("name", CASES)
def test_two_rank_routing_reproduces_the_reference_dump(name):
    import torch.multiprocessing as mp
    c = G.case(name)
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, name, q)) for r in range(2)]
    for p in procs:
        p.start()
    status, parts, matrix = q.get(timeout=240)
    for p in procs:
        p.join(timeout=60)
    assert status == "ok", parts
    assert all(p.exitcode == 0 for p in procs)
    lo = np.concatenate([p[0] for p in parts])
    hi = np.concatenate([p[1] for p in parts])
    ix = np.concatenate([p[2] for p in parts])
    cnt = np.concatenate([p[3] for p in parts])
    keys = set(zip(lo.tolist(), hi.tolist(), ix.tolist()))
    assert len(keys) == len(lo), "a key lives on two ranks"
    if len(G.reads_of(c)) > 1:  # the cut-up single record has its own k-mer multiset (piece borders); whole files are golden
        assert len(lo) == c["nb_kmers"]
        assert G.dump_sha(lo, hi, ix, cnt) == c["dump_sha256"]
    else:
        assert len(lo) > 0 and int(cnt.astype(np.int64).sum()) == c["occurrences"]
    assert min(len(p[0]) for p in parts) > 0.25 * len(lo), "shards are badly unbalanced: %r" % matrix


def test_exchange_plan_offsets():
    from brisk_b200 import exchange_plan
    m = np.array([[1, 2, 3], [4, 5, 6], [7, 8, 9]], np.uint64)
    ss, rs = exchange_plan(m, 1)
    assert ss.tolist() == [0, 4, 9, 15] and rs.tolist() == [0, 2, 7, 15]
    ss, rs = exchange_plan(m, 0)
    assert ss.tolist() == [0, 1, 3, 6] and rs.tolist() == [0, 1, 5, 12]
