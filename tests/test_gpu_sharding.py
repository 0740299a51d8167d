"""GPU parity of the routed path: the radix partition of superkmer records by owner rank and the owner-side apply
(single GPU, virtual ranks), and — when the box has two or more GPUs — the real thing: one process per GPU under
torchrun with the NCCL all-to-all (tests/mgpu_worker.py)."""
import os
import subprocess
import sys

import numpy as np
import pytest

import golden_util as G

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def make_index(c, **kw):
    from brisk_b200 import Brisk, Parameters
    return Brisk(Parameters(c["k"], c["m"], c["b"]), **kw)


@pytest.mark.parametrize("n_ranks", [1, 2, 3, 8, 64])
@pytest.mark.parametrize("name", ["synth_cfg2", "synth_cfg3", "synth_cfg4", "synth_cfg5", "ties_k31_m15", "test_fa_k63"])
def test_partition_groups_records_by_owner(name, n_ranks):
    from brisk_b200 import decode_records, owner_of_bucket
    c = G.case(name)
    idx = make_index(c)
    bases, offs = G.flat(G.reads_of(c))
    flat, rw = idx.enumerate_reads(bases, offs)
    recs, starts = idx.partition_reads(bases, offs, n_ranks)
    assert starts[0] == 0 and starts[-1] == len(recs) == len(flat) == c["superkmers_emitted"]
    assert np.all(np.diff(starts.astype(np.int64)) >= 0)
    # same multiset of records
    key = lambda a: a[np.lexsort(a.T[::-1])]
    assert np.array_equal(key(flat), key(recs))
    # every record sits in the region of the owner of its bucket
    buckets = (recs[:, rw - 1] >> np.uint64(64 - 2 * c["b"])).astype(np.uint64)
    region = np.searchsorted(starts[1:], np.arange(len(recs)), side="right")
    uniq = {int(b): owner_of_bucket(int(b), c["b"], n_ranks) for b in np.unique(buckets)}
    owners = np.array([uniq[int(b)] for b in buckets])
    assert np.array_equal(owners, region)
    if n_ranks in (2, 8) and len(recs) > 20000:
        share = np.diff(starts.astype(np.int64)) / len(recs)
        assert share.max() < 1.25 / n_ranks and share.min() > 0.75 / n_ranks


@pytest.mark.parametrize("name", ["synth_cfg2", "synth_cfg3", "synth_cfg4", "synth_cfg5", "debug_fa_k31"])
def test_virtual_shards_reproduce_the_reference_dump(name):
    """N handles on one GPU play the ranks: partition -> (host swap instead of NCCL) -> owner-side count; the union
    of the shard dumps is the reference's golden dump and the routed query finds every inserted occurrence."""
    c = G.case(name)
    n_ranks = 3
    reads = G.reads_of(c)
    shards = [make_index(c) for _ in range(n_ranks)]
    inbox = [[] for _ in range(n_ranks)]
    for r in range(n_ranks):
        bases, offs = G.flat(reads[r::n_ranks])
        recs, starts = shards[r].partition_reads(bases, offs, n_ranks)
        for d in range(n_ranks):
            inbox[d].append(recs[int(starts[d]):int(starts[d + 1])])
    hits = total = 0
    for d in range(n_ranks):
        got = np.ascontiguousarray(np.concatenate(inbox[d]))
        shards[d].apply_records(got, "count")
        h, t = shards[d].apply_records(got, "get")
        hits += h
        total += t
    parts = [s.dump() for s in shards]
    lo, hi, ix, cnt = (np.concatenate([p[i] for p in parts]) for i in range(4))
    assert len(lo) == c["nb_kmers"]
    assert G.dump_sha(lo, hi, ix, cnt) == c["dump_sha256"]
    assert sum(s.stats()["nb_buckets"] for s in shards) == c["nb_buckets"]
    assert hits == c["occurrences"]


@pytest.mark.parametrize("name", ["synth_cfg3", "synth_cfg5"])
def test_merged_ordered_dumps_of_two_shards_equal_the_unsharded_ordered_dump(name):
    """The shards' ordered dumps, merged on bucket id (brisk_b200.merge_ordered_dumps), are the unsharded index's ordered
    dump element for element: the global order DenseMenuYo::next would walk."""
    from brisk_b200 import merge_ordered_dumps
    c = G.case(name)
    reads = G.reads_of(c)
    whole = make_index(c)
    whole.count_reads(*G.flat(reads))
    want = whole.dump(ordered=True)
    src = make_index(c)
    recs, starts = src.partition_reads(*G.flat(reads), 2)
    parts = []
    for r in range(2):
        shard = make_index(c)
        shard.apply_records(np.ascontiguousarray(recs[int(starts[r]):int(starts[r + 1])]), "count")
        parts.append(shard.dump(ordered=True))
    got = merge_ordered_dumps(whole.params, parts)
    for a, b in zip(got, want):
        assert np.array_equal(a, b)
    st = whole.stats()
    assert st["largest_bucket"] >= 1 and st["largest_bucket"] == int(np.bincount(
        np.unique(__import__("brisk_b200").bucket_ids(whole.params, *want[:3]), return_inverse=True)[1]).max())


@pytest.mark.parametrize("world", [2])
def test_nccl_sharded_index_matches_golden(world):
    import torch
    if torch.cuda.device_count() < world:
        pytest.skip("needs %d GPUs" % world)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world), "--master-addr", "127.0.0.1",
           "--master-port", "29631", os.path.join(ROOT, "tests", "mgpu_worker.py")]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stdout[-3000:] + out.stderr[-3000:]
    assert "MGPU OK" in out.stdout
