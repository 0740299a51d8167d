import sys, numpy as np
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
import golden_util as G
from brisk_b200 import Brisk, Parameters
c = G.case("synth_cfg2")
idx = Brisk(Parameters(c["k"], c["m"], c["b"]))
bases, offs = G.flat(G.reads_of(c)[:200])
flat, rw = idx.enumerate_reads(bases, offs)
for n in (1, 2):
    recs, starts = idx.partition_reads(bases, offs, n)
    z = np.flatnonzero((recs == 0).all(axis=1))
    print("n_ranks", n, "n_recs", len(recs), "starts", starts, "zero rows", len(z), z[:20], z[-5:])
    key = lambda a: a[np.lexsort(a.T[::-1])]
    print("  equal:", np.array_equal(key(flat), key(recs)))
from brisk_b200 import owner_of_bucket
recs, starts = idx.partition_reads(bases, offs, 2)
b = (recs[:, rw-1] >> np.uint64(64 - 2*c["b"]))
own = np.array([owner_of_bucket(int(x), c["b"], 2) for x in b])
print("host owners of output rows:", np.bincount(own, minlength=2), "first rows owners", own[:40])
bf = (flat[:, rw-1] >> np.uint64(64 - 2*c["b"]))
ownf = np.array([owner_of_bucket(int(x), c["b"], 2) for x in bf])
print("host owners of enumerate rows:", np.bincount(ownf, minlength=2))
fs = set(map(tuple, flat.tolist())); rs = set(map(tuple, recs.tolist()))
print("distinct flat", len(fs), "distinct part", len(rs), "missing", len(fs - rs), "extra", len(rs - fs))
miss = np.array(sorted(fs - rs)[:5], dtype=np.uint64); print(miss)
if len(miss): print("owners of missing", [owner_of_bucket(int(x[rw-1]) >> (64-2*c["b"]), c["b"], 2) for x in miss])
