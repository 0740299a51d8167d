"""brisk_b200 — B200-native k-mer insertion path of Brisk (hand-written sm_100a CUDA behind a C ABI).

The product is libbrisk_b200.so (brisk_b200/csrc, include/brisk_b200.h).  This package is the Python host
mirror of the reference's interface for that path: Parameters(k, m, b) and Brisk (insert / get /
count over reads and k-mers, dump, stats), plus the synthetic read generator of the benchmark configs.
Nothing here computes on the CPU: every call goes through the C ABI and fails loudly when the CUDA
library is missing.
"""
from .brisk import (Brisk, BriskError, Parameters, bucket_ids, decode_records, exchange_plan, lib_path,  # noqa: F401
                    merge_ordered_dumps, owner_of_bucket)
