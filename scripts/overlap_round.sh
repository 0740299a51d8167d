#!/bin/bash
# A/B of the host-input overlap (run_reads_overlapped) under the development switches
mkdir -p gpurun_out
while read -r c g t; do
  TAG="count=$c get=$g ctas=$t" BRISK_OVERLAP_COUNT=$c BRISK_OVERLAP_GET=$g BRISK_OVERLAP_CTAS=$t python scripts/e2e_timeline.py 2>&1 | tail -1
done <<'LIST' | tee gpurun_out/ov_timeline4.txt
3 3 2
4 3 2
5 3 2
2,5 3 2
3,6 3 2
3 2 2
3 4 2
3 2,4 2
4 4 2
4 2 2
3,6 2 3
3 3 3
LIST
