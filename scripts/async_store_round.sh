#!/bin/bash
# A/B of BRISK_ASYNC_STORE (DESIGN.md "measured and not enabled"): the count call's store kernel on the side stream beside
# the next call's enumeration, device-resident input (scripts/dev_timeline.py); residency cap of that kernel in CTAs per SM
mkdir -p gpurun_out
while read -r a c m; do
  TAG="async=$a ctas=$c" MODE=$m BRISK_ASYNC_STORE=$a BRISK_ASYNC_CTAS=$c python scripts/dev_timeline.py 2>&1 | tail -1
done <<'LIST' | tee gpurun_out/async_store.txt
0 0 count+query
1 0 count+query
1 1 count+query
1 2 count+query
1 3 count+query
0 0 count
1 0 count
1 2 count
1 3 count
LIST
