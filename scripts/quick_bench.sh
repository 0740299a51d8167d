#!/bin/bash
# one short bench line (no CPU leg, no independent index) + the stage times
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py tests/test_gpu_long_reads.py tests/test_gpu_overlap.py -x -q 2>&1 | tail -2
python bench.py --no-cpu-baseline --no-parity > gpurun_out/quick.json 2>gpurun_out/quick.err
python - <<PY
import json
d=json.load(open("gpurun_out/quick.json"))
print("value %.2f G  %.3f ms | e2e %.2f G %.3f ms" % (d["value"]/1e9, d["ms_per_step"], d["e2e"]["value"]/1e9, d["e2e"]["ms_per_step"]))
print({k: round(v, 3) for k, v in d["stage_ms_per_step"].items()})
PY
