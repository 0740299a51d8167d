#!/bin/bash
# full GPU suite, smoke and the default bench line of the tree as it is
mkdir -p gpurun_out
(time python -m pytest tests -m gpu -x -q) > gpurun_out/gpu_all3.log 2>&1; tail -5 gpurun_out/gpu_all3.log
python __graft_entry__.py smoke > gpurun_out/smoke3.log 2>&1; tail -2 gpurun_out/smoke3.log | cut -c1-200
python bench.py > gpurun_out/final3_n1.json 2> gpurun_out/final3_n1.err
python - <<PY
import json
d=json.load(open("gpurun_out/final3_n1.json"))
print(d["value"]/1e9, d["ms_per_step"], d["e2e"]["value"]/1e9, d["e2e"]["ms_per_step"], d["parity_check"]["ok"], d["gpu_launches"], d["roofline"]["frac"], d["roofline"]["traffic"], d["clocks"])
PY
