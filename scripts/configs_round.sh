#!/bin/bash
# one bench line per synthetic configuration of BASELINE.json on one GPU (profiles/r2_configs.md).  usage: scripts/configs_round.sh tag
TAG=${1:-x}
for CFG in cfg2 cfg4 cfg5 cfg3; do
  python bench.py --config $CFG --steps 8 --warmup 3 > gpurun_out/cfg_${TAG}_$CFG.json 2> gpurun_out/cfg_${TAG}_$CFG.err
  python - <<PY
import json
try:
    d = json.load(open("gpurun_out/cfg_${TAG}_$CFG.json"))
    cb = d.get("cpu_baseline", {})
    print("$CFG", "value %.2f G occ/s" % (d["value"] / 1e9), "ms/step %.3f" % d["ms_per_step"], "e2e %.2f" % (d["e2e"]["value"] / 1e9),
          {k: round(v, 3) for k, v in d["stage_ms_per_step"].items()}, "roofline", d["roofline"]["kernel"][:30], round(d["roofline"]["frac"], 4),
          "parity", d["parity_check"].get("ok"), "cpu %.1f M" % (cb.get("value", 0) / 1e6), cb.get("counter_app", {}).get("t1", {}).get("occurrences_per_s"))
except Exception as e:
    print("$CFG failed", e); print(open("gpurun_out/cfg_${TAG}_$CFG.err").read()[-1500:])
PY
done
