#!/bin/bash
# multi-GPU evidence on one lease of N GPUs: real-NCCL parity worker, concurrent H2D ceiling, bench lines.  usage: scripts/mgpu_round.sh N tag [bench args...]
N=$1; TAG=$2; shift 2
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
$TR --master-port 29631 tests/mgpu_worker.py > gpurun_out/mgpu_${TAG}_n$N.log 2>&1; tail -3 gpurun_out/mgpu_${TAG}_n$N.log
$TR --master-port 29633 scripts/h2d_bw.py > gpurun_out/h2d_${TAG}_n$N.log 2>&1; tail -2 gpurun_out/h2d_${TAG}_n$N.log
for MODE in weak weak-fixed-table strong; do
  $TR --master-port 29632 bench.py --gpus $N --steps 8 --warmup 3 --scaling $MODE "$@" > gpurun_out/bench_${TAG}_n${N}_$MODE.json 2> gpurun_out/bench_${TAG}_n${N}_$MODE.err
  python - <<PY
import json
try:
    d = json.load(open("gpurun_out/bench_${TAG}_n${N}_$MODE.json"))
    print("$MODE", "value %.2f G occ/s" % (d["value"] / 1e9), "ms/step %.3f" % d["ms_per_step"], "e2e %.2f" % (d["e2e"]["value"] / 1e9), {k: round(v, 3) for k, v in d["stage_ms_per_step"].items()}, "parity", d["parity_check"].get("ok"))
except Exception as e:
    print("$MODE failed", e); print(open("gpurun_out/bench_${TAG}_n${N}_$MODE.err").read()[-1500:])
PY
done
