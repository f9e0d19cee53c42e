#!/bin/bash
# Does the parity pass or the step count change the timed kernel at N ranks?  tools/r2_n2_diag2.sh N tag
N=${1:-2}; TAG=${2:-d}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29519"
i=0
for mode in "--steps 20 --warmup 5" "--steps 20 --warmup 5 --no-parity" "--steps 50 --warmup 10" "--steps 50 --warmup 10 --no-parity" "--steps 20 --warmup 5"; do
  i=$((i+1))
  timeout 300 $TR bench.py --gpus $N --no-e2e --no-cpu-baseline --extras none $mode \
      > gpurun_out/diag2_n${N}_${i}_$TAG.json 2> gpurun_out/diag2_n${N}_${i}_$TAG.err
  echo "$mode rc=$?"
  python - <<PY
import json
try:
    d = json.load(open("gpurun_out/diag2_n${N}_${i}_$TAG.json")); r = d["roofline"]
    print("  ms/step", round(d["ms_per_step"], 4), "kernel", r["kernel_ms_dist"], d["clocks"])
except Exception as e:
    print("  no json", e)
PY
done
