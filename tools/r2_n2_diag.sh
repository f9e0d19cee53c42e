#!/bin/bash
# Where does the pooling launch lose time at N ranks?  (under gpurun --gpus N): tools/r2_n2_diag.sh N tag
N=${1:-2}; TAG=${2:-d}
mkdir -p gpurun_out
timeout 300 python tools/route_probe.py --ranks 8 > gpurun_out/route_probe_$TAG.txt 2>&1; echo "route_probe rc=$?"; cat gpurun_out/route_probe_$TAG.txt | tail -4
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29519"
for mode in "--layout sample" "--layout table" "--exchange nccl" "--layout sample --no-rotate"; do
  name=$(echo $mode | tr -d ' -')
  timeout 300 $TR bench.py --gpus $N --steps 50 --warmup 10 --no-e2e --no-cpu-baseline --extras none --no-parity $mode \
      > gpurun_out/diag_n${N}_${name}_$TAG.json 2> gpurun_out/diag_n${N}_${name}_$TAG.err
  echo "$mode rc=$?"
  python - <<PY
import json
try:
    d = json.load(open("gpurun_out/diag_n${N}_${name}_$TAG.json")); r = d["roofline"]
    print("  ms/step", round(d["ms_per_step"], 4), "kernel", r["kernel_ms_dist"])
except Exception as e:
    print("  no json", e)
PY
done
