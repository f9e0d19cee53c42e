#!/bin/bash
# quick N-rank check of the two exchange-heavy workloads: tools/multi_gpu_quick.sh N tag [extra bench args]
N=${1:-2}; TAG=${2:-q}; shift 2
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
for wl in ${WL:-c4tb c2}; do
  timeout 300 $TR bench.py --gpus $N --workload $wl --steps 100 --warmup 10 --no-cpu-baseline --no-e2e "$@" \
      > gpurun_out/${wl}_n${N}_$TAG.json 2> gpurun_out/${wl}_n${N}_$TAG.err
  echo "$wl N=$N rc=$?"; python - <<PY
import json
try:
    d=json.load(open("gpurun_out/${wl}_n${N}_$TAG.json")); r=d["roofline"]
    print(round(d["value"]/1e9,2),"G/s", round(d["ms_per_step"],4),"ms/step kernel",round(r["kernel_ms"],4), r.get("kernel_ms_max_over_ranks"), d["config"].get("placement"))
except Exception as e: print("no json", e)
PY
  grep -i -E "error|Traceback" -A5 gpurun_out/${wl}_n${N}_$TAG.err | head -10
done
