#!/bin/bash
# N-rank record run: c4tb (both slab layouts), c2 (rotation on/off, NCCL comparator), c5.  tools/multi_gpu_final.sh N tag
N=${1:-8}; TAG=${2:-fin}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
run() { name=$1; shift; timeout 300 $TR bench.py --gpus $N --steps 100 --warmup 10 --no-cpu-baseline "$@" \
          > gpurun_out/${name}_n${N}_$TAG.json 2> gpurun_out/${name}_n${N}_$TAG.err; echo "$name rc=$?";
        python - <<PY
import json
try:
    d=json.load(open("gpurun_out/${name}_n${N}_$TAG.json")); r=d["roofline"]
    print("   ", round(d["value"]/1e9,2),"G/s", round(d["ms_per_step"],4),"ms/step kernel",round(r["kernel_ms"],4), r.get("kernel_ms_max_over_ranks"), "e2e", round(d.get("e2e",{}).get("value",0)/1e9,2))
except Exception as e: print("no json", e)
PY
}
run c4tb --workload c4tb
run c4tb_table --workload c4tb --layout table --no-e2e
run c2 --workload c2
run c2_norot --workload c2 --no-rotate --no-e2e
run c2_nccl --workload c2 --exchange nccl --no-e2e
run c5 --workload c5 --no-e2e
