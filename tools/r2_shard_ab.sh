#!/bin/bash
# One GPU, c4tb with forced row cuts (every part on this GPU: each cut table's index list is read once per part):
# bags per CTA of the compacting shard kernel, against the scan-everything path and the uncut placement.
TAG=${1:-sh}; mkdir -p gpurun_out; : > gpurun_out/shard_ab_$TAG.txt
run() { local name=$1 envs=$2; shift 2
  env $envs timeout 600 python bench.py --gpus 1 --workload c4tb --steps 30 --warmup 5 --no-e2e --no-cpu-baseline "$@" > gpurun_out/shard_$TAG.json 2> gpurun_out/shard_$TAG.err
  python - "$name" <<PY >> gpurun_out/shard_ab_$TAG.txt
import json, sys
try:
    d = json.load(open("gpurun_out/shard_$TAG.json")); r = d["roofline"]
    print("%-28s %.2f G/s  ms/step %.4f  kernel %.4f  parity %s  split %s" % (sys.argv[1], d["value"] / 1e9, d["ms_per_step"], r["kernel_ms"],
          d["parity"].get("bit_exact"), d["run"]["placement"].get("row_split_tables")))
except Exception as e:
    print(sys.argv[1], "no json", e)
PY
}
run uncut "EMB_X=0"
for m in 1 2 4 8; do run "rowcut compact mult=$m" "EMB_SHARD_MULT=$m" --max-table-bytes 3000000000; done
run "rowcut scan" "EMB_SHARD_COMPACT=0" --max-table-bytes 3000000000
run "rowcut compact mult=4 1.3e9" "EMB_SHARD_MULT=4" --max-table-bytes 1300000000
run "rowcut scan 1.3e9" "EMB_SHARD_COMPACT=0" --max-table-bytes 1300000000
cat gpurun_out/shard_ab_$TAG.txt
