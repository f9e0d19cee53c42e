#!/bin/bash
# c4tb (Terabyte cardinalities, 48 GB, global batch 65536) at N ranks: the planner's own placement against forced
# row cuts (tables above --max-table-bytes are cut by rows), the latter with and without the compacting shard kernel.
#   tools/r2_c4tb.sh N tag      (under gpurun --gpus N)
N=${1:-2}; TAG=${2:-c4}
mkdir -p gpurun_out
if [ $N -eq 1 ]; then TR="python"; else
  TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29523"; fi
run() {  # name, env, args...
  local name=$1 envs=$2; shift 2
  env $envs timeout 600 $TR bench.py --gpus $N --workload c4tb --steps 50 --warmup 10 --no-e2e --no-cpu-baseline "$@" \
      > gpurun_out/c4tb_n${N}_${name}_$TAG.json 2> gpurun_out/c4tb_n${N}_${name}_$TAG.err
  echo "$name rc=$?"
  python - <<PY
import json
try:
    d = json.load(open("gpurun_out/c4tb_n${N}_${name}_$TAG.json")); r = d["roofline"]
    print("  %.2f G/s  ms/step %.4f  kernel %.4f (max over ranks %.4f)  parity %s  split %s  sliced %s" % (
        d["value"] / 1e9, d["ms_per_step"], r["kernel_ms"], r["kernel_ms_max_over_ranks"], d["parity"].get("bit_exact"),
        d["run"]["placement"].get("row_split_tables"),
        d["run"]["placement"].get("bag_sliced_replicas")))
except Exception as e:
    print("  no json", e)
PY
  grep -i -E "error|Traceback" -A6 gpurun_out/c4tb_n${N}_${name}_$TAG.err | head -12
}
if [ "${QUICK:-0}" = 1 ]; then  # the one line the planner's own placement (bench.py's c4tb extra) does not give
  run rowcut2_compact "EMB_SHARD_COMPACT=1" --max-table-bytes 6000000000
  exit 0
fi
run planner "EMB_X=0"
run rowcut_compact "EMB_SHARD_COMPACT=1" --max-table-bytes 3000000000
run rowcut_scan "EMB_SHARD_COMPACT=0" --max-table-bytes 3000000000
run rowcut2_compact "EMB_SHARD_COMPACT=1" --max-table-bytes 6000000000
run rowcut2_scan "EMB_SHARD_COMPACT=0" --max-table-bytes 6000000000
