#!/bin/bash
# Why is the C2 pooling launch slow at N ranks?  One variation per line, kernel only.  tools/r2_n8_diag.sh N tag
N=${1:-8}; TAG=${2:-d8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
nvidia-smi topo -m > gpurun_out/topo_$TAG.txt 2>&1
nvidia-smi nvlink -s > gpurun_out/nvlink_$TAG.txt 2>&1
i=0
run() { i=$((i+1)); local envs=$1; shift
  env $envs timeout 200 $TR bench.py --gpus $N --steps 30 --warmup 5 --no-e2e --no-cpu-baseline --extras none --no-parity "$@" \
      > gpurun_out/diag8_${i}_$TAG.json 2> gpurun_out/diag8_${i}_$TAG.err
  python - "$envs $*" <<PY
import json, sys
try:
    d = json.load(open("gpurun_out/diag8_${i}_$TAG.json")); r = d["roofline"]
    print("%-44s ms/step %.4f  kernel %.4f (max %.4f) p10 %.4f p90 %.4f" % (sys.argv[1], d["ms_per_step"], r["kernel_ms"],
          r["kernel_ms_max_over_ranks"], r["kernel_ms_dist"]["p10"], r["kernel_ms_dist"]["p90"]))
except Exception as e:
    print(sys.argv[1], "no json", e)
PY
}
run "EMB_X=0"
run "EMB_PDL=0"
run "EMB_X=0" --layout table
run "EMB_X=0" --exchange nccl
run "EMB_X=0" --variant 1
run "EMB_X=0" --no-rotate
run "EMB_X=0" --workload c4
