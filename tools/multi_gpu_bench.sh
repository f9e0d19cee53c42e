#!/bin/bash
# bench.py at N ranks for the three multi-GPU workloads (under gpurun --gpus N): tools/multi_gpu_bench.sh N [tag]
N=${1:-2}; TAG=${2:-r1}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
mkdir -p gpurun_out
for wl in c4tb c2 c5; do
  timeout 600 $TR bench.py --gpus $N --workload $wl --steps 100 --warmup 10 --no-cpu-baseline \
      > gpurun_out/${wl}_n${N}_$TAG.json 2> gpurun_out/${wl}_n${N}_$TAG.err
  echo "$wl N=$N rc=$?"; cat gpurun_out/${wl}_n${N}_$TAG.json | cut -c1-1800; grep -i -E "error|Traceback" -A5 gpurun_out/${wl}_n${N}_$TAG.err | head -20
done
timeout 600 $TR bench.py --gpus $N --workload c2 --exchange nccl --steps 100 --warmup 10 --no-cpu-baseline --no-e2e \
    > gpurun_out/c2nccl_n${N}_$TAG.json 2> gpurun_out/c2nccl_n${N}_$TAG.err
echo "c2 nccl N=$N rc=$?"; cut -c1-400 gpurun_out/c2nccl_n${N}_$TAG.json
