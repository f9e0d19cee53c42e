#!/bin/bash
# C2 weak-scaling curve on ONE box (under gpurun --gpus NMAX): tools/r2_scale.sh NMAX tag [extra bench args]
# N = 1, 2, 4, ... NMAX back to back, kernel only (no e2e / cpu baseline / extras), as the driver's SCALE run does.
NMAX=${1:-2}; TAG=${2:-s}; shift 2
mkdir -p gpurun_out
for N in 1 2 4 8; do
  [ $N -gt $NMAX ] && break
  if [ $N -eq 1 ]; then RUN="python"; else
    RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521"; fi
  timeout 600 $RUN bench.py --gpus $N --steps 50 --warmup 10 --no-e2e --no-cpu-baseline --extras none "$@" \
      > gpurun_out/scale_n${N}_$TAG.json 2> gpurun_out/scale_n${N}_$TAG.err
  echo "N=$N rc=$?"
done
python - <<PY
import json
v1 = None
for n in (1, 2, 4, 8):
    try:
        d = json.load(open("gpurun_out/scale_n%d_$TAG.json" % n))
    except Exception:
        continue
    r = d["roofline"]
    v1 = v1 or d["value"]
    print("N=%d  %.2f G/s  eff %.3f  ms/step %.4f  kernel %.4f (max over ranks %.4f)  parity %s" % (
        n, d["value"] / 1e9, d["value"] / (n * v1), d["ms_per_step"], r["kernel_ms"], r["kernel_ms_max_over_ranks"],
        d["parity"].get("bit_exact")))
PY
