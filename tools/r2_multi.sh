#!/bin/bash
# Round-2 multi-GPU check (under gpurun --gpus N): tools/r2_multi.sh N tag
#   1. tests/test_gpu_multi.py (peer-slab exchange, ShardedTables at world 1..N, batches in flight)
#   2. bench.py exactly as the driver launches it at N ranks (C2 + the c4tb / c5 extras, parity on every rank)
N=${1:-2}; TAG=${2:-r2}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -x -q > gpurun_out/pytest_multi_n${N}_$TAG.log 2>&1
echo "pytest multi rc=$?"; tail -4 gpurun_out/pytest_multi_n${N}_$TAG.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
t0=$(date +%s)
timeout 900 $TR bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_n${N}_$TAG.json 2> gpurun_out/bench_n${N}_$TAG.err
echo "bench N=$N rc=$? wall=$(( $(date +%s) - t0 ))s"
python - <<PY
import json
try:
    d = json.load(open("gpurun_out/bench_n${N}_$TAG.json"))
    r = d["roofline"]
    print("c2", round(d["value"] / 1e9, 2), "G/s", round(d["ms_per_step"], 4), "ms/step, kernel", round(r["kernel_ms"], 4),
          "frac", round(r["frac"], 3), "parity", d.get("parity", {}).get("bit_exact"), "e2e ms", d["e2e"].get("ms_per_step"))
    for k, x in (d.get("extra") or {}).items():
        if isinstance(x, dict) and "value" in x:
            print(k, round(x["value"] / 1e9, 2), "G/s", round(x["ms_per_step"], 4), "ms/step parity",
                  x.get("parity", {}).get("bit_exact"), "e2e", (x.get("e2e") or {}).get("ms_per_step"))
        else:
            print(k, str(x)[:300])
except Exception as e:
    print("no json:", e)
PY
grep -i -E "error|Traceback" -A8 gpurun_out/bench_n${N}_$TAG.err | head -40
# comparator: the same steps without launch overlap (EMB_Q_OVERLAP off), C2 only
timeout 300 $TR bench.py --gpus $N --steps 20 --warmup 5 --no-overlap --no-e2e --no-cpu-baseline --extras none \
    > gpurun_out/bench_n${N}_noov_$TAG.json 2> gpurun_out/bench_n${N}_noov_$TAG.err
python - <<PY
import json
try:
    d = json.load(open("gpurun_out/bench_n${N}_noov_$TAG.json")); r = d["roofline"]
    print("c2 --no-overlap", round(d["value"] / 1e9, 2), "G/s", round(d["ms_per_step"], 4), "ms/step, kernel", round(r["kernel_ms"], 4))
except Exception as e:
    print("no json:", e)
PY
