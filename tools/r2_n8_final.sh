#!/bin/bash
# 8-GPU confirmation: same-box scaling curve (kernel only), then bench.py exactly as the driver launches it.
TAG=${1:-f8}
bash tools/r2_scale.sh 8 $TAG
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29517"
timeout 900 $TR bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/bench_n8_$TAG.json 2> gpurun_out/bench_n8_$TAG.err
echo "bench N=8 rc=$?"
python - <<PY
import json
d = json.load(open("gpurun_out/bench_n8_$TAG.json")); r = d["roofline"]
print("c2", round(d["value"] / 1e9, 2), "G/s", round(d["ms_per_step"], 4), "ms/step, kernel", round(r["kernel_ms"], 4),
      "parity", d["parity"].get("bit_exact"), "e2e ms", d["e2e"].get("ms_per_step"), "floor", d["e2e"].get("floor_ms"))
for k, x in (d.get("extra") or {}).items():
    if isinstance(x, dict) and "value" in x:
        print(k, round(x["value"] / 1e9, 2), "G/s", round(x["ms_per_step"], 4), "ms/step parity", x.get("parity", {}).get("bit_exact"),
              "e2e", (x.get("e2e") or {}).get("ms_per_step"), "floor", (x.get("e2e") or {}).get("floor_ms"))
PY
grep -i -E "error|Traceback" -A8 gpurun_out/bench_n8_$TAG.err | head -20
