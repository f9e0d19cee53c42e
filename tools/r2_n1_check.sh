#!/bin/bash
# One-GPU round-2 check: tools/r2_n1_check.sh tag   (GPU tests, bench with / without launch overlap, host-path sweep)
TAG=${1:-r2}
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_$TAG.log 2>&1
echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu_$TAG.log
for mode in "" "--no-overlap"; do
  name=bench${mode//-/}
  timeout 600 python bench.py --steps 20 --warmup 5 $mode > gpurun_out/${name}_$TAG.json 2> gpurun_out/${name}_$TAG.err
  echo "bench $mode rc=$?"
  python - <<PY
import json
try:
    d = json.load(open("gpurun_out/${name}_$TAG.json")); r = d["roofline"]
    print("  value", round(d["value"] / 1e9, 2), "ms/step", round(d["ms_per_step"], 4), "kernel", round(r["kernel_ms"], 4), "frac", round(r["frac"], 3),
          "e2e", round(d["e2e"]["ms_per_step"], 3), "floor", round(d["e2e"]["floor_ms"], 3), "parity", d["parity"]["bit_exact"])
    for k, x in (d.get("extra") or {}).items():
        if isinstance(x, dict) and "value" in x:
            print("  ", k, round(x["value"] / 1e9, 2), "G/s", round(x["ms_per_step"], 4), "parity", x.get("parity", {}).get("bit_exact"),
                  "e2e", (x.get("e2e") or {}).get("ms_per_step"), "floor", (x.get("e2e") or {}).get("floor_ms"))
except Exception as e:
    print("  no json", e)
PY
done
for wl in c2 c3 c5; do
  timeout 300 python tools/e2e_sweep.py --quick --workload $wl >> gpurun_out/e2e_quick_$TAG.txt 2>&1; echo "e2e sweep $wl rc=$?"
done
cat gpurun_out/e2e_quick_$TAG.txt | cut -c1-230
