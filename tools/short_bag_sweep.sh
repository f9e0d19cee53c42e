#!/bin/bash
# C3 lookup shape (one index per bag): the short-bag kernel's knobs against the stretch kernel, same box, same run.
# Usage: tools/short_bag_sweep.sh <tag>   (under gpurun, one GPU)
TAG=${1:-r3}; OUT=gpurun_out; mkdir -p $OUT
B="python bench.py --workload c3 --steps 100 --warmup 10 --no-cpu-baseline --extras none --no-e2e"
run() { # name, env...
  local name=$1; shift
  env "$@" $B > $OUT/c3_${name}_$TAG.json 2> $OUT/c3_${name}_$TAG.err || echo "$name FAILED"
}
run stretch EMB_SHORT_BAGS=0
for tpw in 1 2 4; do run short_t${tpw} EMB_SHORT_TPW=$tpw; done
run stretch2 EMB_SHORT_BAGS=0
python - <<PY
import json, glob
for f in sorted(glob.glob("$OUT/c3_*_$TAG.json")):
    try:
        r = json.load(open(f)); ro = r["roofline"]; k = ro["kernel_ms_dist"]
        print(f.split("/")[-1], "step_ms %.4f" % r["ms_per_step"], "kernel_ms mean %.4f median %.4f p10 %.4f" % (k["mean"], k["median"], k["p10"]),
              "frac %.3f" % ro["frac"], "bit_exact", r["parity"].get("bit_exact"))
    except Exception as e:
        print(f, "unreadable", e)
PY
