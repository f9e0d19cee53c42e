#!/bin/bash
# Driver-style bench lines at several N on one box (default extras: c4tb, c4rc, c5, c3 at N>1).  Usage: tools/r3_scale.sh <tag> N...
TAG=$1; shift; OUT=gpurun_out; mkdir -p $OUT
for N in "$@"; do
  if [ "$N" = 1 ]; then CMD="python bench.py --steps 20 --warmup 5"
  else CMD="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29600+N)) bench.py --gpus $N --steps 20 --warmup 5"; fi
  ( time $CMD ) > $OUT/bench_${TAG}_n$N.json 2> $OUT/bench_${TAG}_n$N.err; echo "N=$N rc=$?"
  python - <<PY
import json
try:
    r = json.load(open("$OUT/bench_${TAG}_n$N.json"))
    print("N=$N value %.4g ms %.4f parity %s" % (r["value"], r["ms_per_step"], r["parity"].get("bit_exact")))
    for k, v in r.get("extra", {}).items():
        if "value" in v:
            print("  ", k, "value %.4g ms %.4f" % (v["value"], v["ms_per_step"]), "bit_exact", v["parity"].get("bit_exact"),
                  "split", v.get("placement", {}).get("row_split_tables"), "per-rank ms", v["roofline"].get("kernel_ms_per_rank"))
except Exception as e:
    print("N=$N unreadable:", e)
PY
  tail -3 $OUT/bench_${TAG}_n$N.err
done
