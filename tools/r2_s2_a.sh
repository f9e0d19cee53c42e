#!/bin/bash
# Round 2, second session, one-GPU check of the new pieces: emb_lookup_ex_async (two host calls in flight), hot-row
# tables (emb_table_set_hot_rows / emb_table_remap), then the whole GPU suite, a short bench and the hot-row A/B with
# its ncu hit-rate pass.  Usage: tools/r2_s2_a.sh <tag>
set -u
TAG=${1:-s2a}; OUT=gpurun_out; mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,memory.total --format=csv > $OUT/smi_$TAG.txt 2>&1
( time timeout 300 python -m pytest tests/test_gpu_async.py tests/test_gpu_hot.py -m gpu -q ) > $OUT/pytest_new_$TAG.log 2>&1
echo "pytest new rc=$?"; tail -15 $OUT/pytest_new_$TAG.log
( time timeout 400 python bench.py --steps 20 --warmup 5 --extras none --no-cpu-baseline ) > $OUT/bench_quick_$TAG.json 2> $OUT/bench_quick_$TAG.err
echo "bench rc=$?"; tail -3 $OUT/bench_quick_$TAG.err
python - <<PY
import json
try:
    d = json.load(open("$OUT/bench_quick_$TAG.json")); r = d["roofline"]; e = d["e2e"]
    print("  value", round(d["value"] / 1e9, 2), "ms/step", round(d["ms_per_step"], 4), "kernel", round(r["kernel_ms"], 4), "frac", round(r["frac"], 3),
          "parity", d["parity"]["bit_exact"])
    print("  e2e", round(e["ms_per_step"], 3), "in flight", e["calls_in_flight"], "async", round(e["two_calls_in_flight"]["ms_per_step"], 3),
          "sync", round(e["synchronous_call"]["ms_per_step"], 3), "floor", round(e["floor_ms"], 3))
except Exception as ex:
    print("  no json", ex)
PY
HOT=$OUT/l2_hot_$TAG.txt; : > $HOT
run() { echo "## $*" >> $HOT; timeout 420 "$@" >> $HOT 2>&1; echo "rc=$?" >> $HOT; }
run python tools/l2_hot_ab.py --tag big_dim128 --steps 30
grep -E "^\||^###|rc=|Error|error" $HOT | cut -c1-200
( time timeout 600 python -m pytest tests -m gpu -q --deselect tests/test_gpu_async.py --deselect tests/test_gpu_hot.py ) > $OUT/pytest_all_$TAG.log 2>&1
echo "pytest all rc=$?"; tail -8 $OUT/pytest_all_$TAG.log
run python tools/l2_hot_ab.py --tag c5_as_benched --rows 500000 --arms quick --steps 30
run python tools/l2_hot_ab.py --tag narrow_dim16 --tables 26 --rows 4000000 --dim 16 --batch 65536 --pool 10 --arms quick --steps 30
grep -E "^\||^###|rc=|Error|error" $HOT | cut -c1-200
timeout 420 ncu --metrics lts__t_sector_hit_rate.pct,dram__bytes_read.sum,gpu__time_duration.sum --clock-control none \
    -k regex:bag_sum_stretch --csv --log-file $OUT/l2_hot_ncu_$TAG.csv \
    python tools/l2_hot_ab.py --tag ncu --arms quick --hot-mb 64 --steps 2 > $OUT/l2_hot_ncu_$TAG.log 2>&1
echo "ncu rc=$?"; tail -12 $OUT/l2_hot_ncu_$TAG.log | cut -c1-200
