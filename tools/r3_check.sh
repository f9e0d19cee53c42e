#!/bin/bash
# Round-3 one-GPU check: GPU tests, smoke, both bench arms as the driver runs them, ncu launch list + one full capture
# of the C2 kernel and of the C3 (short-bag) kernel.  Usage: tools/r3_check.sh <tag>   (under gpurun)
set -u
TAG=${1:-r3}; OUT=gpurun_out; mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,memory.total --format=csv > $OUT/smi_$TAG.txt 2>&1
( time timeout 1200 python -m pytest tests -m gpu -x -q -rs ) > $OUT/pytest_$TAG.log 2>&1; echo "pytest rc=$?"; tail -9 $OUT/pytest_$TAG.log
python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke_$TAG.log 2>&1; echo "smoke rc=$?"; tail -2 $OUT/smoke_$TAG.log
( time python bench.py --steps 20 --warmup 5 ) > $OUT/bench_$TAG.json 2> $OUT/bench_$TAG.err; echo "bench rc=$?"; tail -4 $OUT/bench_$TAG.err
( time python bench.py --impl reference --steps 20 --warmup 5 ) > $OUT/bench_ref_$TAG.json 2> $OUT/bench_ref_$TAG.err; echo "reference rc=$?"; cut -c1-300 $OUT/bench_ref_$TAG.json
SHORT="python bench.py --steps 6 --warmup 3 --no-e2e --no-cpu-baseline --no-parity --extras none"
$SHORT > $OUT/plain_$TAG.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches_$TAG.csv $SHORT > $OUT/ncu_launch_$TAG.log 2>&1
echo "ncu launches rc=$?"
ncu --set full --clock-control none --import-source on -k regex:bag_sum -s 3 -c 2 -f -o $OUT/prof_c2_$TAG $SHORT > $OUT/ncu_full_c2_$TAG.log 2>&1
echo "ncu full c2 rc=$?"
ncu --set full --clock-control none --import-source on -k regex:bag_sum -s 3 -c 2 -f -o $OUT/prof_c3_$TAG $SHORT --workload c3 > $OUT/ncu_full_c3_$TAG.log 2>&1
echo "ncu full c3 rc=$?"
