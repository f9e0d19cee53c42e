#!/bin/bash
# One gpurun call on one GPU: GPU tests, smoke, both bench arms as the driver runs them, then the two ncu passes of
# the recipe (/opt/skills/guides/B200_PROFILING.md).  Usage: tools/gpu_check.sh <tag> [bench args...]
set -u
TAG=${1:-r2}; shift || true
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,memory.total --format=csv > $OUT/smi_$TAG.txt 2>&1
( time timeout 1200 python -m pytest tests -m gpu -x -q ) > $OUT/pytest_$TAG.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/pytest_$TAG.log
tail -8 $OUT/pytest_$TAG.log
python __graft_entry__.py smoke > $OUT/smoke_$TAG.log 2>&1; echo "smoke rc=$?"; tail -3 $OUT/smoke_$TAG.log
( time python bench.py --steps 20 --warmup 5 "$@" ) > $OUT/bench_$TAG.json 2> $OUT/bench_$TAG.err; echo "bench rc=$?"; cut -c1-1500 $OUT/bench_$TAG.json; tail -5 $OUT/bench_$TAG.err
( time python bench.py --impl reference --steps 20 --warmup 5 "$@" ) > $OUT/bench_ref_$TAG.json 2> $OUT/bench_ref_$TAG.err; echo "reference rc=$?"; cut -c1-600 $OUT/bench_ref_$TAG.json; tail -3 $OUT/bench_ref_$TAG.err
( time python bench.py --steps 20 --warmup 5 --dtype f32 --no-cpu-baseline --extras none ) > $OUT/bench_f32_$TAG.json 2> $OUT/bench_f32_$TAG.err; echo "bench f32 rc=$?"; cut -c1-400 $OUT/bench_f32_$TAG.json
SHORT="python bench.py --steps 6 --warmup 3 --no-e2e --no-cpu-baseline --no-parity --extras none $*"
$SHORT > $OUT/plain_$TAG.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches_$TAG.csv $SHORT > $OUT/ncu_launch_$TAG.log 2>&1
echo "ncu launches rc=$?"
ncu --set full --clock-control none --import-source on -k regex:bag_sum -s 3 -c 2 -f -o $OUT/prof_$TAG $SHORT > $OUT/ncu_full_$TAG.log 2>&1
echo "ncu full rc=$?"; tail -3 $OUT/ncu_full_$TAG.log
