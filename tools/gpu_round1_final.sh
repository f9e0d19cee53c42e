#!/bin/bash
# One gpurun call: GPU tests, smoke, the default kernel on every workload shape, small-call latencies, the bench line, the
# two ncu passes of the recipe (/opt/skills/guides/B200_PROFILING.md).  Usage: tools/gpu_round1_final.sh <tag>
set -u
TAG=${1:-r1z}
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,memory.total --format=csv > $OUT/smi_$TAG.txt 2>&1
( time python -m pytest tests -m gpu -x -q ) > $OUT/pytest_$TAG.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/pytest_$TAG.log
tail -6 $OUT/pytest_$TAG.log
python __graft_entry__.py smoke > $OUT/smoke_$TAG.log 2>&1; echo "smoke rc=$?"; tail -2 $OUT/smoke_$TAG.log
: > $OUT/ab_$TAG.txt
for wl in c2 c3 c4 c5; do
  echo "workload=$wl default" >> $OUT/ab_$TAG.txt
  python tools/sweep.py --workload $wl --steps 60 --tunings 0,0,0 >> $OUT/ab_$TAG.txt 2>&1
done
cut -c1-200 $OUT/ab_$TAG.txt
python tools/small_call_latency.py > $OUT/small_$TAG.jsonl 2> $OUT/small_$TAG.err; echo "small rc=$?"; cut -c1-400 $OUT/small_$TAG.jsonl
python bench.py > $OUT/bench_$TAG.json 2> $OUT/bench_$TAG.err; echo "bench rc=$?"; cat $OUT/bench_$TAG.json; tail -5 $OUT/bench_$TAG.err
SHORT="python bench.py --steps 6 --warmup 3 --no-e2e --no-cpu-baseline"
$SHORT > $OUT/plain_$TAG.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches_$TAG.csv $SHORT > $OUT/ncu_launch_$TAG.log 2>&1
echo "ncu launches rc=$?"
ncu --set full --clock-control none --import-source on -k regex:bag_sum -s 3 -c 2 -f -o $OUT/prof_$TAG $SHORT > $OUT/ncu_full_$TAG.log 2>&1
echo "ncu full rc=$?"; tail -3 $OUT/ncu_full_$TAG.log
