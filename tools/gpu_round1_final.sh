#!/bin/bash
# One gpurun call: GPU tests, smoke, small-call latencies, descriptor-search A/B on every workload
# shape, the bench line.   Usage: tools/gpu_round1_final.sh <tag>
set -u
TAG=${1:-r1z}
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,memory.total --format=csv > $OUT/smi_$TAG.txt 2>&1
( time python -m pytest tests -m gpu -x -q ) > $OUT/pytest_$TAG.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/pytest_$TAG.log
tail -8 $OUT/pytest_$TAG.log
python __graft_entry__.py smoke > $OUT/smoke_$TAG.log 2>&1; echo "smoke rc=$?"; tail -3 $OUT/smoke_$TAG.log
python tools/small_call_latency.py > $OUT/small_$TAG.jsonl 2> $OUT/small_$TAG.err; echo "small rc=$?"; cat $OUT/small_$TAG.jsonl; tail -3 $OUT/small_$TAG.err
: > $OUT/uni_$TAG.txt
for w in c2 c3 c4 c5; do
  for u in 0 1 0 1; do
    echo "workload=$w EMB_UNIFORM_CHUNKS=$u" >> $OUT/uni_$TAG.txt
    EMB_UNIFORM_CHUNKS=$u python tools/sweep.py --workload $w --steps 60 --tunings 0,0,0 >> $OUT/uni_$TAG.txt 2>&1
  done
done
cat $OUT/uni_$TAG.txt | cut -c1-260
python bench.py > $OUT/bench_$TAG.json 2> $OUT/bench_$TAG.err; echo "bench rc=$?"; cat $OUT/bench_$TAG.json; tail -5 $OUT/bench_$TAG.err
