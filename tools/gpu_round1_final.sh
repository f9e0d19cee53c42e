#!/bin/bash
# One gpurun call: GPU tests, smoke, small-call latencies (descriptor cache on/off), the bench line.
# Usage: tools/gpu_round1_final.sh <tag>
set -u
TAG=${1:-r1z}
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,memory.total --format=csv > $OUT/smi_$TAG.txt 2>&1
( time python -m pytest tests -m gpu -x -q ) > $OUT/pytest_$TAG.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/pytest_$TAG.log
tail -8 $OUT/pytest_$TAG.log
python __graft_entry__.py smoke > $OUT/smoke_$TAG.log 2>&1; echo "smoke rc=$?"; tail -3 $OUT/smoke_$TAG.log
python tools/small_call_latency.py > $OUT/small_$TAG.jsonl 2> $OUT/small_$TAG.err; echo "small rc=$?"; cat $OUT/small_$TAG.jsonl; tail -3 $OUT/small_$TAG.err
EMB_DESC_CACHE=0 python tools/small_call_latency.py --presets c1 toy rm2 --max-rows 1000000 > $OUT/small_nocache_$TAG.jsonl 2> $OUT/small_nocache_$TAG.err; echo "small(nocache) rc=$?"; cat $OUT/small_nocache_$TAG.jsonl
python bench.py > $OUT/bench_$TAG.json 2> $OUT/bench_$TAG.err; echo "bench rc=$?"; cat $OUT/bench_$TAG.json; tail -5 $OUT/bench_$TAG.err
