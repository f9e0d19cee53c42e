#!/bin/bash
# A/B of one environment knob on the workload shapes (default tuning), two rounds each.
# Usage: tools/ab_env.sh <tag> <ENV_NAME> [workloads...]
set -u
TAG=$1; KNOB=$2; shift 2
WL=${*:-c2 c3 c4 c5}
OUT=gpurun_out; mkdir -p $OUT
( time python -m pytest tests -m gpu -x -q ) > $OUT/pytest_$TAG.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/pytest_$TAG.log
tail -5 $OUT/pytest_$TAG.log
: > $OUT/ab_$TAG.txt
for w in $WL; do
  for u in 0 1 0 1; do
    echo "workload=$w $KNOB=$u" >> $OUT/ab_$TAG.txt
    env $KNOB=$u python tools/sweep.py --workload $w --steps 60 --tunings 0,0,0 >> $OUT/ab_$TAG.txt 2>&1
  done
done
cut -c1-230 $OUT/ab_$TAG.txt
