#!/bin/bash
# One-warp CTAs (EMB_CTA_THREADS=32) of the balanced kernel against the 64-thread default, dim-16 shapes.
OUT=gpurun_out; TAG=${1:-cta32}; mkdir -p $OUT
( time python -m pytest tests -m gpu -x -q ) > $OUT/pytest_$TAG.log 2>&1; echo "pytest rc=$?"; tail -4 $OUT/pytest_$TAG.log
: > $OUT/$TAG.txt
for wl in c2 c3; do
  echo "== $wl default" >> $OUT/$TAG.txt
  python tools/sweep.py --workload $wl --steps 60 --tunings 0,0,0 >> $OUT/$TAG.txt 2>&1
  echo "== $wl threads=32" >> $OUT/$TAG.txt
  EMB_CTA_THREADS=32 python tools/sweep.py --workload $wl --steps 60 --tunings 2,4,1 2,4,2 2,4,4 2,8,1 2,8,2 2,8,4 >> $OUT/$TAG.txt 2>&1
done
cut -c1-250 $OUT/$TAG.txt
