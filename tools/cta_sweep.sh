#!/bin/bash
# CTA size of the balanced kernel (EMB_CTA_THREADS) x workload; default tuning otherwise, plus mult variants
for wl in c2 c3 c4 c5; do for t in 256 128 64; do
  echo "== $wl threads=$t"
  EMB_CTA_THREADS=$t timeout 300 python tools/sweep.py --workload $wl --tunings 0,0,0 2,0,1 2,0,2 2,0,4 2>&1 | cut -c60-215
done; done
