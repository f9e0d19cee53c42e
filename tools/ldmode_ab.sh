#!/bin/bash
# L2 priority of the row loads: default build against EMB_LD_MODE=4/5/6 builds (lib/ldN/emblib.so), two rounds.
OUT=gpurun_out; TAG=${1:-ldab}; mkdir -p $OUT; : > $OUT/$TAG.txt
for wl in c2 c3 c4 c5; do for rep in 1 2; do for m in 0 4 5 6; do
  echo "workload=$wl ld_mode=$m" >> $OUT/$TAG.txt
  if [ $m = 0 ]; then python tools/sweep.py --workload $wl --steps 60 --tunings 0,0,0 >> $OUT/$TAG.txt 2>&1
  else EMBLIB_SO=pim-embedding-lookup_b200/lib/ld$m/emblib.so python tools/sweep.py --workload $wl --steps 60 --tunings 0,0,0 >> $OUT/$TAG.txt 2>&1; fi
done; done; done
cut -c1-200 $OUT/$TAG.txt
