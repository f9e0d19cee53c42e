#!/bin/bash
# Round 2, first GPU call: parity of the stretch kernel (variant 6), then the A/B against the round-1 kernels on
# the four workload shapes.  tools/sweep.py times every tuning on the same tables and index batches and checks
# that the bits agree.  Usage: tools/r2_stretch_ab.sh <tag>
set -u
TAG=${1:-r2a}; OUT=gpurun_out; mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/smi_$TAG.txt 2>&1
( time timeout 900 python -m pytest tests -m gpu -x -q ) > $OUT/pytest_$TAG.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/pytest_$TAG.log
tail -8 $OUT/pytest_$TAG.log
: > $OUT/ab_$TAG.txt
run() { echo "## $*" >> $OUT/ab_$TAG.txt; timeout 300 "$@" >> $OUT/ab_$TAG.txt 2>&1; }
# 2,0,0 = round-1 default (one-warp balanced kernel); 6,* = stretch kernel (K, bags multiplier)
run python tools/sweep.py --workload c2 --steps 60 --tunings 2,0,0 6,8,2 6,8,1 6,8,4 6,4,2 6,4,4 2,0,0 6,8,2
run env EMB_STRETCH_STAGE=1 python tools/sweep.py --workload c2 --steps 60 --tunings 6,8,2 6,8,4 6,4,4
run python tools/sweep.py --workload c3 --steps 60 --tunings 2,0,0 6,4,4 6,8,4 6,4,2 6,8,8 6,4,8
run python tools/sweep.py --workload c4 --steps 40 --tunings 2,0,0 6,8,4 6,8,2 6,8,8 6,4,4
run env EMB_STRETCH_STAGE=1 python tools/sweep.py --workload c4 --steps 40 --tunings 6,8,4 6,8,8
run python tools/sweep.py --workload c5 --steps 40 --tunings 2,0,0 6,8,4 6,8,2 6,8,8 6,8,1
run env EMB_STRETCH_STAGE=1 python tools/sweep.py --workload c5 --steps 40 --tunings 6,8,4 6,8,8
# tables that fit L2 (the round-1 warp-autonomous kernel's territory)
run python tools/sweep.py --workload c2 --rows 131072 --steps 60 --tunings 3,8,0 2,0,0 6,8,2 6,8,4
cut -c1-260 $OUT/ab_$TAG.txt
