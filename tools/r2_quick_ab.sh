#!/bin/bash
# Quick A/B on one GPU: stretch-kernel parity tests, then sweeps (tools/sweep.py times every tuning on the same
# tables / index batches and checks that the bits agree).  Usage: tools/r2_quick_ab.sh <tag>
set -u
TAG=${1:-r2b}; OUT=gpurun_out; mkdir -p $OUT
( time timeout 900 python -m pytest tests/test_gpu_stretch.py tests/test_gpu_parity.py tests/test_gpu_device.py -m gpu -x -q ) > $OUT/pytest_$TAG.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/pytest_$TAG.log
tail -6 $OUT/pytest_$TAG.log
: > $OUT/ab_$TAG.txt
run() { echo "## $*" >> $OUT/ab_$TAG.txt; timeout 300 "$@" >> $OUT/ab_$TAG.txt 2>&1; }
run python tools/sweep.py --workload c2 --steps 60 --tunings 0,0,0 6,8,2 6,4,2 6,8,4 6,8,1 1,0,0 6,8,2
run python tools/sweep.py --workload c3 --steps 60 --tunings 0,0,0 6,4,4 6,8,4 6,8,8 6,4,8
run python tools/sweep.py --workload c4 --steps 40 --tunings 0,0,0 6,8,4 6,8,2 6,4,4 6,8,8
run python tools/sweep.py --workload c5 --steps 40 --tunings 0,0,0 6,8,4 6,8,2 6,8,1 6,4,2
run env EMB_STRETCH_STAGE=1 python tools/sweep.py --workload c5 --steps 40 --tunings 6,8,2 6,8,1
run python tools/sweep.py --workload c2 --rows 131072 --steps 60 --tunings 0,0,0 6,8,2 6,8,4
python tools/ab_table.py $OUT/ab_$TAG.txt
