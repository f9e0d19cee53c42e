#!/bin/bash
# Hot-row A/B, second pass: two-class (hot set + uniform cold stream) distributions.  Usage: tools/r2_s2_b.sh <tag>
set -u
TAG=${1:-s2b}; OUT=gpurun_out; mkdir -p $OUT
HOT=$OUT/l2_hot_$TAG.txt; : > $HOT
run() { echo "## $*" >> $HOT; timeout 300 "$@" >> $HOT 2>&1; echo "rc=$?" >> $HOT; }
# 32 tables walked one after the other: the window a hot row must survive is its own table's walk
run python tools/l2_hot_ab.py --tag mix_32_tables --dist mix --hot-frac 0.5 --hot-set-mb 48 --hot-mb 48 --steps 20
# two tables, a batch large enough that the cold bytes between two uses of a hot row exceed L2
run python tools/l2_hot_ab.py --tag mix_2_tables_big_batch --dist mix --tables 2 --rows 8000000 --batch 65536 --hot-frac 0.5 \
    --hot-set-mb 160 --hot-mb 160 80 --steps 20
run python tools/l2_hot_ab.py --tag zipf_2_tables_big_batch --tables 2 --rows 8000000 --batch 65536 --hot-mb 80 --arms quick --steps 20
grep -E "^\||^###|rc=|Error|error" $HOT | cut -c1-200
timeout 300 ncu --metrics lts__t_sector_hit_rate.pct,dram__bytes_read.sum,gpu__time_duration.sum --clock-control none \
    -k regex:bag_sum_stretch --csv --log-file $OUT/l2_hot_ncu_$TAG.csv \
    python tools/l2_hot_ab.py --tag ncu --dist mix --tables 2 --rows 8000000 --batch 65536 --hot-frac 0.5 --hot-set-mb 160 \
    --hot-mb 160 --arms quick --steps 2 > $OUT/l2_hot_ncu_$TAG.log 2>&1
echo "ncu rc=$?"; tail -12 $OUT/l2_hot_ncu_$TAG.log | cut -c1-200
