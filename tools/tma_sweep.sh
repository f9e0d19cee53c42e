#!/bin/bash
# LDG.128 vs TMA (1-D bulk copy per row, tile::gather4) for random row gathers; see tools/gather_tma.cu.
# usage (under gpurun): bash tools/tma_sweep.sh > gpurun_out/tma_sweep.txt
B=build/gather_tma
run() { timeout 120 $B "$@" 2>&1 || echo "{\"failed\": \"$*\", \"rc\": $?}"; }
for region in 64 1586; do
  run ldg $region 64
  for cfg in "32 4 3" "32 8 3" "64 4 3" "32 4 6" "64 8 2" "128 4 2"; do
    run bulk $region 64 $cfg
  done
  for cfg in "32 4 3" "64 4 3" "128 4 3" "128 8 2" "32 8 6"; do
    run g4 $region 64 $cfg 1
  done
  run g4 $region 64 32 4 3 4
done
for region in 1586; do
  run ldg $region 256
  run bulk $region 256 32 4 3
  run bulk $region 256 32 6 4
  run g4 $region 256 32 4 3 1
  run ldg $region 512
  run bulk $region 512 32 4 3
  run bulk $region 512 32 3 4
done
