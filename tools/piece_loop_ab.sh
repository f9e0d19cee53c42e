#!/bin/bash
# Round-2 experiments on the balanced kernel's piece loop, prepared but not yet measured (build switches,
# default off; DESIGN.md section 10 item 3a, profiles/r1_bag_sum_balanced_hotspots.md):
#   contig    EMB_CONTIG=1      one contiguous stretch of the staged window per sub-warp
#   bfsearch  EMB_BF_SEARCH=1   branch-free bag search per piece
#   both      both switches (the search is then per stretch; EMB_BF_SEARCH has no effect on it)
#   here (no GPU):   tools/piece_loop_ab.sh build        -> pim-embedding-lookup_b200/lib/<variant>/emblib.so
#   on the GPU box:  tools/piece_loop_ab.sh run <tag>    -> parity tests on every variant build, then the A/B
set -u
HERE=$(cd "$(dirname "$0")/.." && pwd)
CSRC=$HERE/pim-embedding-lookup_b200/csrc
LIBDIR=$HERE/pim-embedding-lookup_b200/lib
flags() { case $1 in contig) echo "-DEMB_CONTIG=1";; bfsearch) echo "-DEMB_BF_SEARCH=1";; esac; }
VARIANTS="contig bfsearch"
case "${1:-}" in
build)
  for v in $VARIANTS; do
    mkdir -p $LIBDIR/$v
    nvcc -O3 -std=c++17 -lineinfo -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC -ccbin g++ $(flags $v) \
         -shared -Xlinker -soname,emb_host -cudart static -o $LIBDIR/$v/emblib.so $CSRC/emb_b200.cu && echo "built $LIBDIR/$v/emblib.so" &
  done; wait ;;
run)
  TAG=${2:-pieces}; OUT=$HERE/gpurun_out; mkdir -p $OUT; cd $HERE
  for v in $VARIANTS; do
    EMBLIB_SO=$LIBDIR/$v/emblib.so python -m pytest tests/test_gpu_parity.py tests/test_gpu_device.py -m gpu -x -q > $OUT/pytest_${TAG}_$v.log 2>&1
    echo "parity on the $v build rc=$?"; tail -2 $OUT/pytest_${TAG}_$v.log
  done
  : > $OUT/$TAG.txt
  for wl in c2 c3 c4 c5; do for rep in 1 2; do
    echo "workload=$wl build=default" >> $OUT/$TAG.txt
    python tools/sweep.py --workload $wl --steps 60 --tunings 0,0,0 >> $OUT/$TAG.txt 2>&1
    for v in $VARIANTS; do
      echo "workload=$wl build=$v" >> $OUT/$TAG.txt
      EMBLIB_SO=$LIBDIR/$v/emblib.so python tools/sweep.py --workload $wl --steps 60 --tunings 0,0,0 >> $OUT/$TAG.txt 2>&1
    done
  done; done
  cut -c1-200 $OUT/$TAG.txt ;;
*) echo "usage: $0 build | run <tag>"; exit 2 ;;
esac
