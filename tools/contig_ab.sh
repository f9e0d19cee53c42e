#!/bin/bash
# Round-2 experiment, prepared but not yet measured: balanced kernel with ONE contiguous stretch of the staged
# window per sub-warp (build switch EMB_CONTIG=1) against the default round-robin pieces.
#   here (no GPU):   tools/contig_ab.sh build          -> pim-embedding-lookup_b200/lib/contig/emblib.so
#   on the GPU box:  tools/contig_ab.sh run <tag>      -> parity tests on the variant build, then the A/B
set -u
HERE=$(cd "$(dirname "$0")/.." && pwd)
CSRC=$HERE/pim-embedding-lookup_b200/csrc
LIB=$HERE/pim-embedding-lookup_b200/lib/contig/emblib.so
case "${1:-}" in
build)
  mkdir -p "$(dirname $LIB)"
  nvcc -O3 -std=c++17 -lineinfo -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC -ccbin g++ -DEMB_CONTIG=1 \
       -shared -Xlinker -soname,emb_host -cudart static -o $LIB $CSRC/emb_b200.cu && echo "built $LIB" ;;
run)
  TAG=${2:-contig}; OUT=$HERE/gpurun_out; mkdir -p $OUT; cd $HERE
  EMBLIB_SO=$LIB python -m pytest tests/test_gpu_parity.py tests/test_gpu_device.py -m gpu -x -q > $OUT/pytest_$TAG.log 2>&1
  echo "parity on the EMB_CONTIG=1 build rc=$?"; tail -3 $OUT/pytest_$TAG.log
  : > $OUT/$TAG.txt
  for wl in c2 c3 c4 c5; do for rep in 1 2; do
    echo "workload=$wl contig=0" >> $OUT/$TAG.txt
    python tools/sweep.py --workload $wl --steps 60 --tunings 0,0,0 >> $OUT/$TAG.txt 2>&1
    echo "workload=$wl contig=1" >> $OUT/$TAG.txt
    EMBLIB_SO=$LIB python tools/sweep.py --workload $wl --steps 60 --tunings 0,0,0 >> $OUT/$TAG.txt 2>&1
  done; done
  cut -c1-200 $OUT/$TAG.txt ;;
*) echo "usage: $0 build | run <tag>"; exit 2 ;;
esac
