#!/bin/bash
# Single validation shot for the pipelined kernel (variant 5): its parity tests, a C2 sweep against the
# default, the main suite under the default and under EMB_DEFAULT_VARIANT=5.
OUT=gpurun_out; mkdir -p $OUT
python -m pytest tests/test_gpu_pipe.py -m gpu -q > $OUT/pipe_tests.log 2>&1; echo "pipe tests rc=$?"; tail -3 $OUT/pipe_tests.log
python tools/sweep.py --workload c2 --steps 60 --tunings 0,0,0 5,8,2 5,8,4 5,8,8 > $OUT/pipe_sweep.txt 2>&1; echo "sweep rc=$?"; cut -c1-220 $OUT/pipe_sweep.txt
python -m pytest tests -m gpu -x -q --ignore tests/test_gpu_pipe.py > $OUT/pipe_main.log 2>&1; echo "main suite rc=$?"; tail -2 $OUT/pipe_main.log
EMB_DEFAULT_VARIANT=5 python -m pytest tests/test_gpu_parity.py tests/test_gpu_device.py tests/test_gpu_shim.py -m gpu -x -q > $OUT/pipe_forced.log 2>&1; echo "forced-default suite rc=$?"; tail -2 $OUT/pipe_forced.log
EMB_PIPE_BAGS=3 python tools/sweep.py --workload c2 --steps 60 --tunings 0,0,0 5,8,2 5,8,4 5,8,8 > $OUT/pipe_sweep3.txt 2>&1; echo "sweep3 rc=$?"; cut -c1-220 $OUT/pipe_sweep3.txt
EMB_PIPE_BAGS=3 python -m pytest tests/test_gpu_pipe.py -m gpu -q > $OUT/pipe_tests3.log 2>&1; echo "pipe tests (3) rc=$?"; tail -3 $OUT/pipe_tests3.log
