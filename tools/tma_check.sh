#!/bin/bash
# variant 4 (TMA gather4 pooling kernel): parity tests, then sweeps against the default on c2/c3/c4/c5 and small tables
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_device.py -x -q 2>&1 | tail -4
for wl in c2 c3 c4 c5; do timeout 300 python tools/sweep.py --workload $wl --tunings 0,0,0 4,0,0; done
timeout 300 python tools/sweep.py --workload c2 --rows 125000 --tunings 0,0,0 4,0,0
