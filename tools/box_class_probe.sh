#!/bin/bash
# What kind of box is this?  The pool's B200 boxes fall into two classes on the C2 pooling launch (0.2215 vs 0.2441 ms,
# same binary).  Prints clocks / ECC / power settings, the box's device-to-device copy bandwidth (the way
# MEASURED_PEAKS.json measures it) and the C2 launch time, so that the two can be told apart.  Usage: tools/box_class_probe.sh <tag>
TAG=${1:-probe}; OUT=gpurun_out; mkdir -p $OUT
{
nvidia-smi --query-gpu=name,serial,vbios_version,driver_version,clocks.max.sm,clocks.max.mem,clocks.mem,ecc.mode.current,power.limit,power.max_limit,memory.total,pcie.link.gen.current,pcie.link.width.current --format=csv
nvidia-smi -q -d CLOCK,PERFORMANCE | grep -E "Graphics|SM|Memory|Performance State|Slowdown|Clocks Event" | head -30
python - <<'PY'
import torch, json
n = 1 << 30
a = torch.empty(n, dtype=torch.bfloat16, device="cuda"); b = torch.empty_like(a)
best = 0
for _ in range(10):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); b.copy_(a); e1.record(); torch.cuda.synchronize()
    best = max(best, 2 * n * 2 / (e0.elapsed_time(e1) * 1e-3) / 1e9)
print(json.dumps({"copy_GBps_best_of_10": round(best, 1), "mem_clock_under_load": None}))
PY
python tools/sweep.py --workload c2 --steps 60 --tunings 0,0,0
python tools/sweep.py --workload c4 --steps 40 --tunings 0,0,0
} > $OUT/box_class_$TAG.txt 2>&1
cut -c1-330 $OUT/box_class_$TAG.txt
