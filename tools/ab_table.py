#!/usr/bin/env python
"""Readable table of a tools/sweep.py log (one JSON line per tuning, '## command' lines between runs)."""
import json
import sys

for line in open(sys.argv[1]):
    line = line.strip()
    if line.startswith("##"):
        print(line)
        continue
    try:
        d = json.loads(line)
    except Exception:  # noqa: BLE001
        if line:
            print("   ", line[:200])
        continue
    print(f"  v{d['variant']} K{d['unroll']} m{d['mult']} grid {d['grid']:7d}  med {d['kernel_ms_median']:.4f} min {d['kernel_ms_min']:.4f} ms "
          f"{d['lookups_per_s'] / 1e9:7.2f} G lookups/s  frac {d['frac_of_measured_peak']:.3f}  same_bits={d['same_bits_as_first']}")
