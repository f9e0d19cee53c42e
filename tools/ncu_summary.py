#!/usr/bin/env python
"""Summarise ncu outputs brought back in gpurun_out/ into small text files under profiles/.

    python tools/ncu_summary.py launches gpurun_out/launches_TAG.csv  > profiles/rN_launches.txt
    python tools/ncu_summary.py full gpurun_out/prof_TAG.ncu-rep      > profiles/rN_<kernel>_full.txt
"""
import collections
import csv
import io
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__shared_mem_per_block_static", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
    "launch__occupancy_limit_warps", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__bytes_read.sum.per_second",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__cycles_active.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sector_hit_rate.pct", "lts__t_sectors_srcunit_tex_op_read.sum", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__t_sector_hit_rate.pct", "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum",
    "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum",
    "smsp__sass_average_data_bytes_per_sector_mem_global_op_ld.pct", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smsp__inst_executed_op_shared_atom.sum",
]


def launches(path):
    rows = list(csv.reader(open(path)))
    hdr = next(i for i, r in enumerate(rows) if r and r[0] == "ID")
    h = rows[hdr]
    d = collections.OrderedDict()
    for r in rows[hdr + 1:]:
        if len(r) < len(h):
            continue
        rec = dict(zip(h, r))
        d.setdefault(rec["Kernel Name"], []).append(float(rec["Metric Value"].replace(",", "")))
        unit = rec["Metric Unit"]
    tot = sum(sum(v) for v in d.values())
    print(f"# ncu --metrics gpu__time_duration.sum --clock-control none  ({path})")
    print(f"# per-launch times are cold-cache and serialised: compare SHARES.  total = {tot:.0f} {unit}")
    for k, v in sorted(d.items(), key=lambda kv: -sum(kv[1])):
        print(f"{sum(v) / tot * 100:6.2f}%  n={len(v):4d}  mean={sum(v) / len(v):12.0f} {unit}  min={min(v):10.0f}  max={max(v):10.0f}  {k[:150]}")


def full(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    h, units = rows[0], rows[1]
    print(f"# ncu --set full --clock-control none --import-source on  ({path})")
    for r in rows[2:]:
        rec = dict(zip(h, r))
        print(f"\n== {rec.get('Kernel Name', '')[:150]}  (ID {rec.get('ID')})")
        for k in KEYS:
            if k in rec:
                print(f"  {k:85s} {rec[k]:>16s} {units[h.index(k)]}")


if __name__ == "__main__":
    {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2])
