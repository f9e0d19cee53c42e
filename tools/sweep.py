#!/usr/bin/env python
"""Kernel-variant sweep on one GPU: same tables and index batches, every (variant, unroll, mult)
tuning timed with CUDA events.  Prints one JSON line per tuning.  Measurement aid for
DESIGN.md's kernel choices; bench.py is the contract.

    python tools/sweep.py [--workload c2] [--steps 30] [--tunings 1,4,0 2,8,1 ...]
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="c2")
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--tunings", nargs="*", default=["1,4,0", "1,8,0", "2,4,1", "2,8,1", "2,4,2", "2,8,2", "2,4,4",
                                                     "2,8,4"])
    ap.add_argument("--rows", type=int, default=0, help="override rows per table")
    ap.add_argument("--tables", type=int, default=0, help="override the table count")
    args = ap.parse_args()
    import torch

    from pim_embedding_lookup_b200 import emblib
    w = dict(bench.WORKLOADS[args.workload])
    if args.rows:
        w["R"] = args.rows
        w.pop("rows", None)
    if args.tables:
        w["T"] = args.tables
    T, R, D, B = w["T"], w["R"], w["D"], w["B"]
    lib = emblib.load()
    emblib.configure(T, D, 0, 0, device=0)
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    gen = torch.Generator(device="cuda").manual_seed(1234)
    lim = int(1e9 / np.sqrt(R))
    for t in range(T):
        Rt = w["rows"][t] if "rows" in w and not args.rows else R
        a = torch.randint(-lim, lim + 1, (Rt, D), dtype=torch.int32, device="cuda", generator=gen)
        emblib.check(lib.emb_populate_table_device(t, Rt, a.data_ptr(), 0, 1, stream.cuda_stream))
        del a
    batches = [bench.make_batch(w, seed=123 + k) for k in range(2)]
    dev = [([torch.from_numpy(i.view(np.int32)).cuda() for i in idx], [torch.from_numpy(o.view(np.int32)).cuda() for o in off])
           for idx, off in batches]
    out = torch.empty((T, B, D), dtype=torch.float32, device="cuda")
    ref = None
    peak, _ = bench.load_peaks()
    for tn in args.tunings:
        v, k, m = (int(x) for x in tn.split(","))
        emblib.check(lib.emb_set_tuning(v, k, m))
        plans = [emblib.Plan([dict(table_id=t, indices=di[t], offsets=do[t], out=out[t]) for t in range(T)])
                 for di, do in dev]
        for i in range(5):
            plans[i % 2].launch(stream)
        torch.cuda.synchronize()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
        for i in range(args.steps):
            ev[i][0].record(stream)
            plans[i % 2].launch(stream)
            ev[i][1].record(stream)
        torch.cuda.synchronize()
        ms = sorted(a.elapsed_time(b) for a, b in ev)
        # parity between tunings: same bits as the first tuning on batch (steps-1) % 2
        chk = out.view(torch.int32).to(torch.int64).sum().item()
        if ref is None:
            ref = chk
        nl = np.mean([sum(len(i) for i in b[0]) for b in batches])
        by = bench.algorithmic_bytes(nl, T * B, D)
        med = ms[len(ms) // 2]
        print(json.dumps({"workload": args.workload, "rows": R, "tables": T, "variant": v, "unroll": k, "mult": m, "grid": plans[0].grid,
                          "kernel_ms_median": round(med, 4), "kernel_ms_min": round(ms[0], 4),
                          "GBps": round(by / med / 1e6, 1), "frac_of_measured_peak": round(by / med / 1e6 / peak, 4),
                          "lookups_per_s": nl / med * 1e3, "same_bits_as_first": chk == ref}), flush=True)
        for p in plans:
            p.close()
    emblib.shutdown()


if __name__ == "__main__":
    main()
