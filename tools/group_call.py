#!/usr/bin/env python
"""The host-pointer call on a DEVICE GROUP (emb_configure_devices): C2 from page-locked host arrays through
emb_lookup_ex (one call at a time) and emb_lookup_ex_async (two in flight), whole tables dealt over 1, 2, ... GPUs of
the box.  One process; every GPU moves its own tables' bytes over its own PCIe link.  Prints one JSON line per group
size.  Measurement aid; bench.py is the contract.

    python tools/group_call.py [--gpus 1 2 4] [--steps 30] [--workload c2]
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, nargs="*", default=[1, 2])
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--workload", default="c2")
    args = ap.parse_args()
    import torch

    from pim_embedding_lookup_b200 import emblib
    w = bench.WORKLOADS[args.workload]
    T, R, D, B = w["T"], w["R"], w["D"], w["B"]
    lib = emblib.load()
    batches = [bench.make_batch(w, seed=123 + k) for k in range(2)]
    lookups = [sum(len(i) for i in b[0]) for b in batches]
    rows_t = w.get("rows", [R] * T)
    check_t = [0, T - 1]
    ref_bits = None
    for n in args.gpus:
        if n > torch.cuda.device_count():
            print(json.dumps({"gpus": n, "skipped": f"{torch.cuda.device_count()} visible"}), flush=True)
            continue
        emblib.shutdown()
        emblib.configure_devices(T, D, list(range(n)))
        for t in range(T):
            lim = int(1e9 / np.sqrt(rows_t[t]))
            with torch.cuda.device(lib.emb_table_device(t) if lib.emb_table_device(t) >= 0 else t % n):
                gen = torch.Generator(device="cuda").manual_seed(1234 + t)
                a = torch.randint(-lim, lim + 1, (rows_t[t], D), dtype=torch.int32, device="cuda", generator=gen)
                torch.cuda.synchronize()
                emblib.check(lib.emb_populate_table_device(t, rows_t[t], a.data_ptr(), 0, 1, None))
                del a
        # host arrays: one arena per kind and per call, as bench.py's e2e
        calls = []
        for idx, off in batches:
            arena = bench.pinned_arena(lib, sum(4 * len(i) for i in idx) + sum(4 * len(o) for o in off) + 64)
            pos, hi, ho = 0, [], []
            for src, dst in ((idx, hi), (off, ho)):
                for a_ in src:
                    v, pos = bench.carve_u32(arena, pos, len(a_))
                    v[:] = a_
                    dst.append(v)
            ra = bench.pinned_arena(lib, T * B * D * 4 + 64)
            res = [ra[t * B * D * 4:(t + 1) * B * D * 4].view(np.float32).reshape(B, D) for t in range(T)]
            calls.append(emblib.HostCall(hi, ho, res))

        def sync_calls(m):
            for k in range(3):
                calls[k % 2]()
            t0 = time.perf_counter()
            for k in range(m):
                calls[k % 2]()
            return (time.perf_counter() - t0) / m * 1e3

        def in_flight(m):
            def run(c):
                calls[0].start()
                for k in range(c):
                    if k + 1 < c:
                        calls[(k + 1) % 2].start()
                    calls[k % 2].wait()
            run(4)
            t0 = time.perf_counter()
            run(m)
            return (time.perf_counter() - t0) / m * 1e3

        ms = sync_calls(args.steps)
        tm = emblib.last_timings()
        bits = np.concatenate([calls[(args.steps - 1) % 2].results[t].view(np.uint32).ravel() for t in check_t])
        if ref_bits is None:
            ref_bits = bits.copy()  # the same tables (seeded per table) and batches on every group size: same bits
        same = bool(np.array_equal(bits, ref_bits))
        ms2 = in_flight(args.steps)
        print(json.dumps({"workload": args.workload, "gpus": n, "group": lib.emb_group_size(),
                          "sync_ms_per_call": round(ms, 4), "two_in_flight_ms_per_call": round(ms2, 4),
                          "G_lookups_per_s_sync": round(np.mean(lookups) / ms / 1e6, 3),
                          "G_lookups_per_s_in_flight": round(np.mean(lookups) / ms2 / 1e6, 3),
                          "launches_per_call": int(tm.launches), "h2d_bytes": int(tm.h2d_bytes), "d2h_bytes": int(tm.d2h_bytes),
                          "slowest_gpu_ms": {"h2d": round(tm.h2d_ms, 3), "kernel": round(tm.kernel_ms, 3), "d2h": round(tm.d2h_ms, 3)},
                          "index_errors": int(lib.emb_check_indices()), "same_bits_as_first_group_size": same}), flush=True)
    emblib.shutdown()


if __name__ == "__main__":
    main()
