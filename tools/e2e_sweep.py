#!/usr/bin/env python
"""Host-pointer call (emb_lookup_ex) on C2 through every PCIe path x group count x group shape, one GPU.
Page-locked arrays carved from one arena per kind (adjacent arrays -> merged copies) and, for comparison,
separately allocated arrays (one copy per table).  Prints one JSON line per combination; measurement aid for
DESIGN.md section 6, bench.py is the contract.

    python tools/e2e_sweep.py [--steps 20]
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
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--workload", default="c2")
    ap.add_argument("--in-flight", action="store_true",
                    help="arena arrays, copies both ways: emb_lookup_ex one call at a time against emb_lookup_ex_async "
                         "with two calls in flight, per group count")
    ap.add_argument("--quick", action="store_true",
                    help="arena arrays only: the default path (EMB_HOST_AUTO) against forced copies / mapped results")
    args = ap.parse_args()
    import torch

    from pim_embedding_lookup_b200 import emblib
    from pim_embedding_lookup_b200.dputypes import EMB_HOST_AUTO, EMB_HOST_COPY, EMB_HOST_MAPPED_OUT
    w = bench.WORKLOADS[args.workload]
    T, R, D, B = w["T"], w["R"], w["D"], w["B"]
    lib = emblib.load()
    emblib.configure(T, D, 0, 0, device=0)
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    gen = torch.Generator(device="cuda").manual_seed(1234)
    rows_t = w.get("rows", [R] * T)
    for t in range(T):
        lim = int(1e9 / np.sqrt(rows_t[t]))
        a = torch.randint(-lim, lim + 1, (rows_t[t], D), dtype=torch.int32, device="cuda", generator=gen)
        emblib.check(lib.emb_populate_table_device(t, rows_t[t], a.data_ptr(), 0, 1, stream.cuda_stream))
        del a
    batches = [bench.make_batch(w, seed=123 + k) for k in range(2)]
    lookups = [sum(len(i) for i in b[0]) for b in batches]

    def arrays(packed):
        host = []
        for idx, off in batches:
            if packed:
                arena = bench.pinned_arena(lib, sum(4 * len(i) for i in idx) + sum(4 * len(o) for o in off) + 64)
                pos, hi, ho = 0, [], []
                for src, dst in ((idx, hi), (off, ho)):
                    for a_ in src:
                        v, pos = bench.carve_u32(arena, pos, len(a_))
                        v[:] = a_
                        dst.append(v)
            else:
                hi = [emblib.pinned_array(len(i)) for i in idx]
                ho = [emblib.pinned_array(len(o)) for o in off]
                for a_, b_ in zip(hi + ho, idx + off):
                    a_[:] = b_
            host.append((hi, ho))
        if packed:
            ra = bench.pinned_arena(lib, T * B * D * 4 + 64)
            res = [ra[t * B * D * 4:(t + 1) * B * D * 4].view(np.float32).reshape(B, D) for t in range(T)]
        else:
            res = [emblib.pinned_array((B, D), np.float32) for _ in range(T)]
        return [emblib.HostCall(hi, ho, res) for hi, ho in host]

    def timed(calls, n):
        for k in range(3):
            calls[k % 2]()
        t0 = time.perf_counter()
        for k in range(n):
            calls[k % 2]()
        dt = (time.perf_counter() - t0) / n
        return dt * 1e3, emblib.last_timings()

    if args.in_flight:
        calls = arrays(True)
        # the second call in flight needs result arrays of its own
        ra2 = bench.pinned_arena(lib, T * B * D * 4 + 64)
        res2 = [ra2[t * B * D * 4:(t + 1) * B * D * 4].view(np.float32).reshape(B, D) for t in range(T)]
        calls[1] = emblib.HostCall(calls[1].indices, calls[1].offsets, res2)

        def timed_in_flight(n):
            def run(m):
                calls[0].start()
                for k in range(m):
                    if k + 1 < m:
                        calls[(k + 1) % 2].start()
                    calls[k % 2].wait()
            run(4)
            t0 = time.perf_counter()
            run(n)
            return (time.perf_counter() - t0) / n * 1e3

        emblib.set_host_path(EMB_HOST_AUTO)
        for groups in (0, 1, 2, 3, 4, 6, 8, 12):
            os.environ["EMB_GROUPS"] = str(groups)
            ms, tm = timed(calls, args.steps)
            ms2 = timed_in_flight(args.steps)
            print(json.dumps({"workload": args.workload, "groups": groups, "launches_per_call": int(tm.launches),
                              "sync_ms_per_call": round(ms, 4), "two_in_flight_ms_per_call": round(ms2, 4),
                              "G_lookups_per_s_in_flight": round(np.mean(lookups) / ms2 / 1e6, 3)}), flush=True)
        os.environ["EMB_GROUPS"] = "0"
        emblib.shutdown()
        return
    if args.quick:
        calls = arrays(True)
        for path, pname in ((EMB_HOST_AUTO, "auto"), (EMB_HOST_MAPPED_OUT, "mapped_out"), (EMB_HOST_COPY, "copy")):
            emblib.set_host_path(path)
            for groups in (0,) if path == EMB_HOST_AUTO else (0, 2, 4, 8):
                os.environ["EMB_GROUPS"] = str(groups)
                ms, tm = timed(calls, args.steps)
                print(json.dumps({"workload": args.workload, "path": pname, "groups": groups,
                                  "ms_per_call": round(ms, 4), "G_lookups_per_s": round(np.mean(lookups) / ms / 1e6, 3),
                                  "launches": int(tm.launches), "h2d_ms": round(tm.h2d_ms, 3),
                                  "kernel_ms": round(tm.kernel_ms, 3), "d2h_ms": round(tm.d2h_ms, 3),
                                  "chosen_path": int(tm.path), "h2d_bytes": int(tm.h2d_bytes),
                                  "d2h_bytes": int(tm.d2h_bytes)}), flush=True)
        emblib.set_host_path(EMB_HOST_AUTO)
        emblib.shutdown()
        return
    for packed in (True, False):
        calls = arrays(packed)
        for path, pname in ((EMB_HOST_MAPPED_OUT, "mapped_out"), (EMB_HOST_COPY, "copy")):
            emblib.set_host_path(path)
            for shape in (0, 1, 2):
                for groups in (0, 4, 6, 8, 12):
                    os.environ["EMB_GROUP_SHAPE"] = str(shape)
                    os.environ["EMB_GROUPS"] = str(groups)
                    ms, tm = timed(calls, args.steps)
                    print(json.dumps({"arrays": "one arena" if packed else "one allocation per table", "path": pname,
                                      "shape": shape, "groups": groups, "ms_per_call": round(ms, 4),
                                      "G_lookups_per_s": round(np.mean(lookups) / ms / 1e6, 3),
                                      "launches": int(tm.launches), "h2d_ms": round(tm.h2d_ms, 3),
                                      "kernel_ms": round(tm.kernel_ms, 3), "d2h_ms": round(tm.d2h_ms, 3)}), flush=True)
        emblib.set_host_path(EMB_HOST_AUTO)
    emblib.shutdown()


if __name__ == "__main__":
    main()
