#!/usr/bin/env python
"""Latency of ONE lookup() call at the reference's own (small) shapes -- BASELINE.json configs[0]
(Kaggle arch, 26 tables, dim 16, mini-batch 1, one index per lookup) and the presets
the reference's run scripts ship (`toy` upmem/run.sh:93-101, `kaggle`
run.sh:40-45, RM2 upmem/r.sh:6-10).  These calls are launch-bound, not bandwidth-bound: what counts
is host time per call, so every flow is timed with time.perf_counter around the synchronous call:

  cpu      dlrm_s_pytorch.py:483-488 apply_emb: nn.EmbeddingBag(mode="sum") per table (fp32), all
           host cores and one core
  pageable lookup() on plain NumPy arrays (what dlrm_dpu_pytorch.py:484-498 passes) -> bounce path
  pinned   lookup() on page-locked arrays (emb_host_alloc) -> mapped path (no copies at all)
  device   emb_plan_launch on device-resident inputs, CUDA events (kernel only)

Every flow's result is compared with the CPU float path (<= 1e-6 absolute: the fixed-point tables
hold 1e-9 steps) and the two host flows with each other bit for bit.
Prints one JSON line per preset.  Measurement aid; bench.py stays the contract."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

KAGGLE = [1460, 583, 10131227, 2202608, 305, 24, 12517, 633, 3, 93145, 5683, 8351593, 3194, 27, 14992, 5461306, 10,
          5652, 2173, 4, 7046547, 18, 15, 286181, 105, 142572]  # PIM-dlrm-new/tools/visualize.py:949
PRESETS = {
    "c1": dict(rows=KAGGLE, D=16, B=1, P=1, what="C1: Kaggle cardinalities, dim 16, mini-batch 1, 1 index per lookup"),
    "kaggle": dict(rows=KAGGLE, D=16, B=512, P=1, what="run.sh kaggle_env: Kaggle cardinalities, dim 16, 512 bags"),
    "toy": dict(rows=[50_000] * 9, D=64, B=64, P=32, what="run.sh toy preset: 9 tables x 50k rows x dim 64, 64 bags x 32"),
    "rm2": dict(rows=[125_000] * 32, D=64, B=64, P=120, what="r.sh RM2: 32 tables x 125k rows x dim 64, 64 bags x 120"),
}


def pct(xs, q):
    xs = sorted(xs)
    return xs[min(len(xs) - 1, int(q * len(xs)))]


def time_calls(fn, n, warm=20):
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(n):
        t0 = time.perf_counter()
        fn()
        ts.append((time.perf_counter() - t0) * 1e6)
    return {"median_us": round(pct(ts, 0.5), 2), "p10_us": round(pct(ts, 0.1), 2), "p90_us": round(pct(ts, 0.9), 2)}


def run(name, p, calls, max_rows):
    import torch

    from pim_embedding_lookup_b200 import emblib
    rows = [min(r, max_rows) if max_rows else r for r in p["rows"]]
    T, D, B, P = len(rows), p["D"], p["B"], p["P"]
    rng = np.random.default_rng(123)
    emblib.configure(T, D, B, P, device=0)
    emb = []
    for t, n in enumerate(rows):  # create_emb, dlrm_dpu_pytorch.py:391-393
        W = rng.uniform(-np.sqrt(1 / n), np.sqrt(1 / n), size=(n, D)).astype(np.float32)
        emb.append(torch.nn.EmbeddingBag.from_pretrained(torch.from_numpy(W), mode="sum", freeze=True))
        emblib.populate_table_f32(t, W)
    idx = [rng.integers(0, n, size=B * P).astype(np.uint32) for n in rows]
    off = [(np.arange(B) * P).astype(np.uint32) for _ in rows]
    lS_i = [torch.from_numpy(i.astype(np.int64)) for i in idx]
    lS_o = [torch.from_numpy(o.astype(np.int64)) for o in off]

    out = {"preset": name, "what": p["what"], "tables": T, "dim": D, "bags": B, "pooling": P,
           "lookups_per_call": T * B * P, "calls": calls}
    if rows != p["rows"]:
        out["rows_capped_at"] = max_rows

    def apply_emb():
        with torch.no_grad():
            return [emb[k](lS_i[k], lS_o[k]) for k in range(T)]
    ref = [r.numpy() for r in apply_emb()]
    ncores = len(os.sched_getaffinity(0))
    torch.set_num_threads(ncores)
    out["cpu_all_cores"] = dict(time_calls(apply_emb, calls), cores=ncores)
    torch.set_num_threads(1)
    out["cpu_one_core"] = dict(time_calls(apply_emb, calls), cores=1)
    torch.set_num_threads(ncores)

    # the reference's call on the arrays its caller owns (plain NumPy memory)
    call = emblib.HostCall(idx, off)
    res_pageable = [r.copy() for r in call()]
    out["lookup_pageable"] = dict(time_calls(call, calls), path=int(emblib.last_timings().path),
                                  kernel_us=round(emblib.last_timings().kernel_ms * 1e3, 2))
    # the same on page-locked arrays
    p_idx = [emblib.pinned_array(i.shape, np.uint32) for i in idx]
    p_off = [emblib.pinned_array(o.shape, np.uint32) for o in off]
    p_res = [emblib.pinned_array((B, D), np.float32) for _ in rows]
    for t in range(T):
        p_idx[t][:] = idx[t]
        p_off[t][:] = off[t]
    pcall = emblib.HostCall(p_idx, p_off, p_res)
    res_pinned = [r.copy() for r in pcall()]
    out["lookup_pinned"] = dict(time_calls(pcall, calls), path=int(emblib.last_timings().path),
                                kernel_us=round(emblib.last_timings().kernel_ms * 1e3, 2))

    # device-resident plan: kernel only
    stream = torch.cuda.Stream()
    d_idx = [torch.from_numpy(i.view(np.int32)).cuda() for i in idx]
    d_off = [torch.from_numpy(o.view(np.int32)).cuda() for o in off]
    d_out = torch.zeros((T, B, D), dtype=torch.float32, device="cuda")
    plan = emblib.Plan([dict(table_id=t, indices=d_idx[t], offsets=d_off[t], out=d_out[t]) for t in range(T)])
    for _ in range(10):
        plan.launch(stream)
    torch.cuda.synchronize()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(calls)]
    for a, b in ev:
        a.record(stream)
        plan.launch(stream)
        b.record(stream)
    torch.cuda.synchronize()
    ks = [a.elapsed_time(b) * 1e3 for a, b in ev]
    out["plan_launch_kernel"] = {"median_us": round(pct(ks, 0.5), 2), "p10_us": round(pct(ks, 0.1), 2),
                                 "p90_us": round(pct(ks, 0.9), 2), "grid": plan.grid}
    res_dev = d_out.cpu().numpy()
    plan.close()

    out["max_abs_diff_vs_cpu_float"] = float(max(np.abs(res_pageable[t] - ref[t]).max() for t in range(T)))
    out["pageable_pinned_device_bits_equal"] = bool(all(
        np.array_equal(res_pageable[t].view(np.uint32), res_pinned[t].view(np.uint32)) and
        np.array_equal(res_pageable[t].view(np.uint32), res_dev[t].view(np.uint32)) for t in range(T)))
    out["speedup_vs_cpu_all_cores"] = round(out["cpu_all_cores"]["median_us"] / out["lookup_pageable"]["median_us"], 2)
    for a in p_idx + p_off + p_res:
        emblib.free_pinned(a)
    emblib.shutdown()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--presets", nargs="*", default=list(PRESETS))
    ap.add_argument("--calls", type=int, default=300)
    ap.add_argument("--max-rows", type=int, default=0, help="cap rows per table (0 = the preset's cardinalities)")
    args = ap.parse_args()
    for name in args.presets:
        print(json.dumps(run(name, PRESETS[name], args.calls, args.max_rows)), flush=True)


if __name__ == "__main__":
    main()
