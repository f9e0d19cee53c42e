#!/usr/bin/env python
"""Where does the multi-GPU pooling launch lose its ~6 % (C2, 0.276 ms alone vs 0.291-0.296 ms at
N=2..8)?  One GPU, same tables and batches, only the OUTPUT addressing changes:

  tbd    [T][B][D]            one contiguous [B, D] result per table (the reference's layout)
  btd    [B][T][D]            sample-major rows of stride T*D (the exchange layout), one slab
  slabs  N x [B/N][T][D]      the exchange layout cut into N per-rank slabs, all LOCAL memory

If btd/slabs are as slow as the N-GPU runs the layout is the cost; if they match tbd it is the
NVLink stores.  Measurement aid for DESIGN.md section 8.   python tools/route_probe.py [--ranks 8]
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
    ap.add_argument("--ranks", type=int, default=8)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--wide", type=int, nargs="*", default=[],
                    help="extra modes: sample-major rows as wide as this many tables (the stride an N-rank exchange has: "
                         "N x T tables), only the first T columns written")
    args = ap.parse_args()
    import torch

    from pim_embedding_lookup_b200 import emblib
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
    dev = [([torch.from_numpy(i.view(np.int32)).cuda() for i in idx], [torch.from_numpy(o.view(np.int32)).cuda() for o in off])
           for idx, off in batches]
    N = args.ranks
    per = B // N
    tbd = torch.empty((T, B, D), dtype=torch.float32, device="cuda")
    btd = torch.empty((B, T, D), dtype=torch.float32, device="cuda")
    slabs = [torch.empty((per, T, D), dtype=torch.float32, device="cuda") for _ in range(N)]

    wide = {f"wide{k}": torch.empty((B // min(8, max(1, k // T)), k, D), dtype=torch.float32, device="cuda") for k in args.wide}

    def queries(mode, di, do):
        qs = []
        for t in range(T):
            if mode in wide:  # rows of a [B/N][k][D] slab: the bags wrap around it (addresses, not values, matter here)
                buf = wide[mode]
                nb = buf.shape[0]
                qs.append(dict(table_id=t, indices=di[t], offsets=do[t], out=None, out_stride=buf.shape[1] * D,
                               bags_per_rank=nb, out_peer=[buf[:, t, :]] * (B // nb)))
            elif mode == "tbd":
                qs.append(dict(table_id=t, indices=di[t], offsets=do[t], out=tbd[t]))
            elif mode == "btd":
                qs.append(dict(table_id=t, indices=di[t], offsets=do[t], out=btd[:, t, :], out_stride=T * D))
            else:
                qs.append(dict(table_id=t, indices=di[t], offsets=do[t], out=None, out_stride=T * D, bags_per_rank=per,
                               out_peer=[s[:, t, :] for s in slabs]))
        return qs

    nl = np.mean([sum(len(i) for i in b[0]) for b in batches])
    ref = None
    for mode in ["tbd", "btd", "slabs"] + list(wide) + ["tbd"]:
        plans = [emblib.Plan(queries(mode, di, do)) for di, do in dev]
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
        if mode in wide:  # later "ranks" overwrite earlier ones: nothing to compare
            chk = ref
        else:
            got = {"tbd": tbd.permute(1, 0, 2), "btd": btd, "slabs": torch.cat(slabs, 0)}[mode]
            chk = got.contiguous().view(torch.int32).to(torch.int64).sum().item()
        ref = chk if ref is None else ref
        print(json.dumps({"mode": mode, "ranks": N, "kernel_ms_median": round(ms[len(ms) // 2], 4),
                          "kernel_ms_min": round(ms[0], 4), "G_lookups_per_s": round(nl / ms[len(ms) // 2] / 1e6, 2),
                          "same_bits": chk == ref}), flush=True)
        for p in plans:
            p.close()
    emblib.shutdown()


if __name__ == "__main__":
    main()
