#!/usr/bin/env python
"""Hot-row tables under a skewed index distribution, on and off, on one GPU (north_star: "an L2
persisting-access window keeps hot rows resident under skewed index distributions").

Workload: T tables x R rows x dim D, batch B, fixed pooling P, row popularity ~ rank^-alpha over a random permutation
of the row ids (the bench's C5 recipe, truncated to R rows), generated on the device.  Sized so that the hot set does
NOT simply fit in the 126 MB L2 (default: 32 x 4M x dim 128 = 64 GB of tables).

Arms, same tables and index batches (every arm's pooled rows must have the bits of the first arm's):
  plain                  the default kernel, rows where the caller put them
  remap                  emb_table_remap with the order estimated from a SEPARATE sample of the distribution
                         (emb_hot_order on its counts); the kernel translates the caller's ids (4 more bytes per lookup)
  remap+pre              the caller translated its ids itself (emb_remap_indices once per batch, EMB_Q_REMAPPED)
  ... keep/stream/both   emb_table_set_hot_rows with EMB_HOT_KEEP / EMB_HOT_STREAM_COLD / both, the hot prefix sized
                         so that all tables' prefixes together take --hot-mb megabytes

Prints one JSON line per arm (kernel ms = median of --steps launches timed with CUDA events, launches rotate over two
index batches) and a markdown table at the end.  Measurement aid; bench.py is the contract.

    python tools/l2_hot_ab.py [--tables 32 --rows 4000000 --dim 128 --batch 4096 --pool 80 --alpha 1.05]
                              [--hot-mb 32 64 96] [--steps 30] [--arms all|quick] [--tag name]
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def zipf_cdf(torch, R, alpha, device):
    w = torch.arange(1, R + 1, dtype=torch.float64, device=device).pow_(-alpha)
    cdf = torch.cumsum(w, 0)
    return cdf / cdf[-1]


def draw_ranks(torch, cdf, n, gen):
    """n popularity ranks (0 = the most popular row) ~ rank^-alpha, truncated to the table."""
    u = torch.rand(n, dtype=torch.float64, device=cdf.device, generator=gen)
    return torch.searchsorted(cdf, u).clamp_(max=cdf.numel() - 1)


def mix_cdf(torch, R, hot_rows, hot_frac, device):
    """Two classes: a fraction hot_frac of the lookups falls uniformly on the first hot_rows ranks, the rest uniformly
    on all the others -- a hot set with NO reuse between its members and a cold stream with none at all, the case in
    which LRU-like replacement lets the stream push the hot set out once the cold bytes between two uses of a hot
    row approach the cache size."""
    w = torch.empty(R, dtype=torch.float64, device=device)
    w[:hot_rows] = hot_frac / hot_rows
    w[hot_rows:] = (1.0 - hot_frac) / (R - hot_rows)
    cdf = torch.cumsum(w, 0)
    return cdf / cdf[-1]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tables", type=int, default=32)
    ap.add_argument("--rows", type=int, default=4_000_000)
    ap.add_argument("--dim", type=int, default=128)
    ap.add_argument("--batch", type=int, default=4096)
    ap.add_argument("--pool", type=int, default=80)
    ap.add_argument("--alpha", type=float, default=1.05)
    ap.add_argument("--dist", default="zipf", choices=["zipf", "mix"],
                    help="zipf: rank^-alpha; mix: --hot-frac of the lookups uniformly on a hot set of --hot-set-mb "
                         "(all tables together), the rest uniformly on the other rows")
    ap.add_argument("--hot-frac", type=float, default=0.5)
    ap.add_argument("--hot-set-mb", type=float, default=48.0)
    ap.add_argument("--hot-mb", type=float, nargs="*", default=[32.0, 64.0, 96.0])
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--sample-batches", type=int, default=4, help="batches of the separate sample the order is estimated from")
    ap.add_argument("--arms", default="all", choices=["all", "quick"])
    ap.add_argument("--tag", default="")
    args = ap.parse_args()
    import torch

    from pim_embedding_lookup_b200 import emblib
    from pim_embedding_lookup_b200.dputypes import EMB_HOT_KEEP, EMB_HOT_STREAM_COLD, EMB_Q_REMAPPED

    T, R, D, B, P = args.tables, args.rows, args.dim, args.batch, args.pool
    lib = emblib.load()
    emblib.configure(T, D, 0, 0, device=0)
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    gen = torch.Generator(device="cuda").manual_seed(4321)
    lim = int(1e9 / np.sqrt(R))
    for t in range(T):
        a = torch.randint(-lim, lim + 1, (R, D), dtype=torch.int32, device="cuda", generator=gen)
        emblib.check(lib.emb_populate_table_device(t, R, a.data_ptr(), 0, 1, stream.cuda_stream))
        del a
    torch.cuda.empty_cache()

    if args.dist == "zipf":
        cdf = zipf_cdf(torch, R, args.alpha, "cuda")
        dist_name = f"rank^-{args.alpha:g}"
    else:
        hs_rows = max(1, int(args.hot_set_mb * 1e6 / (T * D * 4)))
        cdf = mix_cdf(torch, R, hs_rows, args.hot_frac, "cuda")
        dist_name = (f"{args.hot_frac:g} of the lookups uniform on a hot set of {hs_rows} rows per table "
                     f"({args.hot_set_mb:g} MB in all), the rest uniform on the other rows")
    perms = [torch.randperm(R, device="cuda", generator=gen) for _ in range(T)]
    n = B * P
    off = (torch.arange(B, dtype=torch.int64, device="cuda") * P).to(torch.int32)
    batches = [[perms[t][draw_ranks(torch, cdf, n, gen)].to(torch.int32) for t in range(T)] for _ in range(2)]
    # the order is estimated from OTHER draws of the same distribution, as a deployment would from yesterday's traffic
    new_of_old = []
    for t in range(T):
        counts = torch.zeros(R, dtype=torch.int64, device="cuda")
        for _ in range(args.sample_batches):
            counts += torch.bincount(perms[t][draw_ranks(torch, cdf, n, gen)], minlength=R)
        if t == 0:  # the library's own helper on one table; a device argsort (same order: stable, descending) for the rest
            m0 = emblib.hot_order(counts.cpu().numpy().astype(np.uint64))
        order = torch.argsort(counts, descending=True, stable=True)
        m = torch.empty(R, dtype=torch.int64, device="cuda")
        m[order] = torch.arange(R, dtype=torch.int64, device="cuda")
        m = m.to(torch.int32).cpu().numpy().view(np.uint32)
        if t == 0:
            assert np.array_equal(m, m0), "emb_hot_order disagrees with a stable descending argsort"
        new_of_old.append(m)
        del counts, order
    # how much of the traffic the prefix holds, by construction of the sample
    torch.cuda.synchronize()

    out = torch.empty((T, B, D), dtype=torch.float32, device="cuda")
    results, ref_bits = [], [None]
    lookups = T * n
    alg_bytes = (lookups + T * B) * (4 * D + 4)

    def measure(name, idx_sets, flags, note):
        plans = [emblib.Plan([dict(table_id=t, indices=idx[t], offsets=off, out=out[t], flags=flags) for t in range(T)])
                 for idx in idx_sets]
        for i in range(4):
            plans[i % 2].launch(stream)
        torch.cuda.synchronize()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
        for i in range(args.steps):
            ev[i][0].record(stream)
            plans[i % 2].launch(stream)
            ev[i][1].record(stream)
        torch.cuda.synchronize()
        ms = sorted(a.elapsed_time(b) for a, b in ev)
        plans[0].launch(stream)  # parity: batch 0's rows, every arm
        torch.cuda.synchronize()
        chk = int(out.view(torch.int32).to(torch.int64).sum().item())
        if ref_bits[0] is None:
            ref_bits[0] = (chk, out.clone())
        same = chk == ref_bits[0][0] and bool(torch.equal(out.view(torch.int32), ref_bits[0][1].view(torch.int32)))
        for p in plans:
            p.close()
        med = ms[len(ms) // 2]
        rec = {"arm": name, "kernel_ms_median": round(med, 4), "kernel_ms_min": round(ms[0], 4),
               "lookups_per_s": lookups / med * 1e3, "algorithmic_GBps": round(alg_bytes / med / 1e6, 1),
               "same_bits_as_plain": same, "index_errors": int(lib.emb_check_indices()), "note": note}
        print(json.dumps(rec), flush=True)
        results.append(rec)
        if not same:
            raise SystemExit(f"PARITY FAILURE in arm {name}")
        return med

    measure("plain", batches, 0, "default kernel, rows in the caller's order")
    for t in range(T):
        emblib.remap_table(t, new_of_old[t])
    measure("remap", batches, 0, "rows reordered hot-first; the kernel translates the caller's ids")
    pre = [[b[t].clone() for t in range(T)] for b in batches]
    for b in pre:
        for t in range(T):
            emblib.remap_indices(t, b[t], stream)
    torch.cuda.synchronize()
    measure("remap+pre", pre, EMB_Q_REMAPPED, "rows reordered hot-first; ids translated by the caller")
    # the cost of the load FORM alone: one hot row per table, so that every load carries a policy operand
    # (evict_normal for all but that row) and nothing else changes
    for t in range(T):
        emblib.set_hot_rows(t, 1, EMB_HOT_KEEP)
    measure("remap+pre+form", pre, EMB_Q_REMAPPED, "every row load carries an L2 policy operand, all of them evict_normal")
    modes = [("keep", EMB_HOT_KEEP), ("stream", EMB_HOT_STREAM_COLD), ("both", EMB_HOT_KEEP | EMB_HOT_STREAM_COLD)]
    if args.arms == "quick":
        modes = modes[2:]
    for mb in args.hot_mb:
        hot_rows = max(1, int(mb * 1e6 / (T * D * 4)))
        for mname, mode in modes:
            for t in range(T):
                emblib.set_hot_rows(t, hot_rows, mode)
            measure(f"remap+pre+{mname}@{mb:g}MB", pre, EMB_Q_REMAPPED,
                    f"hot prefix {hot_rows} rows per table ({mb:g} MB over all tables), mode {mname}")
        if args.arms == "all" or mb == args.hot_mb[-1]:
            measure(f"remap+{modes[-1][0]}@{mb:g}MB", batches, 0,
                    f"the same with the kernel translating the ids, mode {modes[-1][0]}")
    base = results[0]["kernel_ms_median"]
    cfg = f"{T} x {R} rows x dim {D}, batch {B}, pooling {P}, {dist_name}; tables {T * R * D * 4 / 1e9:.1f} GB"
    print(f"\n### {args.tag or 'l2_hot_ab'}: {cfg}\n")
    print("| arm | kernel ms (median) | vs plain | G lookups/s | algorithmic GB/s | same bits |")
    print("|---|---|---|---|---|---|")
    for r in results:
        print(f"| {r['arm']} | {r['kernel_ms_median']:.4f} | {r['kernel_ms_median'] / base:.3f} | "
              f"{r['lookups_per_s'] / 1e9:.2f} | {r['algorithmic_GBps']:.0f} | {r['same_bits_as_plain']} |")
    emblib.shutdown()


if __name__ == "__main__":
    main()
