#!/usr/bin/env python
"""BASELINE.json configs[2]: end-to-end DLRM inference samples/s, Kaggle arch
(sparse dim 16, bot 13-512-256-64-16, top 367-512-256-1), 26 tables with the Kaggle cardinalities
(PIM-dlrm-new/tools/visualize.py:949), batch 39292 (README.md:14), one index per lookup, uniform
synthetic indices, random-init weights.  The MLPs and the interaction are plain torch (out of scope
for this repo); what is compared is how the embedding lookups reach them:

  host      the reference's flow (dlrm_dpu_pytorch.py:1028-1066): host index arrays -> lookup() through
            the EmbeddingBag kwarg shim -> host result tensors -> .cuda() -> interaction -> top MLP
  device    SURVEY 8f-2: indices copied to the GPU, DeviceBags pools straight into the [B, 27, 16]
            interaction input, nothing else crosses PCIe
  cpu       dlrm_s_pytorch.py flow: everything on the host cores (nn.EmbeddingBag per table)

Prints one JSON line.  Measurement aid; bench.py stays the contract."""
import argparse
import ctypes as C
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

LN_EMB = [1460, 583, 10131227, 2202608, 305, 24, 12517, 633, 3, 93145, 5683, 8351593, 3194, 27, 14992, 5461306, 10,
          5652, 2173, 4, 7046547, 18, 15, 286181, 105, 142572]
LN_BOT = [13, 512, 256, 64, 16]
LN_TOP = [367, 512, 256, 1]


def mlp(ln, sigmoid_last):
    import torch
    layers = []
    for i in range(len(ln) - 1):
        layers.append(torch.nn.Linear(ln[i], ln[i + 1]))
        layers.append(torch.nn.Sigmoid() if (sigmoid_last and i == len(ln) - 2) else torch.nn.ReLU())
    return torch.nn.Sequential(*layers)


def interact(x, ly):
    """dot interaction, dlrm_s_pytorch.py interact_features: x [B,D], ly [B,T,D] -> [B, D + (T+1)T/2]"""
    import torch
    B, D = x.shape
    Tm = torch.cat([x.unsqueeze(1), ly], dim=1)
    Z = torch.bmm(Tm, Tm.transpose(1, 2))
    n = Tm.shape[1]
    li, lj = torch.tril_indices(n, n, offset=-1, device=x.device)
    return torch.cat([x, Z[:, li, lj]], dim=1)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=39292)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--cpu-steps", type=int, default=3)
    args = ap.parse_args()
    import torch

    import pim_embedding_lookup_b200 as pel
    from pim_embedding_lookup_b200 import emblib, embedding_bag
    torch.manual_seed(0)
    rng = np.random.default_rng(123)
    T, D, B = len(LN_EMB), 16, args.batch
    lib = emblib.load()
    emblib.configure(T, D, B, 1, device=0)
    embedding_bag.install()
    dev = torch.device("cuda", 0)

    # model (create_emb: W ~ U(-1/sqrt(n), 1/sqrt(n)); export_emb on the GPU)
    bot_c, top_c = mlp(LN_BOT, False), mlp(LN_TOP, True)
    bot_g, top_g = mlp(LN_BOT, False).to(dev), mlp(LN_TOP, True).to(dev)
    bot_g.load_state_dict(bot_c.state_dict())
    top_g.load_state_dict(top_c.state_dict())
    emb_cpu = []
    t0 = time.perf_counter()
    for k, n in enumerate(LN_EMB):
        W = rng.uniform(-np.sqrt(1 / n), np.sqrt(1 / n), size=(n, D)).astype(np.float32)
        emb_cpu.append(torch.nn.EmbeddingBag.from_pretrained(torch.from_numpy(W), mode="sum", freeze=True))
        emblib.populate_table_f32(k, W)
    populate_s = time.perf_counter() - t0

    # one batch (RandomDataset: dense ~ U[0,1), P=1 uniform indices)
    X = torch.from_numpy(rng.random((B, 13), dtype=np.float32)).pin_memory()
    lS_i = [torch.from_numpy(rng.integers(0, n, size=B).astype(np.int64)) for n in LN_EMB]
    lS_o = [torch.arange(B, dtype=torch.int64) for _ in LN_EMB]

    # ---- cpu flow
    torch.set_num_threads(len(os.sched_getaffinity(0)))

    def fwd_cpu():
        with torch.no_grad():
            x = bot_c(X)
            ly = torch.stack([emb_cpu[k](lS_i[k], lS_o[k]) for k in range(T)], dim=1)
            return top_c(interact(x, ly))
    p_cpu = fwd_cpu()
    t0 = time.perf_counter()
    for _ in range(args.cpu_steps):
        fwd_cpu()
    cpu_ms = (time.perf_counter() - t0) / args.cpu_steps * 1e3

    # ---- host flow (the reference's): page-locked uint32 index arrays + registered result tensors
    h_idx = [emblib.pinned_array(B, np.uint32) for _ in range(T)]
    h_off = [emblib.pinned_array(B, np.uint32) for _ in range(T)]
    for k in range(T):
        h_idx[k][:] = lS_i[k].numpy()
        h_off[k][:] = lS_o[k].numpy()
    res = torch.zeros((T, B, D), dtype=torch.float32).pin_memory()
    ind_p = (C.c_uint64 * T)(*[a.ctypes.data for a in h_idx])
    off_p = (C.c_uint64 * T)(*[a.ctypes.data for a in h_off])
    res_p = (C.c_uint64 * T)(*[res[k].data_ptr() for k in range(T)])
    lookup_ms = []

    def fwd_host():
        with torch.no_grad():
            x = bot_g(X.to(dev, non_blocking=True))
            t1 = time.perf_counter()
            emb_cpu[0](lS_i[0], lS_o[0], num_of_tables=T, dpu_set_ptr=0, use_dpu=True,
                       final_results_ptr=C.addressof(res_p), indices_ptr=C.addressof(ind_p),
                       offsets_ptr=C.addressof(off_p), latency_print=0)
            lookup_ms.append((time.perf_counter() - t1) * 1e3)
            ly = res.to(dev, non_blocking=True).permute(1, 0, 2)
            return top_g(interact(x, ly)).cpu()
    p_host = fwd_host()
    for _ in range(3):
        fwd_host()
    torch.cuda.synchronize()
    lookup_ms.clear()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        fwd_host()
    host_ms = (time.perf_counter() - t0) / args.steps * 1e3
    host_lookup_ms = float(np.mean(lookup_ms))
    host_path = int(emblib.last_timings().path)

    # ---- device flow: int32 indices cross PCIe once, pooled rows never do
    bags = embedding_bag.DeviceBags(T, D)
    i32 = torch.stack([torch.from_numpy(a.view(np.int32)) for a in h_idx]).pin_memory()   # [T, B]
    o32 = torch.stack([torch.from_numpy(a.view(np.int32)) for a in h_off]).pin_memory()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps + 4)]
    it = [0]

    def fwd_dev():
        with torch.no_grad():
            x = bot_g(X.to(dev, non_blocking=True))
            di = i32.to(dev, non_blocking=True)
            do = o32.to(dev, non_blocking=True)
            a, b = ev[it[0] % len(ev)]
            it[0] += 1
            a.record()
            ly = bags([di[k] for k in range(T)], [do[k] for k in range(T)])
            b.record()
            return top_g(interact(x, ly)).cpu()
    p_dev = fwd_dev()
    for _ in range(3):
        fwd_dev()
    torch.cuda.synchronize()
    it[0] = 0
    t0 = time.perf_counter()
    for _ in range(args.steps):
        fwd_dev()
    dev_ms = (time.perf_counter() - t0) / args.steps * 1e3
    torch.cuda.synchronize()
    dev_lookup_ms = float(np.mean([a.elapsed_time(b) for a, b in ev[:min(args.steps, len(ev))]]))

    print(json.dumps({
        "config": "C3: Kaggle arch, 26 Kaggle-cardinality tables x dim 16 (2.16 GB), batch %d, 1 index per lookup" % B,
        "metric": "DLRM inference samples/s (end to end, host batch in, click probabilities on the host)",
        "samples_per_s": {"host_flow": B / host_ms * 1e3, "device_flow": B / dev_ms * 1e3, "cpu_flow": B / cpu_ms * 1e3},
        "ms_per_batch": {"host_flow": host_ms, "device_flow": dev_ms, "cpu_flow": cpu_ms},
        "lookup_ms": {"host_flow(lookup() wall)": host_lookup_ms, "device_flow(kernel, CUDA events)": dev_lookup_ms},
        "host_path": host_path, "cpu_threads": torch.get_num_threads(), "populate_s": populate_s,
        "max_abs_diff_vs_cpu": {"host_flow": float((p_host - p_cpu).abs().max()),
                                "device_flow": float((p_dev - p_cpu).abs().max())},
        "host_vs_device_bits_equal": bool(torch.equal(p_host, p_dev)),
        "version": pel.__version__}))
    embedding_bag.uninstall()
    emblib.shutdown()


if __name__ == "__main__":
    main()
