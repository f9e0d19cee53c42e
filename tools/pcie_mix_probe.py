#!/usr/bin/env python
"""PCIe both-ways probe at C2 byte counts (78 MB in, 109 MB out), page-locked host memory:
(a) copy engines both ways with each direction cut into k copies;
(b) H2D by the copy engine while the pooling kernel writes its results to mapped host memory.
Measurement aid for DESIGN.md section 6; prints JSON lines."""
import ctypes as C
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    import torch

    from pim_embedding_lookup_b200 import emblib
    w = dict(bench.WORKLOADS["c2"])
    T, R, D, B = w["T"], w["R"], w["D"], w["B"]
    lib = emblib.load()
    emblib.configure(T, D, 0, 0, device=0)
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    gen = torch.Generator(device="cuda").manual_seed(1234)
    lim = int(1e9 / np.sqrt(R))
    for t in range(T):
        a = torch.randint(-lim, lim + 1, (R, D), dtype=torch.int32, device="cuda", generator=gen)
        emblib.check(lib.emb_populate_table_device(t, R, a.data_ptr(), 0, 1, stream.cuda_stream))
        del a
    idx, off = bench.make_batch(w, seed=123)
    tot_i, tot_o = sum(len(i) for i in idx), T * B
    h_idx = emblib.pinned_array(tot_i, np.uint32)
    h_off = emblib.pinned_array(tot_o, np.uint32)
    h_out = emblib.pinned_array((tot_o, D), np.float32)
    d_idx = torch.empty(tot_i, dtype=torch.int32, device="cuda")
    d_off = torch.empty(tot_o, dtype=torch.int32, device="cuda")
    d_out = torch.empty((tot_o, D), dtype=torch.float32, device="cuda")
    pos_i = np.concatenate([[0], np.cumsum([len(i) for i in idx])]).astype(np.int64)
    for t in range(T):
        h_idx[pos_i[t]:pos_i[t + 1]] = idx[t]
        h_off[t * B:(t + 1) * B] = off[t]
    d_idx.copy_(torch.from_numpy(h_idx.view(np.int32)))
    d_off.copy_(torch.from_numpy(h_off.view(np.int32)))
    torch.cuda.synchronize()
    nb_in, nb_out = 4 * tot_i, 4 * D * tot_o
    s_in, s_out = torch.cuda.Stream(), torch.cuda.Stream()
    vp = C.c_void_p

    def timed(fn, reps=10):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(reps):
            fn()
        torch.cuda.synchronize()
        return (time.perf_counter() - t0) / reps * 1e3

    def h2d_chunks(k):
        step = (nb_in // k) & ~255
        for i in range(k):
            n = step if i < k - 1 else nb_in - step * (k - 1)
            lib.emb_memcpy_h2d(vp(d_idx.data_ptr() + i * step), vp(h_idx.ctypes.data + i * step), n, vp(s_in.cuda_stream))

    def d2h_chunks(k):
        step = (nb_out // k) & ~255
        for i in range(k):
            n = step if i < k - 1 else nb_out - step * (k - 1)
            lib.emb_memcpy_d2h(vp(h_out.ctypes.data + i * step), vp(d_out.data_ptr() + i * step), n, vp(s_out.cuda_stream))

    for kin, kout in ((1, 1), (2, 2), (4, 4), (8, 8), (26, 26), (52, 26), (1, 26), (26, 1), (104, 104)):
        def f():
            # interleave enqueues so neither stream starves
            h2d_chunks(kin)
            d2h_chunks(kout)
        ms = timed(f)
        print(json.dumps({"what": "copy engines both ways", "h2d_copies": kin, "d2h_copies": kout, "ms": round(ms, 4),
                          "GBps_sum": round((nb_in + nb_out) / ms / 1e6, 1)}), flush=True)

    # (b) kernel writes results to the host while the copy engine brings (other) inputs in
    def plans(groups, out_host):
        ps = []
        per = (T + groups - 1) // groups
        for g0 in range(0, T, per):
            qs = []
            for t in range(g0, min(T, g0 + per)):
                rb = h_out.ctypes.data if out_host else d_out.data_ptr()
                qs.append(dict(table_id=t, indices=int(d_idx.data_ptr() + 4 * pos_i[t]),
                               offsets=int(d_off.data_ptr() + 4 * t * B), out=int(rb + 4 * D * t * B),
                               indices_len=len(idx[t]), nr_batches=B))
            ps.append(emblib.Plan(qs))
        return ps

    d_idx2 = torch.empty(tot_i, dtype=torch.int32, device="cuda")
    for groups in (1, 4, 13, 26):
        ps = plans(groups, True)
        ms_k = timed(lambda: [p.launch(stream) for p in ps])
        for kin in (1, 26):
            def f():
                step = (nb_in // kin) & ~255
                for i in range(kin):
                    n = step if i < kin - 1 else nb_in - step * (kin - 1)
                    lib.emb_memcpy_h2d(vp(d_idx2.data_ptr() + i * step), vp(h_idx.ctypes.data + i * step), n,
                                       vp(s_in.cuda_stream))
                for p in ps:
                    p.launch(stream)
            ms = timed(f)
            print(json.dumps({"what": "copy-engine H2D || kernel writing mapped results", "kernel_launches": groups,
                              "h2d_copies": kin, "ms": round(ms, 4), "kernel_alone_ms": round(ms_k, 4),
                              "GBps_sum": round((nb_in + nb_out) / ms / 1e6, 1)}), flush=True)
        for p in ps:
            p.close()
    emblib.shutdown()


if __name__ == "__main__":
    main()
