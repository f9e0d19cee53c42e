#!/usr/bin/env python
"""Where does the host-pointer lookup() lose time?  C2 shapes, one GPU.  Times the pooling
kernel with its index/offset arrays and its output either in HBM or in page-locked host memory
(read / written over PCIe by the SMs), and the copy engines on the same bytes as one copy per
table vs one copy for all tables.  Measurement aid for DESIGN.md section 6; prints JSON lines.
"""
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
    n_lookups = sum(len(i) for i in idx)
    # one arena per kind, host (page-locked) and device, tables back to back
    tot_i, tot_o = n_lookups, T * B
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

    def plan(in_host, out_host):
        qs = []
        for t in range(T):
            ib = h_idx.ctypes.data if in_host else d_idx.data_ptr()
            ob = h_off.ctypes.data if in_host else d_off.data_ptr()
            rb = h_out.ctypes.data if out_host else d_out.data_ptr()
            qs.append(dict(table_id=t, indices=int(ib + 4 * pos_i[t]), offsets=int(ob + 4 * t * B),
                           out=int(rb + 4 * D * t * B), indices_len=len(idx[t]), nr_batches=B))
        return emblib.Plan(qs)

    def timed(fn, reps=10):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(reps):
            fn()
        torch.cuda.synchronize()
        return (time.perf_counter() - t0) / reps * 1e3

    nb_in, nb_out = 4 * (tot_i + tot_o), 4 * D * tot_o
    for in_host in (False, True):
        for out_host in (False, True):
            p = plan(in_host, out_host)
            ms = timed(lambda: p.launch(stream))
            print(json.dumps({"what": "pooling kernel", "inputs": "host-mapped" if in_host else "hbm",
                              "output": "host-mapped" if out_host else "hbm", "ms": round(ms, 4),
                              "pcie_GBps": round(((nb_in if in_host else 0) + (nb_out if out_host else 0)) / ms / 1e6, 1)}),
                  flush=True)
            p.close()

    # copy engines through the library's own helpers (cudaMemcpyAsync), per table vs whole arena
    s_in, s_out = torch.cuda.Stream(), torch.cuda.Stream()

    def copies(per_table, do_in, do_out):
        def f():
            if do_in:
                if per_table:
                    for t in range(T):
                        lib.emb_memcpy_h2d(C.c_void_p(d_idx.data_ptr() + 4 * int(pos_i[t])),
                                           C.c_void_p(h_idx.ctypes.data + 4 * int(pos_i[t])), 4 * len(idx[t]),
                                           C.c_void_p(s_in.cuda_stream))
                        lib.emb_memcpy_h2d(C.c_void_p(d_off.data_ptr() + 4 * t * B),
                                           C.c_void_p(h_off.ctypes.data + 4 * t * B), 4 * B,
                                           C.c_void_p(s_in.cuda_stream))
                else:
                    lib.emb_memcpy_h2d(C.c_void_p(d_idx.data_ptr()), C.c_void_p(h_idx.ctypes.data), 4 * tot_i,
                                       C.c_void_p(s_in.cuda_stream))
                    lib.emb_memcpy_h2d(C.c_void_p(d_off.data_ptr()), C.c_void_p(h_off.ctypes.data), 4 * tot_o,
                                       C.c_void_p(s_in.cuda_stream))
            if do_out:
                if per_table:
                    for t in range(T):
                        lib.emb_memcpy_d2h(C.c_void_p(h_out.ctypes.data + 4 * D * t * B),
                                           C.c_void_p(d_out.data_ptr() + 4 * D * t * B), 4 * D * B,
                                           C.c_void_p(s_out.cuda_stream))
                else:
                    lib.emb_memcpy_d2h(C.c_void_p(h_out.ctypes.data), C.c_void_p(d_out.data_ptr()), nb_out,
                                       C.c_void_p(s_out.cuda_stream))
        return f

    for per_table in (False, True):
        for do_in, do_out in ((True, False), (False, True), (True, True)):
            t0 = time.perf_counter()
            f = copies(per_table, do_in, do_out)
            ms = timed(f)
            # host time to enqueue alone
            t0 = time.perf_counter()
            f()
            enq = (time.perf_counter() - t0) * 1e3
            torch.cuda.synchronize()
            print(json.dumps({"what": "copy engines", "per_table": per_table, "h2d": do_in, "d2h": do_out,
                              "ms": round(ms, 4), "enqueue_ms": round(enq, 4),
                              "GBps": round(((nb_in if do_in else 0) + (nb_out if do_out else 0)) / ms / 1e6, 1)}),
                  flush=True)
    emblib.shutdown()


if __name__ == "__main__":
    main()
