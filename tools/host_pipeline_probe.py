#!/usr/bin/env python
"""Timeline of bench.run_host_pipeline on one GPU (C2 tables, no exchange): per step the start / end of the copy-in,
the pooling launch and the copy-out, in ms since the first event, next to the serial loop (copy-in, launch, copy-out,
synchronise).  Measurement aid for DESIGN.md section 6.      python tools/host_pipeline_probe.py [--steps 8]"""
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
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--workload", default="c2")
    args = ap.parse_args()
    import torch
    import torch.distributed as dist

    from pim_embedding_lookup_b200 import emblib, exchange
    w = bench.WORKLOADS[args.workload]
    T, R, D, B = w["T"], w["R"], w["D"], w["B"]
    lib = emblib.load()
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    torch.cuda.set_device(rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", rank))
    emblib.configure(T, D, 0, 0, device=rank)
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    gen = torch.Generator(device="cuda").manual_seed(1234)
    rows_t = w.get("rows", [R] * T)
    for t in range(T):
        a = torch.randint(-1000, 1000, (rows_t[t], D), dtype=torch.int32, device="cuda", generator=gen)
        emblib.check(lib.emb_populate_table_device(t, rows_t[t], a.data_ptr(), 0, 1, stream.cuda_stream))
        del a
    batches = [bench.make_batch(w, seed=123 + k) for k in range(2)]
    sets = []
    slabs = exchange.PeerSlabs(B, [T] * world, D, layout="sample", n_buffers=2) if world > 1 else None
    per = B // world
    for kb, (idx, off) in enumerate(batches):
        n_words = sum(len(i) for i in idx) + sum(len(o) for o in off)
        harena = torch.empty(n_words, dtype=torch.int32).pin_memory()
        darena = torch.empty(n_words, dtype=torch.int32, device="cuda")
        out = torch.empty((B, T, D), dtype=torch.float32, device="cuda")
        pos, di, do = 0, [], []
        for src, dst in ((idx, di), (off, do)):
            for a in src:
                harena[pos:pos + len(a)] = torch.from_numpy(a.view(np.int32))
                dst.append(darena[pos:pos + len(a)])
                pos += len(a)
        if slabs is None:
            plan = emblib.Plan([dict(table_id=t, indices=di[t], offsets=do[t], out=out[:, t, :], out_stride=T * D) for t in range(T)])
        else:
            plan = emblib.Plan([dict(table_id=t, indices=di[t], offsets=do[t], out=None, out_stride=slabs.stride, bags_per_rank=per,
                                     out_peer=slabs.out_peer(t, kb), bag_rotate=rank * per) for t in range(T)])
            out = slabs.local_sample_major(kb)
        sets.append((harena, darena, plan, out))
    res = [torch.empty(tuple(sets[0][3].shape), dtype=torch.float32).pin_memory() for _ in range(2)]
    finish = (lambda k: None) if slabs is None else (lambda k: slabs.barrier())
    n = args.steps
    marks = {k: [torch.cuda.Event(enable_timing=True) for _ in range(6)] for k in range(n)}
    cur = torch.cuda.current_stream

    def copy_in(k):
        marks[k][0].record(cur())
        sets[k % 2][1].copy_(sets[k % 2][0], non_blocking=True)
        marks[k][1].record(cur())

    def launch(k):
        marks[k][2].record(cur())
        sets[k % 2][2].launch(stream)
        marks[k][3].record(cur())

    def copy_out(k):
        marks[k][4].record(cur())
        res[k % 2].copy_(sets[k % 2][3], non_blocking=True)
        marks[k][5].record(cur())

    for name in ("serial", "pipelined"):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        if name == "serial":
            for k in range(n):
                copy_in(k), launch(k), finish(k), copy_out(k)
                stream.synchronize()
        else:
            bench.run_host_pipeline(torch, stream, n, copy_in, launch, finish, copy_out)
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) / n * 1e3
        base = marks[0][0]
        tl = [[round(base.elapsed_time(e), 3) for e in marks[k]] for k in range(n)]
        if world > 1:
            dist.barrier()
        if rank == 0:
            print(json.dumps({"mode": name, "world": world, "ms_per_step": round(dt, 3),
                              "timeline_ms [in0,in1,pool0,pool1,out0,out1]": tl}))
    if slabs is not None:
        torch.cuda.synchronize()
        dist.barrier()
        slabs.close()
    emblib.shutdown()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
