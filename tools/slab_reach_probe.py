#!/usr/bin/env python
"""How much peer memory can the pooling launch's result stores be spread over before they slow down?  Two (or more)
ranks, C2 tables, sample-major peer slabs made artificially wide (rows of K column blocks, only the first T written),
single or double buffered: the bytes stored are the same in every line, only the span of peer memory they land in
changes.  torchrun ... tools/slab_reach_probe.py [--widths 52 208 416] """
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
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--widths", type=int, nargs="*", default=[52, 208, 416])
    ap.add_argument("--layouts", nargs="*", default=["sample"])
    args = ap.parse_args()
    import torch
    import torch.distributed as dist

    from pim_embedding_lookup_b200 import emblib, exchange
    w = bench.WORKLOADS["c2"]
    T, R, D, B = w["T"], w["R"], w["D"], w["B"]
    lib = emblib.load()
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", device_id=torch.device("cuda", rank))
    emblib.configure(T, D, 0, 0, device=rank)
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    gen = torch.Generator(device="cuda").manual_seed(1234 + rank)
    for t in range(T):
        a = torch.randint(-1000, 1000, (R, D), dtype=torch.int32, device="cuda", generator=gen)
        emblib.check(lib.emb_populate_table_device(t, R, a.data_ptr(), 0, 1, stream.cuda_stream))
        del a
    batches = [bench.make_batch(w, seed=123 + 1000 * rank + k) for k in range(2)]
    dev = [([torch.from_numpy(i.view(np.int32)).cuda() for i in idx], [torch.from_numpy(o.view(np.int32)).cuda() for o in off])
           for idx, off in batches]
    per = B // world
    for layout in args.layouts:
        for K in args.widths:
            for nbuf in (1, 2):
                slabs = exchange.PeerSlabs(B, None, D, n_buffers=nbuf, n_cols=K, layout=layout)
                plans = [emblib.Plan([dict(table_id=t, indices=di[t], offsets=do[t], out=None, out_stride=slabs.stride,
                                           bags_per_rank=per, out_peer=slabs.out_peer_col(t, kb % nbuf), bag_rotate=rank * per)
                                      for t in range(T)]) for kb, (di, do) in enumerate(dev)]
                for i in range(4):
                    plans[i % 2].launch(stream)
                    slabs.barrier()
                torch.cuda.synchronize()
                dist.barrier()
                ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
                for i in range(args.steps):
                    ev[i][0].record(stream)
                    plans[i % 2].launch(stream)
                    ev[i][1].record(stream)
                    slabs.barrier()
                torch.cuda.synchronize()
                ms = sorted(a.elapsed_time(b) for a, b in ev)
                t = torch.tensor([ms[len(ms) // 2]], device="cuda")
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                if rank == 0:
                    print(json.dumps({"world": world, "layout": layout, "row_blocks": K, "buffers": nbuf,
                                      "slab_MB_per_buffer": round(slabs.slab_bytes / 1e6, 1),
                                      "peer_MB_written_into_per_step": round(slabs.slab_bytes * (world - 1) / 1e6, 1),
                                      "peer_MB_mapped": round(slabs.slab_bytes * nbuf * (world - 1) / 1e6, 1),
                                      "kernel_ms_median_max_over_ranks": round(float(t.item()), 4)}), flush=True)
                for p in plans:
                    p.close()
                dist.barrier()
                slabs.close()
    emblib.shutdown()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
