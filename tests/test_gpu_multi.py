"""Multi-GPU parity of the pooled-output exchange (each case skips on a box with fewer GPUs; run with
`gpurun --gpus N`): ShardedTables at world sizes 1, 2, 4 and 8, and at two GPUs both transports --
peer-slab stores straight from the pooling kernel over NVLink, and NCCL all_to_all_single --
must leave on every rank exactly the oracle's pooled rows of that rank's samples, for all
tables, bit for bit.  Also a row-split table: raw int32 partials from two GPUs added at the
receiver."""
import os
import socket
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _init(rank, world, port, one_gpu):
    """One process per rank.  one_gpu: every rank uses GPU 0 (two processes time-slice it; CUDA IPC, the peer-slab
    stores, the flag barrier of the completion kernel and its reduce jobs work the same between two processes on one
    device as between two devices), and the process group is gloo, because NCCL refuses two ranks on one GPU -- the
    product path needs the group only to hand the IPC handles around."""
    import torch
    import torch.distributed as dist
    dev = 0 if one_gpu else rank
    torch.cuda.set_device(dev)
    if one_gpu:
        dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    else:
        dist.init_process_group("nccl", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world,
                                device_id=torch.device("cuda", rank))
    return dev


def _worker(rank, world, port, tmp, one_gpu=False):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import ctypes as C

    import torch
    import torch.distributed as dist

    import emb_oracle as eo
    from pim_embedding_lookup_b200 import emblib, exchange
    from pim_embedding_lookup_b200.dputypes import EMB_Q_RAW_SUMS
    dev_id = _init(rank, world, port, one_gpu)
    ok = {}
    try:
        D, B, R = 16, 4096, 20000
        tables_per_rank = [3, 2]
        first, t_global = exchange.table_ranges(tables_per_rank)
        rng = np.random.default_rng(7)  # same on both ranks: the global problem
        tabs = [rng.integers(-2**31, 2**31, size=(R, D), dtype=np.int64).astype(np.int32) for _ in range(t_global)]
        idx, off = [], []
        for _ in range(t_global):
            lens = rng.integers(0, 21, size=B)
            off.append(np.concatenate([[0], np.cumsum(lens)[:-1]]).astype(np.uint32))
            idx.append(rng.integers(0, R, size=int(lens.sum())).astype(np.uint32))
        want_all = np.stack(eo.lookup_rows(tabs, idx, off), axis=1)  # [B, T_global, D]
        per = B // world
        want = want_all[rank * per:(rank + 1) * per]

        T_local = tables_per_rank[rank]
        emblib.configure(T_local + 1, D, device=dev_id)
        mine = list(range(first[rank], first[rank] + T_local))
        dev = []
        for lt, t in enumerate(mine):
            emblib.populate_table_i32(lt, tabs[t])
            dev.append((torch.from_numpy(idx[t].view(np.int32)).cuda(), torch.from_numpy(off[t].view(np.int32)).cuda()))
        stream = torch.cuda.current_stream()

        # (1) fused: the kernel stores into every rank's slab, table-major (default) and sample-major
        ok["fused"] = True
        for layout in ("table", "sample"):
            slabs = exchange.PeerSlabs(B, tables_per_rank, D, layout=layout)
            slabs.local().zero_()
            torch.cuda.synchronize()
            dist.barrier()
            plan = emblib.Plan([dict(table_id=lt, indices=dev[lt][0], offsets=dev[lt][1], out=None,
                                     out_stride=slabs.stride, bags_per_rank=slabs.per, out_peer=slabs.out_peer(lt))
                                for lt in range(T_local)])
            plan.launch(stream)
            slabs.barrier()
            torch.cuda.synchronize()
            got = np.ascontiguousarray(slabs.local_sample_major().cpu().numpy())
            ok["fused"] &= bool(np.array_equal(got.view(np.uint32), want.view(np.uint32)))
            plan.close()
            dist.barrier()
            if layout == "table":
                slabs.close()

        # (2) NCCL all_to_all_single of a [N, B/N, T_local, D] send buffer
        send = torch.empty((B, T_local, D), dtype=torch.float32, device="cuda")
        plan = emblib.Plan([dict(table_id=lt, indices=dev[lt][0], offsets=dev[lt][1], out=send[:, lt, :],
                                 out_stride=T_local * D) for lt in range(T_local)])
        plan.launch(stream)
        if one_gpu:  # no NCCL between two ranks on one GPU: the comparator transport is left to the real-GPU case
            ok["nccl"] = True
            torch.cuda.synchronize()
        else:
            got = exchange.nccl_all_to_all(send, B, tables_per_rank, D).cpu().numpy()
            ok["nccl"] = bool(np.array_equal(got.view(np.uint32), want.view(np.uint32)))
        plan.close()

        # (3) table 0 split row-wise over the two ranks: raw int32 partials into the OWNER
        # sample rank's scratch slabs, added there (emb_reduce_partials)
        lo, hi = (0, R // 3) if rank == 0 else (R // 3, R)
        emblib.populate_table_i32(T_local, tabs[0][lo:hi])
        emblib.check(emblib.load().emb_table_set_row_base(T_local, lo))
        parts = exchange.PeerSlabs(B, [1, 1], D, layout="table")  # slab [2, B/N, D]: block r = partial from rank r
        di, do = torch.from_numpy(idx[0].view(np.int32)).cuda(), torch.from_numpy(off[0].view(np.int32)).cuda()
        plan = emblib.Plan([dict(table_id=T_local, indices=di, offsets=do, out=None, out_stride=parts.stride,
                                 bags_per_rank=parts.per, out_peer=parts.out_peer(0), flags=EMB_Q_RAW_SUMS)])
        plan.launch(stream)
        parts.barrier()
        loc = parts.local(dtype="int32")  # [2, B/N, D]: block r = partial from rank r
        res = torch.empty((per, D), dtype=torch.float32, device="cuda")
        arr = (C.c_void_p * 2)(loc[0].data_ptr(), loc[1].data_ptr())
        emblib.check(emblib.load().emb_reduce_partials(arr, 2, parts.stride, res.data_ptr(), D, per, 1,
                                                       emblib._stream(stream)))
        torch.cuda.synchronize()
        ok["rowsplit"] = bool(np.array_equal(res.cpu().numpy().view(np.uint32), want[:, 0, :].view(np.uint32)))
        ok["no_oob"] = emblib.load().emb_check_indices() == 0
        plan.close()
        dist.barrier()
        parts.close()
        slabs.close()
        emblib.shutdown()
    finally:
        open(os.path.join(tmp, f"ok{rank}"), "w").write(repr(ok))
        dist.destroy_process_group()


@pytest.mark.parametrize("one_gpu", [True, False], ids=["two_ranks_on_one_gpu", "two_gpus"])
def test_exchange_two_gpus(tmp_path, one_gpu):
    torch = pytest.importorskip("torch")
    if not one_gpu and torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    mp.spawn(_worker, args=(2, port, str(tmp_path), one_gpu), nprocs=2, join=True)
    for r in range(2):
        got = eval(open(tmp_path / f"ok{r}").read())
        assert got == {"fused": True, "nccl": True, "rowsplit": True, "no_oob": True}, (r, got)


# ---- ShardedTables: planner placement + row-range shards + fused exchange + partial reduction ----

def _sharded_worker(rank, world, port, tmp, one_gpu=False):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import torch
    import torch.distributed as dist

    import emb_oracle as eo
    from pim_embedding_lookup_b200 import emblib
    from pim_embedding_lookup_b200.sharded import ShardedTables
    _init(rank, world, port, one_gpu)
    ok = {}
    try:
        D, B = 16, 2048
        rows = [30000, 17, 5000, 1, 12000, 640, 26000]  # 30000- and 26000-row tables exceed the slab cap below
        rng = np.random.default_rng(11)  # same on every rank: the global problem
        tabs = [rng.integers(-2**31, 2**31, size=(r, D), dtype=np.int64).astype(np.int32) for r in rows]
        batches = []
        for k in range(4):
            idx, off = [], []
            for r in rows:
                lens = rng.integers(0, 13, size=B)
                off.append(np.concatenate([[0], np.cumsum(lens)[:-1]]).astype(np.uint32))
                idx.append(rng.integers(0, r, size=int(lens.sum())).astype(np.uint32))
            batches.append((idx, off))
        per = B // world
        hint = [B * 6] * len(rows)
        hint[4] = B * 60  # planning hint only: a hot small table, which the planner replicates and shares by bags
        st = ShardedTables(rows, D, B, lookups=hint, max_table_bytes=20000 * D * 4,
                           layout="sample" if world == 1 else "table")
        ok["split"] = sorted(st.split) == [0, 6] and all(len(st.plan.of_table(t)) >= 2 for t in (0, 6))
        ok["sliced"] = world == 1 or any(s.sliced for s in st.plan.shards)  # a replica shared by bags
        for slot, s in enumerate(st.slots):
            st.populate_shard(slot, tabs[s.table][s.row_lo:s.row_hi])
        stream = torch.cuda.Stream()
        need = st.tables_needed()
        plans = []
        for k, (idx, off) in enumerate(batches):
            di = {t: torch.from_numpy(idx[t].view(np.int32)).cuda() for t in need}
            do = {t: torch.from_numpy(off[t].view(np.int32)).cuda() for t in need}
            plans.append(st.make_plan(di, do, buffer=k % 2))
        torch.cuda.synchronize()
        dist.barrier()
        good = True
        for k, (idx, off) in enumerate(batches):  # back to back, alternating buffers
            st.lookup(plans[k], stream)
            got = st.result_sample_major(plans[k].buffer)
            stream.synchronize()
            want = np.stack(eo.lookup_rows(tabs, idx, off), axis=1)[rank * per:(rank + 1) * per]
            good &= bool(np.array_equal(got.cpu().numpy().view(np.uint32), want.view(np.uint32)))
        ok["bits"] = good
        # Batches in flight: 8 lookups enqueued back to back with NO host synchronisation in between, each followed
        # by a device-side copy of result(buffer) on the same stream (the consumer).  Ranks run ahead of each other
        # here: a peer stores batch k+1 into buffer (k+1) % 2 while this rank's copy of batch k may still be
        # pending, and the epoch flags of the completion kernel see a peer one epoch ahead.  The rule that makes two
        # buffers enough: the consumer of result(buffer) is enqueued on the SAME stream before the next finish().
        # Run twice: plain plans, then plans made with overlap=True (EMB_Q_OVERLAP: a pooling launch gathers while the
        # previous batch's completion kernel is still waiting for the peers and stores only once it has finished).
        wants = [np.stack(eo.lookup_rows(tabs, idx, off), axis=1)[rank * per:(rank + 1) * per] for idx, off in batches]
        plans_ov = []
        for k, (idx, off) in enumerate(batches):
            di = {t: torch.from_numpy(idx[t].view(np.int32)).cuda() for t in need}
            do = {t: torch.from_numpy(off[t].view(np.int32)).cuda() for t in need}
            plans_ov.append(st.make_plan(di, do, buffer=k % 2, overlap=True))
        torch.cuda.synchronize()
        ok["inflight"] = True
        for which in (plans, plans_ov):
            dist.barrier()
            copies = []
            for k in range(8):
                p = which[k % 4]
                st.lookup(p, stream)
                with torch.cuda.stream(stream):
                    copies.append(st.result_sample_major(p.buffer).clone())
            stream.synchronize()
            ok["inflight"] &= all(bool(np.array_equal(c.cpu().numpy().view(np.uint32), wants[k % 4].view(np.uint32)))
                                  for k, c in enumerate(copies))
        ok["no_oob"] = emblib.load().emb_check_indices() == 0
        dist.barrier()
        for p in plans + plans_ov:
            p.close()
        st.close()
        emblib.shutdown()
    finally:
        open(os.path.join(tmp, f"ok{rank}"), "w").write(repr(ok))
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [1, 2, 4, 8])
def test_sharded_tables(tmp_path, world):
    torch = pytest.importorskip("torch")
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    mp.spawn(_sharded_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    for r in range(world):
        got = eval(open(tmp_path / f"ok{r}").read())
        assert got == {"split": True, "sliced": True, "bits": True, "inflight": True, "no_oob": True}, (r, got)


@pytest.mark.parametrize("world", [2, 4])
def test_sharded_tables_ranks_on_one_gpu(tmp_path, world):
    """The same worker with every rank on GPU 0 (one process per rank): the row-cut tables' partial sums, the peer-slab
    stores, the flag barrier and the reduce jobs of the completion kernel run on a one-GPU box too."""
    pytest.importorskip("torch")
    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    mp.spawn(_sharded_worker, args=(world, port, str(tmp_path), True), nprocs=world, join=True)
    for r in range(world):
        got = eval(open(tmp_path / f"ok{r}").read())
        assert got == {"split": True, "sliced": True, "bits": True, "inflight": True, "no_oob": True}, (r, got)
