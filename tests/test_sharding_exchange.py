"""Host-side multi-GPU logic on CPU: the table planner, row-range index routing, and the
table-parallel -> batch-parallel exchange layout over a world_size-2 gloo group.  The pooled
rows themselves come from the oracle here (this is a test; the product has no CPU path)."""
import os
import sys

import numpy as np
import pytest

import emb_oracle as eo

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

# Criteo Terabyte cardinalities (PIM-dlrm-new/tools/visualize.py:964), C4 in SURVEY.md 8(d)
TERABYTE = [39884406, 39043, 17289, 7420, 20263, 3, 7120, 1543, 63, 38532951, 2953546, 403346, 10, 2208, 11938,
            155, 4, 976, 14, 39979771, 25641295, 39664984, 585935, 12972, 108, 36]


def test_planner_whole_tables_balanced():
    from pim_embedding_lookup_b200 import sharding
    plan = sharding.plan_tables([1_000_000] * 26, 16, 2, bags=65536)
    assert len(plan.shards) == 26 and all(s.whole for s in plan.shards)
    assert sorted(len(plan.of_rank(r)) for r in range(2)) == [13, 13]
    assert plan.imbalance() < 1.001


def test_planner_levels_26_tables_over_8_ranks():
    from pim_embedding_lookup_b200 import sharding
    flat = sharding.plan_tables([1_000_000] * 26, 16, 8, bags=65536, level=False)
    assert flat.imbalance() > 1.2          # 4-vs-3 tables
    # 64 MB tables are replicable: the two left-over tables are shared by bags, 4 replicas each
    plan = sharding.plan_tables([1_000_000] * 26, 16, 8, bags=65536, lookups=[688128] * 26)
    assert plan.imbalance() < 1.01
    sliced = sorted({s.table for s in plan.shards if s.sliced})
    assert len(sliced) == 2 and all(s.whole for s in plan.shards)
    for t in sliced:                        # the slices of a table tile the batch, on distinct ranks
        parts = plan.of_table(t)
        assert parts[0].bag_lo == 0 and parts[-1].bag_hi == 65536
        assert all(a.bag_hi == b.bag_lo for a, b in zip(parts, parts[1:]))
        assert all((s.bag_hi - s.bag_lo) % 256 == 0 for s in parts)
        assert len({s.rank for s in parts}) == len(parts)
    # tables too big to replicate are cut by rows instead
    plan = sharding.plan_tables([1_000_000] * 26, 16, 8, bags=65536, lookups=[688128] * 26, replicate_max_bytes=0)
    assert plan.imbalance() < flat.imbalance() and not any(s.sliced for s in plan.shards)
    for t in range(26):                     # every row of every table is held exactly once
        parts = plan.of_table(t)
        assert parts[0].row_lo == 0 and parts[-1].row_hi == 1_000_000
        assert all(a.row_hi == b.row_lo for a, b in zip(parts, parts[1:]))
        assert len({s.rank for s in parts}) == len(parts)


def test_planner_terabyte_tables_fit_and_split():
    from pim_embedding_lookup_b200 import sharding
    dim = 64
    plan = sharding.plan_tables(TERABYTE, dim, 8, bags=65536, lookups=[65536 * 10] * 26, hbm_budget=12e9,
                                max_table_bytes=6e9)
    assert max(plan.hbm_bytes) <= 12e9
    big = [t for t, r in enumerate(TERABYTE) if r * dim * 4 > 6e9]
    assert big and all(len(plan.of_table(t)) >= 2 for t in big)
    assert sum(s.rows for s in plan.shards) == sum(TERABYTE)
    with pytest.raises(MemoryError):
        sharding.plan_tables(TERABYTE, dim, 2, hbm_budget=12e9, max_table_bytes=6e9)


def test_row_range_partials_add_up_to_the_unsharded_result():
    """What emb_table_set_row_base + EMB_Q_RAW_SUMS + emb_reduce_partials compute, restated
    with the oracle: int32 partial sums over row ranges wrap-add to the full sum."""
    from pim_embedding_lookup_b200 import sharding
    rng = np.random.default_rng(0)
    R, D, B = 5000, 16, 300
    q = rng.integers(-2**31, 2**31, size=(R, D), dtype=np.int64).astype(np.int32)
    lens = rng.integers(0, 30, size=B)
    off = np.concatenate([[0], np.cumsum(lens)[:-1]]).astype(np.uint32)
    off[0] = 1  # ignored: bag 0 starts at 0
    idx = rng.integers(0, R, size=int(lens.sum())).astype(np.uint32)
    _, full = eo.lookup_rows([q], [idx], [off], want_sums=True)
    acc = np.zeros((B, D), dtype=np.uint32)
    for lo, hi in [(0, 1200), (1200, 1201), (1201, R)]:
        bi, bo = sharding.bucket_indices(idx, off, lo, hi)
        assert ((bi >= lo) & (bi < hi)).all() and len(bo) == B
        _, part = eo.lookup_rows([q[lo:hi]], [bi - lo], [bo], want_sums=True)
        acc += part[0].view(np.uint32)
    assert np.array_equal(acc.view(np.int32), full[0])


def test_a2a_split_sizes_follow_the_dlrm_fork():
    from pim_embedding_lookup_b200 import exchange
    with pytest.raises(ValueError):
        exchange.bags_per_rank(64, 3)
    send, recv = exchange.a2a_split_sizes(66, [3, 2, 2], 16, rank=1)  # extend_distributed.py:394-405
    assert send == [22 * 2 * 16] * 3 and recv == [22 * 3 * 16, 22 * 2 * 16, 22 * 2 * 16]
    assert exchange.table_ranges([3, 2, 2]) == ([0, 3, 5], 7)


# ---- world_size 2 over gloo ------------------------------------------------------------------------

def _exchange_worker(rank, world, port, tmp):
    import torch
    import torch.distributed as dist
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import emb_oracle as eo2
    from pim_embedding_lookup_b200 import exchange
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    try:
        D, B, R = 8, 12, 50
        tables_per_rank = [2, 1]
        first, t_global = exchange.table_ranges(tables_per_rank)
        rng = np.random.default_rng(42)  # same stream on both ranks: the global problem
        tabs = [rng.integers(-1000, 1000, size=(R, D)).astype(np.int32) for _ in range(t_global)]
        idx = [rng.integers(0, R, size=B * 3).astype(np.uint32) for _ in range(t_global)]
        off = [(np.arange(B) * 3).astype(np.uint32) for _ in range(t_global)]
        mine = range(first[rank], first[rank] + tables_per_rank[rank])
        pooled = eo2.lookup_rows([tabs[t] for t in mine], [idx[t] for t in mine], [off[t] for t in mine])
        per = B // world
        # send buffer [N, B/N, T_local, D]: what the kernel writes with out_stride = T_local*D
        send = torch.from_numpy(np.stack(pooled, axis=1)).reshape(world, per, tables_per_rank[rank], D).contiguous()
        got = exchange.nccl_all_to_all(send, B, tables_per_rank, D)
        want_all = np.stack(eo2.lookup_rows(tabs, idx, off), axis=1)  # [B, T_global, D]
        want = want_all[rank * per:(rank + 1) * per]
        ok = got.shape == (per, t_global, D) and np.array_equal(got.numpy().view(np.uint32), want.view(np.uint32))
        open(os.path.join(tmp, f"ok{rank}"), "w").write("1" if ok else "0")
    finally:
        dist.destroy_process_group()


def test_all_to_all_layout_world2_gloo(tmp_path):
    import socket

    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    mp.spawn(_exchange_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    assert [open(tmp_path / f"ok{r}").read() for r in range(2)] == ["1", "1"]


def test_partial_columns_of_split_tables_are_dense_and_rank_independent():
    from pim_embedding_lookup_b200 import sharding
    from pim_embedding_lookup_b200.sharded import partial_columns
    plan = sharding.plan_tables(TERABYTE, 64, 8, bags=65536, lookups=[65536 * 10] * 26, hbm_budget=12e9,
                                max_table_bytes=6e9)
    col, first, n = partial_columns(plan)
    split = sorted({s.table for s in plan.shards if not s.whole})
    assert sorted(first) == split and n == sum(len(plan.of_table(t)) for t in split)
    assert sorted(col.values()) == list(range(n))          # every block used exactly once
    for t in split:                                         # a table's parts are consecutive blocks, in part order
        c0, k = first[t]
        assert [col[(t, p)] for p in range(k)] == list(range(c0, c0 + k))
    assert all((s.table, s.part) in col for s in plan.shards if not s.whole)
