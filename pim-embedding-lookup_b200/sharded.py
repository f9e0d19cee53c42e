"""Embedding tables sharded over the GPUs of one box, one process per GPU.

What the reference does with `dpu_id = table_id*NR_COLS + col` (upmem/include/emb_host.h:167:
every (table, column) on its own DPU) and the DLRM fork does with whole tables per rank
(PIM-dlrm-new/dlrm_dpu_pytorch.py:470-482) followed by the `alltoall` butterfly
(extend_distributed.py:389-426), for tables of any size:

* `sharding.plan_tables` places whole tables (table-wise) and contiguous row ranges of big or
  hot tables (row-wise) on the ranks;
* small tables may be replicated and shared by bags instead (each replica pools a slice of the
  batch: data-parallel, nothing to add up afterwards);
* every rank pools the WHOLE global batch for the shards it holds, in ONE launch of the pooling
  kernel; whole tables are dequantised in the kernel's epilogue and stored straight into the
  `[B/N, T, D]` slab of the rank that owns each sample (peer memory over NVLink, the fused
  all-to-all of exchange.PeerSlabs); row-range shards skip the indices outside their range and
  store raw int32 partial sums into a second slab `[B/N, n_parts_total, D]`;
* after the completion barrier each rank wrap-adds the partials of every split table and
  dequantises once (`emb_reduce_partials`): int32 addition is associative, so the result is the
  unsharded one bit for bit whatever the placement.

Inputs: the ranks holding a shard of table t need table t's (indices, offsets) for the global
batch in device memory; a row-range shard gets the full list and filters on the fly.

torch.distributed (NCCL) is plumbing: handle exchange at set-up and the completion barrier.
"""
from __future__ import annotations

import ctypes as C

from . import emblib, exchange, sharding
from .dputypes import EMB_Q_OVERLAP, EMB_Q_RAW_SUMS, EMB_TABLE_FIXED


def partial_columns(plan):
    """Column block of every row-range shard in the partial slab: {(table, part): col}, and the
    per-table (first_col, n_parts) for the reduction.  Deterministic, the same on every rank."""
    col, first = {}, {}
    n = 0
    for t in sorted({s.table for s in plan.shards if not s.whole}):
        k = plan.of_table(t)[0].n_parts
        first[t] = (n, k)
        for p in range(k):
            col[(t, p)] = n + p
        n += k
    return col, first, n


class ShardedTables:
    """rows[t] = rows of global table t; dim = NR_COLS; global_bags = bags per table per call
    (the global mini-batch, divisible by the world size)."""

    def __init__(self, rows, dim, global_bags, lookups=None, group=None, hbm_budget=150e9,
                 max_table_bytes=None, level=True, device=None, n_buffers=2, layout="table", rotate=True,
                 **plan_kwargs):
        import torch
        import torch.distributed as dist
        self.group = group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        self.rows, self.dim, self.bags = list(rows), dim, global_bags
        self.rotate = rotate  # ranks start pooling at different bags: one sender per destination at a time
        self.per = exchange.bags_per_rank(global_bags, self.world)
        self.plan = sharding.plan_tables(self.rows, dim, self.world, lookups=lookups, bags=global_bags,
                                         hbm_budget=hbm_budget, max_table_bytes=max_table_bytes, level=level,
                                         **plan_kwargs)
        self.local = sorted(self.plan.of_rank(self.rank), key=lambda s: (s.table, s.part, s.bag_lo))
        # table slots of this rank: one per distinct row range held (several bag slices of a replica
        # share their slot); slots[i] is a representative shard, slot_of[j] the slot of local shard j
        self.slots, self.slot_of, seen = [], [], {}
        for s in self.local:
            key = (s.table, s.row_lo, s.row_hi)
            if key not in seen:
                seen[key] = len(self.slots)
                self.slots.append(s)
            self.slot_of.append(seen[key])
        self.part_col, self.split, self.n_part_cols = partial_columns(self.plan)
        self.lib = emblib.load()
        dev = torch.cuda.current_device() if device is None else device
        emblib.configure(max(1, len(self.slots)), dim, 0, 0, device=dev)
        T = len(self.rows)
        # Output slabs are double-buffered: a peer may already be storing batch k+1 into this rank's
        # slab while this rank still reads batch k (the barrier of batch k+1 orders batch k+2 after
        # that read, so two buffers are enough).
        self.n_buffers = n_buffers
        self.out = exchange.PeerSlabs(global_bags, None, dim, group=group, n_cols=T, n_buffers=n_buffers,
                                      layout=layout)
        self.parts = (exchange.PeerSlabs(global_bags, None, dim, group=group, n_cols=self.n_part_cols,
                                         n_buffers=n_buffers, layout=layout) if self.n_part_cols else None)
        self._reduce_args = [[] for _ in range(n_buffers)]
        self._finish = [None] * n_buffers
        for b in range(n_buffers):
            jobs = []
            if self.parts is not None:
                for t, (c0, k) in self.split.items():
                    parts = [self.parts.out_peer_col(c0 + p, b)[self.rank] for p in range(k)]
                    dst = self.out.out_peer_col(t, b)[self.rank]
                    jobs.append(dict(parts=parts, out=dst, part_stride=self.parts.stride, out_stride=self.out.stride,
                                     nr_bags=self.per, is_raw_i32=1))
                    arr = (C.c_void_p * k)(*parts)
                    self._reduce_args[b].append((arr, k, self.parts.stride, C.c_void_p(dst), self.out.stride,
                                                 self.per, 1))
            if self.out.barrier_kind == "kernel":
                self._finish[b] = self.out.make_finish(jobs)  # signal + wait + all reductions: one launch

    # ---- tables ------------------------------------------------------------------------------
    def populate_shard(self, slot, data, device_src=False, stream=None):
        """Load table slot `slot` (index into self.slots): rows [row_lo, row_hi) of global table
        slots[slot].table as int32 fixed point, row-major; `data` is a NumPy array or (device_src) a
        CUDA tensor."""
        s = self.slots[slot]
        if device_src:
            assert tuple(data.shape) == (s.rows, self.dim)
            emblib.check(self.lib.emb_populate_table_device(slot, s.rows, data.data_ptr(), 0, EMB_TABLE_FIXED,
                                                            emblib._stream(stream)), "emb_populate_table_device")
        else:
            assert data.shape == (s.rows, self.dim)
            emblib.populate_table_i32(slot, data)
        if not s.whole:
            emblib.check(self.lib.emb_table_set_row_base(slot, s.row_lo), "emb_table_set_row_base")

    def tables_needed(self):
        """Global ids of the tables whose (indices, offsets) this rank must be given."""
        return sorted({s.table for s in self.local})

    # ---- lookups ------------------------------------------------------------------------------
    def queries(self, indices, offsets, buffer=0, overlap=False):
        """Query dicts (emblib.Plan / emblib.lookup_device) for one global batch: indices[t] /
        offsets[t] are CUDA int32/uint32 tensors of table t for every t in tables_needed().  The
        pooled rows land in output buffer `buffer` (alternate 0, 1, ... between consecutive batches).

        overlap=True (EMB_Q_OVERLAP, plans only) is a promise: the index / offset tensors and the tables are not
        written by whatever kernel runs in front of the pooling launch on its stream (they arrive by copy, or were
        filled earlier).  The launch then gathers while the previous batch's pooling tail and completion wait are
        still running and holds back only its stores -- the buffer rule of finish() is unchanged."""
        qs = []
        for j, s in enumerate(self.local):
            slot = self.slot_of[j]
            b_lo, b_hi = (s.bag_lo, s.bag_hi) if s.sliced else (0, self.bags)
            rot = b_lo + (self.rank * self.per) % (b_hi - b_lo) if self.rotate else 0
            if s.whole:  # all rows here (possibly a replica pooling only its slice of the bags): final rows
                q = dict(table_id=slot, indices=indices[s.table], offsets=offsets[s.table], out=None,
                         out_stride=self.out.stride, bags_per_rank=self.per,
                         out_peer=self.out.out_peer_col(s.table, buffer))
                if s.sliced:
                    q.update(bag_first=s.bag_lo, bag_count=s.bag_hi - s.bag_lo)
                q["bag_rotate"] = rot
            else:        # a row range: raw int32 partial sums, added at the receiver
                q = dict(table_id=slot, indices=indices[s.table], offsets=offsets[s.table], out=None,
                         out_stride=self.parts.stride, bags_per_rank=self.per, flags=EMB_Q_RAW_SUMS,
                         out_peer=self.parts.out_peer_col(self.part_col[(s.table, s.part)], buffer), bag_rotate=rot)
            if overlap:
                q["flags"] = q.get("flags", 0) | EMB_Q_OVERLAP
            qs.append(q)
        return qs

    def make_plan(self, indices, offsets, buffer=0, overlap=False):
        p = emblib.Plan(self.queries(indices, offsets, buffer, overlap))
        p.buffer = buffer
        return p

    def finish(self, stream, buffer=0):
        """Completion barrier of the fused exchange, then the partial-sum reduction of the split
        tables, both ordered on `stream` (a torch.cuda.Stream, the one the plan was launched on).
        When it has run, result(buffer) holds this rank's samples.

        Buffer rule (what makes n_buffers = 2 enough): consecutive batches alternate buffers, and whatever
        CONSUMES result(buffer) of batch k is enqueued on `stream` BEFORE finish() of batch k+1.  A peer can only
        start storing batch k+2 (same buffer as k) after its own finish(k+1) has seen this rank's epoch k+1
        signal, which this rank's stream issues after the consumer of batch k.  A consumer on another stream must
        make `stream` wait for it (event) before the next finish()."""
        if self._finish[buffer] is not None:
            self.out.finish(self._finish[buffer], stream)
            return
        self.out.barrier(stream)  # EMB_BARRIER=nccl comparator: all-reduce, then one launch per split table
        st = emblib._stream(stream)
        for arr, k, pstride, dst, ostride, nb, raw in self._reduce_args[buffer]:
            emblib.check(self.lib.emb_reduce_partials(arr, k, pstride, dst, ostride, nb, raw, st),
                         "emb_reduce_partials")

    def lookup(self, plan, stream):
        """One call: pooling launch + exchange + reduction.  Returns result(plan.buffer), valid for work enqueued
        on `stream`; see finish() for the buffer rule when several batches are in flight."""
        plan.launch(stream)
        self.finish(stream, plan.buffer)
        return self.result(plan.buffer)

    def result(self, buffer=0):
        """This rank's pooled rows, torch CUDA tensor fp32 (no copy): [T, B/N, D] for the default layout "table"
        (result()[t] = table t's [B/N, D]: the per-table list interact_features concatenates,
        dlrm_s_pytorch.py:480-488); [B/N, T, D] for layout "sample" (see exchange.PeerSlabs for when that is safe)."""
        return self.out.local(buffer)

    def result_sample_major(self, buffer=0):
        """[B/N, T, D] view of result() whatever the layout."""
        return self.out.local_sample_major(buffer)

    def close(self):
        if self.parts is not None:
            self.parts.close()
            self.parts = None
        if self.out is not None:
            self.out.close()
            self.out = None
