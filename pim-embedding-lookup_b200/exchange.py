"""Pooled-output exchange between the GPUs of one box: table-parallel -> batch-parallel.

Same meaning as the DLRM fork's `alltoall` (PIM-dlrm-new/extend_distributed.py:389-426,
called from dlrm_dpu_pytorch.py:804): rank r pools ALL B samples for ITS tables and every
sample's pooled rows must end up on the data-parallel rank that owns the sample
(samples [r*B/N, (r+1)*B/N) on rank r).

Two transports:

* `PeerSlabs` (fused): every rank owns one output slab [T_global, B/N, D] (or sample-major
  [B/N, T_global, D]) in cudaMalloc memory, exported with CUDA IPC and mapped by all peers.  The
  pooling kernel itself stores bag b of table t into the slab of rank b // (B/N), row b % (B/N) of
  table t (emb_dev_query.bags_per_rank / out_peer, include/emb_b200.h), so the rows cross NVLink
  as they are produced and the only collective left is the completion step (emb_finish_launch:
  flags in peer memory).
* `nccl_all_to_all`: pool into a local [N, B/N, T_local, D] buffer, one
  `all_to_all_single`, comparator and fallback.

One process per GPU; `torch.distributed` (NCCL on GPUs, gloo in the CPU tests) is the
plumbing.  Host-side layout arithmetic lives in free functions so it is testable without
a GPU.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from .dputypes import EmbIpcHandle


def bags_per_rank(global_bags: int, world: int) -> int:
    if global_bags % world:
        raise ValueError(f"global batch {global_bags} must divide over {world} ranks")
    return global_bags // world


def table_ranges(tables_per_rank):
    """tables_per_rank[r] = number of tables rank r owns -> (first_global_table[r], T_global)."""
    first = np.concatenate([[0], np.cumsum(tables_per_rank)]).astype(np.int64)
    return first[:-1].tolist(), int(first[-1])


def a2a_split_sizes(global_bags, tables_per_rank, dim, rank):
    """Element counts for all_to_all_single of a [N, B/N, T_local, D] send buffer
    (extend_distributed.py:394-405: table_split_lengths / batch_split_lengths)."""
    world = len(tables_per_rank)
    per = bags_per_rank(global_bags, world)
    send = [per * tables_per_rank[rank] * dim] * world
    recv = [per * tables_per_rank[r] * dim for r in range(world)]
    return send, recv


def assemble_received(recv_flat, global_bags, tables_per_rank, dim):
    """Received chunks (rank-major, each [B/N, T_r, D]) -> [B/N, T_global, D]."""
    import torch
    world = len(tables_per_rank)
    per = bags_per_rank(global_bags, world)
    parts, pos = [], 0
    for r in range(world):
        n = per * tables_per_rank[r] * dim
        parts.append(recv_flat[pos:pos + n].view(per, tables_per_rank[r], dim))
        pos += n
    return torch.cat(parts, dim=1)


def nccl_all_to_all(send, global_bags, tables_per_rank, dim, group=None):
    """send: [N, B/N, T_local, D] contiguous (slice r goes to rank r).  Returns [B/N, T_global, D]."""
    import torch
    import torch.distributed as dist
    rank = dist.get_rank(group)
    s, r = a2a_split_sizes(global_bags, tables_per_rank, dim, rank)
    recv = torch.empty(sum(r), dtype=send.dtype, device=send.device)
    dist.all_to_all_single(recv, send.reshape(-1), output_split_sizes=r, input_split_sizes=s, group=group)
    return assemble_received(recv, global_bags, tables_per_rank, dim)


class _DevView:
    """Raw device pointer as a __cuda_array_interface__ object (for torch.as_tensor)."""

    def __init__(self, ptr, shape, typestr):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (int(ptr), False),
                                         "version": 2, "strides": None}


class PeerSlabs:
    """Symmetric output slabs mapped across all ranks of `group` (one node, NVLink/NVSwitch)."""

    def __init__(self, global_bags, tables_per_rank, dim, group=None, n_buffers=1, n_cols=None, layout="table"):
        """layout "table" (default): slab = [T_global][B/N][D], one contiguous [B/N, D] block per table -- the list
        of per-table tensors `interact_features` concatenates (dlrm_s_pytorch.py:480-488; on one GPU apply_emb
        returns exactly that list).  layout "sample": slab = [B/N][T_global][D], sample-major rows.

        Why table-major is the default (profiles/r2_scaling.md): a pooling launch walks the tables one after the
        other, so with table-major slabs the rows in flight land in a narrow window of every peer's slab, while with
        sample-major slabs they are 64-byte pieces spread over the WHOLE slab of every peer.  Once the peer memory
        written into that way (all peers, all buffers that alternate) passes ~1 GB the stores thrash the GPU's
        address translation for peer mappings: C2 at 8 ranks, double buffered (7 x 2 x 109 MB) 0.479 ms per launch
        against 0.226 ms table-major and 0.258 ms single buffered; dim-64 tables at 8 ranks (7 x 2 x 436 MB) 18 ms.
        tools/slab_reach_probe.py reproduces it on two GPUs with artificially wide slabs.  Sample-major stays
        available for small slabs (it also costs 3.7 % on one GPU: partial-line writes).
        tables_per_rank[r] = tables rank r owns (consecutive global ids, rank-major), or None with
        n_cols = number of D-wide column blocks per sample when the placement is arbitrary
        (sharded.ShardedTables addresses the blocks itself through out_peer_col)."""
        import torch
        import torch.distributed as dist
        from . import emblib
        self.lib = emblib.load()
        self.group = group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        self.per = bags_per_rank(global_bags, self.world)
        if tables_per_rank is None:
            self.first_table, self.t_global = None, int(n_cols)
        else:
            assert len(tables_per_rank) == self.world
            self.first_table, self.t_global = table_ranges(tables_per_rank)
        self.dim = dim
        if layout not in ("table", "sample"):
            raise ValueError("layout must be 'table' or 'sample'")
        self.layout = layout
        # elements between consecutive samples of one table in a slab
        self.stride = self.t_global * dim if layout == "sample" else dim
        self.n_buffers = n_buffers
        self.slab_bytes = self.per * self.t_global * dim * 4
        # one allocation: n_buffers slabs, then this rank's flag block (one uint32 slot per peer) for the
        # completion kernel (emb_finish_launch)
        self._flag_off = (self.slab_bytes * n_buffers + 255) // 256 * 256
        self._mine = C.c_void_p()
        emblib.check(self.lib.emb_device_alloc(C.byref(self._mine), self._flag_off + 256), "emb_device_alloc")
        zeros = np.zeros(64, dtype=np.uint32)
        emblib.check(self.lib.emb_memcpy_h2d(C.c_void_p(self._mine.value + self._flag_off), zeros.ctypes.data, 256,
                                             None), "emb_memcpy_h2d")  # synchronous: zero before any peer maps it
        h = EmbIpcHandle()
        emblib.check(self.lib.emb_ipc_export(self._mine, C.byref(h)), "emb_ipc_export")
        handles = [None] * self.world
        dist.all_gather_object(handles, bytes(h.bytes), group=group)
        self.base = []
        for r, hb in enumerate(handles):
            if r == self.rank:
                self.base.append(int(self._mine.value))
                continue
            ph = EmbIpcHandle()
            C.memmove(ph.bytes, hb, len(hb))
            p = C.c_void_p()
            emblib.check(self.lib.emb_ipc_open(C.byref(ph), C.byref(p)), "emb_ipc_open")
            self.base.append(int(p.value))
        self._flag = torch.zeros(1, dtype=torch.int32, device="cuda")
        # transport of the completion barrier: "kernel" = emb_finish_launch (flags in peer memory over
        # NVLink, one launch), "nccl" = all-reduce of one int (comparator; EMB_BARRIER=nccl)
        import os
        self.barrier_kind = os.environ.get("EMB_BARRIER", "kernel")
        self._epoch = 0
        self._finishers = []
        self._plain = self.make_finish([]) if self.barrier_kind == "kernel" else None

    def out_peer(self, local_table: int, buffer: int = 0):
        """Per-rank destination of this rank's `local_table`: pass as emb_dev_query.out_peer with
        out_stride = self.stride and bags_per_rank = self.per."""
        return self.out_peer_col(self.first_table[self.rank] + local_table, buffer)

    def out_peer_col(self, col_block: int, buffer: int = 0):
        """Per-rank destination of column block `col_block` (elements [col_block*D, (col_block+1)*D) of
        every sample row)."""
        if not 0 <= col_block < self.t_global:
            raise ValueError(f"column block {col_block} outside the slab's {self.t_global}")
        col = col_block * self.dim * 4 * (1 if self.layout == "sample" else self.per)
        return [b + buffer * self.slab_bytes + col for b in self.base]

    def local(self, buffer: int = 0, dtype="float32"):
        """This rank's slab as a torch tensor (no copy): [T_global, B/N, D] for layout "table" (index it
        by table to get the per-table [B/N, D] tensors the DLRM fork's alltoall returns), [B/N, T_global,
        D] for layout "sample"."""
        import torch
        shape = (self.per, self.t_global, self.dim) if self.layout == "sample" else (self.t_global, self.per, self.dim)
        v = _DevView(self.base[self.rank] + buffer * self.slab_bytes, shape, "<f4" if dtype == "float32" else "<i4")
        return torch.as_tensor(v, device="cuda")

    def local_sample_major(self, buffer: int = 0):
        """[B/N, T_global, D] view of this rank's slab whatever the layout (strided for "table")."""
        x = self.local(buffer)
        return x if self.layout == "sample" else x.permute(1, 0, 2)

    def make_finish(self, jobs):
        """Completion object for this slab set: signal + wait + the given reduce jobs in one launch.
        jobs: list of dicts(parts=[device addresses], out=address, part_stride, out_stride, nr_bags,
        is_raw_i32).  All finish objects of one PeerSlabs share its flag block and epoch counter."""
        from .dputypes import EmbReduceJob
        from . import emblib
        arr = (EmbReduceJob * max(1, len(jobs)))()
        for j, q in enumerate(jobs):
            for s_, p in enumerate(q["parts"]):
                arr[j].parts[s_] = int(p)
            arr[j].out = int(q["out"])
            arr[j].n_parts = len(q["parts"])
            arr[j].part_stride = q.get("part_stride", 0)
            arr[j].out_stride = q.get("out_stride", 0)
            arr[j].is_raw_i32 = int(q.get("is_raw_i32", 1))
            arr[j].nr_bags = q["nr_bags"]
        blocks = (C.c_void_p * self.world)(*[b + self._flag_off for b in self.base])
        h = C.c_void_p()
        emblib.check(self.lib.emb_finish_create(blocks, self.world, self.rank, arr if jobs else None, len(jobs),
                                                C.byref(h)), "emb_finish_create")
        self._finishers.append(h)
        return h

    def finish(self, finisher, stream=None):
        """Launch a completion object on `stream` (default: torch's current stream).  Every rank must
        call finish()/barrier() the same number of times in the same order."""
        import torch
        from . import emblib
        self._epoch += 1
        st = stream if stream is not None else torch.cuda.current_stream()
        emblib.check(self.lib.emb_finish_launch(finisher, self._epoch & 0xffffffff, emblib._stream(st)),
                     "emb_finish_launch")

    def barrier(self, stream=None):
        """Stream-ordered completion barrier: when it has run on this rank's stream, every rank's
        pooling kernel (enqueued before its own barrier call) has finished storing."""
        if self._plain is not None:
            self.finish(self._plain, stream)
            return
        import torch
        import torch.distributed as dist
        if stream is None:
            dist.all_reduce(self._flag, group=self.group)
        else:
            with torch.cuda.stream(stream):
                dist.all_reduce(self._flag, group=self.group)

    def close(self):
        for h in self._finishers:
            self.lib.emb_finish_destroy(h)
        self._finishers = []
        for r, b in enumerate(self.base):
            if r != self.rank and b:
                self.lib.emb_ipc_close(C.c_void_p(b))
        if self._mine:
            self.lib.emb_device_free(self._mine)
            self._mine = C.c_void_p()
        self.base = []
