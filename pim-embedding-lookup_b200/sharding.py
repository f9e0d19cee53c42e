"""Placement of embedding tables over the GPUs of one box.

The reference places one (table, column) per DPU (`dpu_id = table_id*NR_COLS + col`,
upmem/include/emb_host.h:167) and the DLRM fork places whole tables per rank
(dlrm_dpu_pytorch.py:470-482, `cuda:k % ndevices` :878).  Column-wise placement multiplies
index traffic by NR_COLS and breaks 16-byte row loads, so here the units are

* whole tables (table-wise), balanced by longest-processing-time-first on a traffic cost, and
* contiguous row ranges of a table (row-wise) when a table exceeds one GPU's budget or is
  so hot that it alone would unbalance the ranks (26 tables over 8 GPUs is 4-vs-3 otherwise).

A row-range shard pools only the indices that fall in its range (the kernel skips the rest,
`emb_table_set_row_base`) and emits raw int32 partial sums; the partials of one table are
added at the receiver (`emb_reduce_partials`), exact in any order for int32.

Pure host logic: no GPU, no torch.
"""
from __future__ import annotations

from dataclasses import dataclass, field


@dataclass(frozen=True)
class Shard:
    table: int      # global table id
    row_lo: int     # first global row held
    row_hi: int     # one past the last
    rank: int
    n_parts: int    # how many shards the table was cut into (1 = whole table)
    part: int       # this shard's position among them

    @property
    def rows(self):
        return self.row_hi - self.row_lo

    @property
    def whole(self):
        return self.n_parts == 1


@dataclass
class Plan:
    world: int
    dim: int
    shards: list = field(default_factory=list)
    load: list = field(default_factory=list)       # traffic cost per rank (bytes per call)
    hbm_bytes: list = field(default_factory=list)  # table bytes per rank

    def of_rank(self, rank):
        return [s for s in self.shards if s.rank == rank]

    def of_table(self, table):
        return sorted((s for s in self.shards if s.table == table), key=lambda s: s.part)

    def imbalance(self):
        """max rank load / mean rank load (1.0 = perfectly level)."""
        m = sum(self.load) / len(self.load)
        return max(self.load) / m if m else 1.0


def shard_cost(lookups, bags, dim, frac=1.0):
    """Bytes one call moves for a shard holding `frac` of a table's rows (uniform indices):
    every index is read (4 B), a fraction `frac` of them fetch a row (4*dim B), every bag writes a
    row (4*dim B) and reads an offset (4 B).  frac=1 is SURVEY.md 8(d)'s formula."""
    return lookups * (4 + frac * 4 * dim) + bags * (4 * dim + 4)


def plan_tables(rows, dim, world, lookups=None, bags=0, hbm_budget=150e9, max_table_bytes=None,
                level=True, tolerance=1.10):
    """rows[t] = rows of table t; lookups[t] = expected indices per call (default: equal).
    hbm_budget: table bytes allowed per GPU.  max_table_bytes: largest slab kept whole
    (default hbm_budget / 2).  level: also cut hot tables until LPT reaches max/mean <= tolerance."""
    T = len(rows)
    if lookups is None:
        lookups = [1.0] * T
    if world < 1 or T < 1:
        raise ValueError("need at least one rank and one table")
    cap = max_table_bytes if max_table_bytes is not None else hbm_budget / 2
    row_bytes = 4 * dim
    parts = [max(1, -(-int(r * row_bytes) // int(cap))) for r in rows]
    for t in range(T):
        if parts[t] > rows[t]:
            raise ValueError(f"table {t}: a single row does not fit the budget")

    def build(parts):
        units = []
        for t in range(T):
            k = parts[t]
            for p in range(k):
                lo, hi = rows[t] * p // k, rows[t] * (p + 1) // k
                frac = (hi - lo) / rows[t]
                units.append((shard_cost(lookups[t], bags, dim, frac), (hi - lo) * row_bytes, t, lo, hi, k, p))
        units.sort(key=lambda u: (-u[0], u[2], u[6]))
        load, mem, out = [0.0] * world, [0.0] * world, []
        for cost, nbytes, t, lo, hi, k, p in units:
            # shards of one table go to distinct ranks when possible (they reuse the index list)
            taken = {s.rank for s in out if s.table == t}
            order = sorted(range(world), key=lambda r: (r in taken and k <= world, load[r], r))
            for r in order:
                if mem[r] + nbytes <= hbm_budget:
                    break
            else:
                raise MemoryError(f"table {t} rows [{lo},{hi}) does not fit: {world} x {hbm_budget:.3g} B is full")
            load[r] += cost
            mem[r] += nbytes
            out.append(Shard(t, lo, hi, r, k, p))
        return Plan(world, dim, out, load, mem)

    plan = build(parts)
    if level and world > 1:
        # cut the costliest still-splittable table in two more pieces until level (or no gain)
        for _ in range(4 * world):
            if plan.imbalance() <= tolerance:
                break
            hot = max(range(world), key=lambda r: plan.load[r])
            cands = [s for s in plan.of_rank(hot) if s.rows >= 2 * 1024]
            if not cands:
                break
            worst = max(cands, key=lambda s: shard_cost(lookups[s.table], bags, dim, s.rows / rows[s.table]))
            trial = list(parts)
            trial[worst.table] = min(world, parts[worst.table] * 2)
            if trial == parts:
                break
            new = build(trial)
            if new.imbalance() >= plan.imbalance() - 1e-9:
                break
            parts, plan = trial, new
    return plan


def bucket_indices(indices, offsets, row_lo, row_hi):
    """Optional host-side routing for a row-range shard: keep only the indices inside
    [row_lo, row_hi) and rebuild the offsets.  (The kernel can also filter on its own; this
    trades host work for PCIe bytes.)  NumPy arrays in, NumPy arrays out."""
    import numpy as np
    indices = np.asarray(indices, dtype=np.uint32)
    offsets = np.asarray(offsets, dtype=np.uint32)
    keep = (indices >= row_lo) & (indices < row_hi)
    csum = np.concatenate([[0], np.cumsum(keep)]).astype(np.uint32)
    starts = offsets.astype(np.int64).copy()
    if len(starts):
        starts[0] = 0  # bag 0 starts at 0 whatever offsets[0] says (emb_dpu_lookup.c:60-63)
    return indices[keep], csum[np.minimum(starts, len(indices))]
