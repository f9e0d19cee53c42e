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
    n_parts: int    # how many row ranges the table was cut into (1 = all rows)
    part: int       # this shard's position among them
    bag_lo: int = 0  # bag slice [bag_lo, bag_hi) of the global batch this shard pools (replicated small
    bag_hi: int = 0  # table, data-parallel); bag_hi == 0 means every bag

    @property
    def rows(self):
        return self.row_hi - self.row_lo

    @property
    def sliced(self):
        return self.bag_hi != 0

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
        return sorted((s for s in self.shards if s.table == table), key=lambda s: (s.part, s.bag_lo))

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
                level=True, tolerance=1.01, min_rows=1024, replicate_max_bytes=256e6, bag_align=256):
    """rows[t] = rows of table t; lookups[t] = expected indices per call (default: equal); bags = bags
    per table per call (the global batch).  hbm_budget: table bytes allowed per GPU.
    max_table_bytes: largest slab kept whole (default hbm_budget / 2).

    Placement is longest-processing-time-first on shard_cost().  With level=True a unit that would
    push the least-loaded rank above the mean load (x tolerance) is cut instead: the rank receives the
    piece that fills it exactly to the mean and the remainder goes back into the queue
    (water-filling), so the ranks end up level with about `world` extra cuts.  Two kinds of cut:

    * a table of at most replicate_max_bytes is REPLICATED on the ranks that share it and cut by bags
      (each replica pools a slice [bag_lo, bag_hi) of the batch, multiples of bag_align): no wasted
      index reads, no partial sums -- the data-parallel placement of small tables;
    * a larger table is cut by ROWS (pieces of at least min_rows): every piece reads the whole index
      list, fetches the rows in its range and emits partial sums that the receiver adds.

    level=False places whole units only."""
    T = len(rows)
    if lookups is None:
        lookups = [1.0] * T
    if world < 1 or T < 1:
        raise ValueError("need at least one rank and one table")
    cap = max_table_bytes if max_table_bytes is not None else hbm_budget / 2
    row_bytes = 4 * dim
    if row_bytes > cap:
        raise ValueError("a single row does not fit the budget")
    do_level = level and world > 1

    def replicable(t):
        return bags >= 2 * bag_align and rows[t] * row_bytes <= min(replicate_max_bytes, cap)

    # unit = (t, row_lo, row_hi, bag_lo, bag_hi); bag range (0, 0) = every bag
    def cost(u):
        t, lo, hi, blo, bhi = u
        if bhi:
            return shard_cost(lookups[t], bags, dim, 1.0) * (bhi - blo) / bags
        return shard_cost(lookups[t], bags, dim, (hi - lo) / rows[t])

    def build(fixed_target):
        queue = []
        for t in range(T):
            k = max(1, -(-int(rows[t] * row_bytes) // int(cap)))
            if k > rows[t]:
                raise ValueError(f"table {t}: a single row does not fit the budget")
            for p in range(k):
                queue.append((t, rows[t] * p // k, rows[t] * (p + 1) // k, 0, 0))
        load, mem = [0.0] * world, [0.0] * world
        placed = []  # (unit, rank)
        held = set()  # (t, row_lo, row_hi, rank): slabs already counted in mem

        def nbytes(u, r):
            return 0 if (u[0], u[1], u[2], r) in held else (u[2] - u[1]) * row_bytes

        def put(u, r):
            mem[r] += nbytes(u, r)
            held.add((u[0], u[1], u[2], r))
            load[r] += cost(u)
            placed.append((u, r))

        while queue:
            # costliest first; among equals the big tables first (whole, each on a rank of its own) and
            # the replicable ones last: what is left for the end can then be cut by bags, the cheap kind
            # of cut, to fill the ranks exactly
            queue.sort(key=lambda u: (-cost(u), replicable(u[0]), -(u[2] - u[1]), u[0], u[1], u[3]))
            u = queue.pop(0)
            t, lo, hi, blo, bhi = u
            n, c = hi - lo, cost(u)
            target = fixed_target if fixed_target is not None else (sum(load) + c + sum(cost(v) for v in queue)) / world
            taken = {r for (v, r) in placed if v[0] == t}
            order = sorted(range(world), key=lambda r: (load[r], r in taken, r))
            if do_level and load[order[0]] + c > target * tolerance:
                room = target - load[order[0]]
                if replicable(t) and n == rows[t]:
                    b0, b1 = (blo, bhi) if bhi else (0, bags)
                    take = int(room / (c / (b1 - b0))) // bag_align * bag_align
                    q = next((q for q in order if mem[q] + nbytes(u, q) <= hbm_budget), None)
                    if q is not None and bag_align <= take <= (b1 - b0) - bag_align:
                        put((t, lo, hi, b0, b0 + take), q)
                        queue.append((t, lo, hi, b0 + take, b1))
                        continue
                elif not bhi and n >= 2 * min_rows:
                    q = next((q for q in order if mem[q] + min_rows * row_bytes <= hbm_budget), None)
                    if q is None:
                        raise MemoryError(f"table {t} rows [{lo},{hi}) does not fit: {world} x {hbm_budget:.3g} B is full")
                    fixed = shard_cost(lookups[t], bags, dim, 0.0)
                    var = (c - fixed) / n  # per row
                    take = int((target - load[q] - fixed) / var) if var > 0 else n
                    take = min(take, int((hbm_budget - mem[q]) // row_bytes), n - min_rows)
                    if take >= min_rows:
                        put((t, lo, lo + take, 0, 0), q)
                        queue.append((t, lo + take, hi, 0, 0))
                        continue
            r = next((q for q in order if mem[q] + nbytes(u, q) <= hbm_budget), None)
            if r is None:
                # does not fit whole anywhere: give the emptiest rank the rows it can hold
                q = max(range(world), key=lambda x: hbm_budget - mem[x])
                take = min(int((hbm_budget - mem[q]) // row_bytes), n)
                if bhi or take < 1 or take == n:
                    raise MemoryError(f"table {t} rows [{lo},{hi}) does not fit: {world} x {hbm_budget:.3g} B is full")
                put((t, lo, lo + take, 0, 0), q)
                queue.append((t, lo + take, hi, 0, 0))
                continue
            put(u, r)
        out = []
        for t in range(T):
            mine = sorted(((v, r) for (v, r) in placed if v[0] == t), key=lambda x: (x[0][1], x[0][3]))
            ranges = sorted({(v[1], v[2]) for (v, _) in mine})
            for (v, r) in mine:
                out.append(Shard(t, v[1], v[2], r, len(ranges), ranges.index((v[1], v[2])), v[3], v[4]))
        return Plan(world, dim, out, load, mem)

    best = build(None)
    if do_level:
        # every cut by rows adds a shard's fixed cost (its index and output traffic), which the running
        # mean cannot foresee: aim the next pass at the mean the last one ended with
        plan = best
        for _ in range(4):
            if best.imbalance() <= tolerance:
                break
            plan = build(sum(plan.load) / world)
            if plan.imbalance() < best.imbalance() - 1e-9:
                best = plan
    return best


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
