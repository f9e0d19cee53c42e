#!/usr/bin/env python
"""bench.py -- embedding lookups/s of the sum-pool hot path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload c2|c3|c4|c5|c4tb]

One "step" = one pass of the hot path over one batch: every local table pools its B bags.
N=1 workload = BASELINE.json configs[1] ("c2"): 26 tables x 1M rows x dim 16, batch 65536,
pooling U{1..20}, uniform indices, int32 fixed-point tables (the DPU representation).
N>1 (torchrun, one rank per GPU): weak scaling -- every rank owns 26 such tables and pools
the whole global batch for them; pooled rows go to the rank that owns each sample
(table-parallel -> batch-parallel, extend_distributed.py:389-426) either straight from the
pooling kernel over NVLink peer memory (--exchange fused, default) or by one NCCL
all_to_all_single (--exchange nccl).

Numbers: `value` = lookups/s with inputs resident in HBM (CUDA events, max over ranks);
`e2e` = the same from HOST buffers: at N=1 through the host-pointer C-ABI call (emb_lookup_ex), at
N>1 through the sharded product path (pinned indices -> HBM, pooling launch, peer-slab exchange,
completion kernel, [B/N, T, D] rows -> pinned host); `roofline` = algorithmic bytes of one pooling
launch / its CUDA-event duration against MEASURED_PEAKS.json; `cpu_baseline` = the reference's CPU
path and the oracle port on this box's host cores, N=1 only; `parity` = every pooled row of one batch
compared bit for bit, on every rank, with a torch index_select / index_add_ restatement computed on the
rank that owns the table (the run exits non-zero on a mismatch); `extra` = the same measurement on the
other multi-GPU configs of BASELINE.json (c4tb: Terabyte cardinalities, table + row-wise sharded,
strong scaling; c5: RM2-like skew) so that they land in the driver's records.
`--impl reference` times the CPU implementation alone (see DESIGN.md "Reference arm").
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "oracle")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

WORKLOADS = {
    # BASELINE.json configs[1]
    "c2": dict(name="embedding-bag microbench: 26 tables x 1M rows x dim 16, batch 65536, pooling U{1..20}, "
                    "uniform indices", T=26, R=1_000_000, D=16, B=65536, pool=(1, 20), dist="uniform"),
    # configs[4], RM2-like: dim 128, pooling 80, Zipf indices
    "c5": dict(name="RM2-like: 32 tables x 500k rows x dim 128, batch 4096, pooling 80, zipf(1.05) indices",
               T=32, R=500_000, D=128, B=4096, pool=(80, 80), dist="zipf"),
    # configs[2] lookup shape: Kaggle cardinalities (tools/visualize.py:949), dim 16, batch 39292, one index per bag
    "c3": dict(name="Kaggle-cardinality tables x dim 16, batch 39292, pooling 1, uniform indices", T=26, R=10131227,
               rows=[1460, 583, 10131227, 2202608, 305, 24, 12517, 633, 3, 93145, 5683, 8351593, 3194, 27, 14992,
                     5461306, 10, 5652, 2173, 4, 7046547, 18, 15, 286181, 105, 142572],
               D=16, B=39292, pool=(1, 1), dist="uniform"),
    # configs[3] shape per GPU (dim 64): uniform, fixed pooling 10
    "c4": dict(name="dim-64 tables: 26 x 1M rows x dim 64, batch 65536, pooling 10, uniform indices",
               T=26, R=1_000_000, D=64, B=65536, pool=(10, 10), dist="uniform"),
    # configs[3]: Criteo Terabyte cardinalities (tools/visualize.py:964), dim 64, 48 GB of
    # tables sharded table-wise + row-wise over the ranks (sharded.ShardedTables), GLOBAL batch 65536
    "c4tb": dict(name="Terabyte-cardinality tables: 26 tables up to 40M rows x dim 64 (48 GB), global batch 65536, "
                      "pooling 10, uniform indices, table+row-wise sharded", T=26, R=39979771,
                 rows=[39884406, 39043, 17289, 7420, 20263, 3, 7120, 1543, 63, 38532951, 2953546, 403346, 10, 2208,
                       11938, 155, 4, 976, 14, 39979771, 25641295, 39664984, 585935, 12972, 108, 36],
                 D=64, B=65536, pool=(10, 10), dist="uniform", sharded=True),
}
# The same tables with every table above 3 GB CUT BY ROWS over the ranks (the only layout for a table that exceeds one
# GPU's HBM; north_star): raw int32 partial sums cross NVLink and the completion kernel reduces them at the receiver.
WORKLOADS["c4rc"] = dict(WORKLOADS["c4tb"], max_table_bytes=3e9,
                         name=WORKLOADS["c4tb"]["name"] + "; tables above 3 GB cut by rows (partial sums reduced at the receiver)")
N_BATCHES = 4  # distinct index batches rotated through the timed steps


def parse_args(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS))
    ap.add_argument("--exchange", default="fused", choices=["fused", "nccl"])
    ap.add_argument("--layout", default="table", choices=["table", "sample"],
                    help="peer-slab layout of the fused exchange: [T][B/N][D] (default: the per-table list the interaction "
                         "concatenates) or [B/N][T][D] (thrashes peer address translation at 8 ranks, profiles/r2_scaling.md)")
    ap.add_argument("--no-rotate", action="store_true",
                    help="fused exchange: every rank walks the destination ranks in the same order (comparator)")
    ap.add_argument("--no-overlap", action="store_true",
                    help="plans without EMB_Q_OVERLAP: a pooling launch reads nothing before the kernel in front of it "
                         "has completed (comparator for the programmatic-dependent-launch overlap)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-unmodified-caller", action="store_true",
                    help="skip e2e.unmodified_caller (lookup() on pageable arrays)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-parity", action="store_true", help="skip the bit-exact check of one batch on every rank")
    ap.add_argument("--extras", default="auto",
                    help="comma-separated workloads measured after the main one and reported under `extra` "
                         "(auto = c4tb,c5 when the main workload is c2; none = skip)")
    ap.add_argument("--dtype", default="int32", choices=["int32", "f32"],
                    help="table element kind: int32 fixed point (the DPU representation) or fp32 (the CPU path)")
    ap.add_argument("--max-table-bytes", type=float, default=0.0,
                    help="sharded workloads: largest slab kept whole (forces row cuts below the 10 GB tables)")
    ap.add_argument("--variant", type=int, default=0, help="kernel variant (emb_set_tuning); 0 = default")
    ap.add_argument("--unroll", type=int, default=0)
    ap.add_argument("--mult", type=int, default=0)
    return ap.parse_args(argv)


# ---------------------------------------------------------------------------------------------
# synthetic inputs: seeded NumPy on the host
# ---------------------------------------------------------------------------------------------

def reference_bag_indices(R, P, ra):
    """One bag of the reference's uniform generator, dlrm_data_pytorch.py:1061-1065 (generate_uniform_input_batch,
    fixed pooling): `r = ra.random(P); sparse_group = np.round(r * (R - 1)).astype(np.int64)` with a seeded
    np.random.RandomState; duplicates are kept (:1065, the np.unique is commented out).  Pinned against the
    reference's own function by tests/test_generator.py.  bench.make_batch below draws with
    default_rng.integers(0, R) instead: same support {0..R-1}, but the reference's rounding gives the two end
    rows half the probability of the others (a 1-in-R effect, irrelevant to throughput), and a vectorised
    draw of 17.9 M indices takes 0.2 s where the per-bag loop takes minutes."""
    r = ra.random(P)
    return np.round(r * (R - 1)).astype(np.int64)


def make_batch(w, seed, n_tables=None):
    """One batch for n_tables tables: lists of uint32 indices [L_t] and offsets [B]."""
    rng = np.random.default_rng(seed)
    T = w["T"] if n_tables is None else n_tables
    lo, hi = w["pool"]
    idx, off = [], []
    for t in range(T):
        R = w["rows"][t] if "rows" in w else w["R"]
        lens = rng.integers(lo, hi + 1, size=w["B"])
        off.append(np.concatenate([[0], np.cumsum(lens)[:-1]]).astype(np.uint32))
        n = int(lens.sum())
        if w["dist"] == "zipf":
            # power-law over a fixed random permutation of the row ids (duplicates kept)
            r = rng.zipf(1.05, size=n)
            r = (r - 1) % R
            perm = np.random.default_rng(99).permutation(R).astype(np.uint32)
            idx.append(perm[r])
        else:
            idx.append(rng.integers(0, R, size=n, dtype=np.int64).astype(np.uint32))
    return idx, off


def make_table_batch(w, t, seed):
    """(indices, offsets) of global table t for one global batch; the same on every rank."""
    rng = np.random.default_rng([seed, t])
    lo, hi = w["pool"]
    lens = rng.integers(lo, hi + 1, size=w["B"])
    off = np.concatenate([[0], np.cumsum(lens)[:-1]]).astype(np.uint32)
    idx = rng.integers(0, w["rows"][t], size=int(lens.sum()), dtype=np.int64).astype(np.uint32)
    return idx, off


def algorithmic_bytes(n_lookups, n_bags, D):
    # per lookup 4D (row) + 4 (index); per bag 4D (output row) + 4 (offset)
    return (n_lookups + n_bags) * (4 * D + 4)


def job_config(w, world, args, sharded):
    """`config` of the JSON line: what the job IS.  Identical in the B200 arm and the reference arm (the driver
    compares them); everything that describes HOW an arm ran it lives in other keys."""
    cfg = {"workload": w["name"], "tables": w["T"], "dim": w["D"], "global_batch": w["B"],
           "pooling": list(w["pool"]), "index_distribution": w["dist"], "dtype": args.dtype,
           "index_batches": N_BATCHES,
           "l2": "no flush between steps: inputs larger than L2 (tables %.2f GB per GPU >> 126 MB, %d rotating index "
                 "batches of %.0f MB); ncu's cold-cache launch time agrees with the warm event time within 1.5 %% "
                 "(profiles/r2_launches.txt)" % (
                     (sum(w.get("rows", [w["R"]] * w["T"])) * w["D"] * 4 / 1e9) / (world if sharded else 1), N_BATCHES,
                     w["T"] * w["B"] * (w["pool"][0] + w["pool"][1]) / 2 * 4 / 1e6),
           "parallelism": (f"table-wise + row-wise over {world} rank(s), global batch fixed (strong scaling)" if sharded else
                           f"table-wise x{world}: every rank owns {w['T']} such tables and pools the whole batch for them "
                           "(weak scaling); pooled rows to the sample's rank")}
    if "rows" in w:
        cfg["rows_per_table"] = "per-table cardinalities, max %d, total %d" % (max(w["rows"]), sum(w["rows"]))
    else:
        cfg["rows_per_table"] = w["R"]
    return cfg


# ---------------------------------------------------------------------------------------------
# clocks during the timed region (NVML, sampled every ~5 ms from a thread)
# ---------------------------------------------------------------------------------------------

class ClockSampler:
    REASONS = {0x1: "gpu_idle", 0x2: "applications_clocks_setting", 0x4: "sw_power_cap", 0x8: "hw_slowdown",
               0x10: "sync_boost", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
               0x80: "hw_power_brake_slowdown", 0x100: "display_clock_setting"}

    def __init__(self, index):
        self.samples, self.reasons, self.max_mhz, self._stop, self._t = [], set(), None, False, None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception as e:  # noqa: BLE001
            self.nv, self.err = None, repr(e)

    def _run(self):
        nv = self.nv
        while not self._stop:
            try:
                util = nv.nvmlDeviceGetUtilizationRates(self.h).gpu
                self.samples.append((nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM), util))
                bits = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for b, name in self.REASONS.items():
                    if bits & b and name != "gpu_idle":
                        self.reasons.add(name)
            except Exception:  # noqa: BLE001
                pass
            time.sleep(0.005)

    def __enter__(self):
        if self.nv:
            self._t = threading.Thread(target=self._run, daemon=True)
            self._t.start()
        return self

    def __exit__(self, *a):
        self._stop = True
        if self._t:
            self._t.join()

    def summary(self):
        if not self.nv:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "error": self.err}
        s = sorted(c for c, _ in self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "sm_mhz_min": s[0] if s else None,
                "reasons": sorted(self.reasons), "samples": len(s),
                "window": "warm-up + timed steps + e2e steps (the timed steps alone last a few ms: one NVML sample)"}


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def dist_summary(ms):
    """mean / median / p10 / p90 of per-launch durations (SURVEY.md 8d asks for the spread, not only the mean)."""
    a = np.sort(np.asarray(ms, dtype=np.float64))
    q = lambda f: float(a[min(len(a) - 1, int(round(f * (len(a) - 1))))])  # noqa: E731
    return {"mean": float(a.mean()), "median": q(0.5), "p10": q(0.1), "p90": q(0.9), "min": float(a[0]), "n": len(a)}


# ---------------------------------------------------------------------------------------------
# CPU side: the oracle port (the checker, here only timed) and the reference's float CPU path
# ---------------------------------------------------------------------------------------------

def time_oracle_port(tables_i32, idx, off, budget_s=12.0, min_reps=2):
    """Fixed-point path (quantised tables, wrapping int32 pool, /1e9) in multi-threaded C."""
    import emb_oracle as eo
    n = sum(len(i) for i in idx)
    eo.lookup_rows(tables_i32, idx, off)  # warm-up (page-in)
    reps, t_tot = 0, 0.0
    while reps < min_reps or (t_tot < budget_s and reps < 50):
        t0 = time.perf_counter()
        eo.lookup_rows(tables_i32, idx, off)
        t_tot += time.perf_counter() - t0
        reps += 1
    return n * reps / t_tot, eo.num_threads(), reps


def host_cores():
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return os.cpu_count() or 1


def torch_bags(tables_f32, idx, off):
    """The reference's CPU path: nn.EmbeddingBag(mode="sum", sparse=True) per table, fp32 tables, int64 indices /
    offsets (dlrm_s_pytorch.py:295,483-488).  Returns a function that runs one pass over all tables."""
    import torch
    bags = [torch.nn.EmbeddingBag.from_pretrained(torch.from_numpy(x), mode="sum", sparse=True, freeze=True)
            for x in tables_f32]
    ti = [torch.from_numpy(i.astype(np.int64)) for i in idx]
    to = [torch.from_numpy(o.astype(np.int64)) for o in off]

    def one():
        with torch.no_grad():
            for k, e in enumerate(bags):
                e(ti[k], to[k])
    return one


def time_passes(one, n_lookups, budget_s, min_reps=2, max_reps=50):
    one()
    reps, t_tot = 0, 0.0
    while reps < min_reps or (t_tot < budget_s and reps < max_reps):
        t0 = time.perf_counter()
        one()
        t_tot += time.perf_counter() - t0
        reps += 1
    return n_lookups * reps / t_tot, reps


def host_tables(w, n_tables, seed=1234):
    """Quantised int32 tables as export_emb would make them (U(-1/sqrt(R), 1/sqrt(R)) * 1e9)."""
    rng = np.random.default_rng(seed)
    rows = w.get("rows", [w["R"]] * w["T"])
    return [rng.integers(-int(1e9 / np.sqrt(rows[t])), int(1e9 / np.sqrt(rows[t])) + 1, size=(rows[t], w["D"]),
                         dtype=np.int32) for t in range(n_tables)]


def run_reference(args, w):
    """--impl reference: the reference's CPU implementation of the path alone (nn.EmbeddingBag(mode=sum) per table
    on stock torch, dlrm_s_pytorch.py:295,483-488), all host threads, the same job `config` as the B200 arm.
    A step is one pass over ALL tables of one rank's batch (26 x 65536 bags for c2, ~25 ms), so nothing is
    sampled for the default workload; the sharded 48 GB workload takes the tables that are cheap to build."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    world = args.gpus
    cfg = job_config(w, world, args, bool(w.get("sharded")))
    scaling = "strong" if w.get("sharded") else "weak"
    T_s = w["T"]
    sample_note = f"all {w['T']} tables of one rank's batch"
    if w.get("sharded"):
        # the 40M-row tables are 10 GB each (plus an fp32 copy): the bounded sample takes the largest
        # table that is cheap to build on the host (2.95M rows, 756 MB) and seven smaller ones
        pick = [10, 22, 11, 1, 4, 2, 23, 14]
        w = dict(w, rows=[w["rows"][t] for t in pick], T=len(pick))
        T_s = len(pick)
        sample_note = f"{T_s} of 26 tables (the 10 GB ones are left out)"
    tabs = host_tables(w, T_s)
    idx, off = make_batch(w, seed=123, n_tables=T_s)
    import torch
    # torchrun exports OMP_NUM_THREADS=1; the reference arm uses every host core it may run on
    torch.set_num_threads(host_cores())
    n = sum(len(i) for i in idx)
    ftabs = [t.astype(np.float32) * np.float32(1e-9) for t in tabs]
    one = torch_bags(ftabs, idx, off)
    for _ in range(max(1, min(args.warmup, 5))):
        one()
    steps = max(1, min(args.steps, 200))
    t0 = time.perf_counter()
    done = 0
    while done < steps and (done < 3 or time.perf_counter() - t0 < 30.0):
        one()
        done += 1
    dt = time.perf_counter() - t0
    steps = done
    val = n * steps / dt
    cores = torch.get_num_threads()
    # one thread, for reference: a quarter of the tables, a few passes
    torch.set_num_threads(1)
    T1 = max(1, T_s // 4)
    one1 = torch_bags(ftabs[:T1], idx[:T1], off[:T1])
    v1, reps1 = time_passes(one1, sum(len(i) for i in idx[:T1]), budget_s=4.0, min_reps=1, max_reps=5)
    torch.set_num_threads(cores)
    pv, pcores, _ = time_oracle_port(tabs[:8], idx[:8], off[:8], budget_s=4.0)
    sample = f"{sample_note} x {w['B']} bags ({n} lookups) per step, {steps} steps"
    print(json.dumps({
        "impl": "reference", "metric": "embedding lookups/sec", "value": val, "unit": "lookups/s",
        "n_gpus": args.gpus, "steps": steps, "warmup": args.warmup, "ms_per_step": dt / steps * 1e3,
        "higher_is_better": True, "scaling": scaling,
        "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": cfg,
        "cpu_baseline": {"value": val, "unit": "lookups/s", "cores": cores, "kind": "port", "sample": sample,
                         "what": "the reference's CPU lookup path restated call for call: nn.EmbeddingBag(mode=sum, "
                                 "sparse=True) per table on stock torch, fp32 tables, int64 indices "
                                 "(dlrm_s_pytorch.py:295,483-488); its DPU path (oracle/_ref) is undefined at this shape "
                                 "(B<=512, T*D<=256; SURVEY.md Q3/Q4/Q13)",
                         "single_thread": {"value": v1, "unit": "lookups/s", "cores": 1,
                                           "sample": f"{T1} tables x {reps1} passes"},
                         "oracle_port": {"value": pv, "unit": "lookups/s", "cores": pcores, "sample": "8 tables",
                                         "what": "fixed-point restatement (oracle/emb_oracle.c), multi-threaded C"}},
        "e2e": {"value": val, "unit": "lookups/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


# ---------------------------------------------------------------------------------------------
# GPU arm helpers
# ---------------------------------------------------------------------------------------------

def pinned_arena(lib, nbytes):
    """nbytes of cudaHostAlloc memory (emb_host_alloc) as a NumPy uint8 array; one arena per call so that the
    arrays of consecutive tables are adjacent in host memory (the library then moves them with one copy per
    group instead of one per table)."""
    p = C.c_void_p()
    if lib.emb_host_alloc(C.byref(p), max(int(nbytes), 64)):
        raise RuntimeError("emb_host_alloc failed")
    return np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), shape=(max(int(nbytes), 64),))


def carve_u32(arena, pos, n):
    """n uint32 at byte offset pos of the arena, packed tightly: the arrays of consecutive tables are adjacent in
    host memory, which lets the library move a group of tables with one copy.  Returns (array, next offset)."""
    return arena[pos:pos + 4 * n].view(np.uint32), pos + 4 * n


def table_tensor(lib, slot, D, dtype):
    """The library's slab of a table slot as a torch tensor [rows, D] (no copy)."""
    import torch

    from pim_embedding_lookup_b200 import emblib, exchange
    rows, kind, ptr = C.c_uint64(), C.c_int(), C.c_void_p()
    emblib.check(lib.emb_table_info(slot, C.byref(rows), C.byref(kind), C.byref(ptr)), "emb_table_info")
    return torch.as_tensor(exchange._DevView(ptr.value, (rows.value, D), "<i4" if dtype == "int32" else "<f4"),
                           device="cuda")


def torch_pooled(tab, idx, off, B, row_lo=0, raw=False):
    """What the path must produce for one table, restated with stock torch ops on the GPU -- none of this
    repo's kernels: gather with index_select, segment sums with index_add_ (int32 adds wrap; fp32 tables are
    added in fp64 and rounded once), then the reference's dequantise (emb_host.h:210) as an fp32 divide.
    tab: [rows, D] holding global rows [row_lo, row_lo + rows); idx / off: uint32 bits carried as int32.
    Returns fp32 [B, D] (raw=True: the int32 sums)."""
    import torch
    idx64 = idx.to(torch.int64) & 0xffffffff
    L = idx64.numel()
    starts = off.to(torch.int64) & 0xffffffff
    starts[0] = 0  # emb_dpu_lookup.c:60-63
    bag_of = torch.searchsorted(starts, torch.arange(L, device=idx.device), right=True) - 1
    rel = idx64 - row_lo
    keep = (rel >= 0) & (rel < tab.shape[0])
    if not bool(keep.all()):
        rel, bag_of = rel[keep], bag_of[keep]
    if tab.dtype == torch.float32:
        sums = torch.zeros((B, tab.shape[1]), dtype=torch.float64, device=tab.device)
        sums.index_add_(0, bag_of, tab.index_select(0, rel).to(torch.float64))
        return sums.to(torch.float32)
    sums = torch.zeros((B, tab.shape[1]), dtype=torch.int32, device=tab.device)
    sums.index_add_(0, bag_of, tab.index_select(0, rel))
    if raw:
        return sums
    return sums.to(torch.float32) / torch.tensor(1e9, dtype=torch.float32, device=tab.device)


def compare_rows(got, want, mag=None):
    """Number of mismatching elements.  Fixed-point path (mag is None): the bits must be equal.  fp32 tables: within
    1e-5 of the fp64-accumulated sum relative to `mag`, the sum of the magnitudes of the bag's rows (north_star's
    reassociation tolerance; an fp32 sum of n terms in any order is within (n-1) * 6e-8 of it)."""
    import torch
    if mag is None:
        return int((got.contiguous().view(torch.int32) != want.contiguous().view(torch.int32)).sum().item())
    return int(((got - want).abs() > 1e-5 * mag + 1e-30).sum().item())


def gather_over_ranks(vals, world):
    """[v0, v1, ...] of this rank -> one list per value holding every rank's figure (rank order)."""
    if world == 1:
        return [[float(v)] for v in vals]
    import torch
    import torch.distributed as dist
    mine = torch.tensor([float(v) for v in vals], dtype=torch.float64, device="cuda")
    every = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(every, mine)
    return [[round(float(e[i]), 6) for e in every] for i in range(len(vals))]


def reduce_over_ranks(vals, world, ops):
    """vals: list of floats; ops: parallel list of 'max' / 'sum' / 'min'.  All-reduce over the ranks."""
    if world == 1:
        return list(vals)
    import torch
    import torch.distributed as dist
    out = []
    t = torch.tensor(vals, dtype=torch.float64, device="cuda")
    for op, name in ((dist.ReduceOp.MAX, "max"), (dist.ReduceOp.SUM, "sum"), (dist.ReduceOp.MIN, "min")):
        if name in ops:
            r = t.clone()
            dist.all_reduce(r, op=op)
            out.append((name, r))
    res = []
    for i, op in enumerate(ops):
        res.append(float(dict(out)[op][i]))
    return res


def measure_copy_floor(h2d_bytes, d2h_bytes, world, reps=8):
    """Plain concurrent pinned H2D + D2H of a step's byte counts, one big copy each way, on every rank at the same
    time: what the box's PCIe gives these bytes whatever the code around them does."""
    import torch
    import torch.distributed as dist
    hin = torch.empty(max(h2d_bytes, 4) // 4, dtype=torch.int32).pin_memory()
    din = torch.empty_like(hin, device="cuda")
    dout = torch.empty(max(d2h_bytes, 4) // 4, dtype=torch.int32, device="cuda")
    hout = torch.empty(dout.shape, dtype=torch.int32).pin_memory()
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()

    def one():
        with torch.cuda.stream(s1):
            din.copy_(hin, non_blocking=True)
        with torch.cuda.stream(s2):
            hout.copy_(dout, non_blocking=True)
        s1.synchronize()
        s2.synchronize()
    for _ in range(2):
        one()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    for _ in range(reps):
        one()
    dt = (time.perf_counter() - t0) / reps
    (dt,) = reduce_over_ranks([dt], world, ["max"])
    return dt * 1e3


# ---------------------------------------------------------------------------------------------
# table-wise arm: every rank owns T tables and pools the whole batch for them (weak scaling)
# ---------------------------------------------------------------------------------------------

def bench_tablewise(args, w, steps, warmup):
    import torch
    import torch.distributed as dist

    from pim_embedding_lookup_b200 import emblib, exchange
    from pim_embedding_lookup_b200.dputypes import (EMB_HOST_AUTO, EMB_HOST_COPY, EMB_HOST_MAPPED, EMB_Q_OVERLAP, EMB_TABLE_FIXED,
                                                    EMB_TABLE_FLOAT)

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    lib = emblib.load()
    T, R, D, B = w["T"], w["R"], w["D"], w["B"]
    is_float = args.dtype == "f32"
    emblib.configure(T, D, 0, 0, device=local)
    emblib.check(lib.emb_set_tuning(args.variant, args.unroll, args.mult))
    stream = torch.cuda.Stream()  # a real stream: events and launches are ordered on it
    torch.cuda.set_stream(stream)

    # tables: int32 fixed point as export_emb makes them (or fp32 U(+-1/sqrt(R))), generated on the device
    gen = torch.Generator(device="cuda").manual_seed(1234 + rank)
    rows_t = w.get("rows", [R] * T)
    for t in range(T):
        lim = int(1e9 / np.sqrt(rows_t[t]))
        a = torch.randint(-lim, lim + 1, (rows_t[t], D), dtype=torch.int32, device="cuda", generator=gen)
        if is_float:
            a = a.to(torch.float32) * 1e-9
        emblib.check(lib.emb_populate_table_device(t, rows_t[t], a.data_ptr(), 1 if is_float else 0,
                                                   EMB_TABLE_FLOAT if is_float else EMB_TABLE_FIXED, stream.cuda_stream))
        del a
    torch.cuda.empty_cache()

    # index batches: every rank pools the whole global batch for its own tables
    batches = [make_batch(w, seed=123 + 1000 * rank + k) for k in range(N_BATCHES)]
    lookups = [sum(len(i) for i in b[0]) for b in batches]
    dev_batches = []
    for idx, off in batches:
        dev_batches.append(([torch.from_numpy(i.view(np.int32)).cuda() for i in idx],
                            [torch.from_numpy(o.view(np.int32)).cuda() for o in off]))

    # outputs + plans (descriptors uploaded once; a step is one kernel launch).  The peer slabs are double
    # buffered: batch k lands in buffer k % 2, so a peer may already store batch k+1 while this rank's consumer
    # still reads batch k -- the configuration a real caller needs (sharded.ShardedTables does the same).
    slabs = None
    if world > 1 and args.exchange == "fused":
        slabs = exchange.PeerSlabs(B, [T] * world, D, layout=args.layout, n_buffers=2)
    per = B // world

    # the index / offset tensors are filled once, before the loop: a launch may gather under its predecessor's tail
    qflags = 0 if args.no_overlap else EMB_Q_OVERLAP

    def make_plan(di, do, buffer):
        qs = []
        keep = None
        if slabs is not None:
            for t in range(T):
                qs.append(dict(table_id=t, indices=di[t], offsets=do[t], out=None, out_stride=slabs.stride,
                               bags_per_rank=per, out_peer=slabs.out_peer(t, buffer), flags=qflags,
                               bag_rotate=0 if args.no_rotate else rank * per))
        elif world > 1:  # nccl: send buffer [N, B/N, T_local, D] == bag-major rows of stride T*D
            keep = torch.empty((B, T, D), dtype=torch.float32, device="cuda")
            for t in range(T):
                qs.append(dict(table_id=t, indices=di[t], offsets=do[t], out=keep[:, t, :], out_stride=T * D, flags=qflags))
        else:  # the reference's layout: one [B, D] result per table (emb_host.h:207-212)
            keep = torch.empty((T, B, D), dtype=torch.float32, device="cuda")
            for t in range(T):
                qs.append(dict(table_id=t, indices=di[t], offsets=do[t], out=keep[t], flags=qflags))
        return emblib.Plan(qs), keep

    plans, keep = [], []
    for k, (di, do) in enumerate(dev_batches):
        p, buf = make_plan(di, do, k % 2)
        plans.append(p)
        keep.append(buf)

    def step(k):
        """One step; returns this rank's result tensor [B/N, T_global, D] (N = 1: [T, B, D])."""
        plans[k % N_BATCHES].launch(stream)
        if slabs is not None:
            slabs.barrier()
            return slabs.local_sample_major((k % N_BATCHES) % 2)
        if world > 1:
            return exchange.nccl_all_to_all(keep[k % N_BATCHES], B, [T] * world, D)
        return keep[k % N_BATCHES]

    def sync_all():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    for k in range(max(warmup, 3)):
        step(k)
    sync_all()
    if lib.emb_check_indices():
        raise RuntimeError("out-of-range index in the synthetic batch")

    # ---- parity: every pooled row of one batch, on every rank, against stock torch ops on the owning rank
    parity = {"checked": False}
    if not args.no_parity:
        kb = 1  # batch 1 lands in slab buffer 1: the double buffering is part of what is checked
        got = step(kb).clone()
        sync_all()
        di, do = dev_batches[kb]
        tabs = [table_tensor(lib, t, D, args.dtype) for t in range(T)]
        def routed(fn):
            """fn(t) -> [B, D] for local table t; returns what this rank must hold: [T, B, D] (N = 1) or, after the
            reference's butterfly (extend_distributed.py:389-426) done with NCCL, [B/N, T_global, D]."""
            if world == 1:
                return torch.stack([fn(t) for t in range(T)])
            send = torch.empty((B, T, D), dtype=torch.float32, device="cuda")
            for t in range(T):
                send[:, t, :] = fn(t)
            return exchange.nccl_all_to_all(send, B, [T] * world, D)
        want = routed(lambda t: torch_pooled(tabs[t], di[t], do[t], B))
        mag = routed(lambda t: torch_pooled(tabs[t].abs(), di[t], do[t], B)) if is_float else None
        bad = compare_rows(got, want, mag)
        del mag
        # checksum of checksums (wrap-around sum of the raw bits of every row this rank received)
        chk_got = int(got.contiguous().view(torch.int32).to(torch.int64).sum().item()) & 0xffffffffffffffff
        chk_want = int(want.contiguous().view(torch.int32).to(torch.int64).sum().item()) & 0xffffffffffffffff
        rows_checked = got.numel() // D
        bad_all, rows_all = reduce_over_ranks([float(bad), float(rows_checked)], world, ["sum", "sum"])
        idx_err = lib.emb_check_indices()
        parity = {"checked": True, "ranks": world, "rows_checked": int(rows_all), "mismatching_elements": int(bad_all),
                  "bit_exact": bool(bad_all == 0) if not is_float else None,
                  "within_1e-5_relative": bool(bad_all == 0) if is_float else None,
                  # fp32 tables are compared within the stated tolerance: their bits, and so the checksum, may differ
                  "checksum_of_checksums_equal": None if is_float else chk_got == chk_want, "index_errors": int(idx_err),
                  "against": "torch index_select + index_add_ + fp32 divide on the rank that owns each table, routed with "
                             "NCCL all_to_all_single (extend_distributed.py:389-426; emb_dpu_lookup.c:106-116, "
                             "emb_host.h:207-212)"}
        del got, want, tabs
        torch.cuda.empty_cache()
        if bad_all or (not is_float and chk_got != chk_want) or idx_err:
            raise SystemExit(f"PARITY FAILURE on rank {rank}: {json.dumps(parity)}")

    # ---- the timed region: EXACTLY K steps back to back, one event on either side, nothing between the launches
    K = steps
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches0 = emblib.launch_count()
    sync_all()
    e0.record(stream)
    for k in range(K):
        plans[k % N_BATCHES].launch(stream)
        if slabs is not None:
            slabs.barrier()
        elif world > 1:
            exchange.nccl_all_to_all(keep[k % N_BATCHES], B, [T] * world, D)
    e1.record(stream)
    sync_all()
    gpu_launches = emblib.launch_count() - launches0
    total_ms = e0.elapsed_time(e1)
    # ---- the pooling kernel alone: the same K steps once more, every pooling launch between two events.  (An event
    # record between two launches also keeps them from overlapping, so this is the duration of an ISOLATED launch;
    # in the timed region above a launch starts under its predecessor's tail, EMB_Q_OVERLAP.)
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    for k in range(K):
        ev[k][0].record(stream)
        plans[k % N_BATCHES].launch(stream)
        ev[k][1].record(stream)
        if slabs is not None:
            slabs.barrier()
        elif world > 1:
            exchange.nccl_all_to_all(keep[k % N_BATCHES], B, [T] * world, D)
    sync_all()
    kdist = dist_summary([a.elapsed_time(b) for a, b in ev])
    kernel_ms = kdist["mean"]
    n_lookups = sum(lookups[k % N_BATCHES] for k in range(K))
    lookups_rank = float(n_lookups) / K
    total_ms, n_lookups, kernel_ms_max = reduce_over_ranks([total_ms, float(n_lookups), kernel_ms], world,
                                                           ["max", "sum", "max"])
    kernel_ms_per_rank, lookups_per_rank = gather_over_ranks([kernel_ms, lookups_rank], world)
    value = n_lookups / (total_ms * 1e-3)

    # roofline of the pooling kernel on this rank (algorithmic bytes / mean launch duration)
    peak, peak_src = load_peaks()
    bytes_per_launch = float(np.mean([algorithmic_bytes(lookups[k % N_BATCHES], T * B, D) for k in range(K)]))
    achieved = bytes_per_launch / (kernel_ms * 1e-3) / 1e9

    # DRAM traffic of the same kernel from the committed ncu capture (per launch), if there is one
    traffic, traffic_src = None, None
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    if world == 1 and os.path.exists(tp) and not is_float:
        rec = json.load(open(tp)).get(args.workload if w is WORKLOADS.get(args.workload) else None)
        if rec:
            traffic = rec["dram_bytes_read_per_launch"] + rec["dram_bytes_write_per_launch"]
            traffic_src = rec["source"]

    # ---- e2e from HOST buffers, copies inside the timed region
    e2e = None
    if not args.no_e2e:
        Ke = max(3, min(K, 50))
        if world == 1:
            e2e = e2e_host_call(args, lib, emblib, batches, lookups, T, B, D, Ke,
                                (EMB_HOST_AUTO, EMB_HOST_COPY, EMB_HOST_MAPPED), rows_t=rows_t)
        else:
            e2e = e2e_sharded_path(args, lib, emblib, exchange, slabs, batches, lookups, T, B, D, per, rank, world,
                                   Ke, stream, sync_all)
        e2e["floor_ms"] = measure_copy_floor(e2e["h2d_bytes_per_step"], e2e["d2h_bytes_per_step"], world)
        e2e["floor_what"] = ("one pinned H2D copy and one D2H copy of the step's bytes running concurrently, on all "
                             f"{world} rank(s) at once (max over ranks)")

    # ---- CPU baseline on this box's host cores (rank 0, N=1 only)
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        import torch as _t
        T_s = T  # every table of the batch, like the --impl reference arm
        tabs = [emblib.table_read(t, 0, rows_t[t], dtype=np.float32 if is_float else np.int32) for t in range(T_s)]
        sidx, soff = batches[0][0][:T_s], batches[0][1][:T_s]
        ftabs = tabs if is_float else [x.astype(np.float32) * np.float32(1e-9) for x in tabs]
        _t.set_num_threads(host_cores())
        ns = sum(len(i) for i in sidx)
        tv, treps = time_passes(torch_bags(ftabs, sidx, soff), ns, budget_s=8.0)
        cores = _t.get_num_threads()
        sample = f"{T_s} of {T} tables x {B} bags ({ns} lookups) x {treps} passes"
        cpu = {"value": tv, "unit": "lookups/s", "cores": cores, "kind": "port", "sample": sample,
               "what": "nn.EmbeddingBag(mode=sum, sparse=True) per table, fp32, stock torch (dlrm_s_pytorch.py:295,483-488)"}
        _t.set_num_threads(1)
        n1 = sum(len(i) for i in sidx[:2])
        v1, r1 = time_passes(torch_bags(ftabs[:2], sidx[:2], soff[:2]), n1, budget_s=3.0, min_reps=1, max_reps=5)
        _t.set_num_threads(cores)
        cpu["single_thread"] = {"value": v1, "unit": "lookups/s", "cores": 1, "sample": f"2 tables x {r1} passes"}
        if not is_float:
            v, pc, reps = time_oracle_port(tabs[:8], sidx[:8], soff[:8], budget_s=5.0)
            cpu["oracle_port"] = {"value": v, "unit": "lookups/s", "cores": pc, "passes": reps, "kind": "port"}

    out = None
    if rank == 0:
        cfg = job_config(w, world, args, False)
        out = {
            "metric": "embedding lookups/sec", "value": value, "unit": "lookups/s", "n_gpus": world, "steps": K,
            "warmup": max(warmup, 3), "ms_per_step": total_ms / K, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": args.dtype, "data": "synthetic", "config": cfg,
            "run": {"lookups_per_step_per_gpu": int(np.mean(lookups)),
                    "exchange": ("none (1 GPU)" if world == 1 else args.exchange +
                                 (f", {args.layout}-major peer slabs, double buffered" if args.exchange == "fused" else "")),
                    "tuning": {"variant": args.variant, "unroll": args.unroll, "mult": args.mult}},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src, "kernel_ms": kernel_ms,
                         "kernel_ms_dist": kdist, "kernel_ms_max_over_ranks": kernel_ms_max,
                         "kernel_ms_per_rank": kernel_ms_per_rank, "lookups_per_launch_per_rank": lookups_per_rank,
                         "kernel_ms_what": "mean duration of one ISOLATED pooling launch: the K steps repeated right "
                                           "after the timed region with an event on either side of every pooling "
                                           "launch; in the timed region itself nothing separates the launches and "
                                           + ("they do not overlap (--no-overlap)" if args.no_overlap else
                                              "each starts gathering under its predecessor's tail (EMB_Q_OVERLAP), so "
                                              "ms_per_step can be below kernel_ms"),
                         "algorithmic_bytes_per_launch": bytes_per_launch,
                         "frac_of_nominal_8000": achieved / 8000.0},
            "parity": parity, "gpu_launches": int(gpu_launches),
        }
        if e2e:
            out["e2e"] = e2e
        if cpu:
            out["cpu_baseline"] = cpu
    sync_all()
    for p in plans:
        p.close()
    if slabs is not None:
        slabs.close()
    emblib.shutdown()
    torch.cuda.empty_cache()
    return out


def e2e_host_call(args, lib, emblib, batches, lookups, T, B, D, Ke, paths, rows_t=None):
    """N = 1: the reference-facing call, emb_lookup_ex on HOST arrays.  The arrays are page-locked and carved out of
    one arena per kind, the way a caller that owns its buffers would allocate them (the reference allocates its
    index / offset / result arrays once, dlrm_dpu_pytorch.py:484-498,1028-1051)."""
    EMB_HOST_AUTO, EMB_HOST_COPY, EMB_HOST_MAPPED = paths
    host = []
    for idx, off in batches[:2]:
        arena = pinned_arena(lib, sum(4 * len(i) + 16 for i in idx) + sum(4 * len(o) + 16 for o in off) + 64)
        pos, hi, ho = 0, [], []
        for i in idx:
            a, pos = carve_u32(arena, pos, len(i))
            a[:] = i
            hi.append(a)
        for o in off:
            a, pos = carve_u32(arena, pos, len(o))
            a[:] = o
            ho.append(a)
        host.append((hi, ho, arena))
    # one result arena per batch in flight (two calls in flight write different arrays)
    res_sets, res_arenas = [], []
    for _ in range(2):
        res_arena = pinned_arena(lib, T * (B * D * 4 + 16) + 64)
        res, pos = [], 0
        for _ in range(T):
            a, pos = carve_u32(res_arena, pos, B * D)
            res.append(a.view(np.float32).reshape(B, D))
        res_sets.append(res)
        res_arenas.append(res_arena)
    # pointer arrays built once per batch, as the reference's caller does
    calls = [emblib.HostCall(hi, ho, res_sets[k]) for k, (hi, ho, _) in enumerate(host)]

    def time_e2e(n_steps):
        for k in range(3):
            calls[k % 2]()
        t0 = time.perf_counter()
        for k in range(n_steps):
            calls[k % 2]()  # emb_lookup_ex: synchronous, the results are in host memory on return
        dt = time.perf_counter() - t0
        tm = emblib.last_timings()
        n_e = sum(lookups[k % 2] for k in range(n_steps))
        return n_e / dt, dt / n_steps * 1e3, tm

    def time_e2e_in_flight(n_steps):
        """The same calls through emb_lookup_ex_async / emb_lookup_wait, TWO in flight: call k+1 is issued before
        call k is waited for, so its copy-in crosses PCIe while call k is pooled and its rows drain.  Every call's
        copies are inside the timed region; the region ends when the last call's rows are in host memory."""
        def run(n):
            calls[0].start()
            for k in range(n):
                if k + 1 < n:
                    calls[(k + 1) % 2].start()
                calls[k % 2].wait()
        run(4)
        t0 = time.perf_counter()
        run(n_steps)
        dt = time.perf_counter() - t0
        n_e = sum(lookups[k % 2] for k in range(n_steps))
        return n_e / dt, dt / n_steps * 1e3

    PATHS = {1: "staged: cudaMemcpyAsync in, kernel, cudaMemcpyAsync out, pipelined over table groups",
             2: "mapped: one launch, the kernel reads indices/offsets and writes pooled rows over PCIe",
             3: "inputs staged by the copy engine, pooled rows written to the host arrays by the kernel's "
                "store epilogue over PCIe, pipelined over table groups"}
    v, ms, tm = time_e2e(Ke)  # default path (EMB_HOST_AUTO) on page-locked arrays
    # the synchronous results are the check for the calls in flight: same batches, same bits -- every table of BOTH
    # calls (each call object has run synchronously by now; its result arrays are zeroed in between)
    want_bits = [[r.view(np.uint32).copy() for r in c.results] for c in calls]
    for c in calls:
        for r in c.results:
            r[:] = 0
    va, msa = time_e2e_in_flight(max(Ke, 2))
    tma = emblib.last_timings()  # of the call completed last in the in-flight run
    for c, wbs in zip(calls, want_bits):
        for t, (r, wb) in enumerate(zip(c.results, wbs)):
            if not np.array_equal(r.view(np.uint32), wb):
                raise SystemExit(f"PARITY FAILURE: emb_lookup_ex_async rows differ from emb_lookup_ex rows (table {t})")

    def stage_times(x):
        return {"path": PATHS[int(x.path)], "h2d": x.h2d_ms, "kernel": x.kernel_ms, "d2h": x.d2h_ms, "total": x.total_ms,
                "launches": int(x.launches)}
    in_flight = va >= v  # the headline is the better of the two ways to call; both are in the line
    tm_head = tma if in_flight else tm
    e2e = {"value": va if in_flight else v, "unit": "lookups/s", "h2d_bytes_per_step": int(tm_head.h2d_bytes),
           "d2h_bytes_per_step": int(tm_head.d2h_bytes), "ms_per_step": msa if in_flight else ms, "steps": Ke,
           "calls_in_flight": 2 if in_flight else 1,
           "call": ("emb_lookup_ex_async + emb_lookup_wait (host pointers, page-locked arrays carved from one arena per "
                    "kind), two calls in flight: call k+1 is issued before call k is waited for; wall clock from the "
                    "first issue to the last call's rows in host memory") if in_flight else
                   "emb_lookup_ex (host pointers, page-locked arrays carved from one arena per kind)",
           "two_calls_in_flight": {"value": va, "ms_per_step": msa, "call": "emb_lookup_ex_async + emb_lookup_wait",
                                   "last_call_ms": stage_times(tma),
                                   "parity": f"all {T} result arrays of both calls in flight equal the synchronous rows"},
           "synchronous_call": {"value": v, "ms_per_step": ms, "call": "emb_lookup_ex, one call at a time on page-locked "
                                "arrays", "last_call_ms": stage_times(tm)},
           "path": PATHS[int(tm_head.path)],
           "last_call_ms": stage_times(tm_head),
           "other_paths": {}}
    for forced in (EMB_HOST_COPY, EMB_HOST_MAPPED):
        emblib.set_host_path(forced)
        v2, ms2, _ = time_e2e(max(3, Ke // 5))
        e2e["other_paths"][PATHS[forced].split(":")[0]] = {"value": v2, "ms_per_step": ms2}
    emblib.set_host_path(EMB_HOST_AUTO)
    if rows_t is not None and not args.no_unmodified_caller:
        e2e["unmodified_caller"] = e2e_unmodified_caller(lib, emblib, rows_t, T, B, D, max(3, Ke // 2))
    return e2e


def e2e_unmodified_caller(lib, emblib, rows_t, T, B, D, n_steps, P=10):
    """What a caller that changes NOTHING gets: the reference-signature lookup() with the fixed shape of a reference
    build (MAX_NR_BATCHES = B, MAX_INDICES_PER_BATCH = P = 10), on plain pageable NumPy arrays, one
    allocation per table and kind, pointer arrays built once (dlrm_dpu_pytorch.py:484-498,1028-1051).  Then the same
    call after the one-line change INTEGRATION.md section 3 recommends: emb_host_register() of the result arrays."""
    import ctypes as C
    emblib.configure(T, D, B, P)  # the -D shape macros of the reference build (upmem/Makefile:69-81)
    rng = np.random.default_rng(77)
    idx = [rng.integers(0, rows_t[t], size=B * P, dtype=np.int64).astype(np.uint32) for t in range(T)]
    off = [(np.arange(B, dtype=np.uint32) * np.uint32(P)) for _ in range(T)]
    res = [np.empty((B, D), dtype=np.float32) for _ in range(T)]
    ip = (emblib.u32p * T)(*[a.ctypes.data_as(emblib.u32p) for a in idx])
    op = (emblib.u32p * T)(*[a.ctypes.data_as(emblib.u32p) for a in off])
    rp = (emblib.f32p * T)(*[a.ctypes.data_as(emblib.f32p) for a in res])
    n_lookups = T * B * P

    def timed():
        for _ in range(3):
            lib.lookup(ip, op, rp, None, 0)
        t0 = time.perf_counter()
        for _ in range(n_steps):
            lib.lookup(ip, op, rp, None, 0)
        dt = (time.perf_counter() - t0) / n_steps
        tm = emblib.last_timings()
        return {"value": n_lookups / dt, "unit": "lookups/s", "ms_per_step": dt * 1e3,
                "last_call_ms": {"path": int(tm.path), "h2d": tm.h2d_ms, "kernel": tm.kernel_ms, "d2h": tm.d2h_ms,
                                 "total": tm.total_ms},
                "h2d_bytes_per_step": int(tm.h2d_bytes), "d2h_bytes_per_step": int(tm.d2h_bytes)}

    out = {"shape": f"lookup(): {T} tables x {B} bags x fixed pooling {P}, pageable NumPy arrays, one per table",
           "lookup_pageable": timed()}
    # parity of the fixed-shape entry point against emb_lookup_ex on the same arrays (page-locked copies)
    want = emblib.lookup_ex(idx, off)
    for t in range(T):
        if not np.array_equal(res[t].view(np.uint32), want[t].view(np.uint32)):
            raise SystemExit(f"PARITY FAILURE: lookup() rows differ from emb_lookup_ex rows (table {t})")
    out["parity"] = f"all {T} result arrays equal emb_lookup_ex on the same arrays"
    for a in res:
        emblib.check(lib.emb_host_register(C.c_void_p(a.ctypes.data), a.nbytes), "emb_host_register")
    try:
        out["lookup_registered"] = timed()
        out["lookup_registered"]["what"] = "the same call after emb_host_register() of the result arrays only"
    finally:
        for a in res:
            lib.emb_host_unregister(C.c_void_p(a.ctypes.data))
    emblib.configure(T, D, 0, 0)
    return out


def run_host_pipeline(torch, stream, n_steps, copy_in, launch, finish, copy_out):
    """N > 1 e2e: n_steps steps from host buffers with TWO steps in flight.  Step k's host->device copy runs on its
    own stream while step k-1 pools and step k-1's / k-2's rows drain to the host on a third stream: both PCIe
    directions busy, as in the one-GPU host call's table groups (a [B/N, T, D] sample-major result cannot be cut by
    tables, so here the overlap is between steps).  Every copy of every step is inside the caller's timed region.
    Ordering: the copy-in of step k waits for the launch that read its device arena two steps ago; finish(k) waits for
    the copy-out of step k-1 (the buffer rule of ShardedTables.finish: the consumer of result(buffer) is ordered
    before the next finish); the host never has more than two steps enqueued."""
    s_in, s_out = torch.cuda.Stream(), torch.cuda.Stream()
    in_done = [torch.cuda.Event() for _ in range(2)]
    comp_done = [torch.cuda.Event() for _ in range(2)]
    out_done = [torch.cuda.Event() for _ in range(2)]
    for k in range(n_steps):
        slot = k % 2
        if k >= 2:
            out_done[slot].synchronize()
            s_in.wait_event(comp_done[slot])
        with torch.cuda.stream(s_in):
            copy_in(k)
            in_done[slot].record(s_in)
        stream.wait_event(in_done[slot])
        launch(k)
        if k >= 1:
            stream.wait_event(out_done[(k - 1) % 2])
        finish(k)
        comp_done[slot].record(stream)
        s_out.wait_event(comp_done[slot])
        with torch.cuda.stream(s_out):
            copy_out(k)
            out_done[slot].record(s_out)
    s_out.synchronize()
    stream.synchronize()


def e2e_sharded_path(args, lib, emblib, exchange, slabs, batches, lookups, T, B, D, per, rank, world, Ke, stream,
                     sync_all):
    """N > 1: the sharded product path from host buffers, per step and per rank: this rank's indices + offsets in
    ONE pinned arena -> one H2D copy into a device arena the plan reads -> pooling launch (rows stored straight
    into the owners' slabs over NVLink) -> completion kernel -> this rank's [B/N, T_global, D] rows -> one D2H copy
    into pinned memory -> synchronise."""
    import torch
    import torch.distributed as dist
    sets = []
    for k, (idx, off) in enumerate(batches[:2]):
        n_words = sum(len(i) for i in idx) + sum(len(o) for o in off)
        harena = torch.empty(n_words, dtype=torch.int32).pin_memory()
        darena = torch.empty(n_words, dtype=torch.int32, device="cuda")
        pos, di, do = 0, [], []
        for a in idx:
            harena[pos:pos + len(a)] = torch.from_numpy(a.view(np.int32))
            di.append(darena[pos:pos + len(a)])
            pos += len(a)
        for a in off:
            harena[pos:pos + len(a)] = torch.from_numpy(a.view(np.int32))
            do.append(darena[pos:pos + len(a)])
            pos += len(a)
        if slabs is not None:
            qs = [dict(table_id=t, indices=di[t], offsets=do[t], out=None, out_stride=slabs.stride, bags_per_rank=per,
                       out_peer=slabs.out_peer(t, k % 2), bag_rotate=0 if args.no_rotate else rank * per)
                  for t in range(T)]
            send = None
        else:
            send = torch.empty((B, T, D), dtype=torch.float32, device="cuda")
            qs = [dict(table_id=t, indices=di[t], offsets=do[t], out=send[:, t, :], out_stride=T * D) for t in range(T)]
        sets.append((harena, darena, emblib.Plan(qs), send))
    res_shape = tuple(slabs.local(0).shape) if slabs is not None else (per, T * world, D)  # the slab as it lies
    res_host = [torch.empty(res_shape, dtype=torch.float32).pin_memory() for _ in range(2)]
    a2a = {}

    def copy_in(k):
        harena, darena, _, _ = sets[k % 2]
        darena.copy_(harena, non_blocking=True)

    def launch(k):
        sets[k % 2][2].launch(stream)

    def finish(k):
        if slabs is not None:
            slabs.barrier()
        else:
            a2a[k % 2] = exchange.nccl_all_to_all(sets[k % 2][3], B, [T] * world, D)

    def copy_out(k):
        r = slabs.local(k % 2) if slabs is not None else a2a[k % 2]
        res_host[k % 2].copy_(r, non_blocking=True)

    run_host_pipeline(torch, stream, 4, copy_in, launch, finish, copy_out)
    sync_all()
    t0 = time.perf_counter()
    run_host_pipeline(torch, stream, Ke, copy_in, launch, finish, copy_out)
    dt = time.perf_counter() - t0
    n_e = sum(lookups[k % 2] for k in range(Ke))
    dt, n_e = reduce_over_ranks([dt, float(n_e)], world, ["max", "sum"])
    for _, _, plan, _ in sets:
        plan.close()
    return {"value": n_e / dt, "unit": "lookups/s", "h2d_bytes_per_step": int(sets[0][0].numel() * 4),
            "d2h_bytes_per_step": int(res_host[0].numel() * 4), "ms_per_step": dt / Ke * 1e3, "steps": Ke,
            "call": "sharded product path per rank: pinned index/offset arena -> HBM (1 copy), pooling launch with peer-slab "
                    "stores over NVLink, completion kernel, [B/N, T_global, D] rows -> pinned host (1 copy); two steps in "
                    "flight (copy-in, pooling and copy-out of neighbouring steps overlap); bytes are per rank"}


# ---------------------------------------------------------------------------------------------
# sharded arm (C4): ONE global problem cut over the ranks -- strong scaling
# ---------------------------------------------------------------------------------------------

def bench_sharded(args, w, steps, warmup):
    import torch
    import torch.distributed as dist

    from pim_embedding_lookup_b200 import emblib, sharding
    from pim_embedding_lookup_b200.sharded import ShardedTables

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    lib = emblib.load()
    rows, D, B = w["rows"], w["D"], w["B"]
    T = len(rows)
    P = (w["pool"][0] + w["pool"][1]) / 2.0
    emblib.check(lib.emb_set_tuning(args.variant, args.unroll, args.mult))
    kw = {}
    if args.max_table_bytes or w.get("max_table_bytes"):
        kw["max_table_bytes"] = args.max_table_bytes or w["max_table_bytes"]
    st = ShardedTables(rows, D, B, lookups=[B * P] * T, hbm_budget=150e9, layout=args.layout,
                       rotate=not args.no_rotate, **kw)
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)

    # local shards: int32 fixed point as export_emb makes them, generated on the device
    gen = torch.Generator(device="cuda").manual_seed(1234 + rank)
    for slot, sh in enumerate(st.slots):
        lim = int(1e9 / np.sqrt(rows[sh.table]))
        a = torch.randint(-lim, lim + 1, (sh.rows, D), dtype=torch.int32, device="cuda", generator=gen)
        st.populate_shard(slot, a, device_src=True, stream=stream)
        del a
    torch.cuda.empty_cache()

    need = st.tables_needed()
    host_batches = [{t: make_table_batch(w, t, 123 + k) for t in need} for k in range(N_BATCHES)]
    plans, dev = [], []
    for k, hb in enumerate(host_batches):
        di = {t: torch.from_numpy(hb[t][0].view(np.int32)).cuda() for t in need}
        do = {t: torch.from_numpy(hb[t][1].view(np.int32)).cuda() for t in need}
        dev.append((di, do))
        plans.append(st.make_plan(di, do, buffer=k % 2, overlap=not args.no_overlap))
    lookups_global = float(T * B * P)  # every (table, index) of the global batch, counted once
    # algorithmic bytes of THIS rank's launch: the indices its shards read, the rows in their ranges, every bag
    # writes one row and reads one offset
    my_bytes = sum(sharding.shard_cost(B * P, B, D, 1.0) * (sh.bag_hi - sh.bag_lo) / B if sh.sliced else
                   sharding.shard_cost(B * P, B, D, sh.rows / rows[sh.table]) for sh in st.local)

    def sync_all():
        torch.cuda.synchronize()
        dist.barrier()
        torch.cuda.synchronize()

    def step(k):
        plans[k % N_BATCHES].launch(stream)
        st.finish(stream, k % 2)

    for k in range(max(warmup, 3)):
        step(k)
    sync_all()
    if lib.emb_check_indices():
        raise RuntimeError("out-of-range index in the synthetic batch")

    # ---- parity: every pooled row of one batch on every rank.  Each rank restates, with stock torch ops on ITS
    # shards, the raw int32 sums they contribute (a row-range shard: only the rows it holds; a bag-sliced replica:
    # only its bags); the contributions are wrap-added across the ranks with an NCCL all-reduce, dequantised once
    # (emb_host.h:210) and compared bit for bit with what the fused exchange + completion kernel left in this
    # rank's [B/N, T, D] slab.
    parity = {"checked": False}
    if not args.no_parity:
        kb = 1
        step(kb)
        sync_all()
        got = st.result_sample_major(kb % 2).clone()
        di, do = dev[kb]
        contrib = torch.zeros((T, B, D), dtype=torch.int32, device="cuda")
        for j, sh in enumerate(st.local):
            tab = table_tensor(lib, st.slot_of[j], D, "int32")
            sums = torch_pooled(tab, di[sh.table], do[sh.table], B, row_lo=sh.row_lo, raw=True)
            lo, hi = (sh.bag_lo, sh.bag_hi) if sh.sliced else (0, B)
            contrib[sh.table, lo:hi] += sums[lo:hi]
            del sums
        dist.all_reduce(contrib, op=dist.ReduceOp.SUM)  # int32: wrap-around, as the reference's adds
        mine = contrib[:, rank * st.per:(rank + 1) * st.per, :]
        want = (mine.to(torch.float32) / torch.tensor(1e9, dtype=torch.float32, device="cuda")).permute(1, 0, 2).contiguous()
        bad = compare_rows(got, want)
        chk_got = int(got.contiguous().view(torch.int32).to(torch.int64).sum().item()) & 0xffffffffffffffff
        chk_want = int(want.view(torch.int32).to(torch.int64).sum().item()) & 0xffffffffffffffff
        bad_all, rows_all, chk_bad = reduce_over_ranks([float(bad), float(got.numel() // D), float(chk_got != chk_want)],
                                                       world, ["sum", "sum", "sum"])
        idx_err = lib.emb_check_indices()
        parity = {"checked": True, "ranks": world, "rows_checked": int(rows_all), "mismatching_elements": int(bad_all),
                  "bit_exact": bool(bad_all == 0), "checksum_of_checksums_equal": chk_bad == 0, "index_errors": int(idx_err),
                  "against": "torch index_select + index_add_ on every rank's own shards (raw int32 contributions), NCCL "
                             "all-reduce (wrap-around sum over shards), one fp32 divide (emb_dpu_lookup.c:106-116, "
                             "emb_host.h:207-212, extend_distributed.py:389-426)"}
        del contrib, mine, want, got
        torch.cuda.empty_cache()
        if bad_all or chk_bad or idx_err:
            raise SystemExit(f"PARITY FAILURE (sharded) on rank {rank}: {json.dumps(parity)}")

    # timed region: K steps back to back; then the same K steps with every pooling launch between two events
    # (an isolated launch -- see bench_tablewise)
    K = steps
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches0 = emblib.launch_count()
    sync_all()
    e0.record(stream)
    for k in range(K):
        step(k)
    e1.record(stream)
    sync_all()
    gpu_launches = emblib.launch_count() - launches0
    total_ms = e0.elapsed_time(e1)
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    for k in range(K):
        ev[k][0].record(stream)
        plans[k % N_BATCHES].launch(stream)
        ev[k][1].record(stream)
        st.finish(stream, k % 2)
    sync_all()
    kdist = dist_summary([a.elapsed_time(b) for a, b in ev])
    kernel_ms = kdist["mean"]
    total_ms_max, kernel_ms_max, agg_gbps = reduce_over_ranks(
        [total_ms, kernel_ms, my_bytes / (kernel_ms * 1e-3) / 1e9], world, ["max", "max", "sum"])
    # every rank's isolated pooling launch and the work the planner gave it: what the cost model is judged by
    my_lookups = sum(B * P * ((sh.bag_hi - sh.bag_lo) / B if sh.sliced else 1.0) for sh in st.local)
    kernel_ms_per_rank, bytes_per_rank, lookups_per_rank = gather_over_ranks([kernel_ms, my_bytes, my_lookups], world)
    value = lookups_global * K / (total_ms_max * 1e-3)
    peak, peak_src = load_peaks()
    achieved = my_bytes / (kernel_ms * 1e-3) / 1e9  # rank 0's pooling launch

    # ---- e2e: per step, this rank's index/offset arrays come from ONE page-locked arena and its [B/N, T, D]
    # result goes back to page-locked host memory, all inside the timed region
    e2e = None
    if not args.no_e2e:
        sets = []
        for k, hb in enumerate(host_batches[:2]):
            n_words = sum(len(hb[t][0]) + len(hb[t][1]) for t in need)
            harena = torch.empty(n_words, dtype=torch.int32).pin_memory()
            darena = torch.empty(n_words, dtype=torch.int32, device="cuda")
            pos, di, do = 0, {}, {}
            for t in need:
                for arr, dct in ((hb[t][0], di), (hb[t][1], do)):
                    harena[pos:pos + len(arr)] = torch.from_numpy(arr.view(np.int32))
                    dct[t] = darena[pos:pos + len(arr)]
                    pos += len(arr)
            sets.append((harena, darena, st.make_plan(di, do, buffer=k % 2)))
        res_host = torch.empty(tuple(st.result(0).shape), dtype=torch.float32).pin_memory()
        h2d, d2h = sets[0][0].numel() * 4, res_host.numel() * 4

        res_hosts = [res_host, torch.empty_like(res_host).pin_memory()]
        steps_e2e = (lambda k: sets[k % 2][1].copy_(sets[k % 2][0], non_blocking=True),
                     lambda k: sets[k % 2][2].launch(stream),
                     lambda k: st.finish(stream, k % 2),
                     lambda k: res_hosts[k % 2].copy_(st.result(k % 2), non_blocking=True))
        run_host_pipeline(torch, stream, 4, *steps_e2e)
        sync_all()
        Ke = max(3, min(K, 50))
        t0 = time.perf_counter()
        run_host_pipeline(torch, stream, Ke, *steps_e2e)
        dt = time.perf_counter() - t0
        (dt,) = reduce_over_ranks([dt], world, ["max"])
        for _, _, p in sets:
            p.close()
        e2e = {"value": lookups_global * Ke / dt, "unit": "lookups/s", "h2d_bytes_per_step": int(h2d),
               "d2h_bytes_per_step": int(d2h), "ms_per_step": dt / Ke * 1e3, "steps": Ke,
               "call": "ShardedTables.lookup per rank: pinned index/offset arena -> HBM (1 copy), pooling launch, peer-slab "
                       "exchange + completion kernel (barrier + partial reduction), [B/N, T, D] rows -> pinned host (1 copy); "
                       "two steps in flight; bytes are rank 0's",
               "floor_ms": measure_copy_floor(h2d, d2h, world)}

    out = None
    if rank == 0:
        split = {str(tt): k for tt, (_, k) in st.split.items()}
        sliced = {}
        for sh in st.plan.shards:
            if sh.sliced:
                sliced[str(sh.table)] = sliced.get(str(sh.table), 0) + 1
        out = {
            "metric": "embedding lookups/sec", "value": value, "unit": "lookups/s", "n_gpus": world, "steps": K,
            "warmup": max(warmup, 3), "ms_per_step": total_ms_max / K, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "int32", "data": "synthetic",
            "config": job_config(w, world, args, True),
            "run": {"lookups_per_step_global": int(lookups_global), "table_bytes_total": int(sum(rows) * D * 4),
                    "exchange": f"fused: peer-slab stores from the pooling kernel ({args.layout}-major slabs, double buffered) + "
                                "one completion kernel (flag barrier over NVLink + partial reduction)",
                    "placement": {"shards_on_rank0": len(st.local), "row_split_tables": split, "bag_sliced_replicas": sliced,
                                  "planner_imbalance": round(st.plan.imbalance(), 4),
                                  "hbm_bytes_per_rank": [int(x) for x in st.plan.hbm_bytes]},
                    "tuning": {"variant": args.variant, "unroll": args.unroll, "mult": args.mult}},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": None, "peak_source": peak_src, "kernel_ms": kernel_ms, "kernel_ms_dist": kdist,
                         "kernel_ms_max_over_ranks": kernel_ms_max, "kernel_ms_per_rank": kernel_ms_per_rank,
                         "kernel_ms_max_over_mean": round(max(kernel_ms_per_rank) / (sum(kernel_ms_per_rank) / world), 4),
                         "algorithmic_bytes_per_rank": bytes_per_rank, "index_positions_read_per_rank": lookups_per_rank,
                         "algorithmic_bytes_per_launch": my_bytes,
                         "aggregate_GBps_all_ranks": agg_gbps, "note": "rank 0's pooling launch; bytes = indices + "
                         "in-range rows + outputs + offsets of its shards (sharding.shard_cost)"},
            "parity": parity, "gpu_launches": int(gpu_launches),
        }
        if e2e:
            out["e2e"] = e2e
    sync_all()
    for p in plans:
        p.close()
    st.close()
    emblib.shutdown()
    torch.cuda.empty_cache()
    return out


def condensed(r):
    """What an `extra` workload contributes to the main JSON line."""
    keep = {k: r[k] for k in ("value", "unit", "ms_per_step", "steps", "scaling", "gpu_launches")}
    # compact: the driver keeps only the tail of the line, and every extra must fit in it
    keep["parity"] = {k: v for k, v in r["parity"].items() if k != "against" and v is not None}
    keep["workload"] = r["config"]["workload"]
    keep["roofline"] = {k: r["roofline"].get(k) for k in ("achieved", "frac", "kernel_ms", "kernel_ms_max_over_ranks",
                                                           "kernel_ms_per_rank", "kernel_ms_max_over_mean",
                                                           "algorithmic_bytes_per_launch", "aggregate_GBps_all_ranks")}
    if "placement" in r.get("run", {}):
        keep["placement"] = r["run"]["placement"]
    if "e2e" in r:
        keep["e2e"] = {k: r["e2e"].get(k) for k in ("value", "ms_per_step", "floor_ms", "h2d_bytes_per_step",
                                                     "d2h_bytes_per_step")}
    return keep


def small_call_c1(calls=200, max_rows=2_000_000):
    """BASELINE.json configs[0] (C1): the reference's --mini-batch-size=1 inference shape -- ONE lookup() call over the
    26 Kaggle tables, one bag of one index each (dlrm_s_pytorch.py:444,492-494) -- as host wall time per synchronous
    call, beside nn.EmbeddingBag per table on the host cores (tools/small_call_latency.py).  Launch-bound, so the
    figure is microseconds per call, not lookups/s."""
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import small_call_latency as scl
    r = scl.run("c1", scl.PRESETS["c1"], calls, max_rows)
    return {"workload": r["what"], "unit": "us per lookup() call (median of %d)" % calls, "higher_is_better": False,
            "lookup_on_numpy_arrays": r["lookup_pageable"], "lookup_on_page_locked_arrays": r["lookup_pinned"],
            "plan_launch_kernel": r["plan_launch_kernel"], "cpu_embedding_bag_all_cores": r["cpu_all_cores"],
            "cpu_embedding_bag_one_core": r["cpu_one_core"], "rows_capped_at": r.get("rows_capped_at"),
            "parity": {"host_and_device_flows_bit_equal": r["pageable_pinned_device_bits_equal"],
                       "max_abs_diff_vs_cpu_float": r["max_abs_diff_vs_cpu_float"]},
            "speedup_vs_cpu_all_cores": r["speedup_vs_cpu_all_cores"]}


def run_b200(args, w):
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torch.distributed.run")
    if not torch.cuda.is_available():
        raise SystemExit("no CUDA device: the B200 arm has no CPU stand-in (use --impl reference for the CPU path)")
    torch.cuda.set_device(local)
    extras = args.extras
    if extras == "auto":
        # Every BASELINE.json config and both element kinds in the one line: c3 = configs[2]'s lookup shape, c1 =
        # configs[0] (one lookup() call at mini-batch 1, N = 1 only), f32 = the default workload on fp32 tables;
        # c4rc (tables cut by rows) needs ranks to cut over.  c3 / f32 / c1 are one-GPU figures (c3's batch of 39292 does
        # not divide over 8 ranks).
        extras = (("c4tb,c5,c3,f32,c1" if world == 1 else "c4tb,c4rc,c5")
                  if (args.workload == "c2" and args.dtype == "int32") else "none")
    extras = [e for e in extras.split(",") if e and e != "none"]
    need_group = world > 1 or bool(w.get("sharded")) or any(WORKLOADS.get(e, {}).get("sharded") for e in extras)
    if need_group:
        if world == 1:  # ShardedTables wants a process group even for one rank
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            os.environ.setdefault("MASTER_PORT", "29533")
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", local))

    def one(wl, steps, warmup):
        return bench_sharded(args, wl, steps, warmup) if wl.get("sharded") else bench_tablewise(args, wl, steps, warmup)

    with ClockSampler(local) as clk:
        out = one(w, args.steps, args.warmup)
        extra = {}
        for name in extras:
            if name in WORKLOADS and WORKLOADS[name]["B"] % world:
                if rank == 0:
                    extra[name] = {"skipped": f"batch {WORKLOADS[name]['B']} does not divide over {world} ranks"}
                continue
            if name == "c1":
                if world == 1:
                    extra[name] = small_call_c1()
                continue
            saved = args.no_cpu_baseline, args.no_unmodified_caller, args.dtype
            args.no_cpu_baseline = args.no_unmodified_caller = True
            if name == "f32":
                args.dtype = "f32"
            try:
                r = one(w if name == "f32" else WORKLOADS[name], max(3, min(args.steps, 50)), min(args.warmup, 5))
            finally:
                args.no_cpu_baseline, args.no_unmodified_caller, args.dtype = saved
            if rank == 0:
                extra[name] = condensed(r)
                if name == "f32":
                    extra[name]["dtype"] = "f32"
    if rank == 0:
        out["clocks"] = clk.summary()
        if extra:
            out["extra"] = extra
        # the headline, its parity block and the launch count first; the extras last (the driver keeps head and tail)
        front = ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
                 "vs_baseline", "dtype", "data", "parity", "gpu_launches")
        out = {**{k: out[k] for k in front if k in out}, **{k: v for k, v in out.items() if k not in front}}
        print(json.dumps(out))
    if need_group:
        dist.destroy_process_group()


def main():
    args = parse_args()
    w = WORKLOADS[args.workload]
    # stdout carries exactly one JSON line: libraries that print banners there (NCCL's version line
    # with NCCL_DEBUG=VERSION) are sent to stderr for the duration of the run
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(real_stdout, "w")
    if args.impl == "reference":
        run_reference(args, w)
    else:
        run_b200(args, w)
    sys.stdout.flush()


if __name__ == "__main__":
    main()
