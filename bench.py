#!/usr/bin/env python
"""bench.py -- embedding lookups/s of the sum-pool hot path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload c2|c4|c5]

One "step" = one pass of the hot path over one batch: every local table pools its B bags.
N=1 workload = BASELINE.json configs[1] ("c2"): 26 tables x 1M rows x dim 16, batch 65536,
pooling U{1..20}, uniform indices, int32 fixed-point tables (the DPU representation).
N>1 (torchrun, one rank per GPU): weak scaling -- every rank owns 26 such tables and pools
the whole global batch for them; pooled rows go to the rank that owns each sample
(table-parallel -> batch-parallel, extend_distributed.py:389-426) either straight from the
pooling kernel over NVLink peer memory (--exchange fused, default) or by one NCCL
all_to_all_single (--exchange nccl).

Numbers: `value` = lookups/s with inputs resident in HBM (CUDA events, max over ranks);
`e2e` = the same through the host-pointer C-ABI call (emb_lookup_ex: pinned host buffers,
H2D + kernel + D2H inside the timed region); `roofline` = algorithmic bytes of one pooling
launch / its mean CUDA-event duration against MEASURED_PEAKS.json; `cpu_baseline` = the
oracle port (multi-threaded C) on this box's host cores, N=1 only.
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
    # configs[3] shape per GPU (dim 64): uniform, fixed pooling 10
    # configs[2] lookup shape: Kaggle cardinalities (tools/visualize.py:949), dim 16, batch 39292, one index per bag
    "c3": dict(name="Kaggle-cardinality tables x dim 16, batch 39292, pooling 1, uniform indices", T=26, R=10131227,
               rows=[1460, 583, 10131227, 2202608, 305, 24, 12517, 633, 3, 93145, 5683, 8351593, 3194, 27, 14992,
                     5461306, 10, 5652, 2173, 4, 7046547, 18, 15, 286181, 105, 142572],
               D=16, B=39292, pool=(1, 1), dist="uniform"),
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
N_BATCHES = 4  # distinct index batches rotated through the timed steps


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS))
    ap.add_argument("--exchange", default="fused", choices=["fused", "nccl"])
    ap.add_argument("--layout", default="sample", choices=["table", "sample"],
                    help="peer-slab layout of the fused exchange: [T][B/N][D] or [B/N][T][D]")
    ap.add_argument("--no-rotate", action="store_true",
                    help="fused exchange: every rank walks the destination ranks in the same order (comparator)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--variant", type=int, default=0, help="kernel variant (emb_set_tuning); 0 = default")
    ap.add_argument("--unroll", type=int, default=0)
    ap.add_argument("--mult", type=int, default=0)
    return ap.parse_args()


# ---------------------------------------------------------------------------------------------
# synthetic inputs: seeded NumPy on the host
# ---------------------------------------------------------------------------------------------

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


def algorithmic_bytes(n_lookups, n_bags, D):
    # per lookup 4D (row) + 4 (index); per bag 4D (output row) + 4 (offset)
    return (n_lookups + n_bags) * (4 * D + 4)


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
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
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
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(s)}


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


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


def time_torch_embeddingbag(tables_f32, idx, off, budget_s=10.0):
    """The reference's CPU path: nn.EmbeddingBag(mode="sum") per table (dlrm_s_pytorch.py:483-488)."""
    import torch
    torch.set_num_threads(host_cores())
    bags = [torch.nn.EmbeddingBag.from_pretrained(torch.from_numpy(w), mode="sum", freeze=True) for w in tables_f32]
    ti = [torch.from_numpy(i.astype(np.int64)) for i in idx]
    to = [torch.from_numpy(o.astype(np.int64)) for o in off]
    n = sum(len(i) for i in idx)

    def one():
        with torch.no_grad():
            for k, e in enumerate(bags):
                e(ti[k], to[k])
    one()
    reps, t_tot = 0, 0.0
    while reps < 2 or (t_tot < budget_s and reps < 50):
        t0 = time.perf_counter()
        one()
        t_tot += time.perf_counter() - t0
        reps += 1
    return n * reps / t_tot, torch.get_num_threads(), reps


def host_tables(w, n_tables, seed=1234):
    """Quantised int32 tables as export_emb would make them (U(-1/sqrt(R), 1/sqrt(R)) * 1e9)."""
    rng = np.random.default_rng(seed)
    rows = w.get("rows", [w["R"]] * w["T"])
    return [rng.integers(-int(1e9 / np.sqrt(rows[t])), int(1e9 / np.sqrt(rows[t])) + 1, size=(rows[t], w["D"]),
                         dtype=np.int32) for t in range(n_tables)]


def run_reference(args, w):
    """--impl reference: the CPU implementation alone, all host threads, bounded sample."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    # bounded sample of the workload: the first T_s tables of one batch (same shapes and
    # distributions as the GPU arm); each step = one pass over the sample
    T_s = min(w["T"], 8)
    if w.get("sharded"):
        # the 40M-row tables are 10 GB each (plus an fp32 copy): the bounded sample takes the largest
        # table that is cheap to build on the host (2.95M rows, 756 MB) and seven smaller ones
        pick = [10, 22, 11, 1, 4, 2, 23, 14]
        w = dict(w, rows=[w["rows"][t] for t in pick], T=len(pick))
    tabs = host_tables(w, T_s)
    idx, off = make_batch(w, seed=123, n_tables=T_s)
    import torch
    # torchrun exports OMP_NUM_THREADS=1; the reference arm uses every host core it may run on
    torch.set_num_threads(host_cores())
    n = sum(len(i) for i in idx)
    # the reference's CPU path: nn.EmbeddingBag(mode="sum", sparse=True) per table, fp32 tables,
    # int64 indices/offsets (dlrm_s_pytorch.py:295,483-488), all host threads
    ftabs = [t.astype(np.float32) * np.float32(1e-9) for t in tabs]
    bags = [torch.nn.EmbeddingBag.from_pretrained(torch.from_numpy(x), mode="sum", sparse=True, freeze=True)
            for x in ftabs]
    ti = [torch.from_numpy(i.astype(np.int64)) for i in idx]
    to = [torch.from_numpy(o.astype(np.int64)) for o in off]

    def one():
        with torch.no_grad():
            for k, e in enumerate(bags):
                e(ti[k], to[k])
    for _ in range(max(1, min(args.warmup, 5))):
        one()
    steps = max(1, min(args.steps, 200))
    t0 = time.perf_counter()
    done = 0
    while done < steps and (done < 3 or time.perf_counter() - t0 < 20.0):
        one()
        done += 1
    dt = time.perf_counter() - t0
    steps = done
    val = n * steps / dt
    pv, pcores, _ = time_oracle_port(tabs, idx, off, budget_s=6.0)
    sample = f"{T_s} of {w['T']} tables x {w['B']} bags ({n} lookups) per step, {steps} steps"
    print(json.dumps({
        "impl": "reference", "metric": "embedding lookups/sec", "value": val, "unit": "lookups/s",
        "n_gpus": args.gpus, "steps": steps, "warmup": args.warmup, "ms_per_step": dt / steps * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": w["name"], "sample": sample},
        "cpu_baseline": {"value": val, "unit": "lookups/s", "cores": torch.get_num_threads(), "kind": "reference",
                         "sample": sample,
                         "what": "the reference's CPU lookup path restated call for call: nn.EmbeddingBag(mode=sum) "
                                 "per table on stock torch (dlrm_s_pytorch.py:295,483-488); its DPU path "
                                 "(oracle/_ref) is undefined at this shape (B<=512, T*D<=256; SURVEY.md Q3/Q4/Q13)",
                         "oracle_port": {"value": pv, "unit": "lookups/s", "cores": pcores,
                                         "what": "fixed-point restatement (oracle/emb_oracle.c), multi-threaded C"}},
        "e2e": {"value": val, "unit": "lookups/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


# ---------------------------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------------------------

def pinned_u32(lib, n):
    """n uint32 of cudaHostAlloc memory as a NumPy array (emb_host_alloc)."""
    p = C.c_void_p()
    rc = lib.emb_host_alloc(C.byref(p), max(4 * n, 4))
    if rc:
        raise RuntimeError("emb_host_alloc failed")
    return np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint32)), shape=(n,))


def run_b200(args, w):
    import torch
    import torch.distributed as dist

    from pim_embedding_lookup_b200 import emblib, exchange
    from pim_embedding_lookup_b200.dputypes import EMB_HOST_AUTO, EMB_HOST_COPY, EMB_HOST_MAPPED, EMB_TABLE_FIXED

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torch.distributed.run")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    lib = emblib.load()
    T, R, D, B = w["T"], w["R"], w["D"], w["B"]
    emblib.configure(T, D, 0, 0, device=local)
    emblib.check(lib.emb_set_tuning(args.variant, args.unroll, args.mult))
    stream = torch.cuda.Stream()  # a real stream: events and launches are ordered on it
    torch.cuda.set_stream(stream)

    # tables: int32 fixed point as export_emb makes them, generated on the device
    gen = torch.Generator(device="cuda").manual_seed(1234 + rank)
    rows_t = w.get("rows", [R] * T)
    for t in range(T):
        lim = int(1e9 / np.sqrt(rows_t[t]))
        a = torch.randint(-lim, lim + 1, (rows_t[t], D), dtype=torch.int32, device="cuda", generator=gen)
        emblib.check(lib.emb_populate_table_device(t, rows_t[t], a.data_ptr(), 0, EMB_TABLE_FIXED, stream.cuda_stream))
        del a
    torch.cuda.empty_cache()

    # index batches: every rank pools the whole global batch for its own tables
    batches = [make_batch(w, seed=123 + 1000 * rank + k) for k in range(N_BATCHES)]
    lookups = [sum(len(i) for i in b[0]) for b in batches]
    dev_batches = []
    for idx, off in batches:
        dev_batches.append(([torch.from_numpy(i.view(np.int32)).cuda() for i in idx],
                            [torch.from_numpy(o.view(np.int32)).cuda() for o in off]))

    # outputs + plans (descriptors uploaded once; a step is one kernel launch)
    slabs = None
    if world > 1 and args.exchange == "fused":
        slabs = exchange.PeerSlabs(B, [T] * world, D, layout=args.layout)
    per = B // world
    plans, keep = [], []
    for k, (di, do) in enumerate(dev_batches):
        qs = []
        if slabs is not None:
            for t in range(T):
                qs.append(dict(table_id=t, indices=di[t], offsets=do[t], out=None, out_stride=slabs.stride,
                               bags_per_rank=per, out_peer=slabs.out_peer(t),
                               bag_rotate=0 if args.no_rotate else rank * per))
        elif world > 1:  # nccl: send buffer [N, B/N, T_local, D] == bag-major rows of stride T*D
            buf = torch.empty((B, T, D), dtype=torch.float32, device="cuda")
            keep.append(buf)
            for t in range(T):
                qs.append(dict(table_id=t, indices=di[t], offsets=do[t], out=buf[:, t, :], out_stride=T * D))
        else:  # the reference's layout: one [B, D] result per table (emb_host.h:207-212)
            buf = torch.empty((T, B, D), dtype=torch.float32, device="cuda")
            keep.append(buf)
            for t in range(T):
                qs.append(dict(table_id=t, indices=di[t], offsets=do[t], out=buf[t]))
        plans.append(emblib.Plan(qs))

    def step(k):
        plans[k % N_BATCHES].launch(stream)
        if slabs is not None:
            slabs.barrier()
        elif world > 1:
            exchange.nccl_all_to_all(keep[k % N_BATCHES], B, [T] * world, D)

    def sync_all():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    for k in range(max(args.warmup, 3)):
        step(k)
    sync_all()
    if lib.emb_check_indices():
        raise RuntimeError("out-of-range index in the synthetic batch")

    K = args.steps
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches0 = emblib.launch_count()
    with ClockSampler(local) as clk:
        sync_all()
        e0.record(stream)
        for k in range(K):
            ev[k][0].record(stream)
            plans[k % N_BATCHES].launch(stream)
            ev[k][1].record(stream)
            if slabs is not None:
                slabs.barrier()
            elif world > 1:
                exchange.nccl_all_to_all(keep[k % N_BATCHES], B, [T] * world, D)
        e1.record(stream)
        sync_all()
    gpu_launches = emblib.launch_count() - launches0
    total_ms = e0.elapsed_time(e1)
    kernel_ms = float(np.mean([a.elapsed_time(b) for a, b in ev]))
    n_lookups = sum(lookups[k % N_BATCHES] for k in range(K))
    t = torch.tensor([total_ms, float(n_lookups), kernel_ms], dtype=torch.float64, device="cuda")
    if world > 1:
        tmax = t.clone()
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        tsum = t.clone()
        dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
        total_ms, n_lookups, kernel_ms = float(tmax[0]), float(tsum[1]), float(tmax[2])
    value = n_lookups / (total_ms * 1e-3)

    # roofline of the pooling kernel on this rank (algorithmic bytes / mean launch duration)
    peak, peak_src = load_peaks()
    bytes_per_launch = float(np.mean([algorithmic_bytes(lookups[k % N_BATCHES], T * B, D) for k in range(K)]))
    achieved = bytes_per_launch / (kernel_ms * 1e-3) / 1e9

    # DRAM traffic of the same kernel from the committed ncu capture (per launch), if there is one
    traffic, traffic_src = None, None
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    if world == 1 and os.path.exists(tp):
        rec = json.load(open(tp)).get(args.workload)
        if rec:
            traffic = rec["dram_bytes_read_per_launch"] + rec["dram_bytes_write_per_launch"]
            traffic_src = rec["source"]

    # ---- e2e: the host-pointer C-ABI call, pinned host buffers, copies inside the timed region
    e2e = None
    if not args.no_e2e:
        host = []
        for idx, off in batches[:2]:
            hi = [pinned_u32(lib, len(i)) for i in idx]
            ho = [pinned_u32(lib, len(o)) for o in off]
            for a, b in zip(hi + ho, idx + off):
                a[:] = b
            host.append((hi, ho))
        res = [pinned_u32(lib, B * D).view(np.float32).reshape(B, D) for _ in range(T)]
        # pointer arrays built once per batch, as the reference's caller does (dlrm_dpu_pytorch.py:1028-1051)
        calls = [emblib.HostCall(hi, ho, res) for hi, ho in host]

        def time_e2e(n_steps):
            for k in range(3):
                calls[k % 2]()
            sync_all()
            t0 = time.perf_counter()
            for k in range(n_steps):
                calls[k % 2]()  # emb_lookup_ex: synchronous, the results are in host memory on return
            dt = time.perf_counter() - t0
            tm = emblib.last_timings()
            n_e = sum(lookups[k % 2] for k in range(n_steps))
            if world > 1:
                te = torch.tensor([dt, float(n_e)], dtype=torch.float64, device="cuda")
                a = te.clone()
                dist.all_reduce(a, op=dist.ReduceOp.MAX)
                b = te.clone()
                dist.all_reduce(b, op=dist.ReduceOp.SUM)
                dt, n_e = float(a[0]), float(b[1])
            return n_e / dt, dt / n_steps * 1e3, tm

        PATHS = {1: "staged: cudaMemcpyAsync in, kernel, cudaMemcpyAsync out, pipelined over table groups",
                 2: "mapped: one launch, the kernel reads indices/offsets and writes pooled rows over PCIe",
                 3: "inputs staged by the copy engine, pooled rows written to the host arrays by the kernel's "
                    "store epilogue over PCIe, pipelined over table groups"}
        Ke = max(3, min(K, 50))
        v, ms, tm = time_e2e(Ke)  # default path (EMB_HOST_AUTO) on page-locked arrays
        e2e = {"value": v, "unit": "lookups/s", "h2d_bytes_per_step": int(tm.h2d_bytes),
               "d2h_bytes_per_step": int(tm.d2h_bytes), "ms_per_step": ms, "steps": Ke,
               "call": "emb_lookup_ex (host pointers in page-locked memory), per rank", "path": PATHS[int(tm.path)],
               "last_call_ms": {"h2d": tm.h2d_ms, "kernel": tm.kernel_ms, "d2h": tm.d2h_ms, "total": tm.total_ms},
               "other_paths": {}}
        for forced in (EMB_HOST_COPY, EMB_HOST_MAPPED):
            emblib.set_host_path(forced)
            v2, ms2, tm2 = time_e2e(max(3, Ke // 5))
            e2e["other_paths"][PATHS[forced].split(":")[0]] = {"value": v2, "ms_per_step": ms2}
        emblib.set_host_path(EMB_HOST_AUTO)

    # ---- CPU baseline on this box's host cores (rank 0, N=1 only)
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        T_s = min(T, 8)
        tabs = [emblib.table_read(t, 0, rows_t[t]) for t in range(T_s)]
        sidx, soff = batches[0][0][:T_s], batches[0][1][:T_s]
        ftabs = [x.astype(np.float32) * np.float32(1e-9) for x in tabs]
        tv, tth, treps = time_torch_embeddingbag(ftabs, sidx, soff)
        sample = f"{T_s} of {T} tables x {B} bags ({sum(len(i) for i in sidx)} lookups) x {treps} passes"
        cpu = {"value": tv, "unit": "lookups/s", "cores": tth, "kind": "reference", "sample": sample,
               "what": "nn.EmbeddingBag(mode=sum) per table, fp32, stock torch (dlrm_s_pytorch.py:295,483-488)"}
        v, cores, reps = time_oracle_port(tabs, sidx, soff, budget_s=6.0)
        cpu["oracle_port"] = {"value": v, "unit": "lookups/s", "cores": cores, "passes": reps, "kind": "port"}

    if rank == 0:
        out = {
            "metric": "embedding lookups/sec", "value": value, "unit": "lookups/s", "n_gpus": world, "steps": K,
            "warmup": max(args.warmup, 3), "ms_per_step": total_ms / K, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "int32", "data": "synthetic",
            "config": {"workload": w["name"], "tables_per_gpu": T, "rows": R, "dim": D, "global_batch": B,
                       "lookups_per_step_per_gpu": int(np.mean(lookups)),
                       "l2": f"inputs larger than L2: tables {T * R * D * 4 / 1e9:.2f} GB per GPU, "
                             f"{N_BATCHES} rotating index batches",
                       "exchange": ("none (1 GPU)" if world == 1 else args.exchange +
                                    (f", {args.layout}-major peer slabs" if args.exchange == "fused" else "")),
                       "tuning": {"variant": args.variant, "unroll": args.unroll, "mult": args.mult},
                       "parallelism": f"table-wise x{world}, pooled rows to the sample's rank"},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src, "kernel_ms": kernel_ms,
                         "algorithmic_bytes_per_launch": bytes_per_launch,
                         "frac_of_nominal_8000": achieved / 8000.0},
            "clocks": clk.summary(), "gpu_launches": int(gpu_launches),
        }
        if e2e:
            out["e2e"] = e2e
        if cpu:
            out["cpu_baseline"] = cpu
        print(json.dumps(out))
    sync_all()
    for p in plans:
        p.close()
    if slabs is not None:
        slabs.close()
    emblib.shutdown()
    if world > 1:
        dist.destroy_process_group()


# ---------------------------------------------------------------------------------------------
# sharded arm (C4): ONE global problem cut over the ranks -- strong scaling
# ---------------------------------------------------------------------------------------------

def make_table_batch(w, t, seed):
    """(indices, offsets) of global table t for one global batch; the same on every rank."""
    rng = np.random.default_rng([seed, t])
    lo, hi = w["pool"]
    lens = rng.integers(lo, hi + 1, size=w["B"])
    off = np.concatenate([[0], np.cumsum(lens)[:-1]]).astype(np.uint32)
    idx = rng.integers(0, w["rows"][t], size=int(lens.sum()), dtype=np.int64).astype(np.uint32)
    return idx, off


def run_sharded(args, w):
    import torch
    import torch.distributed as dist

    from pim_embedding_lookup_b200 import emblib, sharding
    from pim_embedding_lookup_b200.sharded import ShardedTables

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torch.distributed.run")
    torch.cuda.set_device(local)
    if world == 1:  # ShardedTables wants a process group even for one rank
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29533")
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", local))
    lib = emblib.load()
    rows, D, B = w["rows"], w["D"], w["B"]
    T = len(rows)
    P = (w["pool"][0] + w["pool"][1]) / 2.0
    emblib.check(lib.emb_set_tuning(args.variant, args.unroll, args.mult))
    st = ShardedTables(rows, D, B, lookups=[B * P] * T, hbm_budget=150e9, layout=args.layout,
                       rotate=not args.no_rotate)
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
        plans.append(st.make_plan(di, do, buffer=k % 2))
    lookups_global = float(T * B * P)  # every (table, index) of the global batch, counted once
    # algorithmic bytes of THIS rank's launch: every index of its shards' tables is read, a row-range
    # shard fetches its fraction of the rows, every bag writes one row and reads one offset
    my_bytes = sum(sharding.shard_cost(B * P, B, D, 1.0) * (sh.bag_hi - sh.bag_lo) / B if sh.sliced else
                   sharding.shard_cost(B * P, B, D, sh.rows / rows[sh.table]) for sh in st.local)

    def sync_all():
        torch.cuda.synchronize()
        dist.barrier()
        torch.cuda.synchronize()

    def step(k):
        plans[k % N_BATCHES].launch(stream)
        st.finish(stream, k % 2)

    for k in range(max(args.warmup, 3)):
        step(k)
    sync_all()
    if lib.emb_check_indices():
        raise RuntimeError("out-of-range index in the synthetic batch")

    K = args.steps
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches0 = emblib.launch_count()
    with ClockSampler(local) as clk:
        sync_all()
        e0.record(stream)
        for k in range(K):
            ev[k][0].record(stream)
            plans[k % N_BATCHES].launch(stream)
            ev[k][1].record(stream)
            st.finish(stream, k % 2)
        e1.record(stream)
        sync_all()
    gpu_launches = emblib.launch_count() - launches0
    total_ms = e0.elapsed_time(e1)
    kernel_ms = float(np.mean([a.elapsed_time(b) for a, b in ev]))
    t = torch.tensor([total_ms, kernel_ms, my_bytes / (kernel_ms * 1e-3) / 1e9], dtype=torch.float64, device="cuda")
    tmax, tsum = t.clone(), t.clone()
    dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
    total_ms_max, kernel_ms_max, agg_gbps = float(tmax[0]), float(tmax[1]), float(tsum[2])
    value = lookups_global * K / (total_ms_max * 1e-3)
    peak, peak_src = load_peaks()
    achieved = my_bytes / (kernel_ms * 1e-3) / 1e9  # rank 0's pooling launch

    # ---- e2e: per step, this rank's index/offset arrays come from page-locked host memory and its
    # [B/N, T, D] result goes back to page-locked host memory, all inside the timed region
    e2e = None
    if not args.no_e2e:
        pin = [{t: (torch.from_numpy(hb[t][0].view(np.int32)).pin_memory(),
                    torch.from_numpy(hb[t][1].view(np.int32)).pin_memory()) for t in need} for hb in host_batches[:2]]
        res_host = torch.empty(tuple(st.result(0).shape), dtype=torch.float32).pin_memory()
        h2d = sum(a.numel() * 4 + b.numel() * 4 for a, b in pin[0].values())
        d2h = res_host.numel() * 4

        def e2e_step(k):
            di, do = dev[k % 2]
            for tt in need:
                di[tt].copy_(pin[k % 2][tt][0], non_blocking=True)
                do[tt].copy_(pin[k % 2][tt][1], non_blocking=True)
            plans[k % 2].launch(stream)
            st.finish(stream, k % 2)
            res_host.copy_(st.result(k % 2), non_blocking=True)
            stream.synchronize()
        for k in range(3):
            e2e_step(k)
        sync_all()
        Ke = max(3, min(K, 50))
        t0 = time.perf_counter()
        for k in range(Ke):
            e2e_step(k)
        dt = time.perf_counter() - t0
        te = torch.tensor([dt], dtype=torch.float64, device="cuda")
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
        dt = float(te[0])
        e2e = {"value": lookups_global * Ke / dt, "unit": "lookups/s", "h2d_bytes_per_step": int(h2d),
               "d2h_bytes_per_step": int(d2h), "ms_per_step": dt / Ke * 1e3, "steps": Ke,
               "call": "ShardedTables.lookup per rank: pinned host indices/offsets -> HBM, pooling launch, peer-slab "
                       "exchange + barrier, partial reduction, [B/N, T, D] rows -> pinned host; bytes are rank 0's"}

    if rank == 0:
        split = {str(tt): k for tt, (_, k) in st.split.items()}
        sliced = {}
        for sh in st.plan.shards:
            if sh.sliced:
                sliced[str(sh.table)] = sliced.get(str(sh.table), 0) + 1
        out = {
            "metric": "embedding lookups/sec", "value": value, "unit": "lookups/s", "n_gpus": world, "steps": K,
            "warmup": max(args.warmup, 3), "ms_per_step": total_ms_max / K, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "int32", "data": "synthetic",
            "config": {"workload": w["name"], "tables": T, "dim": D, "global_batch": B,
                       "lookups_per_step_global": int(lookups_global),
                       "table_bytes_total": int(sum(rows) * D * 4),
                       "l2": f"inputs larger than L2: {sum(rows) * D * 4 / 1e9:.1f} GB of tables over {world} GPU(s), "
                             f"{N_BATCHES} rotating index batches",
                       "exchange": f"fused: peer-slab stores from the pooling kernel ({args.layout}-major slabs) + one completion "
                                   "kernel (flag barrier over NVLink + partial reduction)",
                       "placement": {"shards_on_rank0": len(st.local), "row_split_tables": split, "bag_sliced_replicas": sliced,
                                     "planner_imbalance": round(st.plan.imbalance(), 4),
                                     "hbm_bytes_per_rank": [int(x) for x in st.plan.hbm_bytes]},
                       "tuning": {"variant": args.variant, "unroll": args.unroll, "mult": args.mult},
                       "parallelism": f"table-wise + row-wise over {world} rank(s), pooled rows to the sample's rank"},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": None, "peak_source": peak_src, "kernel_ms": kernel_ms,
                         "kernel_ms_max_over_ranks": kernel_ms_max, "algorithmic_bytes_per_launch": my_bytes,
                         "aggregate_GBps_all_ranks": agg_gbps, "note": "rank 0's pooling launch; bytes = indices + "
                         "in-range rows + outputs + offsets of its shards (sharding.shard_cost)"},
            "clocks": clk.summary(), "gpu_launches": int(gpu_launches),
        }
        if e2e:
            out["e2e"] = e2e
        print(json.dumps(out))
    sync_all()
    for p in plans:
        p.close()
    st.close()
    emblib.shutdown()
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
    elif w.get("sharded"):
        run_sharded(args, w)
    else:
        run_b200(args, w)
    sys.stdout.flush()


if __name__ == "__main__":
    main()
