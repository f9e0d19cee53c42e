"""ctypes bridge to emblib.so -- the Python side of the drop-in boundary.

Mirrors how the reference binds its C library (PIM-dlrm-new/dlrm_dpu_pytorch.py:117-125,
308-309; upmem/c_test.py:15-16): `CDLL("emblib.so")`, then `populate_mram` / `lookup` with
the same argtypes.  On top of the two reference entry points it declares the extension
entry points of include/emb_b200.h.

There is no fallback: if emblib.so is missing or no CUDA device is visible every call
raises (load()) or the C side prints and exits, as the reference's DPU_ASSERT does.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

from .dputypes import (EMB_MAX_PEERS, EMB_Q_RAW_SUMS, EMB_TABLE_FIXED, EMB_TABLE_FLOAT, DpuRuntimeTotals,
                       EmbDevQuery, EmbIpcHandle, EmbReduceJob, EmbShape, EmbStaticQuery, EmbTimings)

_PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_PKG, "lib", "emblib.so")
_lib = None

u32p = C.POINTER(C.c_uint32)
i32p = C.POINTER(C.c_int32)
f32p = C.POINTER(C.c_float)


class EmbError(RuntimeError):
    pass


def build(verbose: bool = False) -> str:
    """Compile emblib.so (+ the emb_host microbenchmark) for sm_100a with nvcc."""
    r = subprocess.run(["make", "-C", os.path.join(_PKG, "csrc"), "all"], capture_output=True, text=True)
    if verbose or r.returncode:
        print(r.stdout[-4000:], r.stderr[-4000:])
    if r.returncode:
        raise EmbError("building emblib.so failed")
    return LIB_PATH


def _declare(lib):
    # --- the reference's two entry points, argtypes exactly as its callers set them ---
    lib.populate_mram.argtypes = [C.c_uint32, C.c_uint64, C.c_uint32, i32p, C.POINTER(DpuRuntimeTotals)]
    lib.populate_mram.restype = C.c_void_p
    lib.lookup.argtypes = [C.POINTER(u32p), C.POINTER(u32p), C.POINTER(f32p), C.c_void_p, C.c_int64]
    lib.lookup.restype = C.c_void_p
    # --- extensions ---
    sig = {
        "emb_configure": ([C.c_uint32] * 4 + [C.c_int], C.c_int),
        "emb_configure_devices": ([C.c_uint32] * 4 + [C.POINTER(C.c_int), C.c_uint32, u32p], C.c_int),
        "emb_group_size": ([], C.c_int),
        "emb_table_device": ([C.c_uint32], C.c_int),
        "emb_get_shape": ([C.POINTER(EmbShape)], C.c_int),
        "emb_device": ([], C.c_int),
        "emb_shutdown": ([], None),
        "emb_last_error": ([], C.c_char_p),
        "emb_populate_table_i32": ([C.c_uint32, C.c_uint64, i32p], C.c_int),
        "emb_populate_table_f32": ([C.c_uint32, C.c_uint64, f32p, C.c_int], C.c_int),
        "emb_populate_table_device": ([C.c_uint32, C.c_uint64, C.c_void_p, C.c_int, C.c_int, C.c_void_p], C.c_int),
        "emb_table_set_row_base": ([C.c_uint32, C.c_uint64], C.c_int),
        "emb_table_read": ([C.c_uint32, C.c_uint64, C.c_uint64, C.c_void_p], C.c_int),
        "emb_table_info": ([C.c_uint32, C.POINTER(C.c_uint64), C.POINTER(C.c_int), C.POINTER(C.c_void_p)], C.c_int),
        "emb_free_table": ([C.c_uint32], C.c_int),
        "emb_l2_persist": ([C.c_uint32, C.c_uint64, C.c_uint64, C.c_float, C.c_void_p], C.c_int),
        "emb_table_set_hot_rows": ([C.c_uint32, C.c_uint64, C.c_int], C.c_int),
        "emb_table_remap": ([C.c_uint32, u32p], C.c_int),
        "emb_hot_order": ([C.POINTER(C.c_uint64), C.c_uint64, u32p], C.c_int),
        "emb_remap_indices": ([C.c_uint32, C.c_void_p, C.c_uint64, C.c_void_p], C.c_int),
        "emb_lookup_ex": ([C.c_uint32, C.POINTER(u32p), C.POINTER(u32p), u32p, u32p, C.POINTER(f32p)], C.c_int),
        "emb_lookup_ex_async": ([C.c_uint32, C.POINTER(u32p), C.POINTER(u32p), u32p, u32p, C.POINTER(f32p),
                                 C.POINTER(C.c_uint64)], C.c_int),
        "emb_lookup_wait": ([C.c_uint64], C.c_int),
        "emb_plan_create": ([C.POINTER(EmbDevQuery), C.c_uint32, C.POINTER(C.c_void_p)], C.c_int),
        "emb_plan_launch": ([C.c_void_p, C.c_void_p], C.c_int),
        "emb_plan_destroy": ([C.c_void_p], None),
        "emb_plan_bytes": ([C.c_void_p], C.c_uint64),
        "emb_plan_grid": ([C.c_void_p], C.c_uint32),
        "emb_lookup_device": ([C.POINTER(EmbDevQuery), C.c_uint32, C.c_void_p], C.c_int),
        "emb_check_indices": ([], C.c_int),
        "emb_lookup_static": ([C.POINTER(EmbStaticQuery), C.c_uint32, C.c_uint32, C.c_void_p, C.POINTER(C.c_int)],
                              C.c_int),
        "emb_reduce_partials": ([C.POINTER(C.c_void_p), C.c_uint32, C.c_uint32, C.c_void_p, C.c_uint32,
                                 C.c_uint64, C.c_int, C.c_void_p], C.c_int),
        "emb_finish_create": ([C.POINTER(C.c_void_p), C.c_uint32, C.c_uint32, C.POINTER(EmbReduceJob), C.c_uint32,
                               C.POINTER(C.c_void_p)], C.c_int),
        "emb_finish_launch": ([C.c_void_p, C.c_uint32, C.c_void_p], C.c_int),
        "emb_finish_destroy": ([C.c_void_p], None),
        "emb_device_alloc": ([C.POINTER(C.c_void_p), C.c_size_t], C.c_int),
        "emb_device_free": ([C.c_void_p], C.c_int),
        "emb_host_alloc": ([C.POINTER(C.c_void_p), C.c_size_t], C.c_int),
        "emb_host_free": ([C.c_void_p], C.c_int),
        "emb_memcpy_h2d": ([C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p], C.c_int),
        "emb_memcpy_d2h": ([C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p], C.c_int),
        "emb_stream_sync": ([C.c_void_p], C.c_int),
        "emb_ipc_export": ([C.c_void_p, C.POINTER(EmbIpcHandle)], C.c_int),
        "emb_ipc_open": ([C.POINTER(EmbIpcHandle), C.POINTER(C.c_void_p)], C.c_int),
        "emb_ipc_close": ([C.c_void_p], C.c_int),
        "emb_last_timings": ([C.POINTER(EmbTimings)], C.c_int),
        "emb_launch_count": ([], C.c_uint64),
        "emb_set_unroll": ([C.c_int], C.c_int),
        "emb_set_tuning": ([C.c_int, C.c_int, C.c_int], C.c_int),
        "emb_set_host_path": ([C.c_int], C.c_int),
        "emb_host_register": ([C.c_void_p, C.c_size_t], C.c_int),
        "emb_host_unregister": ([C.c_void_p], C.c_int),
    }
    for name, (args, res) in sig.items():
        fn = getattr(lib, name)
        fn.argtypes = args
        fn.restype = res
    return lib


EXPORTS = ("populate_mram", "lookup")


def load(path: str | None = None) -> C.CDLL:
    """CDLL the library (the reference does `CDLL("./emblib.so")` from its build directory)."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    p = path or os.environ.get("EMBLIB_SO", LIB_PATH)
    if not os.path.exists(p):
        raise EmbError(f"{p} not found: run __graft_entry__.build() / make -C pim-embedding-lookup_b200/csrc "
                       "(there is no CPU fallback)")
    lib = _declare(C.CDLL(os.path.abspath(p)))
    if path is None:
        _lib = lib
    return lib


def check(rc: int, what: str = "emblib") -> None:
    if rc != 0:
        msg = load().emb_last_error()
        raise EmbError(f"{what} failed ({rc}): {msg.decode(errors='replace') if msg else ''}")


def _p(a: np.ndarray, ty):
    return a.ctypes.data_as(ty)


def _ptr_array(arrs, ty):
    return (ty * len(arrs))(*[_p(a, ty) for a in arrs])


# ---------------------------------------------------------------------------------------------
# Convenience layer (NumPy in, NumPy out).  Thin: every method is one C call.
# ---------------------------------------------------------------------------------------------

def configure(nr_tables, nr_cols, max_nr_batches=0, max_indices_per_batch=0, device=-1):
    check(load().emb_configure(nr_tables, nr_cols, max_nr_batches, max_indices_per_batch, device), "emb_configure")


def configure_devices(nr_tables, nr_cols, devices, max_nr_batches=0, max_indices_per_batch=0, table_device=None):
    """Device group (emb_configure_devices): whole tables over several GPUs of the box behind lookup()."""
    devs = (C.c_int * len(devices))(*devices)
    td = None
    if table_device is not None:
        keep = np.ascontiguousarray(table_device, dtype=np.uint32)
        assert keep.size == nr_tables
        td = _p(keep, u32p)
    check(load().emb_configure_devices(nr_tables, nr_cols, max_nr_batches, max_indices_per_batch, devs, len(devices), td),
          "emb_configure_devices")


def shutdown():
    load().emb_shutdown()


def shape() -> EmbShape:
    s = EmbShape()
    check(load().emb_get_shape(C.byref(s)), "emb_get_shape")
    return s


def populate_mram(table_id: int, column: np.ndarray, col: int, runtime: DpuRuntimeTotals | None = None) -> int:
    """One (table, column) payload, as dlrm_dpu_pytorch.py:328-332 sends it."""
    column = np.ascontiguousarray(column, dtype=np.int32)
    rt = C.byref(runtime) if runtime is not None else None
    return load().populate_mram(table_id, column.shape[0], col, _p(column, i32p), rt)


def export_emb(tables_f32) -> int:
    """The reference's export_emb (dlrm_dpu_pytorch.py:306-356) for a list of fp32 [R, D]
    tables, minus the per-element Python: quantise + layout run on the GPU."""
    lib = load()
    for k, w in enumerate(tables_f32):
        w = np.ascontiguousarray(w, dtype=np.float32)
        check(lib.emb_populate_table_f32(k, w.shape[0], _p(w, f32p), EMB_TABLE_FIXED), "emb_populate_table_f32")
    return len(tables_f32)


def populate_table_i32(table_id: int, rows_i32: np.ndarray):
    a = np.ascontiguousarray(rows_i32, dtype=np.int32)
    check(load().emb_populate_table_i32(table_id, a.shape[0], _p(a, i32p)), "emb_populate_table_i32")


def populate_table_f32(table_id: int, w: np.ndarray, kind: int = EMB_TABLE_FIXED):
    a = np.ascontiguousarray(w, dtype=np.float32)
    check(load().emb_populate_table_f32(table_id, a.shape[0], _p(a, f32p), kind), "emb_populate_table_f32")


def table_read(table_id: int, row0: int, n: int, dtype=np.int32) -> np.ndarray:
    D = shape().nr_cols
    out = np.empty((n, D), dtype=dtype)
    check(load().emb_table_read(table_id, row0, n, out.ctypes.data), "emb_table_read")
    return out


def hot_order(counts) -> np.ndarray:
    """emb_hot_order: new_of_old permutation that puts the most looked-up rows first (host only)."""
    c = np.ascontiguousarray(counts, dtype=np.uint64)
    out = np.empty(c.size, dtype=np.uint32)
    check(load().emb_hot_order(c.ctypes.data_as(C.POINTER(C.c_uint64)), c.size, _p(out, u32p)), "emb_hot_order")
    return out


def remap_table(table_id: int, new_of_old) -> None:
    """emb_table_remap: reorder the slab; callers keep their row ids."""
    m = np.ascontiguousarray(new_of_old, dtype=np.uint32)
    check(load().emb_table_remap(table_id, _p(m, u32p)), "emb_table_remap")


def set_hot_rows(table_id: int, hot_rows: int, mode: int) -> None:
    check(load().emb_table_set_hot_rows(table_id, hot_rows, mode), "emb_table_set_hot_rows")


def remap_indices(table_id: int, indices_dev, stream=None) -> None:
    """emb_remap_indices on a device tensor of row ids (int32 bits of uint32), in place."""
    check(load().emb_remap_indices(table_id, _addr(indices_dev), _numel(indices_dev), _stream(stream)),
          "emb_remap_indices")


def pinned_array(shape, dtype=np.uint32) -> np.ndarray:
    """NumPy array in page-locked host memory (emb_host_alloc): lookup()/lookup_ex() on such arrays
    take the EMB_HOST_MAPPED path (no staging copies).  Freed with free_pinned()."""
    n = int(np.prod(shape)) * np.dtype(dtype).itemsize
    p = C.c_void_p()
    check(load().emb_host_alloc(C.byref(p), max(n, 16)), "emb_host_alloc")
    buf = (C.c_ubyte * max(n, 16)).from_address(p.value)
    return np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape)


def free_pinned(a: np.ndarray) -> None:
    check(load().emb_host_free(C.c_void_p(a.ctypes.data)), "emb_host_free")


def set_host_path(path: int) -> None:
    check(load().emb_set_host_path(path), "emb_set_host_path")


def lookup(indices, offsets, handle=None, latency_print=0):
    """The reference call: shapes come from the configured NR_TABLES/MAX_NR_BATCHES/...;
    indices[t] uint32 [B*P], offsets[t] uint32 [B].  Returns list of fp32 [B, D]."""
    s = shape()
    T, B, D = s.nr_tables, s.max_nr_batches, s.nr_cols
    idx = [np.ascontiguousarray(i, dtype=np.uint32) for i in indices]
    off = [np.ascontiguousarray(o, dtype=np.uint32) for o in offsets]
    assert len(idx) == T and len(off) == T
    for t in range(T):
        assert idx[t].size >= B * s.max_indices_per_batch and off[t].size >= B
    res = [np.empty((B, D), dtype=np.float32) for _ in range(T)]
    load().lookup(_ptr_array(idx, u32p), _ptr_array(off, u32p), _ptr_array(res, f32p), handle, latency_print)
    return res


def lookup_ex(indices, offsets, results=None):
    """Explicit lengths, variable pooling (emb_lookup_ex).  Returns list of fp32 [B_t, D]."""
    D = shape().nr_cols
    idx = [np.ascontiguousarray(i, dtype=np.uint32) for i in indices]
    off = [np.ascontiguousarray(o, dtype=np.uint32) for o in offsets]
    T = len(idx)
    il = np.asarray([i.size for i in idx], dtype=np.uint32)
    nb = np.asarray([o.size for o in off], dtype=np.uint32)
    res = results if results is not None else [np.empty((int(nb[t]), D), dtype=np.float32) for t in range(T)]
    check(load().emb_lookup_ex(T, _ptr_array(idx, u32p), _ptr_array(off, u32p), _p(il, u32p), _p(nb, u32p),
                               _ptr_array(res, f32p)), "emb_lookup_ex")
    return res


class HostCall:
    """One emb_lookup_ex() call over fixed host arrays with the pointer arrays built once -- what the
    reference's caller prepares per batch (`ind_pointers_c`, `off_pointers_c`, `final_results_ptr`,
    dlrm_dpu_pytorch.py:484-498,1028-1051).  Calling the object is exactly one C call."""

    def __init__(self, indices, offsets, results=None):
        D = shape().nr_cols
        self.indices = [np.ascontiguousarray(i, dtype=np.uint32) for i in indices]
        self.offsets = [np.ascontiguousarray(o, dtype=np.uint32) for o in offsets]
        T = len(self.indices)
        self._il = np.asarray([i.size for i in self.indices], dtype=np.uint32)
        self._nb = np.asarray([o.size for o in self.offsets], dtype=np.uint32)
        self.results = results if results is not None else [np.empty((int(self._nb[t]), D), dtype=np.float32)
                                                           for t in range(T)]
        self._args = (T, _ptr_array(self.indices, u32p), _ptr_array(self.offsets, u32p), _p(self._il, u32p),
                      _p(self._nb, u32p), _ptr_array(self.results, f32p))
        self._fn = load().emb_lookup_ex

    def __call__(self):
        rc = self._fn(*self._args)
        if rc:
            check(rc, "emb_lookup_ex")
        return self.results

    def start(self) -> int:
        """emb_lookup_ex_async: enqueue the call, return its ticket.  The arrays must stay untouched until wait()."""
        tk = C.c_uint64(0)
        rc = load().emb_lookup_ex_async(*self._args, C.byref(tk))
        if rc:
            check(rc, "emb_lookup_ex_async")
        self._ticket = tk.value
        return tk.value

    def wait(self):
        """emb_lookup_wait on the ticket of the last start(); the results are in host memory on return."""
        rc = load().emb_lookup_wait(self._ticket)
        if rc:
            check(rc, "emb_lookup_wait")
        return self.results


def last_timings() -> EmbTimings:
    t = EmbTimings()
    check(load().emb_last_timings(C.byref(t)), "emb_last_timings")
    return t


def launch_count() -> int:
    return int(load().emb_launch_count())


class Plan:
    """Device-resident lookup: descriptors uploaded once, each launch() is one kernel.

    `queries` is a list of dicts with keys table_id, indices, offsets, out (objects with
    .data_ptr(), e.g. torch CUDA tensors, or raw ints), optional indices_len, nr_batches,
    out_stride, flags, bags_per_rank, out_peer, bag_first, bag_count, bag_rotate.  The plan keeps references to the
    tensors."""

    def __init__(self, queries):
        self._keep = queries
        arr, n = _query_array(queries)
        h = C.c_void_p()
        check(load().emb_plan_create(arr, n, C.byref(h)), "emb_plan_create")
        self._h = h
        self.bytes = int(load().emb_plan_bytes(h))
        self.grid = int(load().emb_plan_grid(h))

    def launch(self, stream=None):
        rc = load().emb_plan_launch(self._h, _stream(stream))
        if rc:
            check(rc, "emb_plan_launch")

    def close(self):
        if self._h:
            load().emb_plan_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def _query_array(queries):
    n = len(queries)
    arr = (EmbDevQuery * n)()
    for i, q in enumerate(queries):
        e = arr[i]
        e.table_id = q["table_id"]
        e.indices = _addr(q["indices"])
        e.offsets = _addr(q["offsets"])
        e.out = _addr(q.get("out"))
        e.nr_batches = q["nr_batches"] if "nr_batches" in q else _numel(q["offsets"])
        e.indices_len = q["indices_len"] if "indices_len" in q else _numel(q["indices"])
        e.flags = q.get("flags", 0)
        e.out_stride = q.get("out_stride", 0)
        e.bags_per_rank = q.get("bags_per_rank", 0)
        e.bag_first = q.get("bag_first", 0)
        e.bag_count = q.get("bag_count", 0)
        e.bag_rotate = q.get("bag_rotate", 0)
        for r, p in enumerate(q.get("out_peer", ())):
            e.out_peer[r] = _addr(p)
    return arr, n


def lookup_device(queries, stream=None) -> None:
    """One-shot device-resident lookup (emb_lookup_device): same query dicts as Plan, asynchronous on
    `stream` (a torch.cuda.Stream, a raw handle, or None for the library's stream)."""
    arr, n = _query_array(queries)
    check(load().emb_lookup_device(arr, n, _stream(stream)), "emb_lookup_device")


def lookup_static(queries, nr_cols: int, stream=None, check_range: bool = False):
    """emb_lookup_static: pool caller-owned device arrays without the library allocating anything (for hosts whose
    own allocator owns the GPU, and for loops that cannot take cudaMalloc's device-wide stall).  `queries` is a list of
    dicts with keys table (device array [nr_rows][nr_cols]), nr_rows, indices, offsets, out and optional kind
    (EMB_TABLE_FIXED), flags, nr_batches, indices_len, out_stride; at most EMB_STATIC_MAX_QUERIES of them.
    Asynchronous on `stream` (None = the legacy default stream) unless check_range, which waits and returns whether
    an index fell outside its table.  Calls share one fixed scratch array per GPU: keep them ordered on the GPU."""
    n = len(queries)
    arr = (EmbStaticQuery * max(n, 1))()
    for i, q in enumerate(queries):
        e = arr[i]
        e.table = _addr(q["table"])
        e.nr_rows = q["nr_rows"]
        e.kind = q.get("kind", EMB_TABLE_FIXED)
        e.flags = q.get("flags", 0)
        e.indices = _addr(q["indices"])
        e.offsets = _addr(q["offsets"])
        e.out = _addr(q.get("out"))
        e.nr_batches = q["nr_batches"] if "nr_batches" in q else _numel(q["offsets"])
        e.indices_len = q["indices_len"] if "indices_len" in q else _numel(q["indices"])
        e.out_stride = q.get("out_stride", 0)
    s = None if stream is None else (stream if isinstance(stream, int) else int(stream.cuda_stream))
    oor = C.c_int(0)
    check(load().emb_lookup_static(arr, n, nr_cols, s, C.byref(oor) if check_range else None), "emb_lookup_static")
    return bool(oor.value) if check_range else None


def _addr(x):
    if x is None:
        return None
    if isinstance(x, int):
        return x
    return int(x.data_ptr())


def _numel(x):
    return int(x.numel()) if hasattr(x, "numel") else int(x.size)


def _stream(s):
    if s is None:
        return None
    if isinstance(s, int):
        return s
    h = int(s.cuda_stream)  # torch.cuda.Stream
    # NULL means "the library's own stream" in the C ABI; torch's default stream has handle 0,
    # so name it explicitly: cudaStreamLegacy == (cudaStream_t)0x1
    return h if h else 1
