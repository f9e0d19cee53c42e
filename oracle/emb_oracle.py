"""ctypes/NumPy face of oracle/emb_oracle.c -- the CPU checker.

TEST INFRASTRUCTURE ONLY (see emb_oracle.c's header): imported by tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs, never by the
product package.  Builds libemb_oracle.so on first use (gcc, a second or two).

Every function restates one piece of the reference; citations are in emb_oracle.c.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libemb_oracle.so")
_lib = None

_u32p = C.POINTER(C.c_uint32)
_i32p = C.POINTER(C.c_int32)
_f32p = C.POINTER(C.c_float)


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "emb_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.run(["make", "-s", "-C", _HERE, "oracle"], check=True)
    return _SO


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        _lib = C.CDLL(build())
        _lib.emb_oracle_quantise_one.argtypes = [C.c_float]
        _lib.emb_oracle_quantise_one.restype = C.c_int32
        _lib.emb_oracle_dequant.argtypes = [C.c_int32]
        _lib.emb_oracle_dequant.restype = C.c_float
        _lib.emb_oracle_num_threads.restype = C.c_int
    return _lib


def _ptr(a: np.ndarray, ty):
    return a.ctypes.data_as(ty)


def num_threads() -> int:
    return lib().emb_oracle_num_threads()


def set_num_threads(n: int) -> None:
    lib().emb_oracle_set_num_threads(C.c_int(n))


def quantise_rows(w: np.ndarray) -> np.ndarray:
    """fp32 [R, D] -> int32 [R, D] (dlrm_dpu_pytorch.py:318 + c_int32 wrap :329)."""
    w = np.ascontiguousarray(w, dtype=np.float32)
    q = np.empty(w.shape, dtype=np.int32)
    lib().emb_oracle_quantise_rows(_ptr(w, _f32p), C.c_uint64(w.size), _ptr(q, _i32p))
    return q


def quantise_columns(w: np.ndarray) -> np.ndarray:
    """fp32 [R, D] -> int32 [D, R]: the populate_mram() payloads (dlrm_dpu_pytorch.py:318-320)."""
    w = np.ascontiguousarray(w, dtype=np.float32)
    r, d = w.shape
    q = np.empty((d, r), dtype=np.int32)
    lib().emb_oracle_quantise_columns(_ptr(w, _f32p), C.c_uint64(r), C.c_uint32(d), _ptr(q, _i32p))
    return q


def dequant(s: np.ndarray) -> np.ndarray:
    """int32 -> fp32 exactly as emb_host.h:210 ((float)x / pow(10,9))."""
    s = np.asarray(s, dtype=np.int32)
    return (s.astype(np.float32).astype(np.float64) / np.float64(10.0) ** 9).astype(np.float32)


def pool_column(column, indices, offsets, indices_len=None, nr_batches=None) -> np.ndarray:
    """One DPU: emb_dpu_lookup.c:106-116."""
    column = np.ascontiguousarray(column, dtype=np.int32)
    indices = np.ascontiguousarray(indices, dtype=np.uint32)
    offsets = np.ascontiguousarray(offsets, dtype=np.uint32)
    il = len(indices) if indices_len is None else int(indices_len)
    nb = len(offsets) if nr_batches is None else int(nr_batches)
    res = np.empty(nb, dtype=np.int32)
    lib().emb_oracle_pool_column(_ptr(column, _i32p), _ptr(indices, _u32p), _ptr(offsets, _u32p),
                                 C.c_uint32(il), C.c_uint32(nb), _ptr(res, _i32p))
    return res


def _ptr_array(arrs, ty):
    return (ty * len(arrs))(*[_ptr(a, ty) for a in arrs])


def lookup_rows(tables, indices, offsets, indices_len=None, nr_batches=None, want_sums=False):
    """Whole lookup() call over row-major int32 tables [R_t, D] (intended semantics,
    SURVEY.md Appendix A).  Returns list of fp32 [B_t, D] (and the raw int32 sums)."""
    tables = [np.ascontiguousarray(t, dtype=np.int32) for t in tables]
    indices = [np.ascontiguousarray(i, dtype=np.uint32) for i in indices]
    offsets = [np.ascontiguousarray(o, dtype=np.uint32) for o in offsets]
    T = len(tables)
    D = tables[0].shape[1]
    il = np.asarray([len(i) for i in indices] if indices_len is None else indices_len, dtype=np.uint32)
    nb = np.asarray([len(o) for o in offsets] if nr_batches is None else nr_batches, dtype=np.uint32)
    outs = [np.empty((int(nb[t]), D), dtype=np.float32) for t in range(T)]
    sums = [np.empty((int(nb[t]), D), dtype=np.int32) for t in range(T)] if want_sums else None
    lib().emb_oracle_lookup_rows(
        C.c_uint32(T), C.c_uint32(D), _ptr_array(tables, _i32p), _ptr_array(indices, _u32p),
        _ptr_array(offsets, _u32p), _ptr(il, _u32p), _ptr(nb, _u32p), _ptr_array(outs, _f32p),
        _ptr_array(sums, _i32p) if want_sums else None)
    return (outs, sums) if want_sums else outs


def lookup_columns(cols, indices, offsets, indices_len=None, nr_batches=None):
    """Whole lookup() call over column-major tables cols[t] = int32 [D, R_t] (the DPU layout)."""
    cols = [np.ascontiguousarray(c, dtype=np.int32) for c in cols]
    indices = [np.ascontiguousarray(i, dtype=np.uint32) for i in indices]
    offsets = [np.ascontiguousarray(o, dtype=np.uint32) for o in offsets]
    T = len(cols)
    D = cols[0].shape[0]
    rows = np.asarray([c.shape[1] for c in cols], dtype=np.uint64)
    il = np.asarray([len(i) for i in indices] if indices_len is None else indices_len, dtype=np.uint32)
    nb = np.asarray([len(o) for o in offsets] if nr_batches is None else nr_batches, dtype=np.uint32)
    outs = [np.empty((int(nb[t]), D), dtype=np.float32) for t in range(T)]
    lib().emb_oracle_lookup_columns(
        C.c_uint32(T), C.c_uint32(D), _ptr(rows, C.POINTER(C.c_uint64)), _ptr_array(cols, _i32p),
        _ptr_array(indices, _u32p), _ptr_array(offsets, _u32p), _ptr(il, _u32p), _ptr(nb, _u32p),
        _ptr_array(outs, _f32p))
    return outs


def bag_sum_f32(table, indices, offsets, indices_len=None, nr_batches=None) -> np.ndarray:
    """Float path: nn.EmbeddingBag(mode='sum') restated (sequential fp32 adds in index order)."""
    table = np.ascontiguousarray(table, dtype=np.float32)
    indices = np.ascontiguousarray(indices, dtype=np.uint32)
    offsets = np.ascontiguousarray(offsets, dtype=np.uint32)
    il = len(indices) if indices_len is None else int(indices_len)
    nb = len(offsets) if nr_batches is None else int(nr_batches)
    out = np.empty((nb, table.shape[1]), dtype=np.float32)
    lib().emb_oracle_bag_sum_f32(_ptr(table, _f32p), C.c_uint32(table.shape[1]), _ptr(indices, _u32p),
                                 _ptr(offsets, _u32p), C.c_uint32(il), C.c_uint32(nb), _ptr(out, _f32p))
    return out
