"""ctypes mirrors of the ABI-visible structs of emblib.so.

Part 1 keeps the names, field order and helpers of the reference's bridge module
(upmem/dputypes.py:18-98) so code written against it keeps working; part 2 mirrors the
extension structs of include/emb_b200.h.
"""
import csv
from ctypes import (POINTER, Structure, addressof, c_double, c_float, c_int, c_int32, c_int64, c_long,
                    c_size_t, c_ubyte, c_uint, c_uint32, c_uint64, c_void_p, resize, sizeof)
from enum import IntEnum


class CtypesEnum(IntEnum):
    @classmethod
    def from_param(cls, obj):
        return int(obj)


class DpuRuntimeConfigEnum(CtypesEnum):  # emb_host.h:72-75
    RT_ALL = 0
    RT_LAUNCH = 1


class DpuTimespec(Structure):  # emb_host.h:54-57 (nsec first)
    _fields_ = [("tv_nsec", c_long), ("tv_sec", c_long)]

    def __init__(self, tv_sec=0, tv_nsec=0):
        super().__init__(tv_nsec, tv_sec)

    def as_ms(self, offset=0):
        return ((self.tv_sec * 1000) + (self.tv_nsec / 1e6)) - offset

    def __repr__(self):
        return f"(tv_sec={self.tv_sec}, tv_nsec={self.tv_nsec} (as_ms={self.as_ms()}))"


class DpuRuntimeInterval(Structure):  # emb_host.h:63-66
    _fields_ = [("start", DpuTimespec), ("stop", DpuTimespec)]

    def __init__(self):
        super().__init__(DpuTimespec(), DpuTimespec())

    def __repr__(self):
        return f"(start={self.start}, stop={self.stop})"


class DpuRuntimeGroup(Structure):  # emb_host.h:81-85
    _fields_ = [("in_use", c_uint), ("length", c_uint), ("intervals", POINTER(DpuRuntimeInterval))]

    def __init__(self, in_use=0, length=4, values=None):
        v = (DpuRuntimeInterval * length)() if values is None else values
        super().__init__(in_use, length, v)

    def __repr__(self):
        return f"(in_use={self.in_use}, length={self.length}, intervals={self.intervals})"


class DpuRuntimeTotals(Structure):  # emb_host.h:41-48
    _fields_ = [
        ("execution_time_prepare", c_double),
        ("execution_time_populate_copy_in", c_double),
        ("execution_time_copy_in", c_double),
        ("execution_time_copy_out", c_double),
        ("execution_time_aggregate_result", c_double),
        ("execution_time_launch", c_double),
    ]

    def __init__(self):
        super().__init__(0, 0, 0, 0, 0, 0)


def customresize(array, new_size):
    resize(array, sizeof(array._type_) * new_size)
    return (array._type_ * new_size).from_address(addressof(array))


def write_results(config, runtime_group, headers=("DPU", "Start", "Stop")):
    """`DPU,Start,Stop` interval CSV (dputypes.py:87-98; read by graph/runtime_intervals/main.py)."""
    if not config.runtimes:
        return
    with open(config.runtime_file, "a") as csvfile:
        w = csv.writer(csvfile, delimiter=",", quotechar='"', quoting=csv.QUOTE_MINIMAL)
        if headers:
            w.writerow(list(headers))
        for nr, item in enumerate(runtime_group):
            for i in range(item.in_use):
                w.writerow([nr, f"{item.intervals[i].start.as_ms():.16f}", f"{item.intervals[i].stop.as_ms():.16f}"])


# ---- part 2: include/emb_b200.h extension structs ------------------------------------------

EMB_MAX_PEERS = 8
EMB_TABLE_FIXED = 1
EMB_TABLE_FLOAT = 2
EMB_Q_RAW_SUMS = 1
EMB_Q_OVERLAP = 2  # plans: the launch reads nothing the kernel in front of it writes (include/emb_b200.h)
EMB_Q_REMAPPED = 4  # the query's indices are slab row ids of a reordered table already
EMB_HOT_KEEP, EMB_HOT_STREAM_COLD = 1, 2  # emb_table_set_hot_rows modes
EMB_HOST_AUTO, EMB_HOST_COPY, EMB_HOST_MAPPED, EMB_HOST_MAPPED_OUT, EMB_HOST_BOUNCE = 0, 1, 2, 3, 4
EMB_IPC_HANDLE_BYTES = 64


class QueryLen(Structure):  # emb_types.h:21-24
    _pack_ = 1
    _fields_ = [("indices_len", c_uint32), ("nr_batches", c_uint32)]


class EmbShape(Structure):
    _fields_ = [("nr_tables", c_uint32), ("nr_cols", c_uint32), ("max_nr_batches", c_uint32),
                ("max_indices_per_batch", c_uint32)]


class EmbDevQuery(Structure):
    _fields_ = [("table_id", c_uint32), ("nr_batches", c_uint32), ("indices_len", c_uint32), ("flags", c_uint32),
                ("indices", c_void_p), ("offsets", c_void_p), ("out", c_void_p), ("out_stride", c_uint32),
                ("bags_per_rank", c_uint32), ("out_peer", c_void_p * EMB_MAX_PEERS),
                ("bag_first", c_uint32), ("bag_count", c_uint32), ("bag_rotate", c_uint32), ("reserved_", c_uint32)]


EMB_STATIC_MAX_QUERIES = 64  # emb_lookup_static: queries per call its fixed descriptor array holds


class EmbStaticQuery(Structure):  # include/emb_b200.h emb_static_query
    _fields_ = [("table", c_void_p), ("nr_rows", c_uint64), ("kind", c_uint32), ("flags", c_uint32),
                ("nr_batches", c_uint32), ("indices_len", c_uint32), ("indices", c_void_p), ("offsets", c_void_p),
                ("out", c_void_p), ("out_stride", c_uint32), ("reserved_", c_uint32)]


class EmbReduceJob(Structure):  # include/emb_b200.h emb_reduce_job
    _fields_ = [("parts", c_void_p * EMB_MAX_PEERS), ("out", c_void_p), ("n_parts", c_uint32),
                ("part_stride", c_uint32), ("out_stride", c_uint32), ("is_raw_i32", c_uint32), ("nr_bags", c_uint64)]


class EmbTimings(Structure):
    _fields_ = [("h2d_ms", c_double), ("kernel_ms", c_double), ("d2h_ms", c_double), ("total_ms", c_double),
                ("h2d_bytes", c_uint64), ("d2h_bytes", c_uint64), ("launches", c_uint64), ("path", c_uint64)]


class EmbIpcHandle(Structure):
    _fields_ = [("bytes", c_ubyte * EMB_IPC_HANDLE_BYTES)]


__all__ = [n for n in dir() if not n.startswith("_")]
