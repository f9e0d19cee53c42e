"""Run the reference's OWN host library + DPU program (compiled verbatim from
/root/reference by oracle/Makefile `ref`, against the UPMEM-SDK stand-in in upmem_sim/).

TEST INFRASTRUCTURE ONLY.  Used by oracle/make_golden.py (in the build container, where
/root/reference exists) to produce tests/golden/, and by tests that find a prebuilt
oracle/_ref/<shape>/emblib.so.  Shapes are compile-time macros in the reference
(upmem/Makefile:69-81) and its state is process-global (emb_host.h:30-33), so every
(T, D, B, P) gets its own build directory and every run its own child process.

Domain where the reference is well defined: B even and <= 512 (Q3),
T*D <= 256 (Q13), and T == 1 or D >= 64 (Q4).  The DPU program's result flush (Q3) is
racy on hardware; the simulator runs tasklets one after another with the tasklet that
owns the last bag LAST, which is the schedule under which the program produces its
intended result.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
import sys
import tempfile

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
NR_TASKLETS = 14


def shape_dir(T, D, B, P):
    return os.path.join(_HERE, "_ref", f"T{T}_D{D}_B{B}_P{P}")


def in_domain(T, D, B):
    return B % 2 == 0 and B <= 512 and T * D <= 256 and (T == 1 or D >= 64)


def build(T, D, B, P, reference="/root/reference"):
    """Compile the reference for one shape.  Needs /root/reference (build container only)."""
    d = shape_dir(T, D, B, P)
    if os.path.exists(os.path.join(d, "emblib.so")) and os.path.exists(os.path.join(d, "dpu_prog.so")):
        return d
    if not os.path.isdir(reference):
        raise FileNotFoundError(f"{reference} not present and {d} not prebuilt")
    subprocess.run(["make", "-s", "-C", _HERE, "ref", f"REFERENCE={reference}", f"NR_TABLES={T}",
                    f"NR_COLS={D}", f"MAX_NR_BATCHES={B}", f"MAX_INDICES_PER_BATCH={P}",
                    f"NR_TASKLETS={NR_TASKLETS}"], check=True)
    return d


def _child(npz_in, npz_out):
    z = np.load(npz_in)
    T, D, B, P = (int(z[k]) for k in ("T", "D", "B", "P"))
    d = shape_dir(T, D, B, P)
    lib = C.CDLL(os.path.join(d, "emblib.so"))
    # exact argtypes of the reference's Python caller (dlrm_dpu_pytorch.py:308-309)
    lib.populate_mram.argtypes = [C.c_uint32, C.c_uint64, C.c_uint32, C.POINTER(C.c_int32), C.c_void_p]
    lib.populate_mram.restype = C.c_void_p
    lib.lookup.argtypes = [C.POINTER(C.POINTER(C.c_uint32)), C.POINTER(C.POINTER(C.c_uint32)),
                           C.POINTER(C.POINTER(C.c_float)), C.c_void_p, C.c_int64]
    lib.lookup.restype = C.c_void_p
    owner = (B - 1) % NR_TASKLETS
    order = [t for t in range(NR_TASKLETS) if t != owner] + [owner]
    lib.upmem_sim_set_tasklet_order((C.c_uint * NR_TASKLETS)(*order), C.c_uint(NR_TASKLETS))
    handle = None
    keep = []
    for t in range(T):
        cols = np.ascontiguousarray(z[f"cols{t}"], dtype=np.int32)  # [D, R]
        keep.append(cols)
        for j in range(D):
            handle = lib.populate_mram(t, cols.shape[1], j,
                                       cols[j].ctypes.data_as(C.POINTER(C.c_int32)), None)
    n_calls = int(z["n_calls"])
    outs = {}
    for c in range(n_calls):
        idx = [np.ascontiguousarray(z[f"idx{c}_{t}"], dtype=np.uint32) for t in range(T)]
        off = [np.ascontiguousarray(z[f"off{c}_{t}"], dtype=np.uint32) for t in range(T)]
        for t in range(T):
            assert idx[t].size >= B * P and off[t].size >= B
        res = [np.full((B, D), np.nan, dtype=np.float32) for _ in range(T)]
        ip = (C.POINTER(C.c_uint32) * T)(*[a.ctypes.data_as(C.POINTER(C.c_uint32)) for a in idx])
        op = (C.POINTER(C.c_uint32) * T)(*[a.ctypes.data_as(C.POINTER(C.c_uint32)) for a in off])
        rp = (C.POINTER(C.c_float) * T)(*[a.ctypes.data_as(C.POINTER(C.c_float)) for a in res])
        lib.lookup(ip, op, rp, handle, 0)
        for t in range(T):
            outs[f"out{c}_{t}"] = res[t]
    np.savez(npz_out, **outs)


def run(T, D, B, P, cols, calls):
    """cols: list of T int32 [D, R_t]; calls: list of (indices[T], offsets[T]) with
    indices[t] uint32 [B*P], offsets[t] uint32 [B].  Returns list (per call) of list (per
    table) of fp32 [B, D] -- what the reference's lookup() wrote into final_results."""
    build(T, D, B, P)
    with tempfile.TemporaryDirectory() as tmp:
        fin, fout = os.path.join(tmp, "in.npz"), os.path.join(tmp, "out.npz")
        payload = dict(T=T, D=D, B=B, P=P, n_calls=len(calls))
        for t in range(T):
            payload[f"cols{t}"] = np.asarray(cols[t], dtype=np.int32)
        for c, (idx, off) in enumerate(calls):
            for t in range(T):
                payload[f"idx{c}_{t}"] = np.asarray(idx[t], dtype=np.uint32)
                payload[f"off{c}_{t}"] = np.asarray(off[t], dtype=np.uint32)
        np.savez(fin, **payload)
        subprocess.run([sys.executable, os.path.abspath(__file__), "--child", fin, fout], check=True)
        z = np.load(fout)
        return [[z[f"out{c}_{t}"] for t in range(T)] for c in range(len(calls))]


if __name__ == "__main__":
    if len(sys.argv) == 4 and sys.argv[1] == "--child":
        _child(sys.argv[2], sys.argv[3])
    else:
        sys.exit("usage: ref_sim.py --child in.npz out.npz")
