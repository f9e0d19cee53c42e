"""The EmbeddingBag kwarg seam of the reference's forked PyTorch, on stock torch.

The fork threads seven extra arguments through nn.EmbeddingBag.forward ->
F.embedding_bag -> at::native::embedding_bag and, when `use_dpu` is set, casts three of them
to pointer arrays and calls the C `lookup()` (PIM-Pytorch/torch/nn/modules/sparse.py:355,
393-397; torch/nn/functional.py:2058-2077,2242-2245; aten/src/ATen/native/EmbeddingBag.cpp:
738-778).  `install()` gives stock `torch.nn.EmbeddingBag` the same signature and routes
`use_dpu=True` calls to emblib.so's `lookup()`, so PIM-dlrm-new/dlrm_dpu_pytorch.py:559-570
runs without the fork.  Anything else falls through to torch's own forward.

As in the fork, the results land in the caller's preallocated buffers (`final_results_ptr`
is the address of a c_uint64[T] array of float* -- dlrm_dpu_pytorch.py:484-498) and the
return value is a dummy 1-element tensor (EmbeddingBag.cpp:754-777).
"""
from __future__ import annotations

import ctypes as C

from . import emblib

_u32pp = C.POINTER(C.POINTER(C.c_uint32))
_f32pp = C.POINTER(C.POINTER(C.c_float))
_installed = False


def dpu_lookup(final_results_ptr: int, indices_ptr: int, offsets_ptr: int, dpu_set_ptr: int = 0,
               latency_print: int = 0) -> None:
    """EmbeddingBag.cpp:738-751: reinterpret the three integers as uint32** / uint32** / float**
    and call lookup()."""
    if not (final_results_ptr and indices_ptr and offsets_ptr):
        raise ValueError("use_dpu=True needs final_results_ptr, indices_ptr and offsets_ptr")
    emblib.load().lookup(C.cast(C.c_void_p(indices_ptr), _u32pp), C.cast(C.c_void_p(offsets_ptr), _u32pp),
                         C.cast(C.c_void_p(final_results_ptr), _f32pp), C.c_void_p(dpu_set_ptr or None),
                         int(latency_print))


class DeviceBags:
    """Device-resident hand-off to the MLPs: indices/offsets are torch CUDA tensors,
    the pooled rows come back as ONE CUDA tensor [B, T, D] -- the layout interact_features
    concatenates into (dlrm_dpu_pytorch.py:722-755) -- written by a single launch on torch's current
    stream.  Nothing crosses PCIe; the call is asynchronous like any torch CUDA op.

    indices: int32/uint32 CUDA tensors (int64 tensors are narrowed with .to(torch.int32));
    offsets: bag start positions as in nn.EmbeddingBag (offsets[0] == 0)."""

    def __init__(self, n_tables: int, dim: int):
        self.T, self.D = n_tables, dim

    def __call__(self, lS_i, lS_o, out=None):
        import torch
        T, D = self.T, self.D
        B = int(lS_o[0].numel())
        if out is None:
            out = torch.empty((B, T, D), dtype=torch.float32, device=lS_i[0].device)
        qs, keep = [], []
        for t in range(T):
            i = lS_i[t] if lS_i[t].dtype == torch.int32 else lS_i[t].to(torch.int32)
            o = lS_o[t] if lS_o[t].dtype == torch.int32 else lS_o[t].to(torch.int32)
            keep.append((i, o))
            qs.append(dict(table_id=t, indices=i, offsets=o, out=out[:, t, :], out_stride=T * D))
        emblib.lookup_device(qs, torch.cuda.current_stream())
        for i, o in keep:  # the caching allocator must not hand these out before the launch has run
            i.record_stream(torch.cuda.current_stream())
            o.record_stream(torch.cuda.current_stream())
        return out


def install():
    """Patch torch.nn.EmbeddingBag.forward (idempotent).  Returns the original forward."""
    global _installed
    import torch

    cls = torch.nn.EmbeddingBag
    if _installed:
        return cls._pim_original_forward
    original = cls.forward

    def forward(self, input, offsets=None, per_sample_weights=None, num_of_tables: int = 0, dpu_set_ptr: int = 0,
                use_dpu: bool = False, final_results_ptr: int = 0, indices_ptr: int = 0, offsets_ptr: int = 0,
                latency_print: int = 0):
        if use_dpu:
            if num_of_tables:
                s = emblib.shape()
                if int(num_of_tables) != s.nr_tables:
                    raise ValueError(f"num_of_tables={num_of_tables} but the library is configured for "
                                     f"NR_TABLES={s.nr_tables}")
            dpu_lookup(final_results_ptr, indices_ptr, offsets_ptr, dpu_set_ptr, latency_print)
            return torch.empty(1)
        return original(self, input, offsets, per_sample_weights)

    cls._pim_original_forward = original
    cls.forward = forward
    _installed = True
    return original


def uninstall():
    global _installed
    import torch
    cls = torch.nn.EmbeddingBag
    if _installed:
        cls.forward = cls._pim_original_forward
        del cls._pim_original_forward
        _installed = False
