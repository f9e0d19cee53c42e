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
