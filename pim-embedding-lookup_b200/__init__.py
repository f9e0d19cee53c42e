"""pim-embedding-lookup_b200 -- B200-native backend for the embedding-lookup hot path of
UBC-ECE-Sasha/PIM-Embedding-Lookup (populate_mram / lookup behind emblib.so).

Layout: csrc/ (CUDA kernels + C ABI, built into lib/emblib.so), emblib.py (ctypes bridge),
dputypes.py (struct mirrors, as upmem/dputypes.py), embedding_bag.py (the EmbeddingBag
kwarg seam of the forked PyTorch), sharding.py + exchange.py + sharded.py (tables over the
GPUs of one box: placement, fused pooled-output exchange, the per-rank driver object).
"""
from . import dputypes, emblib  # noqa: F401
from .emblib import EmbError, Plan, build, load  # noqa: F401

__version__ = "0.1.0"
