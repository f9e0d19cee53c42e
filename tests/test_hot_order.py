"""emb_hot_order (include/emb_b200.h): the host-side helper that turns lookup counts into the permutation
emb_table_remap() takes.  Pure host code, so it is checked here without a device."""
import numpy as np
import pytest


@pytest.fixture(scope="module")
def emblib():
    import __graft_entry__ as ge
    from pim_embedding_lookup_b200 import emblib as m
    import os
    if not os.path.exists(m.LIB_PATH):
        ge.build()
    m.load()
    return m


def test_hot_order_sorts_by_count_and_keeps_ties_in_row_order(emblib):
    counts = np.array([3, 9, 0, 9, 1, 3, 0, 7], dtype=np.uint64)
    m = emblib.hot_order(counts)
    # slab order: rows 1, 3 (9 lookups, in row order), 7, then 0, 5 (3 lookups), 4, then 2, 6
    assert list(np.argsort(m)) == [1, 3, 7, 0, 5, 4, 2, 6]
    assert sorted(m.tolist()) == list(range(8))


def test_hot_order_matches_a_stable_argsort_on_a_zipf_sample(emblib):
    rng = np.random.default_rng(1)
    R = 200_000
    sample = (rng.zipf(1.05, size=500_000) - 1) % R
    counts = np.bincount(rng.permutation(R)[sample], minlength=R).astype(np.uint64)
    m = emblib.hot_order(counts)
    order = np.argsort(-counts.astype(np.int64), kind="stable")
    want = np.empty(R, dtype=np.uint32)
    want[order] = np.arange(R, dtype=np.uint32)
    assert np.array_equal(m, want)
    hot = np.argsort(m)[:1000]  # the 1000 rows the prefix holds are the 1000 most frequent
    assert counts[hot].min() >= np.sort(counts)[-1000]


def test_hot_order_rejects_bad_arguments(emblib):
    lib = emblib.load()
    assert lib.emb_hot_order(None, 4, None) != 0
    out = np.empty(1, dtype=np.uint32)
    c = np.zeros(1, dtype=np.uint64)
    import ctypes as C
    assert lib.emb_hot_order(c.ctypes.data_as(C.POINTER(C.c_uint64)), 0, emblib._p(out, emblib.u32p)) != 0
