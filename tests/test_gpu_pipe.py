"""Parity of the pipelined pooling kernel (variant 5, bag_sum_pipe_kernel) with the oracle, through the
C ABI: one warp per CTA walks several sub-chunks of bags, the next sub-chunk's index window is copied
asynchronously while the current one's rows are gathered.  Bit-exact like every int32 path
(emb_dpu_lookup.c:106-128 + emb_host.h:207-212).
"""
import numpy as np
import pytest

import emb_oracle as eo

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")


@pytest.fixture()
def emblib():
    from pim_embedding_lookup_b200 import emblib as m
    m.load()
    m.shutdown()
    yield m
    m.load().emb_set_tuning(0, 0, 0)
    m.shutdown()


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def ragged(rng, R, B, max_len, empty_frac=0.1):
    lens = rng.integers(1, max_len + 1, size=B)
    lens[rng.random(B) < empty_frac] = 0
    off = np.concatenate([[0], np.cumsum(lens)[:-1]]).astype(np.uint32)
    idx = rng.integers(0, R, size=int(lens.sum())).astype(np.uint32)
    return idx, off


@pytest.mark.parametrize("D,nch", [(D, n) for D in (4, 8, 16, 32) for n in (2, 4, 8)] + [(64, 2), (64, 4)])
def test_pipe_kernel_matches_oracle(emblib, D, nch):
    """Ragged bags with empty ones, bag counts that are not multiples of the CTA's bags, one table with a
    single bag.  D = 64 is outside the kernel's row sizes and must fall back to the balanced kernel."""
    rng = np.random.default_rng(1000 + D + nch)
    shapes = [(4096, 5000, 33), (100, 777, 60), (7, 1, 5), (50000, 129, 20)]
    emblib.configure(len(shapes), D)
    assert emblib.load().emb_set_tuning(5, 8, nch) == 0
    tabs, idxs, offs = [], [], []
    for t, (R, B, L) in enumerate(shapes):
        q = rng.integers(-2**31, 2**31, size=(R, D), dtype=np.int64).astype(np.int32)
        emblib.populate_table_i32(t, q)
        i, o = ragged(rng, R, B, L)
        if t == 1:
            o[0] = 3  # Q1: bag 0 starts at 0 whatever offsets[0] says
        tabs.append(q), idxs.append(i), offs.append(o)
    got = emblib.lookup_ex(idxs, offs)
    want = eo.lookup_rows(tabs, idxs, offs)
    for t in range(len(shapes)):
        assert np.array_equal(bits(got[t]), bits(want[t])), t
    assert emblib.load().emb_check_indices() == 0


def test_pipe_kernel_uniform_tables_like_c2(emblib):
    """All tables pool the same batch (descriptor found by division), pooling U{1..20}."""
    rng = np.random.default_rng(2)
    T, D, R, B = 5, 16, 20000, 4099
    emblib.configure(T, D)
    assert emblib.load().emb_set_tuning(5, 0, 0) == 0
    tabs = [rng.integers(-2**31, 2**31, size=(R, D), dtype=np.int64).astype(np.int32) for _ in range(T)]
    for t in range(T):
        emblib.populate_table_i32(t, tabs[t])
    idxs, offs = zip(*[ragged(rng, R, B, 20, empty_frac=0.0) for _ in range(T)])
    got = emblib.lookup_ex(list(idxs), list(offs))
    want = eo.lookup_rows(tabs, list(idxs), list(offs))
    for t in range(T):
        assert np.array_equal(bits(got[t]), bits(want[t])), t


def test_pipe_kernel_long_bags_take_extra_passes(emblib):
    """Index windows longer than a ring buffer (16 entries per bag) are finished in synchronous passes."""
    rng = np.random.default_rng(9)
    emblib.configure(1, 16)
    assert emblib.load().emb_set_tuning(5, 8, 4) == 0
    q = rng.integers(-2**31, 2**31, size=(1000, 16), dtype=np.int64).astype(np.int32)
    emblib.populate_table_i32(0, q)
    lens = np.array([0, 5000, 1, 70000, 3, 0, 12000, 2] + [300] * 40 + [0, 1, 2, 900], dtype=np.int64)
    off = np.concatenate([[0], np.cumsum(lens)[:-1]]).astype(np.uint32)
    idx = rng.integers(0, 1000, size=int(lens.sum())).astype(np.uint32)
    got = emblib.lookup_ex([idx], [off])[0]
    assert np.array_equal(bits(got), bits(eo.lookup_rows([q], [idx], [off])[0]))


def test_pipe_kernel_overlapping_and_backward_bags(emblib):
    """Unsorted offsets (emb_dpu_lookup.c:109-116 pools every bag on its own): the CTA falls back to bag-by-bag."""
    rng = np.random.default_rng(11)
    emblib.configure(1, 16)
    assert emblib.load().emb_set_tuning(5, 8, 2) == 0
    q = rng.integers(-2**31, 2**31, size=(300, 16), dtype=np.int64).astype(np.int32)
    emblib.populate_table_i32(0, q)
    idx = rng.integers(0, 300, size=4000).astype(np.uint32)
    off = rng.integers(0, 4001, size=500).astype(np.uint32)
    got = emblib.lookup_ex([idx], [off])[0]
    assert np.array_equal(bits(got), bits(eo.lookup_rows([q], [idx], [off])[0]))


def test_pipe_kernel_overflow_and_bad_index(emblib):
    emblib.configure(1, 8)
    assert emblib.load().emb_set_tuning(5, 8, 4) == 0
    q = np.full((4, 8), 2**31 - 1, dtype=np.int32)
    q[1] = -2**31
    emblib.populate_table_i32(0, q)
    idx = np.array([0, 0, 0, 1, 1, 0, 1], dtype=np.uint32)
    off = np.array([0, 3, 5], dtype=np.uint32)
    got = emblib.lookup_ex([idx], [off])[0]
    assert np.array_equal(bits(got), bits(eo.lookup_rows([q], [idx], [off])[0]))  # Q12: wraps
    assert emblib.load().emb_check_indices() == 0
    idx[2] = 4  # one row past the table
    emblib.lookup_ex([idx], [off])
    assert emblib.load().emb_check_indices() != 0


def test_pipe_kernel_device_plan_slices_shards_and_slabs(emblib):
    """Device-resident plan: raw sums of two row shards reduce to the unsharded result, a bag slice covers
    its part of the batch only, and bags are routed to per-rank slabs."""
    from pim_embedding_lookup_b200.dputypes import EMB_Q_RAW_SUMS
    rng = np.random.default_rng(21)
    D, R, B = 16, 3000, 1000
    emblib.configure(3, D)
    lib = emblib.load()
    assert lib.emb_set_tuning(5, 8, 4) == 0
    q = rng.integers(-2**31, 2**31, size=(R, D), dtype=np.int64).astype(np.int32)
    idx, off = ragged(rng, R, B, 25)
    want, want_sums = eo.lookup_rows([q], [idx], [off], want_sums=True)
    want, want_sums = want[0], want_sums[0]
    d_idx = torch.from_numpy(idx.view(np.int32)).cuda()
    d_off = torch.from_numpy(off.view(np.int32)).cuda()
    # table 0: whole; tables 1 and 2: the two row halves
    emblib.populate_table_i32(0, q)
    half = R // 2
    emblib.populate_table_i32(1, q[:half])
    emblib.populate_table_i32(2, q[half:])
    emblib.check(lib.emb_table_set_row_base(2, half))
    emblib.check(lib.emb_table_set_row_base(1, 0))
    out = torch.zeros((B, D), dtype=torch.float32, device="cuda")
    p1 = torch.zeros((B, D), dtype=torch.int32, device="cuda")
    p2 = torch.zeros((B, D), dtype=torch.int32, device="cuda")
    sl = torch.full((B, D), -1.0, dtype=torch.float32, device="cuda")
    slabs = [torch.zeros((250, D), dtype=torch.float32, device="cuda") for _ in range(4)]
    plan = emblib.Plan([
        dict(table_id=0, indices=d_idx, offsets=d_off, out=out),
        dict(table_id=1, indices=d_idx, offsets=d_off, out=p1, flags=EMB_Q_RAW_SUMS),
        dict(table_id=2, indices=d_idx, offsets=d_off, out=p2, flags=EMB_Q_RAW_SUMS),
        dict(table_id=0, indices=d_idx, offsets=d_off, out=sl, bag_first=333, bag_count=401),
        dict(table_id=0, indices=d_idx, offsets=d_off, bags_per_rank=250, out_peer=slabs),
    ])
    plan.launch()
    torch.cuda.synchronize()
    plan.close()
    assert np.array_equal(bits(out.cpu().numpy()), bits(want))
    assert np.array_equal((p1 + p2).cpu().numpy(), want_sums)  # int32 adds wrap like the kernel's
    s = sl.cpu().numpy()
    assert np.array_equal(bits(s[333:734]), bits(want[333:734]))
    assert (s[:333] == -1.0).all() and (s[734:] == -1.0).all()
    got = torch.cat(slabs).cpu().numpy()
    assert np.array_equal(bits(got), bits(want))
