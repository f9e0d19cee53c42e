"""Parity of the CUDA path with the oracle and with the reference's own outputs, through the
C ABI (ctypes -> emblib.so).  Bit-exact for the int32 fixed-point path (sums and dequantised
floats); float tables: <= 1e-5 relative (north_star), and in fact bit-exact because the
kernel adds in index order like the CPU kernel does.
"""
import glob
import os

import numpy as np
import pytest

import emb_oracle as eo

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")

GOLD = os.path.join(os.path.dirname(__file__), "golden")
DPU_FILES = sorted(glob.glob(os.path.join(GOLD, "ref_dpu_*.npz")))


@pytest.fixture()
def emblib():
    from pim_embedding_lookup_b200 import emblib as m
    m.load()
    m.shutdown()
    yield m
    m.shutdown()


def u32(a):  # torch has no uint32 arithmetic; carry the bits as int32
    return torch.from_numpy(np.ascontiguousarray(a, dtype=np.uint32).view(np.int32)).cuda()


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def ragged(rng, R, B, max_len, empty_frac=0.1):
    lens = rng.integers(1, max_len + 1, size=B)
    lens[rng.random(B) < empty_frac] = 0
    off = np.concatenate([[0], np.cumsum(lens)[:-1]]).astype(np.uint32)
    idx = rng.integers(0, R, size=int(lens.sum())).astype(np.uint32)
    return idx, off


# ---- the reference's two entry points against the reference's own outputs ------------------

@pytest.mark.parametrize("path", DPU_FILES, ids=[os.path.basename(f)[8:-4] for f in DPU_FILES])
def test_populate_mram_and_lookup_match_reference_outputs(emblib, path):
    z = np.load(path)
    T, D, B, P = (int(z[k]) for k in "TDBP")
    emblib.configure(T, D, B, P)
    h = None
    for t in range(T):
        cols = z[f"cols{t}"]
        for j in range(D):  # one call per (table, column), dlrm_dpu_pytorch.py:328-332
            h = emblib.populate_mram(t, cols[j], j)
        assert np.array_equal(emblib.table_read(t, 0, cols.shape[1]), cols.T)
    for c in range(int(z["n_calls"])):
        idx = [z[f"idx{c}_{t}"] for t in range(T)]
        off = [z[f"off{c}_{t}"] for t in range(T)]
        got = emblib.lookup(idx, off, h)
        for t in range(T):
            assert np.array_equal(bits(got[t]), bits(z[f"out{c}_{t}"])), (c, t)
    assert emblib.load().emb_check_indices() == 0


def test_shape_from_environment_like_the_reference(emblib, monkeypatch):
    # upmem/run.sh exports NR_TABLES/NR_COLS/MAX_NR_BATCHES/MAX_INDICES_PER_BATCH
    z = np.load(os.path.join(GOLD, "ref_dpu_toy_c_test.npz"))
    T, D, B, P = (int(z[k]) for k in "TDBP")
    for k, v in zip(("NR_TABLES", "NR_COLS", "MAX_NR_BATCHES", "MAX_INDICES_PER_BATCH"), (T, D, B, P)):
        monkeypatch.setenv(k, str(v))
    h = None
    for j in range(D):
        h = emblib.populate_mram(0, z["cols0"][j], j)
    s = emblib.shape()
    assert (s.nr_tables, s.nr_cols, s.max_nr_batches, s.max_indices_per_batch) == (T, D, B, P)
    got = emblib.lookup([z["idx0_0"]], [z["off0_0"]], h, latency_print=1)[0]
    assert np.array_equal(bits(got), bits(z["out0_0"]))
    expect = eo.dequant((np.arange(1, 9) * 10).astype(np.int32))  # upmem/c_test.py:40,56
    assert np.array_equal(bits(got), bits(np.tile(expect, (B, 1))))


def test_runtime_totals_are_filled(emblib):
    from pim_embedding_lookup_b200.dputypes import DpuRuntimeTotals
    emblib.configure(1, 4, 2, 2)
    rt = DpuRuntimeTotals()
    emblib.populate_mram(0, np.arange(10, dtype=np.int32), 0, rt)
    assert rt.execution_time_populate_copy_in > 0


# ---- variable pooling / every dimension against the oracle -----------------------------------

@pytest.mark.parametrize("D", [4, 8, 16, 32, 64, 128, 6, 20, 48, 256])
def test_lookup_ex_matches_oracle(emblib, D):
    rng = np.random.default_rng(100 + D)
    T = 3
    emblib.configure(T, D)
    tabs, idxs, offs = [], [], []
    for t in range(T):
        R = [1, 257, 5000][t]
        q = rng.integers(-2**31, 2**31, size=(R, D), dtype=np.int64).astype(np.int32)
        tabs.append(q)
        emblib.populate_table_i32(t, q)
        i, o = ragged(rng, R, [1, 333, 2049][t], 40)
        if t == 1:
            o[0] = 2  # Q1: bag 0 starts at 0 whatever offsets[0] says
        idxs.append(i)
        offs.append(o)
    got = emblib.lookup_ex(idxs, offs)
    want = eo.lookup_rows(tabs, idxs, offs)
    for t in range(T):
        assert np.array_equal(bits(got[t]), bits(want[t])), t
    assert emblib.load().emb_check_indices() == 0


def test_repeated_calls_on_the_same_arrays_follow_their_contents(emblib):
    """The reference's caller allocates its index / offset / result arrays once and refills them per
    batch (dlrm_dpu_pytorch.py:484-498,1028-1051).  The library then skips the descriptor upload
    (same addresses, same shapes); the results must still follow the arrays' CONTENTS, and a call
    with other arrays, other lengths or a re-populated table must not see stale descriptors."""
    rng = np.random.default_rng(77)
    T, D, B = 3, 16, 40
    emblib.configure(T, D)
    tabs = []
    for t in range(T):
        q = rng.integers(-2**31, 2**31, size=(500 + t, D), dtype=np.int64).astype(np.int32)
        tabs.append(q)
        emblib.populate_table_i32(t, q)
    for pinned in (False, True):
        mk = (lambda n, dt: emblib.pinned_array(n, dt)) if pinned else (lambda n, dt: np.zeros(n, dtype=dt))
        idxs = [mk(B * 5, np.uint32) for _ in range(T)]
        offs = [mk(B, np.uint32) for _ in range(T)]
        res = [mk((B, D), np.float32) for _ in range(T)]
        call = emblib.HostCall(idxs, offs, res)
        for rep in range(4):  # same arrays, new contents every time
            for t in range(T):
                lens = rng.integers(0, 6, size=B)
                lens[-1] = 5
                o = np.concatenate([[0], np.cumsum(lens)[:-1]])
                offs[t][:] = o
                idxs[t][:] = rng.integers(0, tabs[t].shape[0], size=B * 5)
            got = call()
            want = eo.lookup_rows(tabs, [np.array(i) for i in idxs], [np.array(o) for o in offs])
            for t in range(T):
                assert np.array_equal(bits(got[t]), bits(want[t])), (pinned, rep, t)
        # other lengths through the same library state
        i2, o2 = zip(*[ragged(rng, tabs[t].shape[0], 17 + t, 9) for t in range(T)])
        got = emblib.lookup_ex(list(i2), list(o2))
        want = eo.lookup_rows(tabs, list(i2), list(o2))
        for t in range(T):
            assert np.array_equal(bits(got[t]), bits(want[t])), (pinned, "ragged", t)
        # a table populated again (other size): the next call on the first arrays sees the new table
        tabs[1] = rng.integers(-2**31, 2**31, size=(900, D), dtype=np.int64).astype(np.int32)
        emblib.populate_table_i32(1, tabs[1])
        got = call()
        want = eo.lookup_rows(tabs, [np.array(i) for i in idxs], [np.array(o) for o in offs])
        for t in range(T):
            assert np.array_equal(bits(got[t]), bits(want[t])), (pinned, "repopulated", t)
        if pinned:
            for a in idxs + offs + res:
                emblib.free_pinned(a)


def test_rm2_preset_on_pageable_arrays_is_bounced_and_tiny_calls_too(emblib):
    """upmem/r.sh:6-10 (RM2 as shipped): 32 tables x dim 64, 64 bags x 120 indices = 1.5 MB per call in plain
    NumPy memory.  That is above the 1 MB floor of the bounce rule but below 128 KB per table, so the
    call is bounced (one launch) instead of staged with 96 copies; and a call of a few KB is bounced
    even from page-locked arrays.  Both bit-exact against the oracle."""
    from pim_embedding_lookup_b200.dputypes import EMB_HOST_BOUNCE
    rng = np.random.default_rng(8)
    T, D, B, P, R = 32, 64, 64, 120, 3000
    emblib.configure(T, D, B, P)
    tabs = [rng.integers(-2**31, 2**31, size=(R, D), dtype=np.int64).astype(np.int32) for _ in range(T)]
    for t in range(T):
        emblib.populate_table_i32(t, tabs[t])
    idxs = [rng.integers(0, R, size=B * P).astype(np.uint32) for _ in range(T)]
    offs = [(np.arange(B) * P).astype(np.uint32) for _ in range(T)]
    got = emblib.lookup(idxs, offs)
    tm = emblib.last_timings()
    assert tm.path == EMB_HOST_BOUNCE and tm.launches == 1
    want = eo.lookup_rows(tabs, idxs, offs)
    for t in range(T):
        assert np.array_equal(bits(got[t]), bits(want[t])), t
    # a tiny call on page-locked arrays: 2 bags of table 0..T-1
    pi = [emblib.pinned_array(6, np.uint32) for _ in range(T)]
    po = [emblib.pinned_array(2, np.uint32) for _ in range(T)]
    pr = [emblib.pinned_array((2, D), np.float32) for _ in range(T)]
    for t in range(T):
        pi[t][:] = rng.integers(0, R, size=6)
        po[t][:] = [0, 4]
    got = emblib.lookup_ex(pi, po, pr)
    assert emblib.last_timings().path == EMB_HOST_BOUNCE
    want = eo.lookup_rows(tabs, [np.array(a) for a in pi], [np.array(a) for a in po])
    for t in range(T):
        assert np.array_equal(bits(got[t]), bits(want[t])), t
    for a in pi + po + pr:
        emblib.free_pinned(a)


def test_c1_shape_mini_batch_one(emblib):
    """BASELINE.json configs[0]: Kaggle arch, 26 tables x dim 16, mini-batch 1, one
    index per lookup -- through populate_table_f32 (export_emb's arithmetic on the GPU) and the
    reference's lookup().  Cardinalities from tools/visualize.py:949, the four multi-million-row
    tables capped at 300k rows so the oracle's quantise stays sub-second (tools/small_call_latency.py
    runs the full cardinalities)."""
    rows = [1460, 583, 10131227, 2202608, 305, 24, 12517, 633, 3, 93145, 5683, 8351593, 3194, 27, 14992, 5461306,
            10, 5652, 2173, 4, 7046547, 18, 15, 286181, 105, 142572]
    rows = [min(r, 300_000) for r in rows]
    rng = np.random.default_rng(123)
    T, D = len(rows), 16
    emblib.configure(T, D, 1, 1)
    W = [rng.uniform(-np.sqrt(1 / n), np.sqrt(1 / n), size=(n, D)).astype(np.float32) for n in rows]
    for t in range(T):
        emblib.populate_table_f32(t, W[t])
    tabs = [eo.quantise_rows(w) for w in W]
    for trial in range(3):
        idxs = [np.array([rng.integers(0, n)], dtype=np.uint32) for n in rows]
        if trial == 2:
            idxs = [np.array([n - 1], dtype=np.uint32) for n in rows]  # last row of every table
        offs = [np.zeros(1, dtype=np.uint32) for _ in rows]
        got = emblib.lookup(idxs, offs)
        want = eo.lookup_rows(tabs, idxs, offs)
        for t in range(T):
            assert got[t].shape == (1, D)
            assert np.array_equal(bits(got[t]), bits(want[t])), (trial, t)
            # and it is the float row the CPU path (dlrm_s_pytorch.py:483-488) would return, to 1e-9 steps
            assert np.abs(got[t][0] - W[t][idxs[t][0]]).max() <= 1e-6


@pytest.mark.parametrize("D", [16, 64, 6])
def test_page_locked_arrays_take_the_mapped_path_and_match(emblib, D):
    """Host arrays in page-locked memory: the pooling kernel reads them and writes the results over
    PCIe itself (EMB_HOST_MAPPED, no staging copies); pageable arrays are staged (EMB_HOST_COPY).
    Both must give the oracle's bits."""
    from pim_embedding_lookup_b200.dputypes import (EMB_HOST_AUTO, EMB_HOST_BOUNCE, EMB_HOST_COPY, EMB_HOST_MAPPED,
                                                    EMB_HOST_MAPPED_OUT)
    rng = np.random.default_rng(31 + D)
    T = 4
    emblib.configure(T, D)
    tabs, idxs, offs, pidx, poff, pres = [], [], [], [], [], []
    for t in range(T):
        R = [3, 1000, 70000, 257][t]
        q = rng.integers(-2**31, 2**31, size=(R, D), dtype=np.int64).astype(np.int32)
        tabs.append(q)
        emblib.populate_table_i32(t, q)
        i, o = ragged(rng, R, [5, 4097, 1, 600][t], 30)
        idxs.append(i)
        offs.append(o)
        a = emblib.pinned_array(max(i.size, 1), np.uint32)[:i.size]
        a[:] = i
        b = emblib.pinned_array(o.size, np.uint32)
        b[:] = o
        pidx.append(a)
        poff.append(b)
        pres.append(emblib.pinned_array((o.size, D), np.float32))
    want = eo.lookup_rows(tabs, idxs, offs)

    def run(path, i, o, r, expect):
        emblib.set_host_path(path)
        if r is not None:
            for x in r:
                x[:] = np.nan
        got = emblib.lookup_ex(i, o, r)
        assert emblib.last_timings().path == expect
        for t in range(T):
            assert np.array_equal(bits(got[t]), bits(want[t])), (path, t)

    try:
        run(EMB_HOST_AUTO, pidx, poff, pres, EMB_HOST_MAPPED)      # small + all page-locked: one launch
        assert emblib.last_timings().launches == 1
        total = sum(i.nbytes + o.nbytes + o.size * D * 4 for i, o in zip(idxs, offs))
        small = total <= (1 << 20)
        # pageable arrays: small calls bounce through the library's page-locked arena, large ones are staged
        run(EMB_HOST_AUTO, idxs, offs, None, EMB_HOST_BOUNCE if small else EMB_HOST_COPY)
        run(EMB_HOST_AUTO, idxs, offs, pres, EMB_HOST_BOUNCE if small else EMB_HOST_MAPPED_OUT)
        run(EMB_HOST_BOUNCE, idxs, offs, None, EMB_HOST_BOUNCE)
        assert emblib.last_timings().launches == 1
        run(EMB_HOST_COPY, idxs, offs, None, EMB_HOST_COPY)
        run(EMB_HOST_MAPPED_OUT, idxs, offs, pres, EMB_HOST_MAPPED_OUT)  # pageable inputs, page-locked results
        run(EMB_HOST_MAPPED_OUT, pidx, poff, pres, EMB_HOST_MAPPED_OUT)
        run(EMB_HOST_COPY, pidx, poff, pres, EMB_HOST_COPY)        # forced staging of page-locked arrays
        emblib.set_host_path(EMB_HOST_MAPPED)                      # forced mapping of pageable arrays: error
        with pytest.raises(Exception):
            emblib.lookup_ex(idxs, offs)
        emblib.set_host_path(EMB_HOST_MAPPED_OUT)
        with pytest.raises(Exception):
            emblib.lookup_ex(pidx, poff)  # results default to pageable NumPy arrays
        # caller-owned pageable results, page-locked in place with emb_host_register
        own = [np.full((o.size, D), np.nan, dtype=np.float32) for o in offs]
        lib = emblib.load()
        for r in own:
            emblib.check(lib.emb_host_register(r.ctypes.data, r.nbytes))
        try:
            run(EMB_HOST_MAPPED_OUT, idxs, offs, own, EMB_HOST_MAPPED_OUT)
        finally:
            for r in own:
                emblib.check(lib.emb_host_unregister(r.ctypes.data))
    finally:
        emblib.set_host_path(EMB_HOST_AUTO)
    assert emblib.load().emb_check_indices() == 0


def test_large_call_pipelines_groups(emblib):
    """Enough bytes for several pipeline groups.  EMB_HOST_AUTO on page-locked arrays copies both ways through the
    copy engines (the faster of the two on C2, profiles/r2_host_path.md); EMB_HOST_MAPPED_OUT stages the inputs and
    lets the kernel write the results in place.  Same bits either way."""
    from pim_embedding_lookup_b200.dputypes import EMB_HOST_AUTO, EMB_HOST_COPY, EMB_HOST_MAPPED_OUT
    rng = np.random.default_rng(5)
    T, D, R, B = 6, 16, 50000, 30000
    emblib.configure(T, D)
    tabs, pidx, poff, pres = [], [], [], []
    for t in range(T):
        q = rng.integers(-2**31, 2**31, size=(R, D), dtype=np.int64).astype(np.int32)
        tabs.append(q)
        emblib.populate_table_i32(t, q)
        i, o = ragged(rng, R, B, 20)
        a = emblib.pinned_array(i.size, np.uint32)
        a[:] = i
        b = emblib.pinned_array(o.size, np.uint32)
        b[:] = o
        pidx.append(a)
        poff.append(b)
        pres.append(emblib.pinned_array((B, D), np.float32))
    want = eo.lookup_rows(tabs, pidx, poff)
    try:
        for forced, expect in ((EMB_HOST_AUTO, EMB_HOST_COPY), (EMB_HOST_MAPPED_OUT, EMB_HOST_MAPPED_OUT)):
            emblib.set_host_path(forced)
            for r in pres:
                r[:] = 0
            got = emblib.lookup_ex(pidx, poff, pres)
            tm = emblib.last_timings()
            assert tm.path == expect and tm.launches > 1, (forced, tm.path, tm.launches)
            for t in range(T):
                assert np.array_equal(bits(got[t]), bits(want[t])), (forced, t)
    finally:
        emblib.set_host_path(EMB_HOST_AUTO)


TUNINGS = [(7, 0, 0), (1, 4, 0), (1, 8, 0), (4, 0, 0), (6, 8, 1), (6, 8, 2), (6, 4, 4), (6, 8, 8), (6, 4, 1), (6, 4, 2), (6, 8, 4)]


@pytest.mark.parametrize("tuning", TUNINGS, ids=[f"v{v}k{k}m{m}" for v, k, m in TUNINGS])
@pytest.mark.parametrize("D", [4, 8, 16, 32, 64, 128])
def test_kernel_variants_agree(emblib, tuning, D):
    """Every kernel variant (bag per sub-warp / TMA gather4 / stretch kernel; rows in flight, bags per CTA)."""
    rng = np.random.default_rng(5)
    emblib.configure(2, D)
    assert emblib.load().emb_set_tuning(*tuning) == 0
    try:
        tabs, idxs, offs = [], [], []
        for t, (R, B, L) in enumerate([(4096, 5000, 33), (100, 777, 300)]):
            q = rng.integers(-2**31, 2**31, size=(R, D), dtype=np.int64).astype(np.int32)
            emblib.populate_table_i32(t, q)
            i, o = ragged(rng, R, B, L)
            tabs.append(q), idxs.append(i), offs.append(o)
        got = emblib.lookup_ex(idxs, offs)
        want = eo.lookup_rows(tabs, idxs, offs)
        for t in range(2):
            assert np.array_equal(bits(got[t]), bits(want[t])), t
    finally:
        emblib.load().emb_set_tuning(0, 0, 0)


@pytest.mark.parametrize("variant", [1, 4, 6, 7])
def test_overlapping_and_backward_bags(emblib, variant):
    """lookup() accepts any offsets: bag i covers [offsets[i], offsets[i+1]) on its own
    (emb_dpu_lookup.c:109-116), so bags may overlap or be empty because they run backwards."""
    rng = np.random.default_rng(11)
    emblib.configure(1, 16)
    assert emblib.load().emb_set_tuning(variant, 0, 0) == 0
    try:
        q = rng.integers(-2**31, 2**31, size=(300, 16), dtype=np.int64).astype(np.int32)
        emblib.populate_table_i32(0, q)
        idx = rng.integers(0, 300, size=4000).astype(np.uint32)
        off = rng.integers(0, 4001, size=500).astype(np.uint32)  # unsorted
        got = emblib.lookup_ex([idx], [off])[0]
        assert np.array_equal(bits(got), bits(eo.lookup_rows([q], [idx], [off])[0]))
    finally:
        emblib.load().emb_set_tuning(0, 0, 0)


def test_long_bags_beyond_the_staging_window(emblib):
    """Bags longer than the CTA's shared-memory index window fall back to direct index loads."""
    rng = np.random.default_rng(9)
    emblib.configure(1, 16)
    q = rng.integers(-2**31, 2**31, size=(1000, 16), dtype=np.int64).astype(np.int32)
    emblib.populate_table_i32(0, q)
    lens = np.array([0, 5000, 1, 70000, 3, 0, 12000, 2], dtype=np.int64)
    off = np.concatenate([[0], np.cumsum(lens)[:-1]]).astype(np.uint32)
    idx = rng.integers(0, 1000, size=int(lens.sum())).astype(np.uint32)
    got = emblib.lookup_ex([idx], [off])[0]
    assert np.array_equal(bits(got), bits(eo.lookup_rows([q], [idx], [off])[0]))


def test_int32_overflow_wraps(emblib):
    # Q12: sums leave int32 and wrap (emb_dpu_lookup.c:114)
    emblib.configure(1, 8)
    q = np.full((4, 8), 2**31 - 1, dtype=np.int32)
    q[1] = -2**31
    emblib.populate_table_i32(0, q)
    idx = np.array([0, 0, 0, 1, 1, 0, 1], dtype=np.uint32)
    off = np.array([0, 3, 5], dtype=np.uint32)
    got = emblib.lookup_ex([idx], [off])[0]
    want = eo.lookup_rows([q], [idx], [off])[0]
    assert np.array_equal(bits(got), bits(want))
    # 3 * (2^31 - 1) mod 2^32 = 2147483645
    assert got[0, 0] == np.float32(2147483645) / np.float32(1e9)


def test_empty_table_of_bags_and_zero_indices(emblib):
    emblib.configure(2, 16)
    q = np.arange(32, dtype=np.int32).reshape(2, 16)
    emblib.populate_table_i32(0, q)
    emblib.populate_table_i32(1, q)
    idx = [np.zeros(0, dtype=np.uint32), np.array([1, 1], dtype=np.uint32)]
    off = [np.zeros(4, dtype=np.uint32), np.array([0, 0, 2], dtype=np.uint32)]
    got = emblib.lookup_ex(idx, off)
    assert np.array_equal(got[0], np.zeros((4, 16), dtype=np.float32))
    want = eo.lookup_rows([q, q], idx, off)
    assert np.array_equal(bits(got[1]), bits(want[1]))


def test_out_of_range_index_is_reported(emblib):
    emblib.configure(1, 16)
    emblib.populate_table_i32(0, np.ones((10, 16), dtype=np.int32))
    got = emblib.lookup_ex([np.array([3, 10, 4], dtype=np.uint32)], [np.array([0], dtype=np.uint32)])[0]
    assert emblib.load().emb_check_indices() == 1   # Q9: flagged, not fetched
    assert emblib.load().emb_check_indices() == 0   # cleared by the read
    assert np.array_equal(bits(got), bits(eo.dequant(np.full((1, 16), 2, dtype=np.int32))))


# ---- table export (quantise) -------------------------------------------------------------------

def test_quantise_on_gpu_matches_reference_expression(emblib):
    z = np.load(os.path.join(GOLD, "quantise.npz"))
    w, cols = z["w"], z["cols"]
    emblib.configure(1, w.shape[1])
    emblib.populate_table_f32(0, w)
    assert np.array_equal(emblib.table_read(0, 0, w.shape[0]), cols.T)


def test_quantise_edge_values(emblib):
    w = np.array([[0.0, -0.0, 0.5e-9, 1.5e-9], [2.5e-9, -0.5e-9, 2.147483647, -2.147483648],
                  [2.2, -2.2, 4.3, 1e10], [np.inf, -np.inf, np.nan, 1e-30]], dtype=np.float32)
    emblib.configure(1, 4)
    emblib.populate_table_f32(0, w)
    assert np.array_equal(emblib.table_read(0, 0, 4), eo.quantise_rows(w))


# ---- float tables (the CPU path, nn.EmbeddingBag(mode="sum")) ------------------------------------

def test_float_tables_match_reference_apply_emb(emblib):
    from pim_embedding_lookup_b200.dputypes import EMB_TABLE_FLOAT
    z = np.load(os.path.join(GOLD, "float_bag.npz"))
    for D in z["dims"]:
        n = int(z["n_tables"])
        emblib.shutdown()
        emblib.configure(n, int(D))
        for k in range(n):
            emblib.populate_table_f32(k, z[f"w{D}_{k}"], EMB_TABLE_FLOAT)
        for variant in (0, 1):  # the default (stretch kernel: reassociated at its cut points) and the index-ordered kernel
            assert emblib.load().emb_set_tuning(variant, 0, 0) == 0
            got = emblib.lookup_ex([z[f"idx{D}_{k}"] for k in range(n)], [z[f"off{D}_{k}"] for k in range(n)])
            for k in range(n):
                # north_star: <= 1e-5 relative (fp32 reassociation) vs the reference's CPU lookup
                np.testing.assert_allclose(got[k], z[f"out{D}_{k}"], rtol=1e-5, atol=1e-7)
                if variant == 1:
                    want = eo.bag_sum_f32(z[f"w{D}_{k}"], z[f"idx{D}_{k}"], z[f"off{D}_{k}"])
                    assert np.array_equal(bits(got[k]), bits(want))  # same order of adds -> same bits
        emblib.load().emb_set_tuning(0, 0, 0)


@pytest.mark.parametrize("D", [16, 32, 64, 128])
def test_tma_kernel_float_tables_row_shards_and_bad_indices(emblib, D):
    """Variant 4 (cp.async.bulk.tensor tile::gather4): fp32 tables summed in index order (bit-equal to
    the ordered kernel), a row shard (indices outside contribute nothing: out-of-bounds TMA
    coordinates are zero-filled), an out-of-range index on a whole table (flagged, contributes
    nothing), empty bags, a window that is not a multiple of the stage size."""
    from pim_embedding_lookup_b200.dputypes import EMB_TABLE_FLOAT
    rng = np.random.default_rng(21)
    R, B = 5000, 1237
    emblib.configure(3, D)
    lib = emblib.load()
    w = rng.uniform(-1, 1, size=(R, D)).astype(np.float32)
    q = rng.integers(-2**31, 2**31, size=(R, D), dtype=np.int64).astype(np.int32)
    emblib.populate_table_f32(0, w, EMB_TABLE_FLOAT)
    emblib.populate_table_i32(1, q)
    lo, hi = 1000, 3500
    emblib.populate_table_i32(2, q[lo:hi])
    emblib.check(lib.emb_table_set_row_base(2, lo))
    i, o = ragged(rng, R, B, 23)
    outs = {}
    for variant in (1, 4):
        assert lib.emb_set_tuning(variant, 0, 0) == 0
        f = torch.empty((B, D), dtype=torch.float32, device="cuda")
        a = torch.empty((B, D), dtype=torch.float32, device="cuda")
        s_ = torch.empty((B, D), dtype=torch.int32, device="cuda")
        plan = emblib.Plan([dict(table_id=0, indices=u32(i), offsets=u32(o), out=f),
                            dict(table_id=1, indices=u32(i), offsets=u32(o), out=a),
                            dict(table_id=2, indices=u32(i), offsets=u32(o), out=s_, flags=1)])
        plan.launch(torch.cuda.current_stream())
        torch.cuda.synchronize()
        outs[variant] = (f.cpu().numpy(), a.cpu().numpy(), s_.cpu().numpy())
        plan.close()
        assert lib.emb_check_indices() == 0
    assert np.array_equal(bits(outs[4][0]), bits(outs[1][0]))  # fp32: same order of additions
    assert np.array_equal(bits(outs[4][1]), bits(eo.lookup_rows([q], [i], [o])[0]))
    keep = (i >= lo) & (i < hi)
    masked = q.copy()
    masked[:lo] = 0
    masked[hi:] = 0
    _, sums = eo.lookup_rows([masked], [i], [o], want_sums=True)
    assert keep.any() and np.array_equal(outs[4][2], sums[0])
    # one bad index on the whole table: flagged, and the bag gets the sum of its other rows
    bad = i.copy()
    bad[7] = R + 5
    a = torch.empty((B, D), dtype=torch.float32, device="cuda")
    plan = emblib.Plan([dict(table_id=1, indices=u32(bad), offsets=u32(o), out=a)])
    plan.launch(torch.cuda.current_stream())
    torch.cuda.synchronize()
    assert lib.emb_check_indices() == 1
    q0 = np.vstack([q, np.zeros((10, D), dtype=np.int32)])
    assert np.array_equal(bits(a.cpu().numpy()), bits(eo.lookup_rows([q0], [bad], [o])[0]))
    plan.close()
    lib.emb_set_tuning(0, 0, 0)
