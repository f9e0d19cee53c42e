"""Hot-row tables (emb_table_set_hot_rows / emb_table_remap / emb_remap_indices, include/emb_b200.h): an L2 policy
per row and a reordered slab change where rows live and which of them L2 keeps -- never a result.  Every case is held
to the oracle on the ORIGINAL tables and the caller's ORIGINAL row ids, bit for bit (emb_dpu_lookup.c:106-116,
emb_host.h:207-212); fp32 tables to the 1e-5 relative north_star states."""
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
    m.shutdown()


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def u32(a):  # torch has no uint32 arithmetic; carry the bits as int32
    return torch.from_numpy(np.ascontiguousarray(a, dtype=np.uint32).view(np.int32)).cuda()


def skewed(rng, R, B, max_len, perm):
    """Ragged bags (some empty) over a power-law row distribution scattered by `perm`."""
    lens = rng.integers(0, max_len + 1, size=B)
    off = np.concatenate([[0], np.cumsum(lens)[:-1]]).astype(np.uint32)
    r = (rng.zipf(1.2, size=int(lens.sum())) - 1) % R
    return perm[r].astype(np.uint32), off


def hot_permutation(emblib, rng, R, perm):
    sample = perm[(rng.zipf(1.2, size=20 * R) - 1) % R]
    return emblib.hot_order(np.bincount(sample, minlength=R))


@pytest.mark.parametrize("mode", [0, 1, 2, 3])
@pytest.mark.parametrize("D", [4, 16, 64, 128])
def test_reordered_hot_tables_give_the_same_rows_through_the_host_call(emblib, D, mode):
    """lookup_ex on the caller's row ids: table 0 reordered + hot prefix, table 1 hot prefix only, table 2 plain."""
    rng = np.random.default_rng(100 + D + mode)
    T, R, B = 3, 5000, 700
    emblib.configure(T, D)
    perm = rng.permutation(R)
    tabs, idx, off = [], [], []
    for t in range(T):
        q = rng.integers(-2**31, 2**31, size=(R, D), dtype=np.int64).astype(np.int32)
        tabs.append(q)
        emblib.populate_table_i32(t, q)
        i, o = skewed(rng, R, B, 30, perm)
        idx.append(i)
        off.append(o)
    m = hot_permutation(emblib, rng, R, perm)
    emblib.remap_table(0, m)
    slab = emblib.table_read(0, 0, R)
    assert np.array_equal(slab[m], tabs[0])  # caller row r lives at slab row m[r]
    emblib.set_hot_rows(0, 400, mode)
    emblib.set_hot_rows(1, 10**9, mode)  # clamped to the table
    want = eo.lookup_rows(tabs, idx, off)
    before = emblib.launch_count()
    got = emblib.lookup_ex(idx, off)
    assert emblib.launch_count() == before + 1
    for t in range(T):
        assert np.array_equal(bits(got[t]), bits(want[t])), t
    assert emblib.load().emb_check_indices() == 0
    # an index outside the table is reported and contributes nothing, reordered table or not
    idx[0][5] = R + 3
    got = emblib.lookup_ex(idx, off)
    assert emblib.load().emb_check_indices() == 1
    t0 = tabs[0].copy()
    want0 = eo.lookup_rows([np.vstack([t0, np.zeros((4, D), dtype=np.int32)])], [idx[0]], [off[0]])[0]
    assert np.array_equal(bits(got[0]), bits(want0))


def test_plans_translate_or_trust_the_caller(emblib):
    """Plans over device arrays: the kernel translates the caller's ids, or the caller translated them itself
    (emb_remap_indices + EMB_Q_REMAPPED).  Overlapping bags take the kernel's bag-by-bag path; fp32 tables too."""
    from pim_embedding_lookup_b200.dputypes import EMB_HOT_KEEP, EMB_HOT_STREAM_COLD, EMB_Q_REMAPPED, EMB_TABLE_FLOAT
    rng = np.random.default_rng(7)
    D, R, B = 64, 20000, 3000
    emblib.configure(3, D)
    perm = rng.permutation(R)
    q = rng.integers(-2**31, 2**31, size=(R, D), dtype=np.int64).astype(np.int32)
    emblib.populate_table_i32(0, q)
    emblib.populate_table_i32(1, q)
    m = hot_permutation(emblib, rng, R, perm)
    for t in (0, 1):
        emblib.remap_table(t, m)
        emblib.set_hot_rows(t, 1500, EMB_HOT_KEEP | EMB_HOT_STREAM_COLD)
    i, o = skewed(rng, R, B, 40, perm)
    want = eo.lookup_rows([q], [i], [o])[0]
    di, do = u32(i), u32(o)
    di_slab = di.clone()
    emblib.remap_indices(1, di_slab, torch.cuda.current_stream())
    assert np.array_equal(di_slab.cpu().numpy().view(np.uint32), m[i])
    outs = [torch.empty((B, D), dtype=torch.float32, device="cuda") for _ in range(2)]
    plan = emblib.Plan([dict(table_id=0, indices=di, offsets=do, out=outs[0]),
                        dict(table_id=1, indices=di_slab, offsets=do, out=outs[1], flags=EMB_Q_REMAPPED)])
    plan.launch(torch.cuda.current_stream())
    torch.cuda.synchronize()
    for k in range(2):
        assert np.array_equal(bits(outs[k].cpu().numpy()), bits(want)), k
    plan.close()
    # one-shot device lookup
    outs[0].zero_()
    emblib.lookup_device([dict(table_id=0, indices=di, offsets=do, out=outs[0])], torch.cuda.current_stream())
    torch.cuda.synchronize()
    assert np.array_equal(bits(outs[0].cpu().numpy()), bits(want))
    # overlapping / backward bags (lookup() permits them): bag b covers [off[b], off[b+1]) whatever the order
    o2 = o.copy()
    o2[10:20] = o2[10:20][::-1]
    want2 = eo.lookup_rows([q], [i], [o2])[0]
    plan = emblib.Plan([dict(table_id=0, indices=di, offsets=u32(o2), out=outs[0])])
    plan.launch(torch.cuda.current_stream())
    torch.cuda.synchronize()
    assert np.array_equal(bits(outs[0].cpu().numpy()), bits(want2))
    plan.close()
    # fp32 table, reordered: within 1e-5 relative of the sequential sum
    w = rng.uniform(-1, 1, size=(R, D)).astype(np.float32)
    emblib.populate_table_f32(2, w, EMB_TABLE_FLOAT)
    emblib.remap_table(2, m)
    emblib.set_hot_rows(2, 1500, EMB_HOT_KEEP)
    plan = emblib.Plan([dict(table_id=2, indices=di, offsets=do, out=outs[1])])
    plan.launch(torch.cuda.current_stream())
    torch.cuda.synchronize()
    ends = np.concatenate([o[1:], [len(i)]]).astype(np.int64)
    ref = np.zeros((B, D), dtype=np.float64)
    mag = np.zeros((B, D), dtype=np.float64)
    for b in range(B):
        rows = w[i[o[b] if b else 0:ends[b]]].astype(np.float64)
        ref[b] = rows.sum(0)
        mag[b] = np.abs(rows).sum(0)
    got = outs[1].cpu().numpy().astype(np.float64)
    assert np.all(np.abs(got - ref) <= 1e-5 * np.maximum(mag, 1e-30))  # tolerance: north_star, <= 1e-5 relative
    plan.close()


def test_rules_around_reordered_tables(emblib):
    from pim_embedding_lookup_b200.dputypes import EMB_HOT_KEEP
    lib = emblib.load()
    rng = np.random.default_rng(9)
    D, R = 16, 3000
    emblib.configure(2, D)
    q = rng.integers(-2**31, 2**31, size=(R, D), dtype=np.int64).astype(np.int32)
    emblib.populate_table_i32(0, q)
    emblib.populate_table_i32(1, q)
    i = rng.integers(0, R, size=4000).astype(np.uint32)
    o = (np.arange(400) * 10).astype(np.uint32)
    di, do = u32(i), u32(o)
    out = torch.empty((400, D), dtype=torch.float32, device="cuda")
    want = eo.lookup_rows([q], [i], [o])[0]
    m = rng.permutation(R).astype(np.uint32)

    bad = m.copy()
    bad[1] = bad[0]
    with pytest.raises(emblib.EmbError, match="not a permutation"):
        emblib.remap_table(0, bad)
    with pytest.raises(emblib.EmbError, match="mode"):
        emblib.set_hot_rows(0, 10, 7)
    old_plan = emblib.Plan([dict(table_id=0, indices=di, offsets=do, out=out)])
    emblib.remap_table(0, m)
    with pytest.raises(emblib.EmbError, match="stale"):  # the slab it pointed into is gone
        old_plan.launch(torch.cuda.current_stream())
    old_plan.close()
    with pytest.raises(emblib.EmbError, match="already reordered"):
        emblib.remap_table(0, m)
    assert lib.emb_table_set_row_base(0, 100) != 0  # a reordered table cannot become a row shard
    assert lib.emb_table_set_row_base(1, 0) == 0
    with pytest.raises(emblib.EmbError, match="row shard"):
        emblib.remap_table(1, m)
    # the index-ordered kernel (variant 1) cannot translate: refuse, do not pool the wrong rows
    emblib.check(lib.emb_set_tuning(1, 0, 0))
    try:
        with pytest.raises(emblib.EmbError, match="stretch kernel|cannot translate"):
            emblib.Plan([dict(table_id=0, indices=di, offsets=do, out=out)])
    finally:
        emblib.check(lib.emb_set_tuning(0, 0, 0))
    # hot prefix on a table that is NOT reordered, variant 1: a hint, silently dropped
    emblib.populate_table_i32(1, q)
    emblib.set_hot_rows(1, 100, EMB_HOT_KEEP)
    emblib.check(lib.emb_set_tuning(1, 0, 0))
    try:
        p = emblib.Plan([dict(table_id=1, indices=di, offsets=do, out=out)])
        p.launch(torch.cuda.current_stream())
        torch.cuda.synchronize()
        assert np.array_equal(bits(out.cpu().numpy()), bits(want))
        p.close()
    finally:
        emblib.check(lib.emb_set_tuning(0, 0, 0))
    # populating the table again restores the caller's order
    emblib.populate_table_i32(0, q)
    assert np.array_equal(emblib.table_read(0, 0, R), q)
    got = emblib.lookup_ex([i, i], [o, o])
    assert np.array_equal(bits(got[0]), bits(want)) and np.array_equal(bits(got[1]), bits(want))
    # a reordering asked for while a kernel that cannot translate is selected is refused up front (not at the first
    # lookup() afterwards, which has no error return)
    emblib.check(lib.emb_set_tuning(1, 0, 0))
    try:
        with pytest.raises(emblib.EmbError, match="cannot translate"):
            emblib.remap_table(0, m)
    finally:
        emblib.check(lib.emb_set_tuning(0, 0, 0))
    got = emblib.lookup_ex([i, i], [o, o])
    assert np.array_equal(bits(got[0]), bits(want))
