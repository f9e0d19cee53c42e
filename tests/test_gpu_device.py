"""Device-resident entry points (plans), row shards, peer slabs and the size-independent
properties at BASELINE.json's full microbenchmark shape.  All calls go through the C ABI."""
import ctypes as C

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


def dev(a, dtype):
    return torch.from_numpy(np.ascontiguousarray(a).view(dtype)).cuda()


def u32(a):  # torch has no uint32 arithmetic; carry the bits as int32
    return dev(np.ascontiguousarray(a, dtype=np.uint32), np.int32)


def ragged(rng, R, B, max_len):
    lens = rng.integers(0, max_len + 1, size=B)
    off = np.concatenate([[0], np.cumsum(lens)[:-1]]).astype(np.uint32)
    idx = rng.integers(0, R, size=int(lens.sum())).astype(np.uint32)
    return idx, off


def test_plan_launch_matches_oracle(emblib):
    rng = np.random.default_rng(1)
    T, D = 4, 16
    emblib.configure(T, D)
    tabs, qs, outs, idxs, offs = [], [], [], [], []
    for t in range(T):
        R = 3000 + 100 * t
        q = rng.integers(-2**31, 2**31, size=(R, D), dtype=np.int64).astype(np.int32)
        tabs.append(q)
        emblib.populate_table_i32(t, q)
        i, o = ragged(rng, R, 1000 + t, 20)
        idxs.append(i)
        offs.append(o)
        out = torch.empty((len(o), D), dtype=torch.float32, device="cuda")
        outs.append(out)
        qs.append(dict(table_id=t, indices=u32(i), offsets=u32(o), out=out))
    plan = emblib.Plan(qs)
    assert plan.bytes == sum((len(i) + len(o)) * (4 * D + 4) for i, o in zip(idxs, offs))
    before = emblib.launch_count()
    plan.launch(torch.cuda.current_stream())
    torch.cuda.synchronize()
    assert emblib.launch_count() == before + 1  # one kernel for all tables
    want = eo.lookup_rows(tabs, idxs, offs)
    for t in range(T):
        assert np.array_equal(bits(outs[t].cpu().numpy()), bits(want[t]))
    plan.close()


def test_raw_sums_and_strided_output(emblib):
    from pim_embedding_lookup_b200.dputypes import EMB_Q_RAW_SUMS
    rng = np.random.default_rng(2)
    D = 32
    emblib.configure(1, D)
    q = rng.integers(-2**31, 2**31, size=(500, D), dtype=np.int64).astype(np.int32)
    emblib.populate_table_i32(0, q)
    i, o = ragged(rng, 500, 300, 9)
    out = torch.zeros((300, 3 * D), dtype=torch.int32, device="cuda")  # [B, T*D]-style interleaved slab
    plan = emblib.Plan([dict(table_id=0, indices=u32(i), offsets=u32(o), out=out[:, D:2 * D],
                             out_stride=3 * D, flags=EMB_Q_RAW_SUMS)])
    plan.launch(torch.cuda.current_stream())
    torch.cuda.synchronize()
    _, sums = eo.lookup_rows([q], [i], [o], want_sums=True)
    got = out.cpu().numpy()
    assert np.array_equal(got[:, D:2 * D], sums[0])
    assert not got[:, :D].any() and not got[:, 2 * D:].any()


def test_row_shards_reduce_to_the_unsharded_result(emblib):
    """Row-wise split: each slot holds a row range, skips foreign indices, writes raw int32
    partials; emb_reduce_partials adds and dequantises.  int32 adds are order-free -> exact."""
    from pim_embedding_lookup_b200.dputypes import EMB_Q_RAW_SUMS
    rng = np.random.default_rng(3)
    D, R, S = 16, 10000, 3
    emblib.configure(S, D)
    q = rng.integers(-2**31, 2**31, size=(R, D), dtype=np.int64).astype(np.int32)
    bounds = [0, 3333, 7000, R]
    i, o = ragged(rng, R, 2000, 25)
    di, do = u32(i), u32(o)
    parts = [torch.empty((len(o), D), dtype=torch.int32, device="cuda") for _ in range(S)]
    qs = []
    for s in range(S):
        emblib.populate_table_i32(s, q[bounds[s]:bounds[s + 1]])
        assert emblib.load().emb_table_set_row_base(s, bounds[s]) == 0
        qs.append(dict(table_id=s, indices=di, offsets=do, out=parts[s], flags=EMB_Q_RAW_SUMS))
    plan = emblib.Plan(qs)
    plan.launch(torch.cuda.current_stream())
    out = torch.empty((len(o), D), dtype=torch.float32, device="cuda")
    arr = (C.c_void_p * S)(*[p.data_ptr() for p in parts])
    rc = emblib.load().emb_reduce_partials(arr, S, D, out.data_ptr(), D, len(o), 1,
                                           torch.cuda.current_stream().cuda_stream)
    assert rc == 0
    torch.cuda.synchronize()
    assert emblib.load().emb_check_indices() == 0  # foreign rows are not errors on a shard
    want = eo.lookup_rows([q], [i], [o])[0]
    assert np.array_equal(bits(out.cpu().numpy()), bits(want))


def test_bags_routed_to_per_rank_slabs(emblib):
    """bags_per_rank: bag b lands in slab b // bags_per_rank (the fused all-to-all's addressing),
    here with all slabs on one GPU."""
    rng = np.random.default_rng(4)
    D, R, B, N = 16, 2000, 1000, 4
    emblib.configure(1, D)
    q = rng.integers(-2**31, 2**31, size=(R, D), dtype=np.int64).astype(np.int32)
    emblib.populate_table_i32(0, q)
    i, o = ragged(rng, R, B, 12)
    per = B // N
    T_all = 3  # receiver layout [B/N, T_all, D]; this table is column block 1
    slabs = [torch.zeros((per, T_all, D), dtype=torch.float32, device="cuda") for _ in range(N)]
    plan = emblib.Plan([dict(table_id=0, indices=u32(i), offsets=u32(o), out=None, out_stride=T_all * D,
                             bags_per_rank=per, out_peer=[s[:, 1, :] for s in slabs])])
    plan.launch(torch.cuda.current_stream())
    torch.cuda.synchronize()
    want = eo.lookup_rows([q], [i], [o])[0]
    for r in range(N):
        got = slabs[r].cpu().numpy()
        assert np.array_equal(bits(got[:, 1, :]), bits(want[r * per:(r + 1) * per]))
        assert not got[:, 0, :].any() and not got[:, 2, :].any()


def test_device_populate_and_l2_window(emblib):
    rng = np.random.default_rng(6)
    D, R = 64, 4096
    emblib.configure(1, D)
    w = rng.uniform(-R ** -0.5, R ** -0.5, size=(R, D)).astype(np.float32)
    dw = torch.from_numpy(w).cuda()
    from pim_embedding_lookup_b200.dputypes import EMB_TABLE_FIXED
    s = torch.cuda.current_stream().cuda_stream
    assert emblib.load().emb_populate_table_device(0, R, dw.data_ptr(), 1, EMB_TABLE_FIXED, s) == 0
    q = eo.quantise_rows(w)
    assert np.array_equal(emblib.table_read(0, 0, R), q)
    assert emblib.load().emb_l2_persist(0, 0, 1024, 1.0, s) == 0
    i, o = ragged(rng, 1024, 512, 80)
    out = torch.empty((512, D), dtype=torch.float32, device="cuda")
    plan = emblib.Plan([dict(table_id=0, indices=u32(i), offsets=u32(o), out=out)])
    plan.launch(torch.cuda.current_stream())
    torch.cuda.synchronize()
    assert emblib.load().emb_l2_persist(0, 0, 0, 0.0, s) == 0
    assert np.array_equal(bits(out.cpu().numpy()), bits(eo.lookup_rows([q], [i], [o])[0]))


# ---- BASELINE.json configs[1] at full size: 26 x 1M x 16, 65536 bags, pooling 1..20 -----------------

def test_full_size_microbench_properties(emblib):
    """Too big for a per-element CPU check in seconds, so: (1) checksum of checksums -- the
    wrap-sum over all bags of the raw int32 pooled rows equals the wrap-sum of the gathered
    rows, computed independently with torch ops; (2) linearity -- pooling (A + B) equals
    pooling A plus pooling B in wrap arithmetic; (3) a random sample of bags against the
    oracle; (4) dequantised output == fp32(raw)/1e9 bit for bit."""
    from pim_embedding_lookup_b200.dputypes import EMB_Q_RAW_SUMS
    T, R, D, B = 26, 1_000_000, 16, 65536
    emblib.configure(2 * T, D)
    g = torch.Generator(device="cuda").manual_seed(123)
    rng = np.random.default_rng(123)
    s = torch.cuda.current_stream()
    tabs, idxs, offs, raws, outs, qs = [], [], [], [], [], []
    for t in range(T):
        a = torch.randint(-2**31, 2**31, (R, D), dtype=torch.int64, device="cuda", generator=g).to(torch.int32)
        tabs.append(a)
        assert emblib.load().emb_populate_table_device(t, R, a.data_ptr(), 0, 1, s.cuda_stream) == 0
        lens = rng.integers(1, 21, size=B)
        off = np.concatenate([[0], np.cumsum(lens)[:-1]]).astype(np.uint32)
        idx = rng.integers(0, R, size=int(lens.sum())).astype(np.uint32)
        idxs.append(idx)
        offs.append(off)
        raws.append(torch.empty((B, D), dtype=torch.int32, device="cuda"))
        outs.append(torch.empty((B, D), dtype=torch.float32, device="cuda"))
        di, do = u32(idx), u32(off)
        qs.append(dict(table_id=t, indices=di, offsets=do, out=raws[t], flags=EMB_Q_RAW_SUMS))
        qs.append(dict(table_id=t, indices=di, offsets=do, out=outs[t]))
    plan = emblib.Plan(qs)
    plan.launch(s)
    torch.cuda.synchronize()
    assert emblib.load().emb_check_indices() == 0
    for t in range(T):
        di = qs[2 * t]["indices"].to(torch.int64) & 0xffffffff
        # (1) checksum of checksums, int64 accumulate then wrap to 32 bits
        want = tabs[t].index_select(0, di).to(torch.int64).sum(0) & 0xffffffff
        got = raws[t].to(torch.int64).sum(0) & 0xffffffff
        assert torch.equal(want, got), t
        # (4) epilogue
        deq = (raws[t].to(torch.float32) / torch.tensor(1e9, dtype=torch.float32, device="cuda"))
        assert torch.equal(deq.view(torch.int32), outs[t].view(torch.int32)), t
    # (3) sampled bags of three tables against the oracle
    for t in (0, 13, 25):
        host_tab = tabs[t].cpu().numpy()
        pick = rng.choice(B - 1, size=512, replace=False)
        sub_idx, sub_off, pos = [], [], 0
        for b in pick:
            seg = idxs[t][offs[t][b]:offs[t][b + 1]]
            sub_off.append(pos)
            sub_idx.append(seg)
            pos += len(seg)
        want = eo.lookup_rows([host_tab], [np.concatenate(sub_idx)], [np.asarray(sub_off, dtype=np.uint32)])[0]
        assert np.array_equal(bits(outs[t][torch.from_numpy(pick).cuda()].cpu().numpy()), bits(want)), t
    plan.close()
    # (2) linearity on table 0: slot T holds B-table, slot T+1 holds A+B (wrapping)
    bt = torch.randint(-2**31, 2**31, (R, D), dtype=torch.int64, device="cuda", generator=g).to(torch.int32)
    ab = (tabs[0] + bt)  # int32 add wraps
    assert emblib.load().emb_populate_table_device(T, R, bt.data_ptr(), 0, 1, s.cuda_stream) == 0
    assert emblib.load().emb_populate_table_device(T + 1, R, ab.data_ptr(), 0, 1, s.cuda_stream) == 0
    rb = torch.empty((B, D), dtype=torch.int32, device="cuda")
    rab = torch.empty((B, D), dtype=torch.int32, device="cuda")
    di, do = qs[0]["indices"], qs[0]["offsets"]
    p2 = emblib.Plan([dict(table_id=T, indices=di, offsets=do, out=rb, flags=EMB_Q_RAW_SUMS),
                      dict(table_id=T + 1, indices=di, offsets=do, out=rab, flags=EMB_Q_RAW_SUMS)])
    p2.launch(s)
    torch.cuda.synchronize()
    assert torch.equal(raws[0] + rb, rab)
    p2.close()


@pytest.mark.parametrize("D,variant", [(16, 1), (6, 0), (16, 4), (64, 4), (128, 4), (16, 6), (8, 6), (64, 6), (128, 6), (4, 6)])
def test_bag_slices_cover_the_batch(emblib, D, variant):
    """bag_first / bag_count: queries that each take a slice of the bags (replicas of a small table on
    several ranks) together produce exactly the whole batch's result; bags outside a slice are not
    touched.  Slices include bag 0 (starts at 0 whatever offsets[0] says) and the last bag (ends at
    indices_len)."""
    rng = np.random.default_rng(9)
    R, B = 1500, 777
    emblib.configure(1, D)
    assert emblib.load().emb_set_tuning(variant, 0, 0) == 0
    q = rng.integers(-2**31, 2**31, size=(R, D), dtype=np.int64).astype(np.int32)
    emblib.populate_table_i32(0, q)
    i, o = ragged(rng, R, B, 15)
    o[0] = 3  # ignored (emb_dpu_lookup.c:60-63)
    want = eo.lookup_rows([q], [i], [o])[0]
    di, do = u32(i), u32(o)
    out = torch.full((B, D), float("nan"), dtype=torch.float32, device="cuda")
    cuts = [0, 1, 64, 65, 300, 776, B]
    # one plan with all slices (one launch) ...
    plan = emblib.Plan([dict(table_id=0, indices=di, offsets=do, out=out, bag_first=a, bag_count=b - a)
                        for a, b in zip(cuts[:-1], cuts[1:])])
    plan.launch(torch.cuda.current_stream())
    torch.cuda.synchronize()
    assert np.array_equal(bits(out.cpu().numpy()), bits(want))
    plan.close()
    # ... and a single slice leaves the other bags alone; bag_count 0 = "to the end"
    out.fill_(float("nan"))
    plan = emblib.Plan([dict(table_id=0, indices=di, offsets=do, out=out, bag_first=300)])
    plan.launch(torch.cuda.current_stream())
    torch.cuda.synchronize()
    got = out.cpu().numpy()
    assert np.isnan(got[:300]).all() and np.array_equal(bits(got[300:]), bits(want[300:]))
    plan.close()
    # bag_rotate changes the order the bags are pooled in, never the result
    for kw in (dict(bag_rotate=300), dict(bag_rotate=B - 1), dict(bag_first=65, bag_count=600, bag_rotate=400)):
        out.fill_(float("nan"))
        plan = emblib.Plan([dict(table_id=0, indices=di, offsets=do, out=out, **kw)])
        plan.launch(torch.cuda.current_stream())
        torch.cuda.synchronize()
        got = out.cpu().numpy()
        a = kw.get("bag_first", 0)
        b = a + kw["bag_count"] if "bag_count" in kw else B
        assert np.array_equal(bits(got[a:b]), bits(want[a:b])), kw
        assert np.isnan(got[:a]).all() and np.isnan(got[b:]).all()
        plan.close()
    with pytest.raises(emblib.EmbError):
        emblib.Plan([dict(table_id=0, indices=di, offsets=do, out=out, bag_first=700, bag_count=100)])
    with pytest.raises(emblib.EmbError):
        emblib.Plan([dict(table_id=0, indices=di, offsets=do, out=out, bag_first=100, bag_count=100, bag_rotate=50)])
    emblib.load().emb_set_tuning(0, 0, 0)
