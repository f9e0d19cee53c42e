"""The reference's DLRM-side call sequence on stock torch, through the kwarg shim
(pim-embedding-lookup_b200/embedding_bag.py) and the two reference entry points:

  export_emb          PIM-dlrm-new/dlrm_dpu_pytorch.py:306-356   quantise, transpose, populate_mram per column
  result prealloc     :484-498                                   float32 [B, D] tensors + c_uint64[T] of addresses
  pointer prep        :1028-1051                                 ctypes uint32 arrays per table + c_uint64[T]
  apply_emb           :559-570                                   emb_l[0](..., use_dpu=True, ...)

Results must equal the oracle bit for bit and the stock nn.EmbeddingBag within the reference's own
validate_result tolerance (1e-6 absolute, upmem/src/load_generator.c:40-65)."""
import ctypes as C

import numpy as np
import pytest

import emb_oracle as eo

pytestmark = pytest.mark.gpu


@pytest.fixture()
def shim(monkeypatch):
    from pim_embedding_lookup_b200 import emblib, embedding_bag
    emblib.load()
    emblib.shutdown()
    embedding_bag.install()
    yield emblib, embedding_bag
    embedding_bag.uninstall()
    emblib.shutdown()


@pytest.mark.parametrize("register_results", [False, True])
def test_dlrm_dpu_call_sequence_on_stock_torch(shim, monkeypatch, register_results):
    import torch
    emblib, _ = shim
    T, D, B, P = 5, 16, 128, 3
    for k, v in dict(NR_TABLES=T, NR_COLS=D, MAX_NR_BATCHES=B, MAX_INDICES_PER_BATCH=P).items():
        monkeypatch.setenv(k, str(v))  # upmem/run.sh exports; dlrm_dpu_pytorch.py:128 reads NR_TABLES
    lib = emblib.load()
    rng = np.random.default_rng(3)
    ln_emb = [1460, 583, 30000, 3, 305]

    # create_emb :359-409 -- W ~ U(-1/sqrt(n), 1/sqrt(n)) fp32, one nn.EmbeddingBag per table
    emb_l = torch.nn.ModuleList()
    for n in ln_emb:
        W = rng.uniform(-np.sqrt(1 / n), np.sqrt(1 / n), size=(n, D)).astype(np.float32)
        e = torch.nn.EmbeddingBag(n, D, mode="sum", sparse=True)
        e.weight.data = torch.tensor(W, requires_grad=True)
        emb_l.append(e)

    # export_emb :306-356
    dpu_set_ptr = None
    for k in range(T):
        tmp = list(emb_l[k].parameters())[0].detach().numpy()
        q = np.rint(tmp * 10**9).astype(int)          # float32 product (:318)
        q = np.ascontiguousarray(np.transpose(q))     # [D, R] (:320)
        for i in range(D):
            col = (C.c_int32 * ln_emb[k])(*q[i])      # wraps mod 2^32 like the reference (:329)
            dpu_set_ptr = lib.populate_mram(k, ln_emb[k], i, col, None)
    assert dpu_set_ptr

    # result prealloc :484-498
    final_results = [torch.zeros(B, D, dtype=torch.float32) for _ in range(T)]
    if register_results:
        for t in final_results:
            emblib.check(lib.emb_host_register(C.c_void_p(t.data_ptr()), t.numel() * 4))
    final_results_ptr = (C.c_uint64 * T)(*[t.data_ptr() for t in final_results])

    # one batch: fixed pooling P, uniform indices (dlrm_data_pytorch.py:1040-1080); pointer prep :1028-1051
    lS_i = [torch.tensor(rng.integers(0, n, size=B * P), dtype=torch.int64) for n in ln_emb]
    lS_o = [torch.arange(0, B * P, P, dtype=torch.int64) for _ in ln_emb]
    ind_store = [(C.c_uint32 * (B * P))(*ind.tolist()) for ind in lS_i]
    off_store = [(C.c_uint32 * B)(*off.tolist()) for off in lS_o]
    ind_pointers_c = (C.c_uint64 * T)(*[C.addressof(a) for a in ind_store])
    off_pointers_c = (C.c_uint64 * T)(*[C.addressof(a) for a in off_store])

    # apply_emb :559-570
    try:
        V = emb_l[0](lS_i[0], lS_o[0], num_of_tables=T, dpu_set_ptr=int(dpu_set_ptr), use_dpu=True,
                     final_results_ptr=C.addressof(final_results_ptr), indices_ptr=C.addressof(ind_pointers_c),
                     offsets_ptr=C.addressof(off_pointers_c), latency_print=0)
        assert V.numel() == 1  # dummy tensor, EmbeddingBag.cpp:754-777
        for k in range(T):
            w = list(emb_l[k].parameters())[0].detach().numpy()
            want = eo.lookup_rows([eo.quantise_rows(w)], [lS_i[k].numpy().astype(np.uint32)],
                                  [lS_o[k].numpy().astype(np.uint32)])[0]
            got = final_results[k].numpy()
            assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), k
            stock = emb_l[k](lS_i[k], lS_o[k]).detach().numpy()  # plain forward still works through the shim
            assert np.max(np.abs(stock - got)) <= 1e-6
    finally:
        if register_results:
            for t in final_results:
                lib.emb_host_unregister(C.c_void_p(t.data_ptr()))


def test_num_of_tables_mismatch_is_an_error(shim, monkeypatch):
    import torch
    emblib, _ = shim
    emblib.configure(2, 4, 2, 1)
    e = torch.nn.EmbeddingBag(4, 4, mode="sum")
    with pytest.raises(ValueError):
        e(torch.zeros(1, dtype=torch.int64), torch.zeros(1, dtype=torch.int64), num_of_tables=3, use_dpu=True,
          final_results_ptr=1, indices_ptr=1, offsets_ptr=1)


def test_device_bags_hand_pooled_rows_to_torch(shim):
    """CUDA index tensors in, one [B, T, D] CUDA tensor out, asynchronous on torch's stream."""
    import torch
    emblib, embedding_bag = shim
    T, D, B = 4, 16, 3000
    emblib.configure(T, D)
    rng = np.random.default_rng(11)
    tabs, idx, off = [], [], []
    for t in range(T):
        R = [7, 100, 50000, 4096][t]
        w = rng.uniform(-R ** -0.5, R ** -0.5, size=(R, D)).astype(np.float32)
        tabs.append(eo.quantise_rows(w))
        emblib.populate_table_f32(t, w)
        lens = rng.integers(0, 9, size=B)
        lens[0] = max(lens[0], 1)
        off.append(np.concatenate([[0], np.cumsum(lens)[:-1]]).astype(np.int64))
        idx.append(rng.integers(0, R, size=int(lens.sum())).astype(np.int64))
    bags = embedding_bag.DeviceBags(T, D)
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        for rep in range(6):  # more calls than descriptor-ring slots, mixed int64 / int32 inputs
            dt = torch.int64 if rep % 2 else torch.int32
            di = [torch.from_numpy(i).to("cuda").to(dt) for i in idx]
            do = [torch.from_numpy(o).to("cuda").to(dt) for o in off]
            out = bags(di, do)
            y = out.sum()  # consumed by a torch op on the same stream, no host sync in between
    s.synchronize()
    want = np.stack(eo.lookup_rows(tabs, [i.astype(np.uint32) for i in idx], [o.astype(np.uint32) for o in off]), axis=1)
    got = out.cpu().numpy()
    assert got.shape == (B, T, D)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
    assert np.isfinite(float(y))
    assert emblib.load().emb_check_indices() == 0
