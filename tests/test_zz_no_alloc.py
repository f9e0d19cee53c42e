"""emb_lookup_static: the lookup for hosts that cannot let this library allocate GPU memory.

Some programs that load emblib.so have no HBM to spare for it: a serving engine (or a framework's caching
allocator) that reserved nearly the whole GPU at start-up, where a library's cudaMalloc fails or eats memory the
host had planned on; or a latency-bound loop that cannot take the device-wide stall of cudaMalloc / cudaFree.  For
them the C ABI has one entry point that pools caller-owned device arrays (tables included) and keeps the two things
the kernel needs besides -- its per-table work list and the out-of-range word -- in a fixed array compiled into the
library, 64 queries per call at most.

What is checked here, always through the C ABI:
  * its output equals the output of the existing device-resident entry point (emb_lookup_device on tables the
    library populated) -- bit for bit on fixed-point tables, within the fp32 tolerance of tests/test_gpu_stretch.py
    on float tables -- on two calls in a row with different shapes, because the shared array still holds the first
    call's descriptors when the second one starts;
  * the out-of-range word is cleared by each call;
  * the size limit is enforced with a message;
  * GPU memory in use does not move across a call.

This file sorts last on purpose (pytest -x stops at the first failure; a failure here must not hide other tests).
It passes alone in a fresh process and after the rest of the suite: it uses nothing the other tests leave behind, and
the static call of the first test runs before the library is configured at all.
"""

import numpy as np
import pytest


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def ragged(rng, R, B, max_len):
    lens = rng.integers(0, max_len + 1, size=B)
    off = np.concatenate([[0], np.cumsum(lens)[:-1]]).astype(np.uint32)
    idx = rng.integers(0, R, size=int(lens.sum())).astype(np.uint32)
    return idx, off, lens


def test_more_queries_than_the_fixed_array_holds_are_refused():
    """No GPU needed: the size check comes before any CUDA call.  The error text names the limit."""
    from pim_embedding_lookup_b200 import emblib
    from pim_embedding_lookup_b200.dputypes import EMB_STATIC_MAX_QUERIES, EmbStaticQuery
    lib = emblib.load()
    n = EMB_STATIC_MAX_QUERIES + 1
    arr = (EmbStaticQuery * n)()
    rc = lib.emb_lookup_static(arr, n, 16, None, None)
    assert rc == -1  # EMB_EINVAL
    msg = lib.emb_last_error().decode()
    assert str(EMB_STATIC_MAX_QUERIES) in msg and "EMB_STATIC_MAX_QUERIES" in msg
    assert lib.emb_lookup_static(None, 0, 16, None, None) == -1


@pytest.fixture()
def emblib():
    from pim_embedding_lookup_b200 import emblib as m
    m.load()
    m.shutdown()
    yield m
    m.shutdown()


def _cuda(torch, a, dtype):
    return torch.from_numpy(np.ascontiguousarray(a).view(dtype)).cuda()


def _case(torch, rng, T, D, B0, max_len, float_tables=False):
    """T tables with their own row counts, bag counts and ragged bags; everything a call needs, on the GPU."""
    c = dict(T=T, D=D, host_tables=[], tables=[], rows=[], idx=[], off=[], lens=[])
    for t in range(T):
        R = 700 + 211 * t
        if float_tables:
            w = rng.uniform(-R ** -0.5, R ** -0.5, size=(R, D)).astype(np.float32)
            c["tables"].append(_cuda(torch, w, np.float32))
        else:
            w = rng.integers(-2**31, 2**31, size=(R, D), dtype=np.int64).astype(np.int32)
            c["tables"].append(_cuda(torch, w, np.int32))
        c["host_tables"].append(w)
        c["rows"].append(R)
        i, o, lens = ragged(rng, R, B0 + 13 * t, max_len)
        c["idx"].append(_cuda(torch, i, np.int32))
        c["off"].append(_cuda(torch, o, np.int32))
        c["lens"].append(lens)
    return c


def _static_call(torch, emblib, c, kind, stream=None, check_range=False):
    # the NaN fill runs on the stream the lookup is enqueued on, so the lookup's stores come after it
    with torch.cuda.stream(stream if stream is not None else torch.cuda.current_stream()):
        outs = [torch.full((c["off"][t].numel(), c["D"]), float("nan"), dtype=torch.float32, device="cuda")
                for t in range(c["T"])]
    qs = [dict(table=c["tables"][t], nr_rows=c["rows"][t], kind=kind, indices=c["idx"][t], offsets=c["off"][t],
               out=outs[t]) for t in range(c["T"])]
    flag = emblib.lookup_static(qs, c["D"], stream=stream, check_range=check_range)
    return outs, flag


def _existing_call(torch, emblib, c, kind):
    """The entry point that has been there all along: tables populated into library-owned slabs, then
    emb_lookup_device (descriptor ring and error word from cudaMalloc)."""
    emblib.shutdown()
    emblib.configure(c["T"], c["D"])
    for t in range(c["T"]):
        if kind == 2:
            emblib.populate_table_f32(t, c["host_tables"][t], kind)
        else:
            emblib.populate_table_i32(t, c["host_tables"][t])
    outs = [torch.full((c["off"][t].numel(), c["D"]), float("nan"), dtype=torch.float32, device="cuda")
            for t in range(c["T"])]
    emblib.lookup_device([dict(table_id=t, indices=c["idx"][t], offsets=c["off"][t], out=outs[t])
                          for t in range(c["T"])], torch.cuda.current_stream())
    torch.cuda.synchronize()
    return outs


@pytest.mark.gpu
def test_static_lookup_equals_the_existing_entry_point_twice_in_a_row(emblib):
    torch = pytest.importorskip("torch")
    rng = np.random.default_rng(20261106)
    # call 1: 20 tables (two set-up launches of 16 + 4 descriptors), dimension 16
    # call 2: 3 tables, other bag counts and rows -- descriptors 3..19 of call 1 are still in the array
    first = _case(torch, rng, T=20, D=16, B0=300, max_len=24)
    second = _case(torch, rng, T=3, D=16, B0=1111, max_len=7)
    torch.cuda.synchronize()

    # The static path first, before the library has been configured or has populated anything.
    got1, _ = _static_call(torch, emblib, first, kind=1)
    got2, _ = _static_call(torch, emblib, second, kind=1)  # same stream: ordered behind call 1
    torch.cuda.synchronize()
    assert emblib.load().emb_device() == -1  # the library itself still holds no device state

    for c, got in ((first, got1), (second, got2)):
        want = _existing_call(torch, emblib, c, kind=1)
        for t in range(c["T"]):
            g, w = got[t].cpu().numpy(), want[t].cpu().numpy()
            assert not np.isnan(g).any()
            assert np.array_equal(bits(g), bits(w)), f"table {t}"  # fixed point: bit-exact


@pytest.mark.gpu
@pytest.mark.parametrize("D", [64, 10])  # 10: the any-dimension kernel
def test_static_lookup_float_tables_and_odd_dimension(emblib, D):
    from pim_embedding_lookup_b200.dputypes import EMB_TABLE_FLOAT
    torch = pytest.importorskip("torch")
    rng = np.random.default_rng(77 + D)
    cases = [_case(torch, rng, T=2, D=D, B0=500, max_len=40, float_tables=True),
             _case(torch, rng, T=5, D=D, B0=90, max_len=12, float_tables=True)]
    s = torch.cuda.Stream()
    torch.cuda.synchronize()
    gots = [_static_call(torch, emblib, c, kind=EMB_TABLE_FLOAT, stream=s)[0] for c in cases]  # two calls in a row
    s.synchronize()
    for c, got in zip(cases, gots):
        want = _existing_call(torch, emblib, c, kind=EMB_TABLE_FLOAT)  # default tuning: the stretch kernel when D allows
        for t in range(c["T"]):
            g, w = got[t].cpu().numpy(), want[t].cpu().numpy()
            # tolerance of tests/test_gpu_stretch.py: 1e-5 of (largest |element| x lookups in the bag)
            scale = np.abs(c["host_tables"][t]).max() * np.maximum(c["lens"][t], 1)[:, None]
            assert (np.abs(g - w) <= 1e-5 * scale).all(), f"table {t}"


@pytest.mark.gpu
def test_static_lookup_clears_its_error_word_and_allocates_nothing(emblib):
    torch = pytest.importorskip("torch")
    rng = np.random.default_rng(5)
    c = _case(torch, rng, T=2, D=32, B0=200, max_len=9)
    good = c["idx"][1].clone()
    bad = good.clone()
    bad[::5] = c["rows"][1] + 3  # rows that do not exist
    torch.cuda.synchronize()

    c["idx"][1] = bad
    _, flag = _static_call(torch, emblib, c, kind=1, check_range=True)
    assert flag is True
    c["idx"][1] = good
    outs = [torch.empty((c["off"][t].numel(), c["D"]), dtype=torch.float32, device="cuda") for t in range(c["T"])]
    qs = [dict(table=c["tables"][t], nr_rows=c["rows"][t], indices=c["idx"][t], offsets=c["off"][t], out=outs[t])
          for t in range(c["T"])]
    torch.cuda.synchronize()
    free_before, _ = torch.cuda.mem_get_info()
    flag = emblib.lookup_static(qs, c["D"], check_range=True)  # kernels already loaded by the call above
    free_after, _ = torch.cuda.mem_get_info()
    assert flag is False  # the word the previous call set did not survive into this one
    assert free_after == free_before
    want = _existing_call(torch, emblib, c, kind=1)
    for t in range(c["T"]):
        assert np.array_equal(bits(outs[t].cpu().numpy()), bits(want[t].cpu().numpy()))
