"""emb_lookup_ex_async / emb_lookup_wait: the host-pointer call with two calls in flight (include/emb_b200.h).
Same arithmetic as lookup() (emb_host.h:234-404, emb_dpu_lookup.c:106-116), so every result is held to the oracle bit
for bit; what is new is the ordering, so the tests keep two calls in flight over different inputs, on every host path
(copies through the copy engines, mapped results, mapped in place, the bounce arena), mix them with synchronous calls
and wait for tickets out of order."""
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


def ragged(rng, R, B, max_len, empty_frac=0.1):
    lens = rng.integers(1, max_len + 1, size=B)
    lens[rng.random(B) < empty_frac] = 0
    off = np.concatenate([[0], np.cumsum(lens)[:-1]]).astype(np.uint32)
    idx = rng.integers(0, R, size=int(lens.sum())).astype(np.uint32)
    return idx, off


def make_calls(emblib, rng, tabs, B, max_len, n_calls, pinned):
    """n_calls HostCall objects over different random batches, each with its own result arrays, + the oracle's rows."""
    T, (R, D) = len(tabs), tabs[0].shape
    calls, wants = [], []
    for _ in range(n_calls):
        idx, off, res = [], [], []
        for _t in range(T):
            i, o = ragged(rng, R, B, max_len)
            if pinned:
                a = emblib.pinned_array(max(i.size, 1), np.uint32)[:i.size]
                a[:] = i
                b = emblib.pinned_array(o.size, np.uint32)
                b[:] = o
                r = emblib.pinned_array((B, D), np.float32)
            else:
                a, b, r = i, o, np.empty((B, D), dtype=np.float32)
            idx.append(a)
            off.append(b)
            res.append(r)
        calls.append(emblib.HostCall(idx, off, res))
        wants.append(eo.lookup_rows(tabs, idx, off))
    return calls, wants


def check(call, want, what):
    for t, w in enumerate(want):
        assert np.array_equal(bits(call.results[t]), bits(w)), (what, t)


SHAPES = {
    # name: (T, D, R, B, max bag, page-locked arrays, forced path or None, path the library must report)
    "copy_groups": (6, 16, 50000, 30000, 20, True, None, "EMB_HOST_COPY"),
    "mapped_out": (6, 16, 50000, 30000, 20, True, "EMB_HOST_MAPPED_OUT", "EMB_HOST_MAPPED_OUT"),
    "pageable_copy": (4, 64, 20000, 20000, 10, False, None, "EMB_HOST_COPY"),
    "mapped_small": (3, 16, 1000, 2000, 8, True, None, "EMB_HOST_MAPPED"),
    "bounced_small": (3, 32, 1000, 300, 8, False, None, "EMB_HOST_BOUNCE"),
}


@pytest.mark.parametrize("shape", sorted(SHAPES))
def test_two_calls_in_flight_match_the_oracle(emblib, shape):
    from pim_embedding_lookup_b200 import dputypes as dt
    T, D, R, B, max_len, pinned, forced, expect = SHAPES[shape]
    rng = np.random.default_rng(11)
    emblib.configure(T, D)
    tabs = []
    for t in range(T):
        q = rng.integers(-2**31, 2**31, size=(R, D), dtype=np.int64).astype(np.int32)
        tabs.append(q)
        emblib.populate_table_i32(t, q)
    calls, wants = make_calls(emblib, rng, tabs, B, max_len, 4, pinned)
    try:
        if forced:
            emblib.set_host_path(getattr(dt, forced))
        for c in calls:
            for r in c.results:
                r[:] = np.float32(-7.0)
        # the documented loop: async(k+1); wait(k); consume k
        n = 9
        calls[0].start()
        for k in range(n):
            if k + 1 < n:
                calls[(k + 1) % 4].start()
            c = calls[k % 4]
            c.wait()
            assert emblib.last_timings().path == getattr(dt, expect)
            check(c, wants[k % 4], (shape, k))
            for r in c.results:  # the next use of these arrays must write every row again
                r[:] = np.float32(-7.0)
    finally:
        emblib.set_host_path(dt.EMB_HOST_AUTO)
    assert emblib.load().emb_check_indices() == 0


def test_tickets_sync_calls_and_a_third_call(emblib):
    """Waiting out of order, waiting twice, a synchronous call while two are in flight (it completes them first), a
    third asynchronous call (it completes the oldest), a ticket that was never issued."""
    lib = emblib.load()
    T, D, R, B = 4, 16, 30000, 20000
    rng = np.random.default_rng(3)
    emblib.configure(T, D)
    tabs = []
    for t in range(T):
        q = rng.integers(-2**31, 2**31, size=(R, D), dtype=np.int64).astype(np.int32)
        tabs.append(q)
        emblib.populate_table_i32(t, q)
    calls, wants = make_calls(emblib, rng, tabs, B, 12, 4, True)

    t0 = calls[0].start()
    t1 = calls[1].start()
    assert t1 == t0 + 1
    emblib.check(lib.emb_lookup_wait(t1))  # the younger one first
    check(calls[1], wants[1], "younger first")
    emblib.check(lib.emb_lookup_wait(t0))
    check(calls[0], wants[0], "older second")
    emblib.check(lib.emb_lookup_wait(t0))  # again: returns at once
    assert lib.emb_lookup_wait(t1 + 5) != 0 and lib.emb_lookup_wait(0) != 0

    for c in calls:
        for r in c.results:
            r[:] = 0
    calls[0].start()
    calls[1].start()
    calls[2]()  # emb_lookup_ex: completes both calls in flight, then runs
    check(calls[2], wants[2], "synchronous call")
    check(calls[0], wants[0], "drained 0")
    check(calls[1], wants[1], "drained 1")
    calls[0].wait()
    calls[1].wait()

    for c in calls:
        for r in c.results:
            r[:] = 0
    a = calls[0].start()
    calls[1].start()
    calls[2].start()  # third call: call `a` completes before its slot is reused
    check(calls[0], wants[0], "displaced by the third call")
    emblib.check(lib.emb_lookup_wait(a))
    calls[3].start()
    calls[2].wait()
    calls[3].wait()
    calls[1].wait()
    for k in range(4):
        check(calls[k], wants[k], ("after all waits", k))
    # shutdown with a call in flight must not hang or crash
    calls[0].start()
    emblib.shutdown()
