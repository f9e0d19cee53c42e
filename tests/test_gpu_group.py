"""Device groups (emb_configure_devices / $EMB_DEVICES, include/emb_b200.h): whole tables over several GPUs behind the
reference's populate_mram() / lookup().  The arithmetic is that of one GPU (emb_dpu_lookup.c:106-116,
emb_host.h:207-212), so every result is held to the oracle bit for bit; what is new is the routing.  A group may name
one GPU twice, so the whole mechanism (contexts, routing by table, per-GPU slots and streams, merged timings) is
exercised on a one-GPU box; the last test needs two real GPUs."""
import os
import subprocess
import sys

import numpy as np
import pytest

import emb_oracle as eo

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture()
def emblib():
    from pim_embedding_lookup_b200 import emblib as m
    m.load()
    m.shutdown()
    yield m
    m.shutdown()


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def ragged(rng, R, B, max_len):
    lens = rng.integers(0, max_len + 1, size=B)
    off = np.concatenate([[0], np.cumsum(lens)[:-1]]).astype(np.uint32)
    idx = rng.integers(0, R, size=int(lens.sum())).astype(np.uint32)
    return idx, off


def make_tables(rng, T, R, D):
    return [rng.integers(-2**31, 2**31, size=(R + 7 * t, D), dtype=np.int64).astype(np.int32) for t in range(T)]


def group_cases():
    n = torch.cuda.device_count()
    cases = [("one_gpu_twice", [0, 0], None), ("three_contexts_with_a_map", [0, 0, 0], [2, 0, 0, 1, 2])]
    if n >= 2:
        cases.append(("two_gpus", [0, 1], None))
    if n >= 4:
        cases.append(("four_gpus", [0, 1, 2, 3], None))
    return cases


@pytest.mark.parametrize("name,devices,table_map", group_cases(), ids=[c[0] for c in group_cases()])
def test_group_lookup_matches_the_oracle(emblib, name, devices, table_map):
    lib = emblib.load()
    rng = np.random.default_rng(21)
    T, D, R, B = 5, 16, 4000, 1500
    emblib.configure_devices(T, D, devices, table_device=table_map)
    assert lib.emb_group_size() == len(devices)
    tabs = make_tables(rng, T, R, D)
    for t in range(T):
        want_dev = devices[table_map[t]] if table_map else devices[t % len(devices)]
        assert lib.emb_table_device(t) == want_dev
        if t % 2:
            emblib.populate_table_i32(t, tabs[t])
        else:  # the reference's per-column upload, dlrm_dpu_pytorch.py:328-332
            for j in range(D):
                emblib.populate_mram(t, tabs[t][:, j], j)
        assert np.array_equal(emblib.table_read(t, 0, tabs[t].shape[0]), tabs[t])  # read back from the GPU that holds it
    idx, off = zip(*[ragged(rng, tabs[t].shape[0], B, 12) for t in range(T)])
    want = eo.lookup_rows(tabs, idx, off)
    got = emblib.lookup_ex(idx, off)
    for t in range(T):
        assert np.array_equal(bits(got[t]), bits(want[t])), t
    tm = emblib.last_timings()
    assert tm.h2d_bytes >= sum(i.size + o.size for i, o in zip(idx, off)) * 4  # summed over the GPUs
    assert tm.d2h_bytes == T * B * D * 4
    used = len(set(table_map)) if table_map else min(len(devices), T)
    assert tm.launches == used  # one pooling launch per GPU that holds a table
    # a larger call (copies through the copy engines on every GPU), twice in flight
    B2 = 40000
    idx2, off2 = zip(*[ragged(rng, tabs[t].shape[0], B2, 20) for t in range(T)])
    idx3, off3 = zip(*[ragged(rng, tabs[t].shape[0], B2, 20) for t in range(T)])
    c2, c3 = emblib.HostCall(idx2, off2), emblib.HostCall(idx3, off3)
    w2, w3 = eo.lookup_rows(tabs, idx2, off2), eo.lookup_rows(tabs, idx3, off3)
    c2.start()
    c3.start()
    c2.wait()
    c3.wait()
    for t in range(T):
        assert np.array_equal(bits(c2.results[t]), bits(w2[t])) and np.array_equal(bits(c3.results[t]), bits(w3[t])), t
    assert lib.emb_check_indices() == 0
    # an index outside its table, in a table that is not on the first GPU, is reported
    bad = [i.copy() for i in idx]
    t_bad = next(t for t in range(T) if (table_map[t] if table_map else t % len(devices)) != 0)
    bad[t_bad][3] = 10**9
    emblib.lookup_ex(bad, off)
    assert lib.emb_check_indices() == 1 and lib.emb_check_indices() == 0
    # the reference's fixed-shape entry point, shapes changed on every GPU of the group by emb_configure
    P = 4
    emblib.configure(T, D, 64, P)
    assert lib.emb_group_size() == len(devices)
    fi = [rng.integers(0, R, size=64 * P).astype(np.uint32) for _ in range(T)]
    fo = [(np.arange(64) * P).astype(np.uint32) for _ in range(T)]
    got = emblib.lookup(fi, fo)
    want = eo.lookup_rows(tabs, fi, fo)
    for t in range(T):
        assert np.array_equal(bits(got[t]), bits(want[t])), t
    # the group cannot change under live tables; after a shutdown one GPU works as before
    with pytest.raises(emblib.EmbError, match="emb_shutdown"):
        emblib.configure_devices(T, D, devices + [0])
    emblib.shutdown()
    emblib.configure(2, D)
    assert lib.emb_group_size() == 1
    emblib.populate_table_i32(0, tabs[0])
    emblib.populate_table_i32(1, tabs[1])
    got = emblib.lookup_ex(idx[:2], off[:2])
    for t in range(2):
        assert np.array_equal(bits(got[t]), bits(eo.lookup_rows(tabs[:2], idx[:2], off[:2])[t]))


def test_device_resident_entry_points_stay_on_the_first_gpu(emblib):
    rng = np.random.default_rng(4)
    T, D, R = 2, 16, 1000
    emblib.configure_devices(T, D, [0, 0])
    tabs = make_tables(rng, T, R, D)
    for t in range(T):
        emblib.populate_table_i32(t, tabs[t])
    i, o = ragged(rng, R, 300, 8)
    di = torch.from_numpy(i.view(np.int32)).cuda()
    do = torch.from_numpy(o.view(np.int32)).cuda()
    out = torch.empty((300, D), dtype=torch.float32, device="cuda")
    p = emblib.Plan([dict(table_id=0, indices=di, offsets=do, out=out)])  # table 0 lives on the first context
    p.launch(torch.cuda.current_stream())
    torch.cuda.synchronize()
    assert np.array_equal(bits(out.cpu().numpy()), bits(eo.lookup_rows([tabs[0]], [i], [o])[0]))
    p.close()
    with pytest.raises(emblib.EmbError, match="not populated"):  # table 1 lives on the second one
        emblib.Plan([dict(table_id=1, indices=di, offsets=do, out=out)])


def test_plan_launches_in_one_thread_beside_group_host_calls_in_another(emblib):
    """A plan carries the context it was made on: emb_plan_launch (which takes no lock) no longer looks at the library's
    current context, which a device-group host call in another thread switches from GPU to GPU while it runs."""
    import threading
    rng = np.random.default_rng(11)
    T, D, R, B = 4, 16, 3000, 2000
    emblib.configure_devices(T, D, [0, 0], table_device=[0, 1, 0, 1])
    tabs = make_tables(rng, T, R, D)
    for t in range(T):
        emblib.populate_table_i32(t, tabs[t])
    idx, off = zip(*[ragged(rng, tabs[t].shape[0], B, 10) for t in range(T)])
    want_host = eo.lookup_rows(tabs, idx, off)
    i, o = ragged(rng, R, 700, 9)
    di = torch.from_numpy(i.view(np.int32)).cuda()
    do = torch.from_numpy(o.view(np.int32)).cuda()
    want_plan = eo.lookup_rows([tabs[0]], [i], [o])[0]
    out = torch.empty((700, D), dtype=torch.float32, device="cuda")
    plan = emblib.Plan([dict(table_id=0, indices=di, offsets=do, out=out)])
    stream = torch.cuda.Stream()
    errors = []

    def host_calls():
        try:
            for _ in range(30):
                got = emblib.lookup_ex(idx, off)
                for t in range(T):
                    if not np.array_equal(bits(got[t]), bits(want_host[t])):
                        errors.append(f"host call, table {t}")
        except Exception as e:  # noqa: BLE001
            errors.append(repr(e))

    def plan_launches():
        try:
            for _ in range(300):
                out.zero_()
                stream.wait_stream(torch.cuda.current_stream())
                plan.launch(stream)
                stream.synchronize()
                if not np.array_equal(bits(out.cpu().numpy()), bits(want_plan)):
                    errors.append("plan launch")
                    break
        except Exception as e:  # noqa: BLE001
            errors.append(repr(e))

    a, b = threading.Thread(target=host_calls), threading.Thread(target=plan_launches)
    a.start(); b.start(); a.join(); b.join()
    plan.close()
    assert not errors, errors[:3]
    assert emblib.load().emb_check_indices() == 0


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_the_callers_current_device_is_left_alone(emblib):
    """A caller whose own current device is not the library's: allocations, copies and launches still land on the
    library's GPU, and the caller finds its device unchanged after every call (single GPU and device group)."""
    rng = np.random.default_rng(12)
    T, D, R, B = 3, 16, 2000, 900
    tabs = make_tables(rng, T, R, D)
    idx, off = zip(*[ragged(rng, tabs[t].shape[0], B, 10) for t in range(T)])
    want = eo.lookup_rows(tabs, idx, off)
    torch.cuda.set_device(1)
    try:
        torch.zeros(1, device="cuda")  # the caller's context on GPU 1 exists
        for devices in (None, [0, 1]):
            emblib.shutdown()
            if devices is None:
                emblib.configure(T, D, device=0)
            else:
                emblib.configure_devices(T, D, devices)
            assert torch.cuda.current_device() == 1
            for t in range(T):
                emblib.populate_table_i32(t, tabs[t])
                assert torch.cuda.current_device() == 1
            for _ in range(2):  # the second call re-uses, the first one allocates
                got = emblib.lookup_ex(idx, off)
                assert torch.cuda.current_device() == 1
                for t in range(T):
                    assert np.array_equal(bits(got[t]), bits(want[t])), (devices, t)
            assert emblib.load().emb_check_indices() == 0
            assert torch.cuda.current_device() == 1
            if devices is None:  # a plan made on GPU 0 launches there although the thread sits on GPU 1
                di = torch.from_numpy(idx[0].view(np.int32)).to("cuda:0")
                do = torch.from_numpy(off[0].view(np.int32)).to("cuda:0")
                out = torch.empty((B, D), dtype=torch.float32, device="cuda:0")
                plan = emblib.Plan([dict(table_id=0, indices=di, offsets=do, out=out)])
                torch.cuda.synchronize(0)
                plan.launch(None)  # the library's own stream on GPU 0
                emblib.load().emb_stream_sync(None)
                assert torch.cuda.current_device() == 1
                assert np.array_equal(bits(out.cpu().numpy()), bits(want[0]))
                plan.close()
    finally:
        torch.cuda.set_device(0)


def test_an_unmodified_caller_gets_the_group_from_the_environment():
    """populate_mram() / lookup() only, shapes AND the group from the environment (EMB_DEVICES beside the reference's
    NR_TABLES ...): what dlrm_dpu_pytorch.py would see on a box with several GPUs."""
    code = """
import numpy as np, sys
sys.path.insert(0, %r); sys.path.insert(0, %r)
import emb_oracle as eo
from pim_embedding_lookup_b200 import emblib
lib = emblib.load()
rng = np.random.default_rng(8)
T, D, B, P, R = 3, 8, 32, 4, 500
tabs = [rng.integers(-2**31, 2**31, size=(R, D), dtype=np.int64).astype(np.int32) for _ in range(T)]
h = None
for t in range(T):
    for j in range(D):
        h = emblib.populate_mram(t, tabs[t][:, j], j)
assert lib.emb_group_size() == 2, lib.emb_group_size()
idx = [rng.integers(0, R, size=B * P).astype(np.uint32) for _ in range(T)]
off = [(np.arange(B) * P).astype(np.uint32) for _ in range(T)]
got = emblib.lookup(idx, off, h, latency_print=1)
want = eo.lookup_rows(tabs, idx, off)
for t in range(T):
    assert np.array_equal(got[t].view(np.uint32), want[t].view(np.uint32)), t
print("GROUP OK")
""" % (ROOT, os.path.join(ROOT, "oracle"))
    env = dict(os.environ, NR_TABLES="3", NR_COLS="8", MAX_NR_BATCHES="32", MAX_INDICES_PER_BATCH="4", EMB_DEVICES="0,0")
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, env=env, cwd=ROOT, timeout=300)
    assert r.returncode == 0 and "GROUP OK" in r.stdout, r.stdout[-1500:] + r.stderr[-1500:]
    assert "C: Dpu launch latency" in r.stdout  # the reference's six lines, emb_host.h:395-402
