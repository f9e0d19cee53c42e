"""Generate tests/golden/*.npz from the REFERENCE ITSELF (build container only).

TEST INFRASTRUCTURE.  Run as `python oracle/make_golden.py`; needs /root/reference.
Three families, each file carrying its inputs and the reference's outputs:

  ref_dpu_*.npz   the reference's own host library + DPU program (upmem/include/emb_host.h,
                  upmem/src/dpu/emb_dpu_lookup.c) compiled verbatim against the UPMEM-SDK
                  stand-in (oracle/ref_sim.py) and called exactly as
                  PIM-dlrm-new/dlrm_dpu_pytorch.py:308-332 / EmbeddingBag.cpp:751 call it.
                  ref_dpu_toy_c_test.npz is the reference's toy known-answer vector
                  (upmem/c_test.py:40,56: 4x8 table, every bag = rows [1,3,2,0]).
  quantise.npz    the table export arithmetic, by executing the reference's expression
                  (dlrm_dpu_pytorch.py:318-320,329) with NumPy + ctypes on seeded weights.
  float_bag.npz   the CPU float path: the reference's DLRM_Net.create_emb/apply_emb imported
                  from /root/reference/PIM-dlrm-new/dlrm_s_pytorch.py (nn.EmbeddingBag sum).

All seeds are fixed; the script is deterministic.
"""
from __future__ import annotations

import contextlib
import ctypes
import io
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = os.path.join(os.path.dirname(HERE), "tests", "golden")
sys.path.insert(0, HERE)
import ref_sim  # noqa: E402


def dpu_case(name, T, D, B, P, rows, seed, kind="uniform", n_calls=2):
    assert ref_sim.in_domain(T, D, B), (T, D, B)
    rng = np.random.default_rng(seed)
    cols = []
    for r in rows:
        if kind == "overflow":  # Q12: sums leave int32 and must wrap
            c = rng.integers(2**30, 2**31 - 1, size=(D, r), dtype=np.int64)
            c *= rng.choice([-1, 1], size=c.shape)
        elif kind == "scaled":  # what export_emb produces: |q| <= 1e9/sqrt(R)
            w = rng.uniform(-np.sqrt(1 / r), np.sqrt(1 / r), size=(r, D)).astype(np.float32)
            c = np.rint(w * (10**9)).astype(int).T
        else:
            c = rng.integers(-2**31, 2**31, size=(D, r), dtype=np.int64)
        cols.append(np.ascontiguousarray(c).astype(np.int32))
    calls = []
    for c in range(n_calls):
        idx, off = [], []
        for r in rows:
            idx.append(rng.integers(0, r, size=B * P).astype(np.uint32))
            if c == 0:  # fixed pooling, offsets = i*P (load_generator.c:88)
                off.append((np.arange(B) * P).astype(np.uint32))
            else:  # ragged: empty bags, offsets[0] != 0 (Q1), last bag runs to INDICES_LEN (Q2)
                o = np.sort(rng.integers(0, B * P + 1, size=B)).astype(np.uint32)
                o[0] = rng.integers(0, 3)
                off.append(o)
        calls.append((idx, off))
    outs = ref_sim.run(T, D, B, P, cols, calls)
    payload = dict(T=T, D=D, B=B, P=P, n_calls=n_calls)
    for t in range(T):
        payload[f"cols{t}"] = cols[t]
    for c, (idx, off) in enumerate(calls):
        for t in range(T):
            payload[f"idx{c}_{t}"] = idx[t]
            payload[f"off{c}_{t}"] = off[t]
            payload[f"out{c}_{t}"] = outs[c][t]
            assert not np.isnan(outs[c][t]).any(), "reference left results unwritten"
    np.savez_compressed(os.path.join(GOLD, f"ref_dpu_{name}.npz"), **payload)
    print("wrote", name)


def toy_c_test():
    # upmem/c_test.py:40 -- row-major 4x8 table; :56 -- every bag looks up rows [1,3,2,0]
    table = np.array([1, 2, 3, 4, 5, 6, 7, 8, 2, 4, 6, 8, 10, 12, 14, 16, 3, 6, 9, 12, 15, 18, 21, 24,
                      4, 8, 12, 16, 20, 24, 28, 32], dtype=np.int32).reshape(4, 8)
    T, D, B, P = 1, 8, 32, 4
    cols = [np.ascontiguousarray(table.T)]
    idx = [np.tile(np.array([1, 3, 2, 0], dtype=np.uint32), B)]
    off = [(np.arange(B) * 4).astype(np.uint32)]
    out = ref_sim.run(T, D, B, P, cols, [(idx, off)])[0][0]
    expect = (np.arange(1, 9) * 10).astype(np.float32)  # column sums 10,20,...,80
    assert np.array_equal(np.rint(out[0].astype(np.float64) * 1e9), expect)
    np.savez_compressed(os.path.join(GOLD, "ref_dpu_toy_c_test.npz"), T=T, D=D, B=B, P=P, n_calls=1,
                        cols0=cols[0], idx0_0=idx[0], off0_0=off[0], out0_0=out)
    print("wrote toy_c_test")


def quantise():
    rng = np.random.default_rng(7)
    parts = [rng.uniform(-1e-3, 1e-3, size=4096), rng.uniform(-1, 1, size=4096),
             rng.uniform(-3, 3, size=2048),  # beyond int32: exercises the c_int32 wrap
             (np.arange(-64, 64) + 0.5) * 1e-9,  # rint ties
             np.array([0.0, -0.0, 1e-10, -1e-10, 2.147483647, -2.147483648, 2.1474836, 4.294967296])]
    w = np.concatenate(parts).astype(np.float32).reshape(-1, 8)
    # --- the reference's expression, dlrm_dpu_pytorch.py:318-320 ---
    tmp_emb = np.rint(w * (10**9)).astype(int)
    tmp_emb = np.transpose(tmp_emb)
    tmp_emb = np.ascontiguousarray(tmp_emb)
    # --- :329 (c_int32 * n)(*col): Python ints -> int32, low 32 bits kept ---
    cols = np.stack([np.frombuffer((ctypes.c_int32 * len(c))(*[int(v) for v in c]), dtype=np.int32)
                     for c in tmp_emb])
    np.savez_compressed(os.path.join(GOLD, "quantise.npz"), w=w, cols=cols)
    print("wrote quantise", w.shape)


def float_bag():
    sys.path.insert(0, "/root/reference/PIM-dlrm-new")
    import torch
    with contextlib.redirect_stdout(io.StringIO()):
        import dlrm_s_pytorch as ref
    np.random.seed(123)
    torch.manual_seed(123)
    ln_emb = np.array([200, 37, 500, 3])
    payload = {}
    for D in (16, 64, 128):
        net = ref.DLRM_Net.__new__(ref.DLRM_Net)
        torch.nn.Module.__init__(net)
        net.qr_flag = net.md_flag = False
        net.qr_threshold = net.md_threshold = 200
        with contextlib.redirect_stdout(io.StringIO()):
            emb_l, v_W_l = net.create_emb(D, ln_emb, None)  # dlrm_s_pytorch.py:295-302
        rng = np.random.default_rng(1000 + D)
        B = 48
        lS_i, lS_o = [], []
        for n in ln_emb:
            lens = rng.integers(0, 21, size=B)
            lens[0] = max(lens[0], 1)
            off = np.concatenate([[0], np.cumsum(lens)[:-1]])
            lS_o.append(torch.tensor(off, dtype=torch.int64))
            lS_i.append(torch.tensor(rng.integers(0, n, size=int(lens.sum())), dtype=torch.int64))
        ref.nbatches = 1 << 30  # module global read by the @emb_timer decorator (:113-127)
        with contextlib.redirect_stdout(io.StringIO()), torch.no_grad():
            ly = net.apply_emb(lS_o, lS_i, emb_l, v_W_l)  # dlrm_s_pytorch.py:427-498
        for k in range(len(ln_emb)):
            payload[f"w{D}_{k}"] = emb_l[k].weight.detach().numpy()
            payload[f"idx{D}_{k}"] = lS_i[k].numpy().astype(np.uint32)
            payload[f"off{D}_{k}"] = lS_o[k].numpy().astype(np.uint32)
            payload[f"out{D}_{k}"] = ly[k].detach().numpy()
    payload["dims"] = np.array([16, 64, 128])
    payload["n_tables"] = len(ln_emb)
    np.savez_compressed(os.path.join(GOLD, "float_bag.npz"), **payload)
    print("wrote float_bag")


if __name__ == "__main__":
    os.makedirs(GOLD, exist_ok=True)
    if sys.argv[1:] == ["float"]:
        float_bag()
        sys.exit(0)
    toy_c_test()
    dpu_case("toy_run_T4_D64", 4, 64, 64, 32, [500, 333, 64, 2], seed=1)          # run.sh toy shape (T capped by Q13)
    dpu_case("kaggle_dim16_T1", 1, 16, 64, 20, [1460], seed=2, kind="scaled")     # D=16, pooling <= 20
    dpu_case("rm2_dim128_T2", 2, 128, 16, 80, [700, 45], seed=3, kind="scaled")   # D=128, P=80
    dpu_case("maxbags_T1_B512", 1, 32, 512, 6, [97], seed=4)                      # B = 512 (Q3 limit)
    dpu_case("overflow_T1", 1, 64, 32, 40, [5], seed=5, kind="overflow")          # Q12 wrap-around
    dpu_case("single_row_T1", 1, 16, 2, 1, [1], seed=6)                           # smallest shape
    quantise()
    float_bag()
