"""The oracle (oracle/emb_oracle.c) pinned against the reference's own outputs.

tests/golden/*.npz were produced by oracle/make_golden.py from the reference itself:
ref_dpu_* by its host library + DPU program compiled verbatim against the UPMEM-SDK
stand-in; quantise by its export expression; float_bag by its DLRM_Net.apply_emb.
Bit-exact for the integer/fixed-point path; the float path is bit-exact here too because
the restatement keeps torch's summation order (tolerance stated where it is not).
"""
import glob
import os

import numpy as np
import pytest

import emb_oracle as eo


def _bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


DPU_FILES = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "ref_dpu_*.npz")))


def test_golden_files_present():
    names = {os.path.basename(f) for f in DPU_FILES}
    assert "ref_dpu_toy_c_test.npz" in names and len(names) >= 6


@pytest.mark.parametrize("path", DPU_FILES, ids=[os.path.basename(f)[8:-4] for f in DPU_FILES])
def test_lookup_matches_reference_dpu_path(path):
    z = np.load(path)
    T, D, B, P = (int(z[k]) for k in "TDBP")
    cols = [z[f"cols{t}"] for t in range(T)]
    rows = [np.ascontiguousarray(c.T) for c in cols]
    for c in range(int(z["n_calls"])):
        idx = [z[f"idx{c}_{t}"] for t in range(T)]
        off = [z[f"off{c}_{t}"] for t in range(T)]
        want = [z[f"out{c}_{t}"] for t in range(T)]
        got_cols = eo.lookup_columns(cols, idx, off)           # DPU layout, emb_dpu_lookup.c statement
        got_rows, sums = eo.lookup_rows(rows, idx, off, want_sums=True)  # load_generator.c statement
        for t in range(T):
            assert want[t].shape == (B, D)
            assert np.array_equal(_bits(got_cols[t]), _bits(want[t]))
            assert np.array_equal(_bits(got_rows[t]), _bits(want[t]))
            assert np.array_equal(_bits(eo.dequant(sums[t])), _bits(want[t]))
            # per-DPU statement on a few columns
            for j in (0, D - 1):
                s = eo.pool_column(cols[t][j], idx[t], off[t])
                assert np.array_equal(s, sums[t][:, j])


def test_toy_known_answer():
    # upmem/c_test.py:40,56 -> per-column sums 10*(j+1), dequantised by 1e9
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "ref_dpu_toy_c_test.npz"))
    out = eo.lookup_columns([z["cols0"]], [z["idx0_0"]], [z["off0_0"]])[0]
    expect = eo.dequant((np.arange(1, 9) * 10).astype(np.int32))
    assert np.array_equal(_bits(out), _bits(np.tile(expect, (32, 1))))
    assert np.array_equal(_bits(out), _bits(z["out0_0"]))


def test_quantise_matches_reference_expression(golden_dir):
    z = np.load(os.path.join(golden_dir, "quantise.npz"))
    w, cols = z["w"], z["cols"]
    assert np.array_equal(eo.quantise_columns(w), cols)
    assert np.array_equal(eo.quantise_rows(w), cols.T)


def test_dequant_is_ieee_fp32_divide():
    rng = np.random.default_rng(0)
    s = np.concatenate([rng.integers(-2**31, 2**31, size=200000), [0, 1, -1, 2**31 - 1, -2**31]]).astype(np.int32)
    a = eo.dequant(s)
    b = s.astype(np.float32) / np.float32(1e9)  # what the CUDA epilogue computes (__fdiv_rn)
    assert np.array_equal(_bits(a), _bits(b))
    c = np.array([eo.lib().emb_oracle_dequant(int(v)) for v in s[:2000]], dtype=np.float32)
    assert np.array_equal(_bits(c), _bits(a[:2000]))


def test_float_path_matches_reference_apply_emb(golden_dir):
    z = np.load(os.path.join(golden_dir, "float_bag.npz"))
    for D in z["dims"]:
        for k in range(int(z["n_tables"])):
            got = eo.bag_sum_f32(z[f"w{D}_{k}"], z[f"idx{D}_{k}"], z[f"off{D}_{k}"])
            want = z[f"out{D}_{k}"]
            # north_star tolerance for float tables: <= 1e-5 relative (fp32 reassociation)
            np.testing.assert_allclose(got, want, rtol=1e-5, atol=1e-7)


def test_threads_do_not_change_results():
    rng = np.random.default_rng(3)
    tab = rng.integers(-2**31, 2**31, size=(1000, 16), dtype=np.int64).astype(np.int32)
    idx = rng.integers(0, 1000, size=50000).astype(np.uint32)
    off = np.sort(rng.integers(0, 50000, size=4096)).astype(np.uint32)
    eo.set_num_threads(1)
    a = eo.lookup_rows([tab], [idx], [off])[0]
    eo.set_num_threads(4)
    b = eo.lookup_rows([tab], [idx], [off])[0]
    eo.set_num_threads(0)
    assert np.array_equal(_bits(a), _bits(b))
