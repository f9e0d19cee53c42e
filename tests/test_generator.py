"""the reference's uniform input generator (PIM-dlrm-new/dlrm_data_pytorch.py:1040-1080,
`round(ra.random(P) * (R - 1))`, duplicates kept, offsets 0, P, 2P, ...).  bench.reference_bag_indices restates
the per-bag expression; here it is held to the reference's own function for one seed when /root/reference is
present (this container), and its statistical difference from bench.make_batch's vectorised draw -- the two end
rows get half the probability -- is checked so that the deviation stated in bench.py stays true."""
import importlib.util
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

REF = "/root/reference/PIM-dlrm-new/dlrm_data_pytorch.py"


def _reference_module():
    sys.path.insert(0, os.path.dirname(REF))
    try:
        spec = importlib.util.spec_from_file_location("ref_dlrm_data_pytorch", REF)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        return mod
    finally:
        sys.path.remove(os.path.dirname(REF))


@pytest.mark.skipif(not os.path.exists(REF), reason="the reference tree is not on this machine")
def test_restatement_matches_the_reference_generator_for_one_seed():
    pytest.importorskip("torch")
    try:
        mod = _reference_module()
    except Exception as e:  # noqa: BLE001  (the reference imports optional packages at module level)
        pytest.skip(f"the reference's data module does not import here: {e!r}")
    ln_emb, n, P, m_den = [1000, 7, 123456], 9, 5, 4
    np.random.seed(123)
    mod.ra = np.random  # the module draws from a global `ra` its driver script seeds (dlrm_dpu_pytorch.py)
    if not hasattr(mod, "generate_dist_input_batch"):
        pytest.skip("generate_dist_input_batch not found")
    # the generator the DPU driver uses for --data-generation=random (uniform branch, :1061-1065)
    out = mod.generate_dist_input_batch(m_den, ln_emb, n, P, True, "uniform", 0, 1, -1, 1)
    offs, idxs = out[1], out[2]
    np.random.seed(123)
    ra = np.random
    ra.rand(n, m_den)  # the dense features are drawn first (dlrm_data_pytorch.py:940)
    for t, R in enumerate(ln_emb):
        want_idx, want_off, pos = [], [], 0
        for _ in range(n):
            g = bench.reference_bag_indices(R, P, ra)
            want_off.append(pos)
            want_idx += g.tolist()
            pos += len(g)
        assert idxs[t].tolist() == want_idx, t
        assert offs[t].tolist() == want_off, t
        assert offs[t].tolist() == [b * P for b in range(n)]  # fixed pooling: offsets 0, P, 2P, ...


def test_end_rows_are_half_as_likely_in_the_reference_expression():
    ra = np.random.RandomState(5)
    R, n = 11, 200000
    ref = np.concatenate([bench.reference_bag_indices(R, 1000, ra) for _ in range(n // 1000)])
    cnt = np.bincount(ref, minlength=R) / len(ref)
    inner = cnt[1:-1].mean()
    assert abs(cnt[0] / inner - 0.5) < 0.05 and abs(cnt[-1] / inner - 0.5) < 0.05
    assert ref.min() == 0 and ref.max() == R - 1
    idx, off = bench.make_batch(dict(T=1, R=R, B=2000, pool=(100, 100), dist="uniform"), seed=1)
    c2 = np.bincount(idx[0], minlength=R) / len(idx[0])
    assert abs(c2[0] / c2[1:-1].mean() - 1.0) < 0.05  # bench.make_batch: every row equally likely
    assert off[0].tolist() == [100 * b for b in range(2000)]
