"""the reference's own DLRM driver, UNMODIFIED, on stock torch + the kwarg shim + emblib.so.

`PIM-dlrm-new/dlrm_dpu_pytorch.py` is executed where it lies (never copied): cwd holds a symlink `emblib.so` (the
script does `CDLL("./emblib.so")`, :117-125), `dputypes` resolves to this package's mirror, and
`embedding_bag.install()` gives stock `nn.EmbeddingBag.forward` the fork's seven kwargs (:559-570).

The reference tree exists only in the build container, which has no GPU; the GPU box has no reference tree.  So here
the run must get through argument parsing, model construction and `export_emb` (:306-332) up to its first library call
and die THERE, loudly, because this library has no CPU path -- which shows that nothing before the device boundary
needs the forked PyTorch or the UPMEM SDK.  What follows the boundary (populate_mram per column, the pointer arrays,
`lookup()` through the patched forward, the dummy return) is replayed call for call on the GPU by
tests/test_gpu_shim.py.  Should a box ever have both a GPU and the reference tree, the same test demands a clean run."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference/PIM-dlrm-new"
DRIVER = os.path.join(REF, "dlrm_dpu_pytorch.py")

RUNNER = """
import os, runpy, sys
sys.path.insert(0, {root!r})
sys.path.insert(0, {pkg!r})      # `import dputypes` -> the package's mirror of upmem/dputypes.py
sys.path.insert(0, {ref!r})      # dlrm_data_pytorch, extend_distributed, tricks, optim ...
import pim_embedding_lookup_b200  # noqa: F401
from pim_embedding_lookup_b200 import embedding_bag
embedding_bag.install()
sys.argv = ["dlrm_dpu_pytorch.py", "--arch-embedding-size=100-100", "--arch-sparse-feature-size=16",
            "--arch-mlp-bot=13-16", "--arch-mlp-top=48-8-1", "--data-generation=random", "--mini-batch-size=4",
            "--num-batches=2", "--num-indices-per-lookup=2", "--num-indices-per-lookup-fixed=True",
            "--inference-only"]
runpy.run_path({driver!r}, run_name="__main__")
print("DRIVER FINISHED")
"""


@pytest.mark.skipif(not os.path.exists(DRIVER), reason="the reference tree is not on this machine")
def test_reference_dlrm_driver_runs_unmodified_up_to_the_device_boundary(tmp_path):
    import torch

    import __graft_entry__ as ge
    from pim_embedding_lookup_b200 import emblib
    if not os.path.exists(emblib.LIB_PATH):
        ge.build()
    os.symlink(emblib.LIB_PATH, tmp_path / "emblib.so")
    (tmp_path / "run.py").write_text(RUNNER.format(root=ROOT, pkg=os.path.join(ROOT, "pim-embedding-lookup_b200"),
                                                   ref=REF, driver=DRIVER))
    env = dict(os.environ, NR_TABLES="2", NR_COLS="16", MAX_NR_BATCHES="4", MAX_INDICES_PER_BATCH="2")
    r = subprocess.run([sys.executable, "run.py"], cwd=tmp_path, env=env, capture_output=True, text=True, timeout=600)
    out = r.stdout + r.stderr
    assert "Using DPU..." in out, out[-2000:]  # the use_dpu branch of the driver (dlrm_dpu_pytorch.py)
    if torch.cuda.is_available():
        assert r.returncode == 0 and "DRIVER FINISHED" in r.stdout, out[-3000:]
    else:
        # export_emb's first populate_mram(): no device, no CPU path -> print and exit(1), as DPU_ASSERT does
        assert r.returncode != 0 and "DRIVER FINISHED" not in r.stdout
        assert "emblib(B200): populate_mram failed" in r.stderr and "no CUDA device" in r.stderr, out[-2000:]
        assert "Traceback" not in r.stderr, out[-3000:]  # it is the library that stops the run, not a Python error
