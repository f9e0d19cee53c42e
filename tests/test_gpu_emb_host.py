"""the C microbenchmark `emb_host` (csrc/load_generator.c, counterpart of
upmem/src/load_generator.c:67-137) run as a process on the GPU: plain C against the C ABI through dlsym,
populate_mram per column + lookup() with the reference's signatures, every result compared inside the binary with
the reference's expression `(float)sum / pow(10,9)` (upmem/include/emb_host.h:210)."""
import json
import os
import subprocess

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "pim-embedding-lookup_b200", "lib", "emb_host")


@pytest.mark.parametrize("shape", [
    dict(tables=9, cols=64, batches=64, pooling=32, rows=50000),    # the reference's toy preset (upmem/run.sh:93-101)
    dict(tables=26, cols=16, batches=512, pooling=8, rows=20000),   # kaggle_env shape (upmem/run.sh:40-45)
    dict(tables=32, cols=64, batches=64, pooling=120, rows=125000),  # RM2 as shipped (upmem/r.sh:6-10)
])
def test_c_microbenchmark_validates_on_the_gpu(shape, tmp_path):
    if not os.path.exists(EXE):
        pytest.fail(f"{EXE} is missing: __graft_entry__.build() makes it")
    args = [EXE]
    for k, v in shape.items():
        args += [f"--{k}", str(v)]
    args += ["--iters", "20"]
    r = subprocess.run(args, capture_output=True, text=True, timeout=300, cwd=str(tmp_path))
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    line = [ln for ln in r.stdout.splitlines() if ln.startswith("{")][-1]
    d = json.loads(line)
    assert d["bench"] == "emb_host" and d["validated"] is True and d["mismatches"] == 0
    assert d["tables"] == shape["tables"] and d["cols"] == shape["cols"] and d["batches"] == shape["batches"]
    assert d["lookup_ms_median"] > 0 and d["lookups_per_s"] > 0
    out = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(out):  # keep the line as evidence
        with open(os.path.join(out, "emb_host_lines.jsonl"), "a") as f:
            f.write(line + "\n")
