"""The drop-in boundary without a GPU: emblib.so loads, exports every symbol that
include/emb_b200.h declares, its structs have the reference's layout, and -- there being no
CPU path -- the product fails loudly when no CUDA device is visible."""
import ctypes as C
import os
import re
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "emb_b200.h")


@pytest.fixture(scope="module")
def lib():
    import __graft_entry__ as ge
    from pim_embedding_lookup_b200 import emblib
    if not os.path.exists(emblib.LIB_PATH):
        ge.build()
    return emblib.load()


def declared_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    src = re.sub(r"^\s*#.*$", "", src, flags=re.M)
    names = re.findall(r"\b([a-z_][a-z0-9_]*)\s*\([^;{}]*\)\s*;", src)
    return sorted(set(n for n in names if n not in ("__attribute__", "packed")))


def test_header_declares_the_reference_entry_points():
    names = declared_functions()
    assert "populate_mram" in names and "lookup" in names  # emb_host.h:136,234
    assert len(names) >= 30


def test_every_declared_symbol_is_exported(lib):
    missing = [n for n in declared_functions() if not hasattr(lib, n)]
    assert not missing, missing


def test_soname_matches_reference():
    from pim_embedding_lookup_b200 import emblib
    out = subprocess.run(["readelf", "-d", emblib.LIB_PATH], capture_output=True, text=True).stdout
    assert "emb_host" in out  # upmem/Makefile:4,54: soname emb_host


def test_library_is_sm100a_only():
    from pim_embedding_lookup_b200 import emblib
    out = subprocess.run(["cuobjdump", "-lelf", emblib.LIB_PATH], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_(\d+a?)", out))
    assert archs == {"100a"}, out


def test_struct_layouts():
    from pim_embedding_lookup_b200 import dputypes as dt
    assert C.sizeof(dt.DpuRuntimeTotals) == 6 * 8          # emb_host.h:41-48
    assert C.sizeof(dt.DpuTimespec) == 2 * C.sizeof(C.c_long)  # emb_host.h:54-57
    assert dt.DpuTimespec.tv_nsec.offset == 0                # nsec first
    assert C.sizeof(dt.DpuRuntimeInterval) == 2 * C.sizeof(dt.DpuTimespec)
    assert C.sizeof(dt.DpuRuntimeGroup) == 16
    assert C.sizeof(dt.QueryLen) == 8                        # emb_types.h:21-24, packed
    assert C.sizeof(dt.EmbShape) == 16
    assert C.sizeof(dt.EmbDevQuery) == 16 + 3 * 8 + 8 + 8 * 8 + 16
    assert C.sizeof(dt.EmbReduceJob) == 8 * 8 + 8 + 4 * 4 + 8
    assert C.sizeof(dt.EmbTimings) == 4 * 8 + 4 * 8
    assert C.sizeof(dt.EmbIpcHandle) == 64


def test_struct_sizes_agree_with_the_c_compiler(tmp_path):
    src = tmp_path / "sz.c"
    src.write_text('#include "emb_b200.h"\n#include <stdio.h>\nint main(void){printf("%zu %zu %zu %zu %zu %zu %zu\\n",'
                   "sizeof(dpu_runtime_totals),sizeof(dpu_runtime_group),sizeof(struct query_len),"
                   "sizeof(emb_dev_query),sizeof(emb_timings),sizeof(emb_shape),sizeof(emb_reduce_job));return 0;}\n")
    exe = tmp_path / "sz"
    subprocess.run(["gcc", "-std=gnu11", "-I", os.path.join(ROOT, "include"), "-o", str(exe), str(src)], check=True)
    got = [int(x) for x in subprocess.run([str(exe)], capture_output=True, text=True).stdout.split()]
    from pim_embedding_lookup_b200 import dputypes as dt
    want = [C.sizeof(x) for x in (dt.DpuRuntimeTotals, dt.DpuRuntimeGroup, dt.QueryLen, dt.EmbDevQuery,
                                  dt.EmbTimings, dt.EmbShape, dt.EmbReduceJob)]
    assert got == want


def test_no_cpu_fallback_without_a_device(lib):
    """No GPU here: configure must report a CUDA error, not quietly run on the host."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is visible")
    from pim_embedding_lookup_b200 import emblib
    with pytest.raises(emblib.EmbError, match="no CUDA device|CUDA"):
        emblib.configure(1, 16, 4, 2)


def test_reference_entry_points_abort_without_a_device():
    """populate_mram()/lookup() have no error return (emb_host.h: DPU_ASSERT aborts); same here."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is visible")
    code = ("import numpy as np, pim_embedding_lookup_b200 as p\n"
            "import os\nos.environ.update(NR_TABLES='1', NR_COLS='4', MAX_NR_BATCHES='2', MAX_INDICES_PER_BATCH='2')\n"
            "p.emblib.populate_mram(0, np.zeros(8, dtype=np.int32), 0)\nprint('survived')\n")
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, cwd=ROOT)
    assert r.returncode != 0 and "survived" not in r.stdout
    assert "emblib(B200): populate_mram failed" in r.stderr


def test_product_package_does_not_import_the_oracle():
    pkg = os.path.join(ROOT, "pim-embedding-lookup_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".c", ".h")):
                text = open(os.path.join(dirpath, f), errors="replace").read()
                assert "emb_oracle" not in text and "oracle/" not in text, os.path.join(dirpath, f)


def test_runtime_interval_csv_is_what_the_reference_plots(tmp_path):
    """`DPU,Start,Stop` rows (upmem/dputypes.py:87-98) that graph/runtime_intervals/main.py reads with pandas:
    column 0 the unit, columns 1 and 2 start / stop in milliseconds."""
    import types

    import pandas as pd

    from pim_embedding_lookup_b200 import dputypes as dt
    grp = (dt.DpuRuntimeGroup * 2)()
    keep = []
    for n, g in enumerate(grp):
        g.in_use = 2
        g.length = 2
        iv = (dt.DpuRuntimeInterval * 2)()
        keep.append(iv)
        for i in range(2):
            iv[i].start.tv_sec, iv[i].start.tv_nsec = 1 + n, 500_000 * i
            iv[i].stop.tv_sec, iv[i].stop.tv_nsec = 1 + n, 500_000 * i + 250_000
        g.intervals = iv
    cfg = types.SimpleNamespace(runtimes=True, runtime_file=str(tmp_path / "runtimes.csv"))
    dt.write_results(cfg, grp)
    data = pd.read_csv(cfg.runtime_file)
    assert list(data.columns) == ["DPU", "Start", "Stop"] and len(data) == 4
    assert list(data["DPU"]) == [0, 0, 1, 1]
    assert data["Start"][1] == pytest.approx(1000.5) and data["Stop"][1] == pytest.approx(1000.75)
    assert (data["Stop"] > data["Start"]).all()
    dt.write_results(types.SimpleNamespace(runtimes=False, runtime_file=str(tmp_path / "none.csv")), grp)
    assert not os.path.exists(tmp_path / "none.csv")
