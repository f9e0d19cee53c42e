"""The kernels divide by 1e9 without the division routine (emb_kernels.cuh dequant()); tests/check_dequant.c
proves on the CPU, for every int32, that the sequence gives the bits of the reference's
`(float)s / pow(10,9)` (upmem/include/emb_host.h:210)."""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))


def test_fast_dequantise_equals_the_reference_expression_for_every_int32(tmp_path):
    exe = str(tmp_path / "check_dequant")
    subprocess.run(["gcc", "-O2", "-mfma", "-fopenmp", "-ffp-contract=off", "-o", exe,
                    os.path.join(HERE, "check_dequant.c"), "-lm"], check=True)
    r = subprocess.run([exe], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "reciprocal 0x3089705f fast_mismatches 0 fp32_divide_mismatches 0" in r.stdout
