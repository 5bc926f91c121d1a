"""tests/cuda/mathcheck.cu on the GPU: the guard-free division / square root / normalisation
helpers of csrc/dc_fastmath.cuh and csrc/dc_pair.cuh against the IEEE intrinsics
(__fsqrt_rn exhaustively over the admissible range, __fdiv_rn on 8.6e9 operand pairs per
mode) and the acos polynomial against double precision (all floats in [-1, 1])."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "tests", "cuda", "bin", "mathcheck")


@pytest.mark.gpu
def test_mathcheck_exact_division_sqrt_and_acos_error():
    assert os.path.exists(BIN), "tests/cuda/bin/mathcheck missing: __graft_entry__.build() makes it"
    r = subprocess.run([BIN], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    lines = r.stdout.strip().splitlines()
    assert lines[-1] == "MATHCHECK OK", r.stdout
    for l in lines:
        if "mismatches=" in l:
            assert l.rstrip().endswith("mismatches=0"), l
        if "max_rel=" in l:
            assert float(l.split("max_rel=")[1]) < 1.2e-6, l
