"""GPU: the image-level cases of the reference's doctest suites
(src/unittest/conversion/test_conversion_*.cpp), ported line by line onto the
drop-in headers (depth-conversions_b200/host/tools/ref_unittests.cpp), against
the reference's own golden PNGs and thresholds."""
import os
import shutil
import subprocess

import pytest

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
EXE = os.path.join(os.path.dirname(HERE), "depth-conversions_b200", "bin", "ref_unittests")
FIX = os.path.join(HERE, "golden", "ref_fixtures", "unittest_conversion")


def test_reference_unit_test_cases(tmp_path):
    assert os.path.exists(EXE), "ref_unittests is not built"
    work = tmp_path / "conversion"
    shutil.copytree(FIX, work)
    r = subprocess.run([EXE, str(work)], capture_output=True, text=True, timeout=600)
    print(r.stdout)
    assert r.returncode == 0, r.stdout[-4000:] + r.stderr[-2000:]
    assert " 0 failed" in r.stdout
    assert r.stdout.count("[ ok ]") >= 18
    # the images the cases write, as the reference's do
    for name in ("test_vertical.png", "test_vertical_parallel.png", "test_flexion.png",
                 "test_mean_unclamped.png", "test_max_curve_double.png"):
        assert (work / name).exists(), name
