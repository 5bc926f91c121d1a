"""GPU: the reference's own shell tests (src/test/depth2x/test_depth2x*.sh, copied
verbatim under tests/golden/ref_fixtures/depth2x/) run UNMODIFIED against this
build's depth2x, on the reference's real Kinect / laser-scanner frames; every PNG
they produce is then compared pixel by pixel with the compiled reference
(oracle/_ref; the C restatement where that library is absent)."""
import os
import re
import shutil
import subprocess

import numpy as np
import pytest

from oracle_lib import Oracle, RefLib, ref_available

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))
FIX = os.path.join(HERE, "golden", "ref_fixtures")
DEPTH2X = os.path.join(os.path.dirname(HERE), "depth-conversions_b200", "bin", "depth2x")
SCRIPTS = ["test_depth2x", "test_depth2x_bearing", "test_depth2x_flexion",
           "test_depth2x_gaussian_curvature", "test_depth2x_mean_curvature",
           "test_depth2x_max_curve", "test_depth2x_range", "test_depth2x_scale"]

KINECT = dict(w=960, h=540, fx=float(np.float32(519.226)), fy=float(np.float32(522.23)),
              cx=float(np.float32(479.462)), cy=float(np.float32(272.737)))
LASER = dict(w=1799, h=397, tmin=float(np.float32(0.87)), tmax=float(np.float32(2.27)))

# output name -> (oracle feature id, clamp of the LAST command of the script that
# writes that name). Feature ids: oracle_lib F_*.
FEATS = {"flexion": 5, "range": 0, "gauss-curv": 6, "mean-curv": 7, "max-curve": 8}
DIRS = {"horizontal": 1, "vertical": 2, "diagonal": 3, "anti-diagonal": 4}


def read16(path):
    import cv2
    img = cv2.imread(path, cv2.IMREAD_UNCHANGED)
    assert img is not None and img.dtype == np.uint16 and img.ndim == 2, path
    return img


@pytest.fixture(scope="module")
def checker():
    return RefLib() if ref_available() else None


def expected(checker, oracle, name, workdir):
    """The image the reference writes under `name`, or None if the name is not one
    of the conversions (or cannot be checked without oracle/_ref)."""
    m = re.fullmatch(r"batch-bearing-(laserscan-)?(\d+)-([a-z-]+)\.png", name)
    if m:
        feat, laser, idx, lo, hi = DIRS[m.group(3)], m.group(1), m.group(2), -20.0, 20.0
        src = f"laserscan-{idx}-depth.png" if laser else f"data{idx}-depth.png"
        in_kind = 0
    else:
        m = re.fullmatch(r"batch-(flexion-from-range|flexion|range|gauss-curv|mean-curv|max-curve|scale)-"
                         r"(laserscan-)?(\d+)\.png", name)
        if not m:
            return None
        kind, laser, idx = m.group(1), m.group(2), m.group(3)
        src = f"laserscan-{idx}-depth.png" if laser else f"data{idx}-depth.png"
        in_kind, lo, hi = 0, -20.0, 20.0
        if kind == "scale":
            d = read16(os.path.join(workdir, src))
            scale, offset = (8.0, 0.0) if len(idx) == 1 else (1.0, 5000.0)
            return (checker or oracle).depth_scaling(d, scale, offset), True
        if kind == "flexion-from-range":
            src, in_kind, kind = f"batch-range-{idx}.png", 1, "flexion"
        # the third command of the curvature scripts rewrites indices 0 / 1 with bounds
        if kind == "gauss-curv" and not laser and len(idx) == 1:
            lo, hi = -1.0, 1.0
        if kind == "mean-curv" and not laser and len(idx) == 1:
            lo, hi = -10.0, 10.0
        feat = FEATS[kind]
    d = read16(os.path.join(workdir, src))
    exact = feat == 0  # range: IEEE-only arithmetic (the 16-bit flexion kernel: +-1 LSB)
    if laser:
        if checker is None:
            return None
        q = np.empty(d.shape, np.uint16)
        checker.lib.ref_cli_frame_equi(feat, d.ctypes.data, LASER["w"], LASER["h"], LASER["tmin"],
                                       LASER["tmax"], lo, hi, None, q.ctypes.data)
        return q, exact
    q = (checker or oracle).cli_frame(feat, in_kind, d, KINECT, lo, hi)[-1]
    return q, exact


@pytest.mark.parametrize("script", SCRIPTS)
def test_reference_shell_script(script, tmp_path, checker, oracle):
    assert os.path.exists(DEPTH2X), "depth2x is not built"
    work = tmp_path / "depth2x"
    shutil.copytree(os.path.join(FIX, "depth2x"), work)
    env = dict(os.environ, TERM="dumb")
    r = subprocess.run(["sh", f"{script}.sh", DEPTH2X, os.path.join(FIX, "log_helpers.sh")],
                       cwd=work, env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, f"{script}.sh failed:\n{r.stdout[-3000:]}\n{r.stderr[-3000:]}"
    assert "Test successful!" in r.stderr or script == "test_depth2x"
    produced = sorted(f for f in os.listdir(work) if f.startswith("batch-") and f.endswith(".png"))
    if script != "test_depth2x":
        assert produced, "the script wrote no image"
    checked = 0
    for name in produced:
        exp = expected(checker, oracle, name, str(work))
        if exp is None:
            continue
        want, exact = exp
        got = read16(os.path.join(work, name))
        assert got.shape == want.shape, name
        diff = np.abs(got.astype(np.int32) - want.astype(np.int32))
        tol = 0 if exact else 1
        assert diff.max() <= tol, f"{name}: off by {diff.max()} LSB at {(diff > tol).sum()} pixels"
        checked += 1
    if script != "test_depth2x":
        assert checked >= 5, f"only {checked} images of {produced} were compared"
    print(f"{script}: {len(produced)} images written, {checked} compared with the reference")
