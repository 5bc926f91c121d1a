"""GPU: the median pre-filter against OpenCV itself:
preprocess::median_blur is cv::medianBlur (src/include/sens_loc/preprocess/
filter.h:143-157); cv2 (the same library, newer version) is in this image. The
median of an odd window with a replicated border is unique: bit exact."""
import cv2
import numpy as np
import pytest

from oracle_lib import F_FLEXION, kinect_scaled
from parity import assert_u16_parity

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("k", [3, 5])
@pytest.mark.parametrize("w,h", [(960, 540), (61, 37), (5, 3), (2, 7), (1, 1)])
def test_median_u16_matches_opencv(gpu_ctx, oracle, w, h, k):
    frames = np.stack([oracle.synth_depth(0xD2C000A0 + k, f, w, h) for f in range(3)])
    rs = np.random.RandomState(w * h + k)
    frames[2] = rs.randint(0, 65536, size=(h, w)).astype(np.uint16)  # full value range
    gpu_ctx.set_frame_size(w, h)
    got = gpu_ctx.median(frames, k)
    for i in range(3):
        want = cv2.medianBlur(frames[i], k)
        assert np.array_equal(got[i], want), (w, h, k, i, int(np.sum(got[i] != want)))


@pytest.mark.parametrize("k", [3, 5])
def test_median_f32_matches_opencv(gpu_ctx, k):
    w, h = 203, 77
    rs = np.random.RandomState(k)
    img = (rs.rand(h, w) * 5000).astype(np.float32)
    img[rs.rand(h, w) < 0.3] = 0
    gpu_ctx.set_frame_size(w, h)
    got = gpu_ctx.median(img, k)
    assert np.array_equal(got, cv2.medianBlur(img, k))


def test_fused_prefilter_equals_filter_then_convert(pkg, gpu_ctx, oracle):
    """dc_job.median_ksize: what `depth_filter median-blur` + `depth2x flexion`
    compute with a PNG in between (transformation_step.py:80-96)."""
    w, h = 320, 200
    k = kinect_scaled(w, h)
    frames = np.stack([oracle.synth_depth(0xD2C000A9, f, w, h) for f in range(70)])
    gpu_ctx.set_pinhole(w, h, k["fx"], k["fy"], k["cx"], k["cy"])
    fused = gpu_ctx.convert(pkg.DC_FLEXION, frames, pkg.DC_IN_U16_DEPTH, pkg.DC_OUT_U16, median=5)[0]
    for i in (0, 33, 69):
        filtered = cv2.medianBlur(frames[i], 5)
        want = oracle.cli_frame(F_FLEXION, 0, filtered, k)[-1]
        assert_u16_parity("flexion", fused[i], want)
    with pytest.raises(pkg.DcError):
        gpu_ctx.convert(pkg.DC_FLEXION, frames[:1], pkg.DC_IN_U16_DEPTH, pkg.DC_OUT_U16, median=7)
    with pytest.raises(pkg.DcError):
        gpu_ctx.convert(pkg.DC_FLEXION, frames[:1].astype(np.float32), pkg.DC_IN_F32_RANGE,
                        pkg.DC_OUT_F32, median=3)


def test_depth2x_median_flag_and_cpp_header(tmp_path, oracle):
    """depth2x --median-blur 5 flexion == depth_filter median-blur -d 5 | depth2x flexion."""
    import os
    import subprocess
    here = os.path.dirname(os.path.abspath(__file__))
    depth2x = os.path.join(os.path.dirname(here), "depth-conversions_b200", "bin", "depth2x")
    w, h = 320, 200
    k = kinect_scaled(w, h)
    (tmp_path / "k.txt").write_text(
        f"{w} {h}\n{k['fx']!r} 0.0 {k['cx']!r}\n0.0 {k['fy']!r} {k['cy']!r}\n0.0 0.0 1.0\n")
    f = np.float32
    kp = dict(w=w, h=h, fx=float(f(repr(k["fx"]))), fy=float(f(repr(k["fy"]))),
              cx=float(f(repr(k["cx"]))), cy=float(f(repr(k["cy"]))))
    frames = [oracle.synth_depth(0xD2C000AA, i, w, h) for i in range(3)]
    for i, fr in enumerate(frames):
        cv2.imwrite(str(tmp_path / f"d{i}.png"), fr)
    r = subprocess.run([depth2x, "-c", "k.txt", "-i", "d{}.png", "-s", "0", "-e", "2", "--median-blur", "5",
                        "flexion", "-o", "f{}.png"], cwd=tmp_path, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    for i, fr in enumerate(frames):
        got = cv2.imread(str(tmp_path / f"f{i}.png"), cv2.IMREAD_UNCHANGED)
        want = oracle.cli_frame(F_FLEXION, 0, cv2.medianBlur(fr, 5), kp)[-1]
        assert_u16_parity("flexion", got, want)
    r = subprocess.run([depth2x, "-c", "k.txt", "-i", "d{}.png", "-s", "0", "-e", "0", "--median-blur", "4",
                        "flexion", "-o", "g{}.png"], cwd=tmp_path, capture_output=True, text=True, timeout=300)
    assert r.returncode == 1 and "--median-blur" in r.stderr
