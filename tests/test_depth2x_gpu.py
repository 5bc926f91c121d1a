"""GPU: the depth2x tool and the C++ drop-in API end to end (PNG in, PNG out),
pixel-compared with the oracle. The scenarios are the reference's shell tests
src/test/depth2x/test_depth2x_{flexion,bearing,range,scale,max_curve,
gaussian_curvature,mean_curvature}.sh (which check exit codes and file existence
only) on synthetic frames."""
import os
import subprocess

import cv2
import numpy as np
import pytest

from oracle_lib import (F_BEARING_H, F_FLEXION, F_GAUSS, F_MAX_CURVE, F_MEAN, F_RANGE, kinect_scaled)
from parity import assert_float_parity, assert_u16_parity

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
BIN = os.path.join(os.path.dirname(HERE), "depth-conversions_b200", "bin")
DEPTH2X = os.path.join(BIN, "depth2x")
SELFTEST = os.path.join(BIN, "host_selftest")
W, H = 960, 540


def run(args, cwd):
    return subprocess.run([DEPTH2X] + args, cwd=cwd, capture_output=True, text=True, timeout=600)


def intrinsic_text(k):
    return (f"{k['w']} {k['h']}\n{k['fx']!r} 0.000000 {k['cx']!r}\n0.000000 {k['fy']!r} {k['cy']!r}\n"
            "0.000000 0.000000 1.000000\n")


def parsed(k):
    """The intrinsic as depth2x sees it: the text parsed to float."""
    f = np.float32
    return dict(w=k["w"], h=k["h"], fx=float(f(repr(k["fx"]))), fy=float(f(repr(k["fy"]))),
                cx=float(f(repr(k["cx"]))), cy=float(f(repr(k["cy"]))))


@pytest.fixture(scope="module")
def scene(tmp_path_factory, oracle):
    """data0..data4 + data002..data006 (006 is 480x270: the size-mismatch case of
    test_depth2x_range.sh:78-93)."""
    d = tmp_path_factory.mktemp("depth2x")
    k = kinect_scaled(W, H)
    (d / "kinect_intrinsic.txt").write_text(intrinsic_text(k))
    frames = {}
    for i in range(5):
        frames[i] = oracle.synth_depth(0xD2C00001, i, W, H)
        cv2.imwrite(str(d / f"data{i}-depth.png"), frames[i])
    for i in range(2, 6):
        cv2.imwrite(str(d / f"data{i:03d}-depth.png"), frames.get(i, oracle.synth_depth(0xD2C00001, i, W, H)))
    frames[5] = oracle.synth_depth(0xD2C00001, 5, W, H)
    cv2.imwrite(str(d / "data006-depth.png"), oracle.synth_depth(0xD2C00001, 6, 480, 270))
    return str(d), parsed(k), frames


def read(cwd, name):
    img = cv2.imread(os.path.join(cwd, name), cv2.IMREAD_UNCHANGED)
    assert img is not None and img.dtype == np.uint16, name
    return img


def test_flexion_patterns_and_missing_index(scene, oracle):
    cwd, k, frames = scene
    base = ["-c", "kinect_intrinsic.txt", "--type", "pinhole-depth"]
    # test_depth2x_flexion.sh: "{}" pattern, flags before the sub-command
    r = run(base + ["-i", "data{}-depth.png", "-s", "0", "-e", "1", "flexion", "-o", "flexion-{}.png"], cwd)
    assert r.returncode == 0, r.stderr
    assert "Processing 2 images took" in r.stderr
    for i in (0, 1):
        assert_u16_parity("flexion", read(cwd, f"flexion-{i}.png"),
                          oracle.cli_frame(F_FLEXION, 0, frames[i], k)[2])
    # "{:03d}" pattern, reversed range (parallel_processing.h:31-34), sub-command first
    r = run(["flexion", "-o", "batch-flexion-{:03d}.png"] + base +
            ["-i", "data{:03d}-depth.png", "--start", "4", "--end", "2"], cwd)
    assert r.returncode == 0, r.stderr
    for i in (2, 3, 4):
        assert_u16_parity("flexion", read(cwd, f"batch-flexion-{i:03d}.png"),
                          oracle.cli_frame(F_FLEXION, 0, frames[i], k)[2])
    # test_depth2x.sh:62-65: indices without files fail, the others are written
    r = run(base + ["-i", "data{}-depth.png", "-s", "3", "-e", "7", "flexion", "-o", "part-{}.png"], cwd)
    assert r.returncode == 1
    for i in (5, 6, 7):
        assert f'Could not process index "{i}"' in r.stderr
    assert "Encountered 3 problematic files" in r.stderr
    assert os.path.exists(os.path.join(cwd, "part-3.png")) and not os.path.exists(os.path.join(cwd, "part-5.png"))


def test_bearing_all_directions(scene, oracle):
    cwd, k, frames = scene
    r = run(["bearing", "-c", "kinect_intrinsic.txt", "-i", "data{}-depth.png", "-s", "0", "-e", "2",
             "--horizontal", "bh-{}.png", "--vertical", "bv-{}.png", "--diagonal", "bd-{}.png",
             "--anti-diagonal", "ba-{}.png", "--batch-size", "2"], cwd)
    assert r.returncode == 0, r.stderr
    for i in range(3):
        for d, n in enumerate(("bh", "bv", "bd", "ba")):
            assert_u16_parity("bearing", read(cwd, f"{n}-{i}.png"),
                              oracle.cli_frame(F_BEARING_H + d, 0, frames[i], k)[2])
    # a subset of the directions writes only those (converter_bearing.h.inl:13-26)
    r = run(["bearing", "-c", "kinect_intrinsic.txt", "-i", "data{}-depth.png", "-s", "0", "-e", "0",
             "--diagonal", "only-d-{}.png"], cwd)
    assert r.returncode == 0 and os.path.exists(os.path.join(cwd, "only-d-0.png"))


def test_range_roundtrip_and_size_mismatch(scene, oracle):
    cwd, k, frames = scene
    r = run(["-c", "kinect_intrinsic.txt", "-i", "data{:03d}-depth.png", "--start", "2", "--end", "4",
             "range", "--output", "batch-range-{:03d}.png"], cwd)
    assert r.returncode == 0, r.stderr
    for i in (2, 3, 4):
        assert_u16_parity("range", read(cwd, f"batch-range-{i:03d}.png"),
                          oracle.cli_frame(F_RANGE, 0, frames[i], k)[2])
    # test_depth2x_range.sh:58-76: range images as --type pinhole-range input
    r = run(["-c", "kinect_intrinsic.txt", "-i", "batch-range-{:03d}.png", "--start", "2", "--end", "4",
             "--type", "pinhole-range", "flexion", "--output", "ffr-{:03d}.png"], cwd)
    assert r.returncode == 0, r.stderr
    rng16 = read(cwd, "batch-range-003.png")
    assert_u16_parity("flexion", read(cwd, "ffr-003.png"), oracle.cli_frame(F_FLEXION, 1, rng16, k)[2])
    # :78-93: data006 is 480x270: index 5 works, index 6 fails, exit code 1
    r = run(["-c", "kinect_intrinsic.txt", "-i", "data{:03d}-depth.png", "--start", "5", "--end", "6",
             "range", "--output", "mm-range-{:03d}.png"], cwd)
    assert r.returncode == 1 and 'Could not process index "6"' in r.stderr
    assert os.path.exists(os.path.join(cwd, "mm-range-005.png"))
    assert not os.path.exists(os.path.join(cwd, "mm-range-006.png"))


def test_curvature_max_curve_scale(scene, oracle):
    cwd, k, frames = scene
    common = ["-c", "kinect_intrinsic.txt", "-i", "data{}-depth.png", "-s", "0", "-e", "1"]
    assert run(common + ["gauss-curvature", "-o", "g-{}.png", "-l", "-1.0", "-u", "1.0"], cwd).returncode == 0
    assert run(common + ["mean-curvature", "-o", "m-{}.png"], cwd).returncode == 0
    assert run(common + ["max-curve", "-o", "mc-{}.png"], cwd).returncode == 0
    for i in (0, 1):
        assert_u16_parity("gauss", read(cwd, f"g-{i}.png"),
                          oracle.cli_frame(F_GAUSS, 0, frames[i], k, -1.0, 1.0)[2])
        assert_u16_parity("mean", read(cwd, f"m-{i}.png"), oracle.cli_frame(F_MEAN, 0, frames[i], k)[2])
        assert_u16_parity("max_curve", read(cwd, f"mc-{i}.png"),
                          oracle.cli_frame(F_MAX_CURVE, 0, frames[i], k)[2])
    # scale needs no intrinsic and takes any frame size (test_depth2x_scale.sh)
    r = run(["scale", "-i", "data{:03d}-depth.png", "-s", "5", "-e", "6", "-o", "sc-{:03d}.png",
             "--factor", "8.0", "--delta", "3"], cwd)
    assert r.returncode == 0, r.stderr
    for i in (5, 6):
        src = read(cwd, f"data{i:03d}-depth.png")
        assert np.array_equal(read(cwd, f"sc-{i:03d}.png"), oracle.depth_scaling(src, 8.0, 3.0))


def test_equirectangular_model(tmp_path, oracle, pkg, gpu_ctx):
    """-m equirectangular (test_depth2x_range.sh:95-111 and friends): the range
    conversion is the identity; features are checked against the C ABI."""
    w, h = 360, 90
    (tmp_path / "laser_intrinsic.txt").write_text(f"{w} {h}\n0.87 2.27\n")
    frame = oracle.synth_depth(0xD2C00007, 0, w, h)
    cv2.imwrite(str(tmp_path / "laserscan-0-depth.png"), frame)
    base = ["-m", "equirectangular", "-c", "laser_intrinsic.txt", "-i", "laserscan-{}-depth.png",
            "-s", "0", "-e", "0"]
    assert run(base + ["range", "--output", "r-{}.png"], str(tmp_path)).returncode == 0
    assert np.array_equal(read(str(tmp_path), "r-0.png"), frame)
    assert run(base + ["flexion", "--output", "f-{}.png"], str(tmp_path)).returncode == 0
    gpu_ctx.set_equirectangular(w, h, float(np.float32(0.87)), float(np.float32(2.27)))
    want = gpu_ctx.convert(pkg.DC_FLEXION, frame, pkg.DC_IN_U16_DEPTH, pkg.DC_OUT_U16)[0]
    assert np.array_equal(read(str(tmp_path), "f-0.png"), want)


def test_cpp_library_api(tmp_path, oracle):
    """sens_loc::conversion::* called from C++ (host/tools/host_selftest.cpp)."""
    w, h = 160, 96
    k = kinect_scaled(w, h)
    (tmp_path / "k.txt").write_text(intrinsic_text(k))
    k = parsed(k)
    depth = oracle.synth_depth(0xD2C00008, 0, w, h)
    cv2.imwrite(str(tmp_path / "d.png"), depth)
    r = subprocess.run([SELFTEST, "api", str(tmp_path / "k.txt"), str(tmp_path / "d.png"), str(tmp_path)],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr

    def f32(name):
        return np.fromfile(str(tmp_path / f"{name}.f32"), np.float32).reshape(h, w)

    rng = oracle.laserscan(depth, k)
    assert_float_parity("range", f32("range"), rng)
    names = {"bearing_h": F_BEARING_H, "bearing_v": F_BEARING_H + 1, "bearing_d": F_BEARING_H + 2,
             "bearing_a": F_BEARING_H + 3, "flexion": F_FLEXION, "gauss": F_GAUSS, "mean": F_MEAN,
             "max_curve": F_MAX_CURVE}
    for n, f in names.items():
        assert_float_parity("bearing" if n.startswith("bearing") else n, f32(n), oracle.feature(f, rng, k))
    assert np.array_equal(f32("bearing_h_fused"), f32("bearing_h"))
    assert np.array_equal(f32("bearing_d_fused"), f32("bearing_d"))
    # the stand-alone quantisers on the GPU's own float images
    for n, f in (("range", F_RANGE), ("bearing_h", F_BEARING_H), ("flexion", F_FLEXION), ("gauss", F_GAUSS),
                 ("mean", F_MEAN), ("max_curve", F_MAX_CURVE)):
        got = cv2.imread(str(tmp_path / f"{n}.png"), cv2.IMREAD_UNCHANGED)
        assert_u16_parity(n, got, oracle.quantise(f, f32(n), rng), exact=False)
    assert np.array_equal(cv2.imread(str(tmp_path / "scaled.png"), cv2.IMREAD_UNCHANGED),
                          oracle.depth_scaling(depth, 8.0, 3.0))
    # preprocess::median_blur (preprocess/filter.h:143-157) == cv::medianBlur
    assert np.array_equal(cv2.imread(str(tmp_path / "median5.png"), cv2.IMREAD_UNCHANGED),
                          cv2.medianBlur(depth, 5))
    assert np.array_equal(f32("median3"), cv2.medianBlur(f32("range"), 3))
    # preprocess::gaussian_blur / bilateral_filter (preprocess/filter.h:49-82, 125-136)
    r32 = f32("range")
    span = float(r32.max() - r32.min())
    want = cv2.GaussianBlur(r32, (5, 5), 1.5)
    assert np.all(np.abs(f32("gaussblur") - want) <= 1e-6 * np.maximum(np.abs(want), 1.0))
    got = cv2.imread(str(tmp_path / "gaussblur.png"), cv2.IMREAD_UNCHANGED).astype(np.int32)
    assert np.abs(got - cv2.GaussianBlur(depth, (3, 3), 0).astype(np.int32)).max() <= 1
    # an int proximity is the diameter, sigma_space 0 -> 1 (OpenCV 4.3.0); a double is sigma_space
    assert np.abs(f32("bilateral_d") - cv2.bilateralFilter(r32, 5, 30.0, 1.0)).max() <= 2e-5 * span
    assert np.abs(f32("bilateral_s") - cv2.bilateralFilter(r32, 0, 30.0, 2.0)).max() <= 2e-5 * span
    got = cv2.imread(str(tmp_path / "bilateral.png"), cv2.IMREAD_UNCHANGED).astype(np.int32)
    want = np.clip(np.rint(cv2.bilateralFilter(depth.astype(np.float32), 0, 30.0, 2.0)), 0, 65535)
    assert np.abs(got - want).max() <= 1


def test_devices_sharding_in_one_process(scene, oracle, pkg):
    """depth2x --devices 2: one context per GPU in ONE process, frames sharded by
    index. max-curve and flexion at 960x540 take kernels that need
    per-device function attributes (54 KB of dynamic shared memory). Needs two GPUs."""
    if pkg.lib().dc_device_count() < 2:
        pytest.skip("one GPU on this box (run with gpurun --gpus 2)")
    cwd, k, frames = scene
    base = ["-c", "kinect_intrinsic.txt", "--type", "pinhole-depth", "--devices", "2",
            "-i", "data{}-depth.png", "-s", "0", "-e", "4"]
    for sub, feat, out in (("max-curve", F_MAX_CURVE, "dev2-mc-{}.png"), ("flexion", F_FLEXION, "dev2-fl-{}.png")):
        r = run(base + [sub, "-o", out], cwd)
        assert r.returncode == 0, r.stderr
        for i in range(5):
            assert_u16_parity(sub, read(cwd, out.format(i)), oracle.cli_frame(feat, 0, frames[i], k)[2])
