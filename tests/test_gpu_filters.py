"""GPU: preprocess::gaussian_blur and preprocess::bilateral_filter (SURVEY.md section 8,
row f4; src/include/sens_loc/preprocess/filter.h:49-82, 125-136) through the C ABI against
OpenCV itself (cv2, the same library, newer version) and against the numpy restatement
oracle/filters_oracle.py. Not bit exact against OpenCV (its accumulation order is build
dependent): Gaussian 1e-6 of the value (u16: +-1), bilateral 2e-5 of the frame's value
range; against the restatement, which accumulates in the kernels' order, far tighter."""
import os
import sys

import cv2
import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import filters_oracle as fo  # noqa: E402

pytestmark = pytest.mark.gpu
KINECT = os.path.join(ROOT, "tests", "golden", "ref_fixtures", "depth2x", "data0-depth.png")


def random_frames(w, h, n=3, seed=0):
    rs = np.random.RandomState(seed + w * h)
    a = (rs.rand(n, h, w) * 3000 + 500).astype(np.float32)
    a[rs.rand(n, h, w) < 0.2] = 0
    return a


@pytest.mark.parametrize("w,h", [(960, 540), (203, 77), (64, 32), (65, 33), (5, 3), (1, 1)])
@pytest.mark.parametrize("kw,kh,sx,sy", [(3, 3, 0, 0), (5, 5, 0, 0), (9, 9, 0, 0), (5, 5, 1.5, 0), (7, 3, 2.0, 1.0),
                                         (0, 0, 1.0, 0), (31, 31, 6.0, 0)])
def test_gaussian_blur_matches_opencv(gpu_ctx, w, h, kw, kh, sx, sy):
    if (w, h) == (960, 540):
        fr = np.stack([cv2.imread(KINECT, cv2.IMREAD_UNCHANGED).astype(np.float32)] * 2)
        fr[1] = random_frames(w, h, 1)[0]
    else:
        fr = random_frames(w, h)
    gpu_ctx.set_frame_size(w, h)
    got = gpu_ctx.gaussian_blur(fr, (kw, kh), sx, sy)
    got16 = gpu_ctx.gaussian_blur(fr.astype(np.uint16), (kw, kh), sx, sy)
    assert got.dtype == np.float32 and got16.dtype == np.uint16
    for i in range(fr.shape[0]):
        want = cv2.GaussianBlur(fr[i], (kw, kh), sx, sigmaY=sy)
        assert np.all(np.abs(got[i] - want) <= 1e-6 * np.maximum(np.abs(want), 1.0)), (w, h, kw, kh, i)
        want16 = cv2.GaussianBlur(fr[i].astype(np.uint16), (kw, kh), sx, sigmaY=sy).astype(np.int32)
        assert np.abs(got16[i].astype(np.int32) - want16).max() <= 1
        if w * h <= 203 * 77:  # the restatement sums in the kernel's order: bit exact
            assert np.array_equal(got[i], fo.gaussian_blur_f32(fr[i], kw, kh, sx, sy))


@pytest.mark.parametrize("w,h", [(960, 540), (203, 77), (33, 32), (5, 3)])
@pytest.mark.parametrize("d,sc,ss", [(5, 20.0, 1.0), (5, 150.0, 2.0), (9, 50.0, 3.0), (3, 10.0, 0.8), (0, 30.0, 2.0),
                                     (31, 500.0, 8.0)])
def test_bilateral_matches_opencv(gpu_ctx, w, h, d, sc, ss):
    if (w, h) == (960, 540):
        fr = np.stack([cv2.imread(KINECT, cv2.IMREAD_UNCHANGED).astype(np.float32)] * 2)
        fr[1] = random_frames(w, h, 1)[0]
    else:
        fr = random_frames(w, h)
    gpu_ctx.set_frame_size(w, h)
    got = gpu_ctx.bilateral(fr, d, sc, ss)
    got16 = gpu_ctx.bilateral(fr.astype(np.uint16), d, sc, ss)
    for i in range(fr.shape[0]):
        span = float(fr[i].max() - fr[i].min())
        want = cv2.bilateralFilter(fr[i], d, sc, ss)
        assert np.abs(got[i] - want).max() <= 2e-5 * span, (w, h, d, sc, ss, i)
        # u16 frames: converted to float, filtered, converted back (math::convert)
        f16 = fr[i].astype(np.uint16).astype(np.float32)
        want16 = np.clip(np.rint(cv2.bilateralFilter(f16, d, sc, ss)), 0, 65535).astype(np.int32)
        assert np.abs(got16[i].astype(np.int32) - want16).max() <= 1
        if w * h <= 203 * 77:
            ref = fo.bilateral_f32(fr[i], d, sc, ss)
            assert np.abs(got[i] - ref).max() <= 1e-6 * span


def test_bilateral_flat_frame_defaults_and_errors(pkg, gpu_ctx):
    w, h = 40, 24
    gpu_ctx.set_frame_size(w, h)
    fr = random_frames(w, h, 2)
    fr[1] = 777.0  # max - min < FLT_EPSILON: the frame is copied
    got = gpu_ctx.bilateral(fr, 5, 25.0, 1.5)
    assert np.array_equal(got[1], fr[1])
    assert np.abs(got[0] - cv2.bilateralFilter(fr[0], 5, 25.0, 1.5)).max() <= 2e-5 * 3500
    # sigma <= 0 means 1 (OpenCV 4.3.0, the version the reference pins)
    assert np.array_equal(gpu_ctx.bilateral(fr[:1], 5, 0.0, 0.0), gpu_ctx.bilateral(fr[:1], 5, 1.0, 1.0))
    with pytest.raises(pkg.DcError):
        gpu_ctx.bilateral(fr, 33, 10.0, 1.0)       # radius 16
    with pytest.raises(pkg.DcError):
        gpu_ctx.gaussian_blur(fr, (4, 3), 1.0)     # even size
    with pytest.raises(pkg.DcError):
        gpu_ctx.gaussian_blur(fr, (0, 0), 0.0)     # no size, no sigma
    with pytest.raises(pkg.DcError):
        gpu_ctx.gaussian_blur(fr, (33, 3), 1.0)    # beyond 31


def test_filter_then_convert_on_the_device(pkg, gpu_ctx, oracle):
    """bilateral filter -> flexion without leaving the device (what `depth_filter bilateral`
    followed by `depth2x flexion --type pinhole-range` computes with a PNG in between,
    scripts/analysis/transformation_step.py:80-96), on the caller's stream."""
    import torch
    from oracle_lib import F_FLEXION, kinect_scaled
    w, h = 320, 200
    k = kinect_scaled(w, h)
    gpu_ctx.set_pinhole(w, h, k["fx"], k["fy"], k["cx"], k["cy"])
    depth = np.stack([oracle.synth_depth(0xD2C000B1, f, w, h) for f in range(4)])
    rng = np.stack([oracle.laserscan(d, k) for d in depth]).astype(np.float32)
    d_in = torch.from_numpy(rng).cuda()
    d_flt = torch.empty_like(d_in)
    d_out = torch.empty_like(d_in)
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        gpu_ctx._check(gpu_ctx._L.dc_bilateral_device(gpu_ctx._h, pkg.DC_IN_F32_RANGE, d_in.data_ptr(), 4, 5,
                                                      30.0, 1.5, d_flt.data_ptr(), st.cuda_stream))
        job = gpu_ctx.make_job(pkg.DC_FLEXION, pkg.DC_IN_F32_RANGE, pkg.DC_OUT_F32, 4, d_flt.data_ptr(),
                               [d_out.data_ptr()])
        gpu_ctx.convert_device(job, st.cuda_stream)
    st.synchronize()
    flt = d_flt.cpu().numpy()
    for i in range(4):
        assert np.abs(flt[i] - cv2.bilateralFilter(rng[i], 5, 30.0, 1.5)).max() <= 2e-5 * float(rng[i].max())
        want = oracle.feature(F_FLEXION, flt[i], k)
        from parity import assert_float_parity
        assert_float_parity("flexion", d_out[i].cpu().numpy(), want)
