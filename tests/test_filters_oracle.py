"""CPU: oracle/filters_oracle.py (the numpy restatement of cv::GaussianBlur and of the
CV_32F cv::bilateralFilter) pinned against the cv2 of this image
on random frames and on a real Kinect frame of the reference's fixtures. Tolerances: the
filters are float sums whose order / FMA use is OpenCV-build dependent (Gaussian 1e-6 of the
value, u16 +-1; bilateral 2e-5 of the frame's value range)."""
import os
import sys

import cv2
import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import filters_oracle as fo  # noqa: E402

KINECT = os.path.join(ROOT, "tests", "golden", "ref_fixtures", "depth2x", "data0-depth.png")


def frames():
    rs = np.random.RandomState(7)
    a = (rs.rand(61, 83) * 3000 + 500).astype(np.float32)
    a[rs.rand(*a.shape) < 0.2] = 0
    k = cv2.imread(KINECT, cv2.IMREAD_UNCHANGED)[100:260, 300:520].astype(np.float32)
    return [a, k]


@pytest.mark.parametrize("n", [1, 3, 5, 7, 9, 11, 15, 31])
@pytest.mark.parametrize("sigma", [0.0, 0.7, 1.5, 4.0])
def test_gaussian_kernel(n, sigma):
    want = cv2.getGaussianKernel(n, sigma, cv2.CV_32F).ravel()
    assert np.array_equal(fo.gaussian_kernel(n, sigma), want)


@pytest.mark.parametrize("kw,kh,sx,sy", [(3, 3, 0, 0), (5, 5, 0, 0), (9, 9, 0, 0), (5, 5, 1.5, 0), (7, 3, 2.0, 1.0),
                                         (0, 0, 1.0, 0), (1, 5, 0, 0), (11, 11, 0, 0), (31, 31, 6.0, 0)])
def test_gaussian_blur(kw, kh, sx, sy):
    for img in frames():
        want = cv2.GaussianBlur(img, (kw, kh), sx, sigmaY=sy)
        got = fo.gaussian_blur_f32(img, kw, kh, sx, sy)
        assert np.all(np.abs(got - want) <= 1e-6 * np.maximum(np.abs(want), 1.0))
        i16 = img.astype(np.uint16)
        want16 = cv2.GaussianBlur(i16, (kw, kh), sx, sigmaY=sy).astype(np.int32)
        got16 = np.clip(np.rint(fo.gaussian_blur_f32(i16.astype(np.float32), kw, kh, sx, sy)), 0, 65535)
        assert np.abs(got16 - want16).max() <= 1


@pytest.mark.parametrize("d,sc,ss", [(5, 20.0, 1.0), (5, 150.0, 2.0), (9, 50.0, 3.0), (3, 10.0, 0.8),
                                     (7, 2000.0, 5.0), (0, 30.0, 2.0), (0, 100.0, 3.0)])
def test_bilateral(d, sc, ss):
    for img in frames():
        want = cv2.bilateralFilter(img, d, sc, ss)
        got = fo.bilateral_f32(img, d, sc, ss)
        assert np.abs(got - want).max() <= 2e-5 * float(img.max() - img.min())


def test_bilateral_flat_frame_and_sigma_defaults():
    flat = np.full((9, 11), 1234.0, np.float32)
    assert np.array_equal(fo.bilateral_f32(flat, 5, 10.0, 1.0), flat)
    img = frames()[0]
    # sigma <= 0 means 1 (OpenCV 4.3.0, the reference's pin)
    assert np.array_equal(fo.bilateral_f32(img, 5, 0.0, 0.0), fo.bilateral_f32(img, 5, 1.0, 1.0))
