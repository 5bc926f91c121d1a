"""TEST INFRASTRUCTURE (oracle): numpy restatement of the two OpenCV filters behind
sens_loc::preprocess::gaussian_blur and ::bilateral_filter
(/root/reference/src/include/sens_loc/preprocess/filter.h:49-82, 125-136).

The arithmetic lives in OpenCV (pinned 4.3.0 by the reference, cmake/use_opencv.cmake:17,
source not vendored): this file restates the published algorithm of
cv::GaussianBlur (separable, kernel of cv::getGaussianKernel, BORDER_REFLECT_101) and of
cv::bilateralFilter for CV_32F single-channel images (colour weights from a 4096-bin
interpolated exp table over the value range of the image, spatial weights inside the
circle of the radius, centre pixel with weight 1), and tests/test_filters_oracle.py pins it
against the cv2 of this image (4.13) on the reference's fixture frames and random images.
Only tests/ may import this."""
import numpy as np


# fixed kernels for sigma <= 0 (getGaussianKernelBitExact; the n = 9 row is in the cv2 of
# this image and, as far as its history shows, in every version that has that function)
SMALL_TABLES = {
    1: [1.0],
    3: [0.25, 0.5, 0.25],
    5: [0.0625, 0.25, 0.375, 0.25, 0.0625],
    7: [0.03125, 0.109375, 0.21875, 0.28125, 0.21875, 0.109375, 0.03125],
    9: [4 / 256, 13 / 256, 30 / 256, 51 / 256, 60 / 256, 51 / 256, 30 / 256, 13 / 256, 4 / 256],
}


def gaussian_kernel(n, sigma):
    """cv::getGaussianKernel(n, sigma, CV_32F): exp(-x^2 / (2 sigma^2)) normalised to sum 1,
    computed in double, stored as float; sigma <= 0 -> the fixed table for n <= 9, else
    sigma = 0.3*((n-1)*0.5 - 1) + 0.8."""
    if sigma <= 0 and n in SMALL_TABLES:
        return np.array(SMALL_TABLES[n], dtype=np.float32)
    if sigma <= 0:
        sigma = 0.3 * ((n - 1) * 0.5 - 1) + 0.8
    x = np.arange(n, dtype=np.float64) - (n - 1) * 0.5
    k = np.exp(-0.5 * x * x / (sigma * sigma))
    return (k / k.sum()).astype(np.float32)


def gaussian_ksize(ksize, sigma, is_u8=False):
    """cv::GaussianBlur's automatic kernel size for ksize <= 0."""
    if ksize > 0:
        return ksize
    return int(np.rint(sigma * (3 if is_u8 else 4) * 2 + 1)) | 1


def reflect101(i, n):
    """cv::borderInterpolate(i, n, BORDER_REFLECT_101)"""
    if n == 1:
        return 0
    while i < 0 or i >= n:
        i = -i if i < 0 else 2 * (n - 1) - i
    return i


def _pad101(img, r):
    h, w = img.shape
    ys = [reflect101(i, h) for i in range(-r, h + r)]
    xs = [reflect101(i, w) for i in range(-r, w + r)]
    return img[np.ix_(ys, xs)]


def gaussian_blur_f32(img, kw, kh, sigma_x, sigma_y=0.0):
    """float image -> float image; rows first, then columns, float accumulation in the
    symmetric order k[c] x[c] + sum_i k[c+i] (x[c+i] + x[c-i])."""
    img = np.asarray(img, dtype=np.float32)
    if sigma_y <= 0:
        sigma_y = sigma_x
    kw = gaussian_ksize(kw, sigma_x)
    kh = gaussian_ksize(kh, sigma_y)
    kx, ky = gaussian_kernel(kw, sigma_x), gaussian_kernel(kh, sigma_y)
    rx, ry = kw // 2, kh // 2
    h, w = img.shape
    p = _pad101(img, max(rx, ry))
    o = max(rx, ry)
    rows = p[:, o:o + w] * kx[rx]
    for i in range(1, rx + 1):
        rows = rows + kx[rx + i] * (p[:, o + i:o + i + w] + p[:, o - i:o - i + w])
    out = rows[o:o + h] * ky[ry]
    for i in range(1, ry + 1):
        out = out + ky[ry + i] * (rows[o + i:o + i + h] + rows[o - i:o - i + h])
    return out.astype(np.float32)


def bilateral_f32(img, d, sigma_color, sigma_space):
    """cv::bilateralFilter on a CV_32FC1 image, BORDER_DEFAULT (no NaN in the input).
    sigma <= 0 -> 1 as OpenCV 4.3.0 (the version the reference pins) does; the cv2 4.13 of
    this image returns the input unchanged for sigma_space <= 0 instead, so tests pass
    explicit positive sigmas to cv2."""
    img = np.asarray(img, dtype=np.float32)
    if sigma_color <= 0:
        sigma_color = 1.0
    if sigma_space <= 0:
        sigma_space = 1.0
    gc = -0.5 / (sigma_color * sigma_color)
    gs = -0.5 / (sigma_space * sigma_space)
    radius = int(np.rint(sigma_space * 1.5)) if d <= 0 else d // 2
    radius = max(radius, 1)
    mn, mx = float(img.min()), float(img.max())
    if abs(mn - mx) < np.finfo(np.float32).eps:
        return img.copy()
    bins = 1 << 12
    length = np.float32(mx - mn)
    scale_index = np.float32(bins) / length
    lut = np.zeros(bins + 2, dtype=np.float32)
    last = np.float32(1.0)
    for i in range(bins + 2):
        if last > 0:
            val = float(i) / float(scale_index)
            lut[i] = np.float32(np.exp(val * val * gc))
            last = lut[i]
    h, w = img.shape
    p = _pad101(img, radius)
    s = np.zeros((h, w), dtype=np.float32)
    ws = np.zeros((h, w), dtype=np.float32)
    for i in range(-radius, radius + 1):
        for j in range(-radius, radius + 1):
            r = np.sqrt(float(i) * i + float(j) * j)
            if r > radius or (i == 0 and j == 0):
                continue
            sw = np.float32(np.exp(r * r * gs))
            v = p[radius + i:radius + i + h, radius + j:radius + j + w]
            alpha = np.abs(v - img) * scale_index
            idx = np.floor(alpha).astype(np.int32)
            alpha = alpha - idx.astype(np.float32)
            wgt = sw * (lut[idx] + alpha * (lut[idx + 1] - lut[idx]))
            s = s + v * wgt
            ws = ws + wgt
    return ((s + img) / (ws + np.float32(1.0))).astype(np.float32)
