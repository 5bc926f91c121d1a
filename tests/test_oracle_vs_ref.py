"""CPU: pins the oracle. (1) the plain-C restatement is bit-identical to the
reference's own headers compiled here (oracle/_ref); (2) that compiled reference
reproduces the reference's checked-in golden PNGs (only where /root/reference
is mounted, i.e. in the build container)."""
import os

import numpy as np
import pytest

from oracle_lib import FEATURE_NAMES, KINECT, RefLib, kinect_scaled, ref_available
from parity import bits

pytestmark = pytest.mark.skipif(not ref_available(), reason="oracle/_ref not built")
UNIT = "/root/reference/src/unittest/conversion/"


@pytest.fixture(scope="module")
def ref():
    return RefLib()


@pytest.mark.parametrize("w,h,seed", [(320, 240, 0xD2C00002), (202, 131, 0xD2C00003)])
def test_restatement_bit_identical_to_compiled_reference(oracle, ref, w, h, seed):
    k = kinect_scaled(w, h)
    d = oracle.synth_depth(seed, 2, w, h)
    for kind in (0, 1):
        r0 = oracle.laserscan(d, k) if kind == 0 else oracle.u16_to_f32(d)
        r1 = ref.laserscan(d, k) if kind == 0 else d.astype(np.float32)
        assert np.array_equal(bits(r0), bits(r1))
        for f in range(1, 9):
            a, b = oracle.feature(f, r0, k), ref.feature(f, r1, k)
            assert np.array_equal(bits(a), bits(b)), FEATURE_NAMES[f]
            qa, qb = oracle.quantise(f, a, r0, -3.0, 7.5), ref.quantise(f, b, r1, -3.0, 7.5)
            assert np.array_equal(qa, qb), FEATURE_NAMES[f]
        assert np.array_equal(oracle.quantise(0, r0, r0), ref.quantise(0, r1, r1))
    for s, o in ((8.0, 0.0), (1.0, 10000.0), (0.37, -12.5)):
        assert np.array_equal(oracle.depth_scaling(d, s, o), ref.depth_scaling(d, s, o))


def test_cli_batch_matches_single_frames(oracle, ref):
    """The taskflow frame-parallel driver (the CPU baseline) equals per-frame calls."""
    w, h = 96, 64
    k = kinect_scaled(w, h)
    frames = np.stack([oracle.synth_depth(0xD2C00004, f, w, h) for f in range(5)])
    feats = [1, 2, 3, 4, 5, 8]
    outs, workers = ref.cli_batch(feats, 0, frames, k, threads=3)
    assert workers == 3
    for i, f in enumerate(feats):
        for n in range(frames.shape[0]):
            _, _, q = oracle.cli_frame(f, 0, frames[n], k)
            assert np.array_equal(outs[i][n], q)


@pytest.mark.skipif(not os.path.isdir(UNIT), reason="/root/reference not mounted")
def test_compiled_reference_reproduces_reference_goldens(oracle, ref):
    """Double path + unit-test intrinsic (src/unittest/conversion/intrinsic.h:8-19)
    against the PNGs the reference's own tests compare with
    (test_conversion_{laser,bearing,flexion,curvature,max_curve}.cpp). The
    reference accepts average_pixel_error < 0.5 .. 5; here: <= 1 LSB per pixel."""
    import cv2
    d = cv2.imread(UNIT + "data0-depth.png", cv2.IMREAD_UNCHANGED)
    k = dict(w=960, h=540, fx=519.226, fy=479.462, cx=522.23, cy=272.737)
    rng = ref.laserscan(d, k, np.float64)
    g = cv2.imread(UNIT + "data0-depth-laser.png", cv2.IMREAD_UNCHANGED)
    q = np.clip(np.rint(rng), 0, 65535).astype(np.uint16)
    assert np.abs(q.astype(int) - g.astype(int)).max() <= 1
    double_goldens = {2: "bearing-vertical", 3: "bearing-diagonal", 5: "flexion-reference",
                      6: "gauss-reference", 7: "mean-reference", 8: "max-curve-double"}
    for f, name in double_goldens.items():
        feat = ref.feature(f, rng, k)
        g = cv2.imread(UNIT + name + ".png", cv2.IMREAD_UNCHANGED)
        q = ref.convert_bearing_u8(feat) if g.dtype == np.uint8 else ref.quantise(f, feat, rng)
        assert np.abs(q.astype(int) - g.astype(int)).max() <= 1, name
    # float-path goldens: the reference's own thresholds (rounding noise, K5)
    kf = {**k, **{n: float(np.float32(k[n])) for n in ("fx", "fy", "cx", "cy")}}
    r32 = ref.laserscan(d, kf)
    g = cv2.imread(UNIT + "bearing-horizontal.png", cv2.IMREAD_UNCHANGED)
    q = ref.quantise(1, ref.feature(1, r32, kf), r32)
    assert oracle.average_pixel_error(q, g) < 1.0   # test_conversion_bearing.cpp:175
    g = cv2.imread(UNIT + "bearing-antidiagonal.png", cv2.IMREAD_UNCHANGED)
    q8 = ref.convert_bearing_u8(ref.feature(4, r32, kf))
    ape = np.linalg.norm(q8.astype(np.float64) - g.astype(np.float64)) / g.size
    assert ape < 0.5                                 # test_conversion_bearing.cpp:158
    # pathological frame must not crash (test_conversion_bearing_pathological.cpp:14-21)
    d = cv2.imread(UNIT + "crashing-file.png", cv2.IMREAD_UNCHANGED)
    r = ref.laserscan(d, kf)
    ref.feature(3, r, kf)
    for f in range(1, 9):
        assert np.array_equal(bits(oracle.feature(f, r, kf)), bits(ref.feature(f, r, kf)))
