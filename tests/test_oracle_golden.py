"""CPU: the plain-C restatement (oracle/dc_oracle.c) against the committed
golden vectors, which were produced by the reference's own headers compiled in
the build container (tests/golden/make_golden.py). Bit-exact."""
import glob
import os

import numpy as np
import pytest

from oracle_lib import FEATURE_NAMES
from parity import bits

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = sorted(os.path.basename(p) for p in glob.glob(os.path.join(GOLD, "*x*.npz")))


def intr(z):
    w, h, fx, fy, cx, cy = z["intrinsic"]
    return dict(w=int(w), h=int(h), fx=float(fx), fy=float(fy), cx=float(cx), cy=float(cy))


@pytest.mark.parametrize("case", CASES)
def test_restatement_matches_golden(oracle, case):
    z = np.load(os.path.join(GOLD, case))
    k, depth = intr(z), z["depth"]
    for kind in (0, 1):
        if f"k{kind}_range" not in z:
            continue
        rng = oracle.laserscan(depth, k) if kind == 0 else oracle.u16_to_f32(depth)
        assert np.array_equal(bits(rng), bits(z[f"k{kind}_range"]))
        assert np.array_equal(oracle.quantise(0, rng, rng), z[f"k{kind}_range_u16"])
        for f in range(1, 9):
            n = FEATURE_NAMES[f]
            feat = oracle.feature(f, rng, k)
            assert np.array_equal(bits(feat), bits(z[f"k{kind}_{n}"])), (case, kind, n)
            q = oracle.quantise(f, feat, rng, -20.0, 20.0)
            assert np.array_equal(q, z[f"k{kind}_{n}_u16"]), (case, kind, n, "u16")


def test_known_answer_scalars(oracle):
    from oracle_lib import KINECT
    z = np.load(os.path.join(GOLD, "kat_scalars.npz"))
    k = {**KINECT, **{n: float(np.float32(KINECT[n])) for n in ("fx", "fy", "cx", "cy")}}
    for (u, v), ray in zip(z["pts"], z["rays"]):
        assert np.array_equal(bits(oracle.pixel_to_sphere(k, int(u), int(v))), bits(ray))
    for p, phi in zip(z["pairs"], z["phis"]):
        got = np.float32(oracle.phi(k, *[int(x) for x in p]))
        assert got.view(np.uint32) == np.float32(phi).view(np.uint32)
    for (b, c, ca), want in zip(z["ba_in"], z["ba"]):
        got = np.float32(oracle.bearing_angle(float(b), float(c), float(ca)))
        assert got.view(np.uint32) == np.float32(want).view(np.uint32)


def test_reference_unit_test_expectations(oracle):
    """Closed-form expectations of the reference's own unit tests, evaluated on
    the float restatement (src/unittest/math/test_triangles.cpp:12-77,
    camera_models/test_pinhole.cpp:19-38, conversion/test_conversion_max_curve.cpp:28-66)."""
    import math
    ba = oracle.bearing_angle
    assert ba(1.0, 1.0, math.cos(math.radians(60))) == pytest.approx(math.radians(60), rel=1e-5)
    assert math.degrees(ba(5.0, 5.0, math.cos(math.radians(45)))) == pytest.approx(67.5, rel=1e-5)
    assert math.degrees(ba(2.4, 6.7, math.cos(math.radians(44)))) == pytest.approx(117.468, rel=1e-4)
    assert ba(2.0, 1.0, math.cos(math.radians(0.05))) < math.radians(90)
    assert ba(1.0, 2.0, math.cos(math.radians(0.05))) > math.radians(90)
    assert ba(1.0, 2.0, 0.5) == pytest.approx(math.radians(90), rel=1e-5)
    assert ba(1.0, math.sqrt(2.0), math.cos(math.radians(45))) == pytest.approx(math.radians(90), rel=1e-4)
    from oracle_lib import KINECT
    k = dict(KINECT, fy=479.462, cx=522.23)  # src/unittest/conversion/intrinsic.h:14-18
    phi = oracle.phi
    assert phi(k, 0, 1, 1, 1) == phi(k, 1, 1, 0, 1)
    assert phi(k, 1, 0, 1, 1) == phi(k, 1, 1, 1, 0)
    assert phi(k, 1, 1, 0, 0) == phi(k, 0, 0, 1, 1)
    assert phi(k, 1, 0, 0, 1) == phi(k, 0, 1, 1, 0)
    # The reference checks strict monotonicity in DOUBLE; in float (the CLI
    # precision) neighbouring rays at the image corner hit the 1 - 10*eps clamp
    # (SURVEY.md K4), so only the weak ordering survives.
    assert phi(k, 0, 0, 1, 1) >= phi(k, 0, 0, 1, 0)
    assert phi(k, 0, 0, 1, 1) >= phi(k, 0, 0, 0, 1)
    assert phi(k, 0, 0, 1, 0) < phi(k, 480, 0, 481, 0)
    assert phi(k, 900, 0, 901, 0) < phi(k, 480, 0, 481, 0)
    assert phi(k, 480, 270, 481, 271) > phi(k, 480, 270, 481, 270)


def test_bearing_border_and_prior_rule(oracle):
    """Iteration ranges / prior pixel per direction
    (src/unittest/conversion/test_conversion_bearing.cpp:17-76)."""
    from oracle_lib import kinect_scaled
    k = kinect_scaled(42, 42)
    rng = np.full((42, 42), 1000.0, np.float32)
    rng[20, 20] = 1200.0
    h = oracle.feature(1, rng, k)
    v = oracle.feature(2, rng, k)
    d = oracle.feature(3, rng, k)
    a = oracle.feature(4, rng, k)
    assert (h[:, 0] == 0).all() and (h[:, 1:] != 0).all()
    assert (v[0, :] == 0).all() and (v[1:, :] != 0).all()
    assert (d[0, :] == 0).all() and (d[:, 0] == 0).all() and (d[1:, 1:] != 0).all()
    assert (a[41, :] == 0).all() and (a[:, 0] == 0).all() and (a[:41, 1:] != 0).all()
    # the bump at (20,20) is the prior of (21,20) hor, (20,21) ver, (21,21) dia, (21,19) anti
    flat = h[5, 5]
    assert h[20, 21] != flat and v[21, 20] != flat and d[21, 21] != flat and a[19, 21] != flat


def test_synth_depth_statistics(oracle):
    d = oracle.synth_depth(0xD2C00001, 7, 960, 540)
    zero = (d == 0).mean()
    assert 0.2 < zero < 0.4
    nz = d[d != 0]
    assert nz.min() >= 796 and nz.max() <= 2504
