"""GPU: the stand-alone quantisers of the library API for every PixelType the
reference instantiates (convert_bearing / convert_flexion / convert_max_curve
<Real, PixelType>, scaling_factor of conversion/util.h:71-79) and
curvature_to_image with missing clamps (min / max of the image, a device
reduction), against the compiled reference (oracle/_ref)."""
import numpy as np
import pytest

from oracle_lib import (F_BEARING_D, F_FLEXION, F_MAX_CURVE, F_MEAN, RefLib, kinect_scaled,
                        ref_available)

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not ref_available(), reason="oracle/_ref not built")]


@pytest.fixture(scope="module")
def ref():
    return RefLib()


@pytest.fixture(scope="module")
def features(oracle, ref):
    w, h = 192, 108
    k = kinect_scaled(w, h)
    d = oracle.synth_depth(0xD2C00031, 0, w, h)
    rng = ref.laserscan(d, k)
    return k, rng, {f: ref.feature(f, rng, k) for f in (F_BEARING_D, F_FLEXION, F_MAX_CURVE, F_MEAN)}


@pytest.mark.parametrize("real", [np.float32, np.float64])
@pytest.mark.parametrize("pixel", [0, 1, 2, 3])
def test_bounded_quantisers_every_pixel_type(pkg, gpu_ctx, ref, features, pixel, real):
    _, _, feats = features
    quants = {F_BEARING_D: pkg.DC_Q_BEARING, F_FLEXION: pkg.DC_Q_FLEXION, F_MAX_CURVE: pkg.DC_Q_MAX_CURVE}
    for f, q in quants.items():
        img = feats[f].astype(real)
        got = gpu_ctx.quantise_ex(q, img, pixel=pixel)
        want = ref.quantise_px(f, img, pixel)
        assert got.dtype == want.dtype
        d = np.abs(got.astype(np.int32) - want.astype(np.int32))
        assert d.max() <= 1, (f, pixel, int(d.max()))
        assert (d == 0).mean() > 0.999


@pytest.mark.parametrize("real", [np.float32, np.float64])
@pytest.mark.parametrize("lo,hi", [(None, None), (-3.0, None), (None, 2.5), (-20.0, 20.0)])
def test_curvature_to_image_with_missing_clamps(pkg, gpu_ctx, ref, features, lo, hi, real):
    _, rng, feats = features
    mean = feats[F_MEAN].astype(real)
    want = ref.curvature_to_image(mean, rng.astype(real), lo, hi)
    got = gpu_ctx.quantise_ex(pkg.DC_Q_CURVATURE, mean, mask=rng.astype(real), lo=lo, hi=hi)
    d = np.abs(got.astype(np.int32) - want.astype(np.int32))
    assert d.max() <= 1, int(d.max())


def test_quantiser_argument_errors(pkg, gpu_ctx, features):
    _, rng, feats = features
    rc, _ = gpu_ctx.quantise_ex_raw(pkg.DC_Q_CURVATURE, feats[F_MEAN], mask=rng, pixel=pkg.DC_PX_U8,
                                    lo=-1.0, hi=1.0)
    assert rc == pkg.DC_ERR_UNSUPPORTED
    const = np.full((8, 8), 3.0, np.float32)
    rc, _ = gpu_ctx.quantise_ex_raw(pkg.DC_Q_CURVATURE, const, mask=const)
    assert rc == pkg.DC_ERR_INVALID  # min == max: math::scale's precondition (scaling.h:36)
    rc, _ = gpu_ctx.quantise_ex_raw(pkg.DC_Q_CURVATURE, const)
    assert rc == pkg.DC_ERR_INVALID  # no mask
    rc, _ = gpu_ctx.quantise_ex_raw(pkg.DC_Q_BEARING, const, pixel=7)
    assert rc == pkg.DC_ERR_INVALID
