"""GPU: the Real = double instantiation against the
reference's own templates compiled with Real = double (oracle/_ref, built by
oracle/Makefile; the built library travels to the GPU box). Tables come from the
same host libm as the reference's, so everything but the final acos / pow
(device libm, <= 2 ulp) is bit exact: tolerance 1e-12 relative, zero masks
bit-equal, quantised images +-1 LSB."""
import numpy as np
import pytest

from oracle_lib import (F_BEARING_H, F_FLEXION, F_GAUSS, F_MAX_CURVE, F_MEAN, RefLib, kinect_scaled,
                        ref_available)

pytestmark = pytest.mark.gpu

# src/unittest/conversion/intrinsic.h:8-26: the camera of the reference's unit tests
UNIT_K = dict(w=960, h=540, fx=519.226, fy=522.23, cx=479.462, cy=272.737)


@pytest.fixture(scope="module")
def ref():
    if not ref_available():
        pytest.skip("oracle/_ref is not built")
    return RefLib()


def close(name, got, want, rel=1e-12):
    assert got.shape == want.shape and got.dtype == np.float64, name
    assert np.array_equal(got == 0, want == 0), f"{name}: zero mask differs"
    assert np.array_equal(np.isnan(got), np.isnan(want)), f"{name}: NaN mask differs"
    ok = ~np.isnan(want)
    err = np.abs(got[ok] - want[ok])
    tol = rel * np.maximum(np.abs(want[ok]), 1e-300)
    bad = err > tol
    assert not bad.any(), f"{name}: {bad.sum()} pixels off, max rel {np.max(err / np.maximum(np.abs(want[ok]), 1e-300)):.3e}"
    return float(np.mean(got[ok] == want[ok]))


@pytest.mark.parametrize("w,h", [(960, 540), (61, 37)])
def test_double_features_match_compiled_reference(pkg, gpu_ctx, oracle, ref, w, h):
    k = dict(UNIT_K, w=w, h=h) if (w, h) == (960, 540) else kinect_scaled(w, h)
    frames = np.stack([oracle.synth_depth(0xD2C000F3, f, w, h) for f in range(3)])
    gpu_ctx.set_pinhole_f64(w, h, k["fx"], k["fy"], k["cx"], k["cy"])
    rng_ref = np.stack([ref.laserscan(f, k, np.float64) for f in frames])
    rng = gpu_ctx.convert(pkg.DC_RANGE, frames, pkg.DC_IN_U16_DEPTH, pkg.DC_OUT_F64)[0]
    assert np.array_equal(rng, rng_ref), "laserscan<double> must be bit exact"
    # from the double range image (the library API) and straight from u16 depth
    for in_kind, src in ((pkg.DC_IN_F64_RANGE, rng_ref), (pkg.DC_IN_U16_DEPTH, frames)):
        b = gpu_ctx.convert(pkg.DC_BEARING, src, in_kind, pkg.DC_OUT_F64, want=(True,) * 4)
        fl = gpu_ctx.convert(pkg.DC_FLEXION, src, in_kind, pkg.DC_OUT_F64)[0]
        cv = gpu_ctx.convert(pkg.DC_CURVATURE, src, in_kind, pkg.DC_OUT_F64, want=(True, True))
        mc = gpu_ctx.convert(pkg.DC_MAX_CURVE, src, in_kind, pkg.DC_OUT_F64, want=(True, True))
        for i in range(frames.shape[0]):
            for d in range(4):
                close(f"bearing{d}", b[d][i], ref.feature(F_BEARING_H + d, rng_ref[i], k))
            assert np.array_equal(fl[i], ref.feature(F_FLEXION, rng_ref[i], k)), "flexion<double> bit exact"
            want_g = ref.feature(F_GAUSS, rng_ref[i], k)
            assert np.array_equal(cv[0][i], want_g), "gaussian<double> bit exact"
            close("mean", cv[1][i], ref.feature(F_MEAN, rng_ref[i], k))
            close("max_curve", mc[0][i], ref.feature(F_MAX_CURVE, rng_ref[i], k))
            assert np.array_equal(mc[1][i], rng_ref[i])


def test_double_quantisers(pkg, gpu_ctx, oracle, ref):
    w, h = 160, 120
    k = kinect_scaled(w, h)
    d = oracle.synth_depth(0xD2C000F4, 0, w, h)
    rng = ref.laserscan(d, k, np.float64)
    gpu_ctx.set_pinhole_f64(w, h, k["fx"], k["fy"], k["cx"], k["cy"])
    for f, q in ((F_BEARING_H, pkg.DC_Q_BEARING), (F_FLEXION, pkg.DC_Q_FLEXION),
                 (F_MAX_CURVE, pkg.DC_Q_MAX_CURVE), (F_GAUSS, pkg.DC_Q_CURVATURE),
                 (F_MEAN, pkg.DC_Q_CURVATURE)):
        feat = ref.feature(f, rng, k)
        want = ref.quantise(f, feat, rng)
        got = gpu_ctx.quantise(q, feat, mask=rng if q == pkg.DC_Q_CURVATURE else None)
        diff = np.abs(got.astype(np.int32) - want.astype(np.int32))
        assert diff.max() <= 1, (f, int(diff.max()))
        assert np.mean(diff == 0) > 0.999


def test_double_equirectangular(pkg, gpu_ctx, oracle, ref):
    w, h = 128, 64
    tmin, tmax = 0.87, 2.27
    d = oracle.synth_depth(0xD2C000F5, 0, w, h)
    rng = d.astype(np.float64)
    gpu_ctx.set_equirectangular_f64(w, h, tmin, tmax)
    out = np.empty((h, w), np.float64)
    import ctypes as C
    for f, feat, slot in ((F_FLEXION, pkg.DC_FLEXION, 0), (F_BEARING_H, pkg.DC_BEARING, 0),
                          (F_MAX_CURVE, pkg.DC_MAX_CURVE, 0)):
        ref.lib.ref_feature_equi_f64.argtypes = [C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_double,
                                                 C.c_double, C.c_void_p]
        ref.lib.ref_feature_equi_f64(f, rng.ctypes.data, w, h, tmin, tmax, out.ctypes.data)
        want = [True] * (slot + 1)
        got = gpu_ctx.convert(feat, rng, pkg.DC_IN_F64_RANGE, pkg.DC_OUT_F64, want=tuple(want))[slot]
        close(str(f), got, out)


def test_double_job_validation(pkg, gpu_ctx):
    gpu_ctx.set_pinhole(64, 48, 60.0, 60.0, 32.0, 24.0)
    f32 = np.ones((48, 64), np.float32)
    with pytest.raises(pkg.DcError):
        gpu_ctx.convert(pkg.DC_FLEXION, f32, pkg.DC_IN_F32_RANGE, pkg.DC_OUT_F64)
    with pytest.raises(pkg.DcError):
        gpu_ctx.convert(pkg.DC_FLEXION, f32.astype(np.float64), pkg.DC_IN_F64_RANGE, pkg.DC_OUT_F32)
    # float camera widened: a double job on a float camera works
    out = gpu_ctx.convert(pkg.DC_FLEXION, f32.astype(np.float64), pkg.DC_IN_F64_RANGE, pkg.DC_OUT_F64)[0]
    assert out.dtype == np.float64 and out.shape == (48, 64)


def test_cpp_library_api_double(tmp_path, oracle, ref):
    """sens_loc::conversion::*<double> called from C++ (host_selftest api64), the
    way src/unittest/conversion/*.cpp instantiate them."""
    import os
    import subprocess

    import cv2
    here = os.path.dirname(os.path.abspath(__file__))
    selftest = os.path.join(os.path.dirname(here), "depth-conversions_b200", "bin", "host_selftest")
    w, h = 160, 96
    k = kinect_scaled(w, h)
    (tmp_path / "k.txt").write_text(
        f"{w} {h}\n{k['fx']!r} 0.0 {k['cx']!r}\n0.0 {k['fy']!r} {k['cy']!r}\n0.0 0.0 1.0\n")
    depth = oracle.synth_depth(0xD2C000F6, 0, w, h)
    cv2.imwrite(str(tmp_path / "d.png"), depth)
    r = subprocess.run([selftest, "api64", str(tmp_path / "k.txt"), str(tmp_path / "d.png"), str(tmp_path)],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr

    def f64(name):
        return np.fromfile(str(tmp_path / f"{name}.f64"), np.float64).reshape(h, w)

    rng = ref.laserscan(depth, k, np.float64)
    assert np.array_equal(f64("range"), rng)
    for n, f in (("bearing_h", F_BEARING_H), ("bearing_a", F_BEARING_H + 3), ("flexion", F_FLEXION),
                 ("gauss", F_GAUSS), ("mean", F_MEAN), ("max_curve", F_MAX_CURVE)):
        close(n, f64(n), ref.feature(f, rng, k))
    for n, f in (("bearing_h64", F_BEARING_H), ("flexion64", F_FLEXION), ("gauss64", F_GAUSS),
                 ("max_curve64", F_MAX_CURVE)):
        got = cv2.imread(str(tmp_path / f"{n}.png"), cv2.IMREAD_UNCHANGED)
        want = ref.quantise(f, f64(n[:-2]), rng)
        assert np.abs(got.astype(np.int32) - want.astype(np.int32)).max() <= 1, n
