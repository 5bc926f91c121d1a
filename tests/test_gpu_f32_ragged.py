"""GPU: the library API's input type (f32 range images, what
sens_loc::conversion::depth_to_*(math::image<float>) sends) and widths that are
even but not a multiple of four run on the guard-free frame-pair kernels; f32
values outside the proven window fall back to the IEEE kernels inside the call."""
import numpy as np
import pytest

from oracle_lib import FEATURE_NAMES, kinect_scaled
from parity import assert_float_parity, assert_u16_parity
from test_gpu_parity import gpu_all_features, set_k

pytestmark = pytest.mark.gpu


def bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


def test_f32_range_input_runs_the_pair_kernels(pkg, gpu_ctx, oracle):
    w, h, n = 960, 540, 3
    k = kinect_scaled(w, h)
    set_k(gpu_ctx, k)
    frames = np.stack([oracle.synth_depth(0xD2C00021, f, w, h) for f in range(n)])
    from_u16 = gpu_all_features(pkg, gpu_ctx, frames, pkg.DC_IN_U16_DEPTH, pkg.DC_OUT_F32)
    u16_u16 = gpu_all_features(pkg, gpu_ctx, frames, pkg.DC_IN_U16_DEPTH, pkg.DC_OUT_U16)
    rng = from_u16["range"]
    before = gpu_ctx.f32_fallbacks
    P = pkg
    gpu_ctx.convert(P.DC_BEARING, rng, P.DC_IN_F32_RANGE, P.DC_OUT_F32, want=(True,) * 4)
    assert gpu_ctx.last_kernel == "dc::pr::pair_kernel<SPEC_TABLES>"
    gpu_ctx.convert(P.DC_MAX_CURVE, rng, P.DC_IN_F32_RANGE, P.DC_OUT_F32)
    assert gpu_ctx.last_kernel == "dc::pr::maxcurve_edge_kernel<SPEC_TABLES>"
    gpu_ctx.convert(P.DC_FLEXION, rng, P.DC_IN_F32_RANGE, P.DC_OUT_F32)
    assert gpu_ctx.last_kernel == "dc::pr::flexion_pair_kernel"
    gpu_ctx.convert(P.DC_CURVATURE, rng, P.DC_IN_F32_RANGE, P.DC_OUT_F32, want=(True, True))
    assert gpu_ctx.last_kernel == "dc::pr::pair_kernel<SPEC_TABLES>"
    from_f32 = gpu_all_features(pkg, gpu_ctx, rng, P.DC_IN_F32_RANGE, P.DC_OUT_F32)
    f32_u16 = gpu_all_features(pkg, gpu_ctx, rng, P.DC_IN_F32_RANGE, P.DC_OUT_U16)
    assert gpu_ctx.f32_fallbacks == before, "in-window input must not take the IEEE fallback"
    # the same kernels behind a different load stage: bit-identical, every feature
    for name in FEATURE_NAMES.values():
        assert np.array_equal(bits(from_f32[name]), bits(from_u16[name])), name
        if name != "range":  # (range of a float range image: round to u16 of the value itself)
            assert np.array_equal(f32_u16[name], u16_u16[name]), name
    for i in range(n):
        r = oracle.laserscan(frames[i], k)
        for f, name in FEATURE_NAMES.items():
            assert_float_parity(name, from_f32[name][i], oracle.feature(f, r, k))


@pytest.mark.parametrize("scale", [2.0 ** -22, 2.0 ** 14, -1.0])
def test_f32_values_outside_the_window_fall_back(pkg, gpu_ctx, oracle, scale):
    """Ranges below 2^-10, above 2^22 or negative: the fast kernels raise the
    guard, the IEEE kernels redo the batch, the result is still the oracle's."""
    w, h = 256, 96
    k = kinect_scaled(w, h)
    set_k(gpu_ctx, k)
    frames = np.stack([oracle.synth_depth(0xD2C00022, f, w, h) for f in range(3)])
    rng = np.stack([oracle.laserscan(f, k) for f in frames])
    bad = rng.copy()
    bad[1] *= np.float32(scale)  # one frame of the batch leaves the window
    before = gpu_ctx.f32_fallbacks
    got = gpu_all_features(pkg, gpu_ctx, bad, pkg.DC_IN_F32_RANGE, pkg.DC_OUT_F32)
    assert gpu_ctx.f32_fallbacks >= before + 4, "bearing, flexion, curvature, max-curve must fall back"
    for i in range(3):
        for f, name in FEATURE_NAMES.items():
            want = oracle.feature(f, bad[i], k)
            if scale < 0 and i == 1:
                # negative ranges are outside the reference's contract; the IEEE
                # kernels still follow its arithmetic wherever that is finite
                ok = np.isfinite(want) & np.isfinite(got[name][i])
                assert np.array_equal(np.isfinite(want), np.isfinite(got[name][i])), name
                assert_float_parity(name, np.where(ok, got[name][i], 0).astype(np.float32),
                                    np.where(ok, want, 0).astype(np.float32), exact=False)
            else:
                assert_float_parity(name, got[name][i], want)
    # the next in-window batch is served by the fast kernels again, without a fallback
    before = gpu_ctx.f32_fallbacks
    again = gpu_all_features(pkg, gpu_ctx, rng, pkg.DC_IN_F32_RANGE, pkg.DC_OUT_F32)
    assert gpu_ctx.f32_fallbacks == before
    for f, name in FEATURE_NAMES.items():
        assert_float_parity(name, again[name][1], oracle.feature(f, rng[1], k))


@pytest.mark.parametrize("w,h", [(322, 203), (66, 50), (6, 5), (2, 3)])
@pytest.mark.parametrize("n", [1, 3])
def test_even_ragged_widths_run_the_pair_kernels(pkg, gpu_ctx, oracle, w, h, n):
    """w % 4 == 2: 4-pixel accesses in two halves (SPEC_HALF), the last column group
    half empty; u16 and f32 input, both output types."""
    k = kinect_scaled(w, h)
    set_k(gpu_ctx, k)
    frames = np.stack([oracle.synth_depth(0xD2C00023, f, w, h) for f in range(n)])
    P = pkg
    f32 = gpu_all_features(pkg, gpu_ctx, frames, P.DC_IN_U16_DEPTH, P.DC_OUT_F32)
    gpu_ctx.convert(P.DC_BEARING, frames, P.DC_IN_U16_DEPTH, P.DC_OUT_U16, want=(True,) * 4)
    assert gpu_ctx.last_kernel == "dc::pr::pair_kernel<SPEC_HALF>"
    gpu_ctx.convert(P.DC_FLEXION, frames, P.DC_IN_U16_DEPTH, P.DC_OUT_U16)
    assert gpu_ctx.last_kernel == "dc::pr::flexion_pair_kernel<LSB>"
    u16 = gpu_all_features(pkg, gpu_ctx, frames, P.DC_IN_U16_DEPTH, P.DC_OUT_U16, clamp=(-5.0, 11.0))
    lib = gpu_all_features(pkg, gpu_ctx, f32["range"], P.DC_IN_F32_RANGE, P.DC_OUT_F32)
    assert gpu_ctx.last_kernel == "dc::pr::pair_kernel<SPEC_HALF>"
    for i in range(n):
        rng = oracle.laserscan(frames[i], k)
        for f, name in FEATURE_NAMES.items():
            want = oracle.feature(f, rng, k)
            assert_float_parity(name, f32[name][i], want)
            assert_float_parity(name, lib[name][i], want)
            assert_u16_parity(name, u16[name][i], oracle.quantise(f, want, rng, -5.0, 11.0))


def test_odd_widths_stay_on_the_scalar_kernels(pkg, gpu_ctx, oracle):
    w, h = 61, 37
    k = kinect_scaled(w, h)
    set_k(gpu_ctx, k)
    frames = np.stack([oracle.synth_depth(0xD2C00024, f, w, h) for f in range(2)])
    f32 = gpu_all_features(pkg, gpu_ctx, frames, pkg.DC_IN_U16_DEPTH, pkg.DC_OUT_F32)
    assert gpu_ctx.last_kernel.startswith("dc::feature_kernel")
    for i in range(2):
        rng = oracle.laserscan(frames[i], k)
        for f, name in FEATURE_NAMES.items():
            assert_float_parity(name, f32[name][i], oracle.feature(f, rng, k))


def test_conversion_right_after_installing_a_camera(pkg, oracle):
    """The per-camera tables are uploaded from pageable memory on the legacy stream
    while the kernels run on non-blocking streams: the first conversion after
    dc_set_pinhole must already see complete tables (small cameras, many times)."""
    w, h = 64, 48
    k = kinect_scaled(w, h)
    frames = np.stack([oracle.synth_depth(0xD2C00025, f, w, h) for f in range(2)])
    want = None
    with pkg.Context(0) as ctx:
        for it in range(40):
            k2 = dict(k, fx=k["fx"] + (it % 2))  # a different camera every time: new tables
            set_k(ctx, k2)
            got = gpu_all_features(pkg, ctx, frames, pkg.DC_IN_U16_DEPTH, pkg.DC_OUT_F32)
            set_k(ctx, k)
            got = gpu_all_features(pkg, ctx, frames, pkg.DC_IN_U16_DEPTH, pkg.DC_OUT_F32)
            if want is None:
                want = got
                rng = oracle.laserscan(frames[0], k)
                for f, name in FEATURE_NAMES.items():
                    assert_float_parity(name, got[name][0], oracle.feature(f, rng, k))
            for name in got:
                assert np.array_equal(bits(got[name]), bits(want[name])), (it, name)
