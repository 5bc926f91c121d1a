"""GPU: the CUDA path, called through the C ABI, against the oracle and the
committed golden vectors. Tolerances: parity.py."""
import glob
import os

import numpy as np
import pytest

from oracle_lib import (F_BEARING_H, F_FLEXION, F_GAUSS, F_MAX_CURVE, F_MEAN,
                        FEATURE_NAMES, kinect_scaled)
from parity import assert_float_parity, assert_u16_parity

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = sorted(os.path.basename(p) for p in glob.glob(os.path.join(GOLD, "*x*.npz")))


def set_k(ctx, k):
    ctx.set_pinhole(k["w"], k["h"], k["fx"], k["fy"], k["cx"], k["cy"])


def gpu_all_features(pkg, ctx, frames, in_kind, out_kind, clamp=(-20.0, 20.0)):
    """Every feature image of the hot path for `frames`, keyed like the goldens."""
    P = pkg
    res = {}
    res["range"] = ctx.convert(P.DC_RANGE, frames, in_kind, out_kind)[0]
    b = ctx.convert(P.DC_BEARING, frames, in_kind, out_kind, want=(True,) * 4)
    res["bearing_h"], res["bearing_v"], res["bearing_d"], res["bearing_a"] = b
    res["flexion"] = ctx.convert(P.DC_FLEXION, frames, in_kind, out_kind)[0]
    res["gauss"], res["mean"] = ctx.convert(P.DC_CURVATURE, frames, in_kind, out_kind,
                                            want=(True, True), clamp=clamp)[:2]
    mc = ctx.convert(P.DC_MAX_CURVE, frames, in_kind, out_kind, want=(True, True))
    res["max_curve"], res["range_fused"] = mc[0], mc[1]
    return res


def intr(z):
    w, h, fx, fy, cx, cy = z["intrinsic"]
    return dict(w=int(w), h=int(h), fx=float(fx), fy=float(fy), cx=float(cx), cy=float(cy))


@pytest.mark.parametrize("case", CASES)
def test_golden_vectors(pkg, gpu_ctx, case):
    z = np.load(os.path.join(GOLD, case))
    k, depth = intr(z), z["depth"]
    set_k(gpu_ctx, k)
    for kind, in_kind in ((0, pkg.DC_IN_U16_DEPTH), (1, pkg.DC_IN_U16_RANGE)):
        if f"k{kind}_range" not in z:
            continue
        f32 = gpu_all_features(pkg, gpu_ctx, depth, in_kind, pkg.DC_OUT_F32)
        u16 = gpu_all_features(pkg, gpu_ctx, depth, in_kind, pkg.DC_OUT_U16)
        for n in FEATURE_NAMES.values():
            assert_float_parity(n, f32[n], z[f"k{kind}_{n}"])
            assert_u16_parity(n, u16[n], z[f"k{kind}_{n}_u16"])
        assert_float_parity("range", f32["range_fused"], z[f"k{kind}_range"])
        assert_u16_parity("range", u16["range_fused"], z[f"k{kind}_range_u16"])
        # library API: float range image in (math::image<float>)
        lib = gpu_all_features(pkg, gpu_ctx, z[f"k{kind}_range"], pkg.DC_IN_F32_RANGE,
                               pkg.DC_OUT_F32)
        for n in FEATURE_NAMES.values():
            assert_float_parity(n, lib[n], z[f"k{kind}_{n}"])


@pytest.mark.parametrize("w,h", [(960, 540), (480, 270), (322, 203)])
def test_batch_against_oracle(pkg, gpu_ctx, oracle, w, h):
    """A batch of synthetic frames (BASELINE config 1/2 size and the 480x270 /
    ragged-width cases), every feature, both output types."""
    k = kinect_scaled(w, h)
    set_k(gpu_ctx, k)
    n = 3
    frames = np.stack([oracle.synth_depth(0xD2C00001, f, w, h) for f in range(n)])
    f32 = gpu_all_features(pkg, gpu_ctx, frames, pkg.DC_IN_U16_DEPTH, pkg.DC_OUT_F32)
    u16 = gpu_all_features(pkg, gpu_ctx, frames, pkg.DC_IN_U16_DEPTH, pkg.DC_OUT_U16,
                           clamp=(-5.0, 11.0))
    stats = {}
    for i in range(n):
        rng = oracle.laserscan(frames[i], k)
        for f, name in FEATURE_NAMES.items():
            want = oracle.feature(f, rng, k)
            eq = assert_float_parity(name, f32[name][i], want)
            stats[name] = min(stats.get(name, 1.0), eq)
            assert_u16_parity(name, u16[name][i], oracle.quantise(f, want, rng, -5.0, 11.0))
    print("bit-equal fraction per feature:", {k_: round(v, 6) for k_, v in stats.items()})
    assert stats["mean"] > 0.99


def test_single_outputs_and_direction_subsets(pkg, gpu_ctx, oracle):
    """bearing with only some directions / curvature with one output: the other
    slots stay untouched (converter_bearing.h.inl:13-26)."""
    w, h = 160, 96
    k = kinect_scaled(w, h)
    set_k(gpu_ctx, k)
    d = oracle.synth_depth(0xD2C00005, 0, w, h)
    rng = oracle.laserscan(d, k)
    for want in [(True, False, False, False), (False, False, False, True),
                 (False, True, True, False)]:
        outs = gpu_ctx.convert(pkg.DC_BEARING, d, pkg.DC_IN_U16_DEPTH, pkg.DC_OUT_F32, want=want)
        for i, on in enumerate(want):
            assert (outs[i] is not None) == on
            if on:
                assert_float_parity("bearing", outs[i], oracle.feature(F_BEARING_H + i, rng, k))
    g, m = gpu_ctx.convert(pkg.DC_CURVATURE, d, pkg.DC_IN_U16_DEPTH, pkg.DC_OUT_F32,
                           want=(True, False))[:2]
    assert m is None
    assert_float_parity("gauss", g, oracle.feature(F_GAUSS, rng, k))
    g, m = gpu_ctx.convert(pkg.DC_CURVATURE, d, pkg.DC_IN_U16_DEPTH, pkg.DC_OUT_F32,
                           want=(False, True))[:2]
    assert g is None
    assert_float_parity("mean", m, oracle.feature(F_MEAN, rng, k))


def test_edge_inputs(pkg, gpu_ctx, oracle):
    """all-invalid frame, all-valid flat frame, maximum depth, single zero."""
    w, h = 64, 48
    k = kinect_scaled(w, h)
    set_k(gpu_ctx, k)
    frames = np.zeros((4, h, w), np.uint16)
    frames[1] = 1500
    frames[2] = 65535
    frames[3] = 1000
    frames[3, 20, 30] = 0
    f32 = gpu_all_features(pkg, gpu_ctx, frames, pkg.DC_IN_U16_DEPTH, pkg.DC_OUT_F32)
    u16 = gpu_all_features(pkg, gpu_ctx, frames, pkg.DC_IN_U16_DEPTH, pkg.DC_OUT_U16)
    for i in range(4):
        rng = oracle.laserscan(frames[i], k)
        for f, name in FEATURE_NAMES.items():
            want = oracle.feature(f, rng, k)
            assert_float_parity(name, f32[name][i], want)
            assert_u16_parity(name, u16[name][i], oracle.quantise(f, want, rng))
    assert not f32["flexion"][0].any() and not f32["max_curve"][0].any()
    # empty batch is a no-op
    job = gpu_ctx.make_job(pkg.DC_FLEXION, pkg.DC_IN_U16_DEPTH, pkg.DC_OUT_U16, 0,
                           frames.ctypes.data, [frames.ctypes.data])
    assert gpu_ctx.convert_host_raw(job) == pkg.DC_OK


def test_scale_and_pinhole_range_roundtrip(pkg, gpu_ctx, oracle):
    """depth2x scale (test_conversion_scaling.cpp:13-39: x8 saturates, +10000
    offset) and range -> `--type pinhole-range` flexion (test_depth2x_range.sh:58-76)."""
    w, h = 96, 64
    k = kinect_scaled(w, h)
    d = oracle.synth_depth(0xD2C00006, 0, w, h)
    d[3, 3] = 60000
    gpu_ctx.set_frame_size(w, h)
    for s, o in ((8.0, 0.0), (1.0, 10000.0), (0.37, -12.5)):
        got = gpu_ctx.convert(pkg.DC_SCALE, d, pkg.DC_IN_U16_RANGE, pkg.DC_OUT_U16,
                              scale=s, offset=o)[0]
        assert np.array_equal(got, oracle.depth_scaling(d, s, o))
    job = gpu_ctx.make_job(pkg.DC_FLEXION, pkg.DC_IN_U16_DEPTH, pkg.DC_OUT_U16, 1,
                           d.ctypes.data, [d.ctypes.data])
    assert gpu_ctx.convert_host_raw(job) == pkg.DC_ERR_NO_CAMERA
    set_k(gpu_ctx, k)
    rng_u16 = gpu_ctx.convert(pkg.DC_RANGE, d, pkg.DC_IN_U16_DEPTH, pkg.DC_OUT_U16)[0]
    fl = gpu_ctx.convert(pkg.DC_FLEXION, rng_u16, pkg.DC_IN_U16_RANGE, pkg.DC_OUT_U16)[0]
    _, _, want = oracle.cli_frame(F_FLEXION, 1, rng_u16, k)
    assert_u16_parity("flexion", fl, want)


def test_bad_jobs_are_rejected(pkg, gpu_ctx):
    P = pkg
    k = kinect_scaled(32, 16)
    set_k(gpu_ctx, k)
    buf = np.zeros((16, 32), np.uint16)
    p = buf.ctypes.data
    J = gpu_ctx.make_job
    assert gpu_ctx.convert_host_raw(J(P.DC_BEARING, 0, 1, 1, p, [])) == P.DC_ERR_INVALID
    assert gpu_ctx.convert_host_raw(J(99, 0, 1, 1, p, [p])) == P.DC_ERR_INVALID
    assert gpu_ctx.convert_host_raw(J(P.DC_FLEXION, 0, 1, 1, None, [p])) == P.DC_ERR_INVALID
    assert gpu_ctx.convert_host_raw(J(P.DC_SCALE, 0, 1, 1, p, [p])) == P.DC_ERR_UNSUPPORTED
    assert gpu_ctx.convert_host_raw(J(P.DC_CURVATURE, 0, 1, 1, p, [p], clamp=(3.0, 3.0))) \
        == P.DC_ERR_INVALID
    L = pkg.lib()
    assert L.dc_set_pinhole(gpu_ctx._h, 0, 16, 1.0, 1.0, 1.0, 1.0) == P.DC_ERR_INVALID
    assert L.dc_set_pinhole(gpu_ctx._h, 32, 16, 1.0, -1.0, 1.0, 1.0) == P.DC_ERR_INVALID
    assert b"pinhole" in L.dc_last_error(gpu_ctx._h)
    assert gpu_ctx.size == (32, 16)


@pytest.mark.parametrize("n", [1, 2, 3, 5])
def test_batch_composition_does_not_change_results(pkg, gpu_ctx, oracle, n):
    """The fast kernels convert two frames per thread (odd batches duplicate the
    last frame): every frame of a batch must equal the same frame converted
    alone, and the u16-input fast path must equal the one-frame kernels that
    the float-input library API uses (bit for bit where the arithmetic is
    IEEE-only)."""
    w, h = 256, 75
    k = kinect_scaled(w, h)
    set_k(gpu_ctx, k)
    frames = np.stack([oracle.synth_depth(0xD2C00009, f, w, h) for f in range(n)])
    together = gpu_all_features(pkg, gpu_ctx, frames, pkg.DC_IN_U16_DEPTH, pkg.DC_OUT_U16)
    together32 = gpu_all_features(pkg, gpu_ctx, frames, pkg.DC_IN_U16_DEPTH, pkg.DC_OUT_F32)
    rng = together32["range"]
    for i in range(n):
        alone = gpu_all_features(pkg, gpu_ctx, frames[i], pkg.DC_IN_U16_DEPTH, pkg.DC_OUT_U16)
        alone32 = gpu_all_features(pkg, gpu_ctx, frames[i], pkg.DC_IN_U16_DEPTH, pkg.DC_OUT_F32)
        scalar32 = gpu_all_features(pkg, gpu_ctx, rng[i], pkg.DC_IN_F32_RANGE, pkg.DC_OUT_F32)
        for name in together:
            assert np.array_equal(together[name][i], alone[name]), name
            assert np.array_equal(together32[name][i].view(np.uint32), alone32[name].view(np.uint32)), name
        for name in ("range", "flexion", "gauss", "range_fused"):
            assert np.array_equal(together32[name][i].view(np.uint32), scalar32[name].view(np.uint32)), name
        for name in ("bearing_h", "bearing_v", "bearing_d", "bearing_a", "max_curve", "mean"):
            assert_float_parity(name, together32[name][i], scalar32[name], exact=False)


@pytest.mark.parametrize("w,h", [(1920, 1080), (4096, 2160)])
def test_full_size_frames_against_oracle(pkg, gpu_ctx, oracle, w, h):
    """BASELINE configs 3 and 4 at their real frame sizes (uniform ray-angle
    tables): two frames, every feature, float and 16 bit."""
    k = kinect_scaled(w, h)
    set_k(gpu_ctx, k)
    frames = np.stack([oracle.synth_depth(0xD2C00003, f, w, h) for f in range(2)])
    f32 = gpu_all_features(pkg, gpu_ctx, frames, pkg.DC_IN_U16_DEPTH, pkg.DC_OUT_F32)
    u16 = gpu_all_features(pkg, gpu_ctx, frames, pkg.DC_IN_U16_DEPTH, pkg.DC_OUT_U16)
    for i in range(2):  # both halves of the frame pair
        rng = oracle.laserscan(frames[i], k)
        for f, name in FEATURE_NAMES.items():
            want = oracle.feature(f, rng, k)
            assert_float_parity(name, f32[name][i], want)
            assert_u16_parity(name, u16[name][i], oracle.quantise(f, want, rng))
        assert_u16_parity("range", u16["range_fused"][i], oracle.quantise(0, rng, rng))


def test_large_batch_properties(pkg, gpu_ctx, oracle):
    """BASELINE config 2 batch size (256 frames of 960x540) through the host API:
    size-independent properties instead of a 256-frame oracle run."""
    w, h, n = 960, 540, 256
    k = kinect_scaled(w, h)
    set_k(gpu_ctx, k)
    base = np.stack([oracle.synth_depth(0xD2C00002, f, w, h) for f in range(4)])
    frames = np.concatenate([base] * (n // 4))
    outs = gpu_ctx.convert(pkg.DC_BEARING, frames, pkg.DC_IN_U16_DEPTH, pkg.DC_OUT_U16, want=(True,) * 4)
    fl = gpu_ctx.convert(pkg.DC_FLEXION, frames, pkg.DC_IN_U16_DEPTH, pkg.DC_OUT_U16)[0]
    for o in outs + [fl]:
        o = o.reshape(n // 4, 4, h, w)
        assert (o == o[0]).all()  # identical frames give identical images wherever they sit
    # every distinct frame against the oracle, each at another place of the batch (with the
    # identity above that covers all 256 frames): both halves of a pair, first / last pair
    for b, at in ((0, 0), (1, 129), (2, 254), (3, 255), (0, 128), (1, 1), (2, 2), (3, 131)):
        rng = oracle.laserscan(base[b], k)
        for d in range(4):
            want = oracle.quantise(F_BEARING_H + d, oracle.feature(F_BEARING_H + d, rng, k), rng)
            assert_u16_parity("bearing", outs[d][at], want)
            # invalid-depth mask (depth_to_bearing.h:204-207): zero depth -> exactly 0
            assert not outs[d][at][base[b] == 0].any()
        assert_u16_parity("flexion", fl[at], oracle.cli_frame(F_FLEXION, 0, base[b], k)[2])
    # border stays 0 (depth_to_flexion.h:156-161)
    assert not fl[:, 0].any() and not fl[:, -1].any() and not fl[:, :, 0].any() and not fl[:, :, -1].any()


@pytest.mark.parametrize("w,h,n", [(4, 3, 1), (8, 37, 2), (64, 75, 3), (132, 36, 5), (260, 110, 2)])
def test_max_curve_shared_edges_small_uniform_cameras(pkg, gpu_ctx, oracle, w, h, n):
    """maxcurve_edge_kernel (one scalar neighbour angle: telephoto intrinsics clamp
    every phi() like the 1080p+ Kinect ones, camera_models/utility.h:39-45) at small,
    ragged sizes: strip seams (36 rows per strip), one thread group per row, odd
    batches, depth and range input, float and 16-bit output, fused range output."""
    k = dict(w=w, h=h, fx=2000.0, fy=2010.0, cx=w * 0.49, cy=h * 0.52)
    set_k(gpu_ctx, k)
    frames = np.stack([oracle.synth_depth(0xD2C00004, f, w, h) for f in range(n)])
    frames[0, h // 2, w // 2] = 0  # an invalid pixel inside every neighbourhood kind
    for in_kind, kind in ((pkg.DC_IN_U16_DEPTH, 0), (pkg.DC_IN_U16_RANGE, 1)):
        f32 = gpu_ctx.convert(pkg.DC_MAX_CURVE, frames, in_kind, pkg.DC_OUT_F32, want=(True, True))
        assert gpu_ctx.last_kernel == "dc::pr::maxcurve_edge_kernel<SPEC_UNIFORM>", gpu_ctx.last_kernel
        u16 = gpu_ctx.convert(pkg.DC_MAX_CURVE, frames, in_kind, pkg.DC_OUT_U16, want=(True, True))
        only = gpu_ctx.convert(pkg.DC_MAX_CURVE, frames, in_kind, pkg.DC_OUT_U16)[0]
        assert np.array_equal(only, u16[0])
        for i in range(n):
            rng = oracle.laserscan(frames[i], k) if kind == 0 else frames[i].astype(np.float32)
            want = oracle.feature(F_MAX_CURVE, rng, k)
            assert_float_parity("max_curve", f32[0][i], want)
            assert_u16_parity("max_curve", u16[0][i], oracle.quantise(F_MAX_CURVE, want, rng))
            assert_float_parity("range", f32[1][i], rng)
            assert_u16_parity("range", u16[1][i], oracle.quantise(0, rng, rng))
            # border stays 0 (depth_to_max_curve.h:101-102)
            assert not f32[0][i][0].any() and not f32[0][i][-1].any()
            assert not f32[0][i][:, 0].any() and not f32[0][i][:, -1].any()


def test_kernel_selection_is_reported(pkg, gpu_ctx, oracle):
    """dc_last_kernel: per-pixel angle tables (960x540 Kinect) take the frame-pair
    kernels with tables, even widths that are not a multiple of 4 their half-access
    form, odd widths the scalar kernel."""
    w, h = 960, 540
    set_k(gpu_ctx, kinect_scaled(w, h))
    frames = np.stack([oracle.synth_depth(0xD2C00002, f, w, h) for f in range(2)])
    gpu_ctx.convert(pkg.DC_BEARING, frames, pkg.DC_IN_U16_DEPTH, pkg.DC_OUT_U16, want=(True,) * 4)
    assert gpu_ctx.last_kernel == "dc::pr::pair_kernel<SPEC_TABLES>"
    gpu_ctx.convert(pkg.DC_MAX_CURVE, frames, pkg.DC_IN_U16_DEPTH, pkg.DC_OUT_U16)
    assert gpu_ctx.last_kernel == "dc::pr::maxcurve_edge_kernel<SPEC_TABLES>"
    gpu_ctx.convert(pkg.DC_FLEXION, frames, pkg.DC_IN_U16_DEPTH, pkg.DC_OUT_U16)
    assert gpu_ctx.last_kernel == "dc::pr::flexion_pair_kernel<LSB>"
    gpu_ctx.convert(pkg.DC_RANGE, frames, pkg.DC_IN_U16_DEPTH, pkg.DC_OUT_U16)
    assert gpu_ctx.last_kernel == "dc::pr::stream_kernel"
    set_k(gpu_ctx, kinect_scaled(322, 203))
    small = np.stack([oracle.synth_depth(0xD2C00002, 0, 322, 203)])
    gpu_ctx.convert(pkg.DC_BEARING, small, pkg.DC_IN_U16_DEPTH, pkg.DC_OUT_U16, want=(True,) * 4)
    assert gpu_ctx.last_kernel == "dc::pr::pair_kernel<SPEC_HALF>"  # even, not a multiple of 4
    set_k(gpu_ctx, kinect_scaled(61, 37))
    odd = np.stack([oracle.synth_depth(0xD2C00002, 0, 61, 37)])
    gpu_ctx.convert(pkg.DC_BEARING, odd, pkg.DC_IN_U16_DEPTH, pkg.DC_OUT_U16, want=(True,) * 4)
    assert gpu_ctx.last_kernel.startswith("dc::feature_kernel")


@pytest.mark.parametrize("w,h", [(360, 90), (128, 64), (61, 37)])
def test_equirectangular_float_against_compiled_reference(pkg, gpu_ctx, oracle, w, h):
    """Row f1 in float: every feature of the equirectangular camera model
    (camera_models/equirectangular.h:43-181) against the reference's own templates
    (oracle/_ref). The laser-scan input is a range image already (conversion/util.h:57-67).
    Non-pinhole angle tables: the table-driven kernels (per-pixel cosines differ between the
    two anti-diagonal angles of max-curve, depth_to_max_curve.h:136-139)."""
    from oracle_lib import RefLib, ref_available
    if not ref_available():
        pytest.skip("oracle/_ref is not built")
    ref = RefLib()
    tmin, tmax = float(np.float32(0.87)), float(np.float32(2.27))
    gpu_ctx.set_equirectangular(w, h, tmin, tmax)
    frames = np.stack([oracle.synth_depth(0xD2C00007, f, w, h) for f in range(3)])
    f32 = gpu_all_features(pkg, gpu_ctx, frames, pkg.DC_IN_U16_RANGE, pkg.DC_OUT_F32)
    lib = gpu_all_features(pkg, gpu_ctx, frames.astype(np.float32), pkg.DC_IN_F32_RANGE, pkg.DC_OUT_F32)
    for i in range(3):
        rng = np.ascontiguousarray(frames[i].astype(np.float32))
        for f, name in FEATURE_NAMES.items():
            if name == "range":
                continue
            want = np.empty((h, w), np.float32)
            ref.lib.ref_feature_equi_f32(f, rng.ctypes.data, w, h, tmin, tmax, want.ctypes.data)
            assert_float_parity(name, f32[name][i], want)
            assert_float_parity(name, lib[name][i], want)


def test_flexion_u16_lsb_kernel_and_exact_switch(pkg, gpu_ctx, oracle, tmp_path):
    """16-bit flexion (what depth2x writes) comes from the kernel that divides by the norms
    with MUFU.RSQ: within +-1 LSB of the reference (in fact < 0.2 LSB before rounding, so
    only pixels at a rounding boundary move). DC_EXACT_U16=1 selects the bit-exact kernel;
    the float image is bit exact either way."""
    import subprocess
    import sys as _sys
    w, h = 960, 540
    k = kinect_scaled(w, h)
    set_k(gpu_ctx, k)
    frames = np.stack([oracle.synth_depth(0xD2C000C1, f, w, h) for f in range(3)])
    got = gpu_ctx.convert(pkg.DC_FLEXION, frames, pkg.DC_IN_U16_DEPTH, pkg.DC_OUT_U16)[0]
    assert gpu_ctx.last_kernel == "dc::pr::flexion_pair_kernel<LSB>"
    f32 = gpu_ctx.convert(pkg.DC_FLEXION, frames, pkg.DC_IN_U16_DEPTH, pkg.DC_OUT_F32)[0]
    assert gpu_ctx.last_kernel == "dc::pr::flexion_pair_kernel"
    moved = 0
    for i in range(3):
        rng = oracle.laserscan(frames[i], k)
        want32 = oracle.feature(F_FLEXION, rng, k)
        assert_float_parity("flexion", f32[i], want32)
        want = oracle.quantise(F_FLEXION, want32, rng)
        d = np.abs(got[i].astype(np.int32) - want.astype(np.int32))
        assert d.max() <= 1
        moved += int((d > 0).sum())
        # before rounding the two kernels are < 0.2 LSB apart: a pixel can only move if the
        # exact value lies that close to a rounding boundary (k + 0.5)
        frac = np.abs((want32.astype(np.float64) * 65535.0) % 1.0 - 0.5)
        assert (frac[d > 0] < 0.2).all()
    assert moved < 0.1 * frames.size
    np.save(tmp_path / "frames.npy", frames)
    code = (
        "import sys, numpy as np; sys.path.insert(0, %r); sys.path.insert(0, %r)\n"
        "from conftest import load_pkg; pkg = load_pkg()\n"
        "fr = np.load(%r)\n"
        "with pkg.Context(0) as c:\n"
        "    c.set_pinhole(%d, %d, %r, %r, %r, %r)\n"
        "    out = c.convert(pkg.DC_FLEXION, fr, pkg.DC_IN_U16_DEPTH, pkg.DC_OUT_U16)[0]\n"
        "    assert c.last_kernel == 'dc::pr::flexion_pair_kernel', c.last_kernel\n"
        "np.save(%r, out)\n"
    ) % (os.path.dirname(os.path.abspath(__file__)), os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
         str(tmp_path / "frames.npy"), w, h, k["fx"], k["fy"], k["cx"], k["cy"], str(tmp_path / "exact.npy"))
    r = subprocess.run([_sys.executable, "-c", code], env=dict(os.environ, DC_EXACT_U16="1"),
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    exact = np.load(tmp_path / "exact.npy")
    for i in range(3):
        rng = oracle.laserscan(frames[i], k)
        assert np.array_equal(exact[i], oracle.quantise(F_FLEXION, oracle.feature(F_FLEXION, rng, k), rng))


@pytest.mark.parametrize("w,h", [(960, 540), (1920, 1080)])
def test_u16_kernels_at_depth_discontinuities(pkg, gpu_ctx, oracle, w, h):
    """The 16-bit bearing / max-curve kernels take nom / sqrt(rad) with an unrefined reciprocal
    (<= 1 ulp off in rare cases): worst where acos is steepest, |ratio| -> 1, i.e. where the
    chord to the neighbour runs along the ray. Frames made of such chords: depth jumps of up to
    60 m between neighbouring pixels in every direction, plus smooth parts and holes. 16-bit
    images within +-1 of the reference's, float images within 1e-5 (per-camera tables at
    960x540, one scalar angle at 1920x1080)."""
    k = kinect_scaled(w, h)
    set_k(gpu_ctx, k)
    rs = np.random.RandomState(w)
    base = oracle.synth_depth(0xD2C000D1, 0, w, h).astype(np.int64)
    frames = []
    for kind in range(4):
        f = base.copy()
        jump = rs.choice([0, 300, 5000, 20000, 60000], size=(h, w), p=[0.5, 0.2, 0.15, 0.1, 0.05])
        if kind == 1:   # vertical stripes: every horizontal / diagonal chord is a jump
            jump = np.where((np.arange(w)[None, :] // 1) % 2 == 0, 0, 40000) + 0 * jump
        elif kind == 2:  # checkerboard
            jump = np.where((np.arange(w)[None, :] + np.arange(h)[:, None]) % 2 == 0, 0, 55000)
        elif kind == 3:  # near and far values next to each other, tiny near values
            f = np.where(rs.rand(h, w) < 0.5, 1 + (f % 7), 65535 - (f % 11))
            jump = 0 * jump
        f = np.where(base == 0, 0, np.clip(f + jump, 0, 65535))
        frames.append(f.astype(np.uint16))
    frames = np.stack(frames)
    b16 = gpu_ctx.convert(pkg.DC_BEARING, frames, pkg.DC_IN_U16_DEPTH, pkg.DC_OUT_U16, want=(True,) * 4)
    b32 = gpu_ctx.convert(pkg.DC_BEARING, frames, pkg.DC_IN_U16_DEPTH, pkg.DC_OUT_F32, want=(True,) * 4)
    m16 = gpu_ctx.convert(pkg.DC_MAX_CURVE, frames, pkg.DC_IN_U16_DEPTH, pkg.DC_OUT_U16)[0]
    fl16 = gpu_ctx.convert(pkg.DC_FLEXION, frames, pkg.DC_IN_U16_DEPTH, pkg.DC_OUT_U16)[0]
    steep = 0
    for i in range(frames.shape[0]):
        rng = oracle.laserscan(frames[i], k)
        for d in range(4):
            want = oracle.feature(F_BEARING_H + d, rng, k)
            assert_float_parity("bearing", b32[d][i], want)
            assert_u16_parity("bearing", b16[d][i], oracle.quantise(F_BEARING_H + d, want, rng))
            steep += int(((want > 0) & ((want < 0.01) | (want > np.pi - 0.01))).sum())
        want = oracle.feature(F_MAX_CURVE, rng, k)
        assert_u16_parity("max_curve", m16[i], oracle.quantise(F_MAX_CURVE, want, rng))
        assert_u16_parity("flexion", fl16[i], oracle.cli_frame(F_FLEXION, 0, frames[i], k)[2])
    assert steep > 10000, steep  # the frames do exercise the steep end of acos
