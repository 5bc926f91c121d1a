"""Comparison helpers shared by the parity tests.

Tolerances (BASELINE.json north_star): float feature values within 1e-5
relative, invalid-depth (exact zero) masks bit-exact, quantised 16-bit images
within +-1 LSB. Features whose arithmetic is IEEE +,-,*,/,sqrt only (range,
flexion, Gaussian curvature) are held to BIT-EXACT equality of their float images; bearing /
max-curve end in acosf and mean curvature in pow(x, 3), where the device and
glibc may differ by a few ulp.
"""
import numpy as np

REL_TOL = 1e-5
BIT_EXACT = {"range", "flexion", "gauss"}
# 16-bit images the KERNELS write: flexion is held to the contract's +-1 LSB, not to
# equality, since the 16-bit flexion kernel divides by the norms with MUFU.RSQ
# (dc_pair_kernels.cuh "flexion for 16-bit output"; DC_EXACT_U16=1 restores equality).
# Its float output stays bit exact (BIT_EXACT above).
U16_EXACT = {"range", "gauss"}


def bits(a):
    a = np.ascontiguousarray(a)
    return a.view(np.uint32) if a.dtype == np.float32 else a


def assert_float_parity(name, got, want, exact=None):
    assert got.shape == want.shape and got.dtype == np.float32 and want.dtype == np.float32
    exact = (name in BIT_EXACT) if exact is None else exact
    # NaN/invalid masks: exact-zero pixels and non-finite pixels must coincide
    assert np.array_equal(got == 0, want == 0), f"{name}: zero mask differs"
    assert np.array_equal(np.isnan(got), np.isnan(want)), f"{name}: NaN mask differs"
    eq = bits(got) == bits(want)
    if exact:
        assert eq.all(), f"{name}: {(~eq).sum()} of {eq.size} pixels not bit-exact"
        return 1.0
    fin = np.isfinite(want)
    g, w = got[fin].astype(np.float64), want[fin].astype(np.float64)
    rel = np.abs(g - w) / np.maximum(np.abs(w), 1e-30)
    rel[w == 0] = 0
    assert rel.size == 0 or rel.max() <= REL_TOL, \
        f"{name}: max relative error {rel.max():.3e} > {REL_TOL}"
    return float(eq.mean())


def assert_u16_parity(name, got, want, exact=None):
    assert got.shape == want.shape and got.dtype == np.uint16 and want.dtype == np.uint16
    exact = (name in U16_EXACT) if exact is None else exact
    d = np.abs(got.astype(np.int32) - want.astype(np.int32))
    if exact:
        assert d.max() == 0, f"{name}: u16 image differs at {(d > 0).sum()} pixels"
    else:
        assert d.max() <= 1, f"{name}: u16 image off by {d.max()} LSB at {(d > 1).sum()} pixels"
    return float((d == 0).mean())
