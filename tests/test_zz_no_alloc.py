"""GPU: dc_curvature_to_image_noalloc (include/depthconv_b200.h), the entry point
for callers that must not allocate: same images as dc_quantise_host_ex with
DC_Q_CURVATURE and no clamps, pixel for pixel (16 bit integers: tolerance 0), on
calls in a row that share the library's fixed scratch. The device buffers are the
CALLER's (torch tensors here); the entry point itself allocates nothing.

The file name sorts last on purpose: pytest -x stops at the
first failure, and a failure here must not hide another test."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def images(seed, n, spread):
    """A curvature-like image (heavy tails, some NaN) and a range image whose
    values below 0.5 blank the result."""
    rs = np.random.RandomState(seed)
    curv = (rs.standard_normal(n) * spread).astype(np.float32)
    curv[rs.randint(0, n, max(1, n // 97))] *= np.float32(50.0)
    curv[rs.randint(0, n, max(1, n // 211))] = np.nan
    rng = rs.uniform(0.0, 6.0, n).astype(np.float32)
    rng[rs.randint(0, n, max(1, n // 13))] = 0.0
    rng[rs.randint(0, n, max(1, n // 17))] = np.float32(0.5)  # cvRound(0.5) = 0: blanked
    return curv, rng


def no_alloc(pkg, curv, rng, prefill=0x5A5A):
    import torch
    n = curv.size
    d_curv = torch.from_numpy(curv).cuda()
    d_rng = torch.from_numpy(rng).cuda()
    d_out = torch.full((n,), prefill, dtype=torch.int16, device="cuda")
    torch.cuda.synchronize()
    rc, msg = pkg.curvature_to_image_noalloc(0, d_curv.data_ptr(), d_rng.data_ptr(), n, d_out.data_ptr())
    return rc, msg, d_out.cpu().numpy().view(np.uint16)


@pytest.mark.parametrize("first,second", [((1000 * 700, 3.0), (37 * 23, 0.01)),
                                          ((4096 * 3 + 5, 0.02), (1920 * 1080, 8.0))])
def test_two_calls_in_a_row_match_the_allocating_entry_point(pkg, gpu_ctx, first, second):
    # the second image has other bounds and another CTA count than the first:
    # anything the first call left in the scratch would show
    for k, (n, spread) in enumerate((first, second, first)):
        curv, rng = images(0xA110C + k, n, spread)
        want = gpu_ctx.quantise_ex(pkg.DC_Q_CURVATURE, curv, mask=rng)
        rc, msg, got = no_alloc(pkg, curv, rng)
        assert rc == pkg.DC_OK, msg
        assert got.dtype == want.dtype and got.shape == want.shape
        assert np.array_equal(got, want), f"call {k}: {(got != want).sum()} of {n} pixels differ"
        assert want.max() > 20000 and (want == 0).sum() > n // 20  # a real image, blanked pixels


def test_maximum_size_and_beyond(pkg, gpu_ctx):
    n = pkg.DC_NOALLOC_MAX_PIXELS
    curv, rng = images(0xB16, n, 1.5)
    want = gpu_ctx.quantise_ex(pkg.DC_Q_CURVATURE, curv, mask=rng)
    rc, msg, got = no_alloc(pkg, curv, rng)
    assert rc == pkg.DC_OK, msg
    assert np.array_equal(got, want)
    # one pixel more than the scratch provides for: refused before any GPU work
    rc, msg = pkg.curvature_to_image_noalloc(0, 1, 1, n + 1, 1)
    assert rc == pkg.DC_ERR_INVALID
    assert "DC_NOALLOC_MAX_PIXELS" in msg and str(n + 1) in msg


def test_errors_leave_the_scratch_usable(pkg, gpu_ctx):
    const = np.full(5000, 3.0, np.float32)
    rc, msg, got = no_alloc(pkg, const, const)
    assert rc == pkg.DC_ERR_INVALID and "minimum" in msg  # as dc_quantise_host_ex: lower < upper
    assert (got == 0x5A5A).all()  # nothing written
    rc, msg, _ = no_alloc(pkg, np.full(64, np.nan, np.float32), np.ones(64, np.float32))
    assert rc == pkg.DC_ERR_INVALID
    rc, msg = pkg.curvature_to_image_noalloc(0, 0, 0, 16, 0)
    assert rc == pkg.DC_ERR_INVALID and "NULL" in msg
    rc, msg = pkg.curvature_to_image_noalloc(4096, 1, 1, 16, 1)
    assert rc == pkg.DC_ERR_NO_DEVICE
    # the status word of the failed calls does not leak into the next one
    curv, rng = images(7, 12345, 2.0)
    rc, msg, got = no_alloc(pkg, curv, rng)
    assert rc == pkg.DC_OK, msg
    assert np.array_equal(got, gpu_ctx.quantise_ex(pkg.DC_Q_CURVATURE, curv, mask=rng))
