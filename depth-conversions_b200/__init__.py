"""ctypes binding of the depthconv_b200 C ABI (include/depthconv_b200.h).

Used by the tests, bench.py and __graft_entry__.py. The directory name carries
a hyphen, so import it with `importlib` (see `tests/conftest.py::load_pkg`) or
via `__graft_entry__.load_pkg()`; the module registers itself as
`depth_conversions_b200`.

There is no fallback: if the CUDA library is missing or no B200 is present,
loading / context creation raises.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("DC_LIB") or os.path.join(HERE, "lib", "libdepthconv_b200.so")
HOST_LIB_PATH = os.path.join(HERE, "lib", "libdc_host.so")
DEPTH2X_PATH = os.path.join(HERE, "bin", "depth2x")

# enums of include/depthconv_b200.h
DC_RANGE, DC_BEARING, DC_FLEXION, DC_CURVATURE, DC_MAX_CURVE, DC_SCALE = range(6)
DC_IN_U16_DEPTH, DC_IN_U16_RANGE, DC_IN_F32_RANGE, DC_IN_F64_RANGE = range(4)
DC_OUT_F32, DC_OUT_U16, DC_OUT_F64 = range(3)
DC_OK, DC_ERR_INVALID, DC_ERR_NO_DEVICE, DC_ERR_CUDA = 0, -1, -2, -3
DC_ERR_NO_CAMERA, DC_ERR_UNSUPPORTED, DC_ERR_NOMEM = -4, -5, -6

EXPORTED_SYMBOLS = [
    "dc_version", "dc_device_count", "dc_create", "dc_destroy", "dc_last_error",
    "dc_set_pinhole", "dc_set_equirectangular", "dc_set_frame_size", "dc_width",
    "dc_height", "dc_convert_device", "dc_convert_host", "dc_synchronize",
    "dc_host_alloc", "dc_host_free", "dc_synth_depth_device", "dc_kernel_launches", "dc_last_kernel",
    "dc_quantise_host", "dc_set_pinhole_f64", "dc_set_equirectangular_f64",
    "dc_quantise_host_f64", "dc_median_host", "dc_median_device", "dc_f32_fallbacks",
    "dc_quantise_host_ex", "dc_curvature_to_image_noalloc",
    "dc_gaussian_blur_host", "dc_gaussian_blur_device", "dc_bilateral_host", "dc_bilateral_device",
]
DC_NOALLOC_MAX_PIXELS = 1 << 24
DC_PX_U16, DC_PX_U8, DC_PX_I8, DC_PX_I16 = range(4)
_PX_DTYPE = {DC_PX_U16: np.uint16, DC_PX_U8: np.uint8, DC_PX_I8: np.int8, DC_PX_I16: np.int16}
DC_Q_RANGE, DC_Q_BEARING, DC_Q_FLEXION, DC_Q_CURVATURE, DC_Q_MAX_CURVE = range(5)


class DcJob(C.Structure):
    _fields_ = [
        ("feature", C.c_int32), ("in_kind", C.c_int32), ("out_kind", C.c_int32),
        ("n_frames", C.c_int32), ("in_", C.c_void_p), ("out", C.c_void_p * 4),
        ("clamp_lo", C.c_float), ("clamp_hi", C.c_float),
        ("scale", C.c_double), ("offset", C.c_double), ("median_ksize", C.c_int32),
    ]


class DcError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"depthconv_b200 error {code}: {msg}")
        self.code = code


_lib = None


def lib():
    """Load libdepthconv_b200.so (raises if it has not been built)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise FileNotFoundError(
                f"{LIB_PATH} is missing: run `make -C {HERE}` "
                "(__graft_entry__.build()); there is no CPU fallback")
        L = C.CDLL(LIB_PATH)
        L.dc_version.restype = C.c_char_p
        L.dc_last_error.restype = C.c_char_p
        L.dc_last_error.argtypes = [C.c_void_p]
        L.dc_create.argtypes = [C.c_int, C.POINTER(C.c_void_p)]
        L.dc_destroy.argtypes = [C.c_void_p]
        L.dc_destroy.restype = None
        L.dc_set_pinhole.argtypes = [C.c_void_p, C.c_int, C.c_int] + [C.c_float] * 4
        L.dc_set_equirectangular.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_float, C.c_float]
        L.dc_set_frame_size.argtypes = [C.c_void_p, C.c_int, C.c_int]
        L.dc_width.argtypes = [C.c_void_p]
        L.dc_height.argtypes = [C.c_void_p]
        L.dc_convert_device.argtypes = [C.c_void_p, C.POINTER(DcJob), C.c_void_p]
        L.dc_convert_host.argtypes = [C.c_void_p, C.POINTER(DcJob)]
        L.dc_synchronize.argtypes = [C.c_void_p]
        L.dc_host_alloc.restype = C.c_void_p
        L.dc_host_alloc.argtypes = [C.c_size_t]
        L.dc_host_free.argtypes = [C.c_void_p]
        L.dc_host_free.restype = None
        L.dc_synth_depth_device.argtypes = [C.c_void_p, C.c_uint64, C.c_int, C.c_int,
                                            C.c_void_p, C.c_void_p]
        L.dc_quantise_host.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_size_t,
                                       C.c_void_p, C.c_float, C.c_float]
        L.dc_set_pinhole_f64.argtypes = [C.c_void_p, C.c_int, C.c_int] + [C.c_double] * 4
        L.dc_set_equirectangular_f64.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_double,
                                                 C.c_double]
        L.dc_quantise_host_f64.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p,
                                           C.c_size_t, C.c_void_p, C.c_double, C.c_double]
        L.dc_median_host.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        L.dc_median_device.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int,
                                       C.c_void_p, C.c_void_p]
        L.dc_gaussian_blur_host.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int,
                                            C.c_double, C.c_double, C.c_void_p]
        L.dc_gaussian_blur_device.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int,
                                              C.c_double, C.c_double, C.c_void_p, C.c_void_p]
        L.dc_bilateral_host.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int,
                                        C.c_double, C.c_double, C.c_void_p]
        L.dc_bilateral_device.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int,
                                          C.c_double, C.c_double, C.c_void_p, C.c_void_p]
        L.dc_kernel_launches.restype = C.c_uint64
        L.dc_kernel_launches.argtypes = [C.c_void_p]
        L.dc_last_kernel.restype = C.c_char_p
        L.dc_last_kernel.argtypes = [C.c_void_p]
        L.dc_quantise_host_ex.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p,
                                          C.c_void_p, C.c_size_t, C.c_void_p,
                                          C.POINTER(C.c_double), C.POINTER(C.c_double)]
        L.dc_f32_fallbacks.restype = C.c_uint64
        L.dc_f32_fallbacks.argtypes = [C.c_void_p]
        L.dc_curvature_to_image_noalloc.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_size_t,
                                                    C.c_void_p, C.c_void_p, C.c_char_p, C.c_size_t]
        _lib = L
    return _lib


def curvature_to_image_noalloc(device, curvature_ptr, range_ptr, n_px, out_ptr, stream=None):
    """dc_curvature_to_image_noalloc on DEVICE pointers (ints): -> (status, message).
    Allocates nothing on the GPU; every call on a GPU shares one fixed scratch."""
    err = C.create_string_buffer(256)
    rc = lib().dc_curvature_to_image_noalloc(device, curvature_ptr, range_ptr, n_px, out_ptr,
                                             stream, err, len(err))
    return rc, err.value.decode()


_IN_DTYPE = {DC_IN_U16_DEPTH: np.uint16, DC_IN_U16_RANGE: np.uint16, DC_IN_F32_RANGE: np.float32,
             DC_IN_F64_RANGE: np.float64}
_OUT_DTYPE = {DC_OUT_F32: np.float32, DC_OUT_U16: np.uint16, DC_OUT_F64: np.float64}


class Context:
    """One context per (thread, GPU)."""

    def __init__(self, device=0):
        self._L = lib()
        h = C.c_void_p()
        rc = self._L.dc_create(device, C.byref(h))
        if rc != DC_OK:
            raise DcError(rc, self._L.dc_last_error(None).decode())
        self._h = h

    def close(self):
        if getattr(self, "_h", None):
            self._L.dc_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _check(self, rc):
        if rc != DC_OK:
            raise DcError(rc, self._L.dc_last_error(self._h).decode())

    def status(self, rc):
        return rc, self._L.dc_last_error(self._h).decode()

    # -- camera ------------------------------------------------------------
    def set_pinhole(self, w, h, fx, fy, cx, cy):
        self._check(self._L.dc_set_pinhole(self._h, w, h, fx, fy, cx, cy))

    def set_pinhole_f64(self, w, h, fx, fy, cx, cy):
        self._check(self._L.dc_set_pinhole_f64(self._h, w, h, fx, fy, cx, cy))

    def set_equirectangular_f64(self, w, h, theta_min, theta_max):
        self._check(self._L.dc_set_equirectangular_f64(self._h, w, h, theta_min, theta_max))

    def set_equirectangular(self, w, h, theta_min, theta_max):
        self._check(self._L.dc_set_equirectangular(self._h, w, h, theta_min, theta_max))

    def set_frame_size(self, w, h):
        self._check(self._L.dc_set_frame_size(self._h, w, h))

    @property
    def size(self):
        return self._L.dc_width(self._h), self._L.dc_height(self._h)

    @property
    def kernel_launches(self):
        return int(self._L.dc_kernel_launches(self._h))

    @property
    def last_kernel(self):
        return self._L.dc_last_kernel(self._h).decode()

    @property
    def f32_fallbacks(self):
        """Batches of f32 range input the IEEE kernels had to redo (dc_f32_fallbacks)."""
        return int(self._L.dc_f32_fallbacks(self._h))

    def synchronize(self):
        self._check(self._L.dc_synchronize(self._h))

    # -- raw calls -----------------------------------------------------------
    def make_job(self, feature, in_kind, out_kind, n_frames, in_ptr, out_ptrs,
                 clamp=(-20.0, 20.0), scale=1.0, offset=0.0, median=0):
        j = DcJob()
        j.feature, j.in_kind, j.out_kind, j.n_frames = feature, in_kind, out_kind, n_frames
        j.in_ = in_ptr
        for k in range(4):
            j.out[k] = out_ptrs[k] if k < len(out_ptrs) and out_ptrs[k] else None
        j.clamp_lo, j.clamp_hi = clamp
        j.scale, j.offset = scale, offset
        j.median_ksize = median
        return j

    def convert_device_raw(self, job, stream=None):
        return self._L.dc_convert_device(self._h, C.byref(job), stream)

    def convert_host_raw(self, job):
        return self._L.dc_convert_host(self._h, C.byref(job))

    def convert_device(self, job, stream=None):
        self._check(self.convert_device_raw(job, stream))

    # -- numpy convenience (host buffers) -----------------------------------
    def convert(self, feature, frames, in_kind, out_kind=DC_OUT_F32, want=(True,),
                clamp=(-20.0, 20.0), scale=1.0, offset=0.0, median=0):
        """frames: (n, h, w) or (h, w) array of the in_kind dtype. `want[k]`
        selects output slot k. Returns a list with an array (same leading shape)
        per wanted slot, None for the others."""
        frames = np.ascontiguousarray(frames, _IN_DTYPE[in_kind])
        single = frames.ndim == 2
        f3 = frames[None] if single else frames
        n = f3.shape[0]
        outs = [np.empty(f3.shape, _OUT_DTYPE[out_kind]) if (k < len(want) and want[k]) else None
                for k in range(4)]
        job = self.make_job(feature, in_kind, out_kind, n, f3.ctypes.data,
                            [o.ctypes.data if o is not None else None for o in outs],
                            clamp, scale, offset, median)
        self._check(self.convert_host_raw(job))
        return [(o[0] if single else o) if o is not None else None for o in outs]

    def median(self, frames, ksize):
        """preprocess::median_blur on (n, h, w) or (h, w) uint16 / float32 frames."""
        frames = np.ascontiguousarray(frames)
        kind = {np.dtype(np.uint16): DC_IN_U16_RANGE, np.dtype(np.float32): DC_IN_F32_RANGE}[frames.dtype]
        n = 1 if frames.ndim == 2 else frames.shape[0]
        out = np.empty_like(frames)
        self._check(self._L.dc_median_host(self._h, kind, frames.ctypes.data, n, ksize,
                                           out.ctypes.data))
        return out

    @staticmethod
    def _filter_args(frames):
        frames = np.ascontiguousarray(frames)
        kind = {np.dtype(np.uint16): DC_IN_U16_RANGE, np.dtype(np.float32): DC_IN_F32_RANGE}[frames.dtype]
        return frames, kind, (1 if frames.ndim == 2 else frames.shape[0]), np.empty_like(frames)

    def gaussian_blur(self, frames, ksize, sigma_x, sigma_y=0.0):
        """preprocess::gaussian_blur on (n, h, w) or (h, w) uint16 / float32 frames;
        ksize = (width, height)."""
        frames, kind, n, out = self._filter_args(frames)
        self._check(self._L.dc_gaussian_blur_host(self._h, kind, frames.ctypes.data, n, ksize[0], ksize[1],
                                                  sigma_x, sigma_y, out.ctypes.data))
        return out

    def bilateral(self, frames, d, sigma_color, sigma_space):
        """preprocess::bilateral_filter on (n, h, w) or (h, w) uint16 / float32 frames."""
        frames, kind, n, out = self._filter_args(frames)
        self._check(self._L.dc_bilateral_host(self._h, kind, frames.ctypes.data, n, d, sigma_color,
                                              sigma_space, out.ctypes.data))
        return out

    def quantise(self, quant, feature, mask=None, clamp=(-20.0, 20.0)):
        """Stand-alone quantiser of the library API: float (or double) image -> u16 image."""
        if np.asarray(feature).dtype == np.float64:
            feature = np.ascontiguousarray(feature, np.float64)
            out = np.empty(feature.shape, np.uint16)
            m = np.ascontiguousarray(mask, np.float64) if mask is not None else None
            self._check(self._L.dc_quantise_host_f64(
                self._h, quant, feature.ctypes.data, m.ctypes.data if m is not None else None,
                feature.size, out.ctypes.data, clamp[0], clamp[1]))
            return out
        feature = np.ascontiguousarray(feature, np.float32)
        out = np.empty(feature.shape, np.uint16)
        m = np.ascontiguousarray(mask, np.float32) if mask is not None else None
        self._check(self._L.dc_quantise_host(self._h, quant, feature.ctypes.data,
                                             m.ctypes.data if m is not None else None,
                                             feature.size, out.ctypes.data, clamp[0], clamp[1]))
        return out

    def quantise_ex_raw(self, quant, feature, mask=None, pixel=DC_PX_U16, lo=None, hi=None):
        """dc_quantise_host_ex: any pixel type; lo / hi None = std::nullopt (curvature:
        the image's own minimum / maximum). Returns (status, image)."""
        feature = np.ascontiguousarray(feature)
        real = DC_OUT_F64 if feature.dtype == np.float64 else DC_OUT_F32
        feature = feature.astype(np.float64 if real == DC_OUT_F64 else np.float32, copy=False)
        m = np.ascontiguousarray(mask, feature.dtype) if mask is not None else None
        out = np.zeros(feature.shape, _PX_DTYPE.get(pixel, np.uint16))
        plo = C.byref(C.c_double(lo)) if lo is not None else None
        phi = C.byref(C.c_double(hi)) if hi is not None else None
        rc = self._L.dc_quantise_host_ex(self._h, quant, pixel, real, feature.ctypes.data,
                                         m.ctypes.data if m is not None else None, feature.size,
                                         out.ctypes.data, plo, phi)
        return rc, out

    def quantise_ex(self, quant, feature, mask=None, pixel=DC_PX_U16, lo=None, hi=None):
        rc, out = self.quantise_ex_raw(quant, feature, mask, pixel, lo, hi)
        self._check(rc)
        return out

    def synth_depth_device(self, seed, frame0, n_frames, out_ptr, stream=None):
        self._check(self._L.dc_synth_depth_device(self._h, seed, frame0, n_frames,
                                                  out_ptr, stream))
