"""ctypes bindings for the CHECKERS under oracle/ (test infrastructure only).

`Oracle`  -> oracle/liboracle.so        plain-C restatement (oracle/dc_oracle.c)
`RefLib`  -> oracle/_ref/liboracle_ref.so  the reference's own headers compiled
             against the cv::Mat shim (oracle/ref_bridge.cpp); present only when
             it was built in a container that mounts /root/reference.
"""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")

# feature ids shared by ref_bridge.cpp and the tests
F_RANGE, F_BEARING_H, F_BEARING_V, F_BEARING_D, F_BEARING_A = 0, 1, 2, 3, 4
F_FLEXION, F_GAUSS, F_MEAN, F_MAX_CURVE = 5, 6, 7, 8
FEATURE_NAMES = {
    F_RANGE: "range", F_BEARING_H: "bearing_h", F_BEARING_V: "bearing_v",
    F_BEARING_D: "bearing_d", F_BEARING_A: "bearing_a", F_FLEXION: "flexion",
    F_GAUSS: "gauss", F_MEAN: "mean", F_MAX_CURVE: "max_curve",
}

KINECT = dict(w=960, h=540, fx=519.226, fy=522.23, cx=479.462, cy=272.737)


def kinect_scaled(w, h):
    """Kinect CLI fixture intrinsic (src/test/depth2x/kinect_intrinsic.txt)
    scaled to w x h, as float32 values (DESIGN.md 'Synthetic input')."""
    f = np.float32
    return dict(w=w, h=h,
                fx=float(f(519.226 * w / 960)), fy=float(f(522.23 * h / 540)),
                cx=float(f(479.462 * w / 960)), cy=float(f(272.737 * h / 540)))


class Pinhole(C.Structure):
    _fields_ = [("w", C.c_int), ("h", C.c_int), ("fx", C.c_float),
                ("fy", C.c_float), ("cx", C.c_float), ("cy", C.c_float)]


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


def build_oracle():
    so = os.path.join(ORACLE_DIR, "liboracle.so")
    src = os.path.join(ORACLE_DIR, "dc_oracle.c")
    if (not os.path.exists(so)) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", ORACLE_DIR, "liboracle.so"],
                              stdout=subprocess.DEVNULL)
    return so


class Oracle:
    """Plain-C restatement. All images are C-contiguous numpy arrays (h, w)."""

    def __init__(self):
        self.lib = C.CDLL(build_oracle())
        L = self.lib
        L.dco_phi.restype = C.c_float
        L.dco_phi.argtypes = [C.c_void_p] + [C.c_int] * 4
        L.dco_bearing_angle.restype = C.c_float
        L.dco_bearing_angle.argtypes = [C.c_float] * 3
        L.dco_average_pixel_error_u16.restype = C.c_double
        L.dco_average_pixel_error_u16.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
        L.dco_synth_depth.argtypes = [C.c_uint64, C.c_int, C.c_int, C.c_int, C.c_void_p]
        L.dco_curvature_to_image_u16.argtypes = [
            C.c_void_p, C.c_void_p, C.c_size_t, C.c_float, C.c_float, C.c_void_p]
        L.dco_depth_scaling_u16.argtypes = [
            C.c_void_p, C.c_size_t, C.c_double, C.c_double, C.c_void_p]
        for n in ("dco_convert_bearing_u16", "dco_convert_flexion_u16",
                  "dco_convert_max_curve_u16", "dco_convert_range_u16",
                  "dco_u16_to_f32"):
            getattr(L, n).argtypes = [C.c_void_p, C.c_size_t, C.c_void_p]

    @staticmethod
    def K(k):
        return Pinhole(k["w"], k["h"], k["fx"], k["fy"], k["cx"], k["cy"])

    def pixel_to_sphere(self, k, u, v):
        s = np.zeros(3, np.float32)
        K = self.K(k)
        self.lib.dco_pixel_to_sphere(C.byref(K), C.c_int(u), C.c_int(v), _ptr(s))
        return s

    def phi(self, k, u1, v1, u2, v2):
        K = self.K(k)
        return self.lib.dco_phi(C.byref(K), u1, v1, u2, v2)

    def bearing_angle(self, b, c, cos_alpha):
        return self.lib.dco_bearing_angle(b, c, cos_alpha)

    def laserscan(self, depth_u16, k):
        assert depth_u16.dtype == np.uint16 and depth_u16.shape == (k["h"], k["w"])
        depth_u16 = np.ascontiguousarray(depth_u16)
        out = np.empty(depth_u16.shape, np.float32)
        K = self.K(k)
        self.lib.dco_laserscan(_ptr(depth_u16), C.byref(K), _ptr(out))
        return out

    def u16_to_f32(self, img):
        img = np.ascontiguousarray(img)
        out = np.empty(img.shape, np.float32)
        self.lib.dco_u16_to_f32(_ptr(img), img.size, _ptr(out))
        return out

    def feature(self, f, range_f32, k):
        assert range_f32.dtype == np.float32 and range_f32.shape == (k["h"], k["w"])
        range_f32 = np.ascontiguousarray(range_f32)
        out = np.empty(range_f32.shape, np.float32)
        K = self.K(k)
        L = self.lib
        if f == F_RANGE:
            return range_f32.copy()
        if F_BEARING_H <= f <= F_BEARING_A:
            L.dco_bearing(C.c_int(f - F_BEARING_H), _ptr(range_f32), C.byref(K), _ptr(out))
        elif f == F_FLEXION:
            L.dco_flexion(_ptr(range_f32), C.byref(K), _ptr(out))
        elif f == F_GAUSS:
            L.dco_gaussian_curvature(_ptr(range_f32), C.byref(K), _ptr(out))
        elif f == F_MEAN:
            L.dco_mean_curvature(_ptr(range_f32), C.byref(K), _ptr(out))
        elif f == F_MAX_CURVE:
            L.dco_max_curve(_ptr(range_f32), C.byref(K), _ptr(out))
        else:
            raise ValueError(f)
        return out

    def quantise(self, f, feat, range_f32, lo=-20.0, hi=20.0):
        feat = np.ascontiguousarray(feat, np.float32)
        out = np.empty(feat.shape, np.uint16)
        L = self.lib
        if f == F_RANGE:
            L.dco_convert_range_u16(_ptr(feat), feat.size, _ptr(out))
        elif F_BEARING_H <= f <= F_BEARING_A:
            L.dco_convert_bearing_u16(_ptr(feat), feat.size, _ptr(out))
        elif f == F_FLEXION:
            L.dco_convert_flexion_u16(_ptr(feat), feat.size, _ptr(out))
        elif f == F_MAX_CURVE:
            L.dco_convert_max_curve_u16(_ptr(feat), feat.size, _ptr(out))
        elif f in (F_GAUSS, F_MEAN):
            rng = np.ascontiguousarray(range_f32, np.float32)
            L.dco_curvature_to_image_u16(_ptr(feat), _ptr(rng), feat.size,
                                         C.c_float(lo), C.c_float(hi), _ptr(out))
        else:
            raise ValueError(f)
        return out

    def depth_scaling(self, depth_u16, scale, offset):
        depth_u16 = np.ascontiguousarray(depth_u16)
        out = np.empty(depth_u16.shape, np.uint16)
        self.lib.dco_depth_scaling_u16(_ptr(depth_u16), depth_u16.size,
                                       scale, offset, _ptr(out))
        return out

    def cli_frame(self, f, in_kind, depth_u16, k, lo=-20.0, hi=20.0):
        """preprocess_depth + depth_to_X + quantiser, as one depth2x index."""
        rng = self.laserscan(depth_u16, k) if in_kind == 0 else self.u16_to_f32(depth_u16)
        feat = self.feature(f, rng, k)
        return rng, feat, self.quantise(f, feat, rng, lo, hi)

    def synth_depth(self, seed, frame, w, h):
        out = np.empty((h, w), np.uint16)
        self.lib.dco_synth_depth(C.c_uint64(seed), frame, w, h, _ptr(out))
        return out

    def average_pixel_error(self, a, b):
        a = np.ascontiguousarray(a, np.uint16)
        b = np.ascontiguousarray(b, np.uint16)
        return self.lib.dco_average_pixel_error_u16(_ptr(a), _ptr(b), a.size)


REF_SO = os.path.join(ORACLE_DIR, "_ref", "liboracle_ref.so")


def ref_available():
    return os.path.exists(REF_SO)


class RefLib:
    """The compiled reference headers (float path + double path)."""

    def __init__(self):
        self.lib = C.CDLL(REF_SO)
        L = self.lib
        L.ref_build_info.restype = C.c_char_p
        f4, d4 = [C.c_float] * 4, [C.c_double] * 4
        L.ref_laserscan_f32.argtypes = [C.c_void_p, C.c_int, C.c_int] + f4 + [C.c_void_p]
        L.ref_laserscan_f64.argtypes = [C.c_void_p, C.c_int, C.c_int] + d4 + [C.c_void_p]
        L.ref_feature_f32.argtypes = [C.c_int, C.c_void_p, C.c_int, C.c_int] + f4 + [C.c_void_p]
        L.ref_feature_f64.argtypes = [C.c_int, C.c_void_p, C.c_int, C.c_int] + d4 + [C.c_void_p]
        L.ref_feature_equi_f32.argtypes = [C.c_int, C.c_void_p, C.c_int, C.c_int,
                                           C.c_float, C.c_float, C.c_void_p]
        L.ref_quantise_f32.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int,
                                       C.c_double, C.c_double, C.c_void_p]
        L.ref_quantise_f64.argtypes = L.ref_quantise_f32.argtypes
        L.ref_convert_bearing_u8_f32.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        L.ref_convert_bearing_u8_f64.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        L.ref_quantise_px_f32.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        L.ref_quantise_px_f64.argtypes = L.ref_quantise_px_f32.argtypes
        L.ref_curvature_to_image_opt_f32.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int,
                                                     C.c_int, C.c_double, C.c_int, C.c_double,
                                                     C.c_void_p]
        L.ref_curvature_to_image_opt_f64.argtypes = L.ref_curvature_to_image_opt_f32.argtypes
        L.ref_phi_f32.restype = C.c_float
        L.ref_phi_f32.argtypes = [C.c_int, C.c_int] + f4 + [C.c_int] * 4
        L.ref_pixel_to_sphere_f32.argtypes = [C.c_int, C.c_int] + f4 + [C.c_int, C.c_int, C.c_void_p]
        L.ref_bearing_angle_f32.restype = C.c_float
        L.ref_bearing_angle_f32.argtypes = [C.c_float] * 3
        L.ref_depth_scaling_u16.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_double,
                                            C.c_double, C.c_void_p]
        L.ref_cli_frame.argtypes = ([C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int] + f4 +
                                    [C.c_double, C.c_double, C.c_void_p, C.c_void_p])
        L.ref_cli_frame_equi.argtypes = [C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_float,
                                         C.c_float, C.c_double, C.c_double, C.c_void_p,
                                         C.c_void_p]
        L.ref_cli_batch.argtypes = ([C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int,
                                     C.c_int, C.c_int] + f4 +
                                    [C.c_double, C.c_double, C.c_void_p, C.c_int])
        L.ref_cli_batch_cyclic.argtypes = ([C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int,
                                            C.c_int, C.c_int, C.c_int] + f4 +
                                           [C.c_double, C.c_double, C.c_void_p, C.c_int])

    @staticmethod
    def _k(k):
        return (k["w"], k["h"], k["fx"], k["fy"], k["cx"], k["cy"])

    def info(self):
        return self.lib.ref_build_info().decode()

    def laserscan(self, depth_u16, k, dtype=np.float32):
        depth_u16 = np.ascontiguousarray(depth_u16)
        out = np.empty(depth_u16.shape, dtype)
        fn = self.lib.ref_laserscan_f32 if dtype == np.float32 else self.lib.ref_laserscan_f64
        fn(_ptr(depth_u16), *self._k(k), _ptr(out))
        return out

    def feature(self, f, rng, k):
        rng = np.ascontiguousarray(rng)
        out = np.empty(rng.shape, rng.dtype)
        fn = self.lib.ref_feature_f32 if rng.dtype == np.float32 else self.lib.ref_feature_f64
        fn(f, _ptr(rng), *self._k(k), _ptr(out))
        return out

    def quantise(self, f, feat, rng, lo=-20.0, hi=20.0):
        feat = np.ascontiguousarray(feat)
        rng = np.ascontiguousarray(rng)
        out = np.empty(feat.shape, np.uint16)
        fn = self.lib.ref_quantise_f32 if feat.dtype == np.float32 else self.lib.ref_quantise_f64
        fn(f, _ptr(feat), _ptr(rng), feat.shape[1], feat.shape[0], lo, hi, _ptr(out))
        return out

    def quantise_px(self, f, feat, pixel):
        """convert_bearing / convert_flexion / convert_max_curve<Real, PixelType>;
        pixel: 0 ushort, 1 uchar, 2 schar, 3 short."""
        feat = np.ascontiguousarray(feat)
        out = np.empty(feat.shape, [np.uint16, np.uint8, np.int8, np.int16][pixel])
        fn = self.lib.ref_quantise_px_f32 if feat.dtype == np.float32 else self.lib.ref_quantise_px_f64
        fn(f, pixel, _ptr(feat), feat.shape[1], feat.shape[0], _ptr(out))
        return out

    def curvature_to_image(self, feat, rng, lo=None, hi=None):
        """curvature_to_image<ushort> with optional clamps (None = std::nullopt)."""
        feat = np.ascontiguousarray(feat)
        rng = np.ascontiguousarray(rng, feat.dtype)
        out = np.empty(feat.shape, np.uint16)
        fn = (self.lib.ref_curvature_to_image_opt_f32 if feat.dtype == np.float32
              else self.lib.ref_curvature_to_image_opt_f64)
        fn(_ptr(feat), _ptr(rng), feat.shape[1], feat.shape[0], int(lo is not None),
           float(lo or 0.0), int(hi is not None), float(hi or 0.0), _ptr(out))
        return out

    def convert_bearing_u8(self, feat):
        feat = np.ascontiguousarray(feat)
        out = np.empty(feat.shape, np.uint8)
        fn = (self.lib.ref_convert_bearing_u8_f32 if feat.dtype == np.float32
              else self.lib.ref_convert_bearing_u8_f64)
        fn(_ptr(feat), feat.shape[1], feat.shape[0], _ptr(out))
        return out

    def phi(self, k, u1, v1, u2, v2):
        return self.lib.ref_phi_f32(*self._k(k), u1, v1, u2, v2)

    def pixel_to_sphere(self, k, u, v):
        s = np.zeros(3, np.float32)
        self.lib.ref_pixel_to_sphere_f32(*self._k(k), u, v, _ptr(s))
        return s

    def bearing_angle(self, b, c, cos_alpha):
        return self.lib.ref_bearing_angle_f32(b, c, cos_alpha)

    def depth_scaling(self, depth_u16, scale, offset):
        depth_u16 = np.ascontiguousarray(depth_u16)
        out = np.empty(depth_u16.shape, np.uint16)
        self.lib.ref_depth_scaling_u16(_ptr(depth_u16), depth_u16.shape[1],
                                       depth_u16.shape[0], scale, offset, _ptr(out))
        return out

    def cli_frame(self, f, in_kind, depth_u16, k, lo=-20.0, hi=20.0):
        depth_u16 = np.ascontiguousarray(depth_u16)
        feat = np.empty(depth_u16.shape, np.float32)
        q = np.empty(depth_u16.shape, np.uint16)
        self.lib.ref_cli_frame(f, in_kind, _ptr(depth_u16), *self._k(k), lo, hi,
                               _ptr(feat), _ptr(q))
        return feat, q

    def cli_batch(self, features, in_kind, frames_u16, k, lo=-20.0, hi=20.0,
                  threads=0, want_outputs=True, n_frames=None):
        """frames_u16: (n, h, w). Returns (list of (n,h,w) u16 arrays, workers).
        n_frames > n re-reads the buffer cyclically (timed baseline; no outputs)."""
        frames_u16 = np.ascontiguousarray(frames_u16)
        n = frames_u16.shape[0]
        if n_frames is not None and n_frames != n:
            assert not want_outputs
            feats = (C.c_int * len(features))(*features)
            workers = self.lib.ref_cli_batch_cyclic(
                C.cast(feats, C.c_void_p), len(features), in_kind, _ptr(frames_u16), n,
                n_frames, *self._k(k), lo, hi, None, threads)
            return [], workers
        feats = (C.c_int * len(features))(*features)
        outs = [np.empty(frames_u16.shape, np.uint16) for _ in features] if want_outputs else []
        if want_outputs:
            arr = (C.c_void_p * len(features))(*[o.ctypes.data for o in outs])
            outs_p = C.cast(arr, C.c_void_p)
        else:
            outs_p = None
        workers = self.lib.ref_cli_batch(C.cast(feats, C.c_void_p), len(features), in_kind,
                                         _ptr(frames_u16), n, *self._k(k), lo, hi,
                                         outs_p, threads)
        return outs, workers
