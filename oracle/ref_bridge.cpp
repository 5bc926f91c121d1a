// TEST INFRASTRUCTURE ONLY (oracle/). Never linked into the product library.
//
// C bridge around the REFERENCE'S OWN headers (compiled from where they lie in
// /root/reference/src/include, never copied) so that tests and the CPU-baseline
// leg of bench.py can call the reference arithmetic through ctypes.
// Build recipe: oracle/Makefile, output oracle/_ref/liboracle_ref.so.
//
// What is bridged (reference file:line):
//   conversion::depth_to_laserscan           conversion/depth_to_laserscan.h:70-93
//   conversion::depth_to_bearing<direction>  conversion/depth_to_bearing.h:221-248
//   conversion::depth_to_flexion             conversion/depth_to_flexion.h:147-167
//   conversion::depth_to_gaussian_curvature  conversion/depth_to_curvature.h:166-184
//   conversion::depth_to_mean_curvature      conversion/depth_to_curvature.h:188-206
//   conversion::depth_to_max_curve           conversion/depth_to_max_curve.h:86-153
//   conversion::convert_bearing/flexion/max_curve, curvature_to_image
//   conversion::depth_scaling                conversion/depth_scaling.h:21-31
//   camera_models::phi                       camera_models/utility.h:15-49
//   the frame-parallel batch driver          apps/util/parallel_processing.h:36-58
//   (tf::Executor + parallel_for(start, end+1, 1) with one task per frame, each
//    task = batch_converter::process_index minus the PNG I/O,
//    apps/util/batch_converter.cpp:12-28 + apps/depth2x/converter_*.h.inl).
#include <sens_loc/camera_models/equirectangular.h>
#include <sens_loc/camera_models/pinhole.h>
#include <sens_loc/camera_models/utility.h>
#include <sens_loc/conversion/depth_scaling.h>
#include <sens_loc/conversion/depth_to_bearing.h>
#include <sens_loc/conversion/depth_to_curvature.h>
#include <sens_loc/conversion/depth_to_flexion.h>
#include <sens_loc/conversion/depth_to_laserscan.h>
#include <sens_loc/conversion/depth_to_max_curve.h>
#include <sens_loc/math/image.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <optional>
#include <taskflow/taskflow.hpp>
#include <thread>

namespace sens_loc::util {
// Declared in sens_loc/util/correctness_util.h:32-34; the definition lives in
// src/lib/util/correctness_util.cpp, which drags in console/rang. The oracle
// only needs the symbol.
[[noreturn]] void unreachable(std::string_view message, std::string_view file,
                              int line) noexcept {
    std::fprintf(stderr, "%.*s:%d: unreachable: %.*s\n", int(file.size()),
                 file.data(), line, int(message.size()), message.data());
    std::exit(1);
}
}  // namespace sens_loc::util

namespace {
using namespace sens_loc;
using camera_models::equirectangular;
using camera_models::pinhole;
using conversion::direction;

struct cam {
    int    model;  // 0 pinhole, 1 equirectangular
    int    w, h;
    double p0, p1, p2, p3;  // pinhole: fx fy cx cy; equirect: theta_min theta_max
};

template <typename Real>
pinhole<Real> make_pinhole(const cam& c) {
    return pinhole<Real>(c.w, c.h, Real(c.p0), Real(c.p1), Real(c.p2), Real(c.p3));
}
template <typename Real>
equirectangular<Real> make_equi(const cam& c) {
    return equirectangular<Real>(c.w, c.h, {Real(c.p0), Real(c.p1)});
}

template <typename T>
math::image<T> wrap(const T* p, int w, int h) {
    cv::Mat m(h, w, math::detail::get_opencv_type<T>(),
              const_cast<void*>(static_cast<const void*>(p)));
    return math::image<T>(m);
}

template <typename T>
void store(const math::image<T>& img, T* out) {
    std::memcpy(out, img.data().data,
                sizeof(T) * std::size_t(img.w()) * std::size_t(img.h()));
}

enum feature {
    F_RANGE = 0,
    F_BEARING_H,
    F_BEARING_V,
    F_BEARING_D,
    F_BEARING_A,
    F_FLEXION,
    F_GAUSS,
    F_MEAN,
    F_MAX_CURVE,
};

template <typename Real, typename Intrinsic>
math::image<Real> run_feature(int f, const math::image<Real>& range,
                              const Intrinsic& K) {
    using namespace conversion;
    switch (f) {
    case F_BEARING_H: return depth_to_bearing<direction::horizontal>(range, K);
    case F_BEARING_V: return depth_to_bearing<direction::vertical>(range, K);
    case F_BEARING_D: return depth_to_bearing<direction::diagonal>(range, K);
    case F_BEARING_A: return depth_to_bearing<direction::antidiagonal>(range, K);
    case F_FLEXION: return depth_to_flexion(range, K);
    case F_GAUSS: return depth_to_gaussian_curvature(range, K);
    case F_MEAN: return depth_to_mean_curvature(range, K);
    case F_MAX_CURVE: return depth_to_max_curve(range, K);
    default: return range;
    }
}

// The quantiser each depth2x converter applies (apps/depth2x/converter_*.h.inl).
template <typename Real>
math::image<ushort> quantise(int f, const math::image<Real>& feat,
                             const math::image<Real>& range, double lo,
                             double hi) {
    using namespace conversion;
    switch (f) {
    case F_BEARING_H:
    case F_BEARING_V:
    case F_BEARING_D:
    case F_BEARING_A: return convert_bearing<Real, ushort>(feat);
    case F_FLEXION: return convert_flexion<ushort>(feat);
    case F_GAUSS:
    case F_MEAN:
        return curvature_to_image<ushort>(feat, range, {Real(lo)}, {Real(hi)});
    case F_MAX_CURVE: return convert_max_curve<ushort>(feat);
    default: {
        // converter_laserscan.h.inl:9-10
        cv::Mat d16(feat.h(), feat.w(), CV_16U);
        feat.data().convertTo(d16, CV_16U);
        return math::image<ushort>(d16);
    }
    }
}

template <typename Real, typename Intrinsic>
int feature_impl(int f, const Real* range, const Intrinsic& K, Real* out) {
    const auto in  = wrap(range, K.w(), K.h());
    const auto res = run_feature<Real>(f, in, K);
    store(res, out);
    return 0;
}

template <typename Real, typename Intrinsic>
int laserscan_impl(const uint16_t* depth, const Intrinsic& K, Real* out) {
    const auto in  = wrap<ushort>(depth, K.w(), K.h());
    const auto res = conversion::depth_to_laserscan<Real, ushort>(in, K);
    store(res, out);
    return 0;
}

// One frame of the CLI path, minus file I/O:
// batch_sensor_converter::preprocess_depth (apps/util/batch_converter.h:150-163)
// then <feature>_converter::process_file (apps/depth2x/converter_*.h.inl).
template <typename Intrinsic>
void cli_frame(int f, int in_kind, const uint16_t* depth, const Intrinsic& K,
               double lo, double hi, float* out_f32, uint16_t* out_u16) {
    const auto in16 = wrap<ushort>(depth, K.w(), K.h());
    const math::image<float> range =
        in_kind == 0 ? conversion::depth_to_laserscan<float, ushort>(in16, K)
                     : math::convert<float>(in16);
    const math::image<float> feat = run_feature<float>(f, range, K);
    if (out_f32 != nullptr)
        store(feat, out_f32);
    if (out_u16 != nullptr) {
        const auto q = quantise<float>(f, feat, range, lo, hi);
        store(q, out_u16);
    }
}
// Every PixelType instantiation of the bounded quantisers (f: a bearing direction,
// F_FLEXION or F_MAX_CURVE; pixel: 0 ushort, 1 uchar, 2 schar, 3 short) and
// curvature_to_image with optional clamps (has_lo / has_hi = 0: std::nullopt, the
// image's own minimum / maximum, depth_to_curvature.h:218-228).
template <typename Real, typename Px>
static int quantise_px_impl(int f, const Real* feat, int w, int h, void* out) {
    using namespace conversion;
    const auto in = wrap(feat, w, h);
    math::image<Px> q = f == F_FLEXION     ? convert_flexion<Px>(in)
                        : f == F_MAX_CURVE ? convert_max_curve<Px>(in)
                                           : convert_bearing<Real, Px>(in);
    store(q, static_cast<Px*>(out));
    return 0;
}
template <typename Real>
static int quantise_px(int f, int pixel, const Real* feat, int w, int h, void* out) {
    switch (pixel) {
    case 1: return quantise_px_impl<Real, uchar>(f, feat, w, h, out);
    case 2: return quantise_px_impl<Real, schar>(f, feat, w, h, out);
    case 3: return quantise_px_impl<Real, short>(f, feat, w, h, out);
    default: return quantise_px_impl<Real, ushort>(f, feat, w, h, out);
    }
}
template <typename Real>
static int curvature_opt(const Real* feat, const Real* range, int w, int h, int has_lo, double lo,
                         int has_hi, double hi, uint16_t* out) {
    const std::optional<Real> a = has_lo ? std::optional<Real>(Real(lo)) : std::nullopt;
    const std::optional<Real> b = has_hi ? std::optional<Real>(Real(hi)) : std::nullopt;
    const auto q = conversion::curvature_to_image<ushort>(wrap(feat, w, h), wrap(range, w, h), a, b);
    store(q, out);
    return 0;
}
}  // namespace

extern "C" {

const char* ref_build_info() {
    return "reference headers @ /root/reference/src/include, Eigen 3.3.7 "
           "(vendored), GSL (vendored), cpp-taskflow 2.3.1 (vendored), cv::Mat "
           "shim; g++ " __VERSION__
           " -O3 -DNDEBUG -msse4.2 -DGSL_THROW_ON_CONTRACT_VIOLATION=1";
}

// ---- per-function entry points, Real = float (the CLI precision) ----------
int ref_laserscan_f32(const uint16_t* depth, int w, int h, float fx, float fy,
                      float cx, float cy, float* out) {
    return laserscan_impl<float>(depth, pinhole<float>(w, h, fx, fy, cx, cy), out);
}
int ref_feature_f32(int f, const float* range, int w, int h, float fx, float fy,
                    float cx, float cy, float* out) {
    return feature_impl<float>(f, range, pinhole<float>(w, h, fx, fy, cx, cy), out);
}
int ref_feature_equi_f32(int f, const float* range, int w, int h,
                         float theta_min, float theta_max, float* out) {
    return feature_impl<float>(
        f, range, equirectangular<float>(w, h, {theta_min, theta_max}), out);
}
int ref_quantise_f32(int f, const float* feat, const float* range, int w, int h,
                     double lo, double hi, uint16_t* out) {
    const auto q = quantise<float>(f, wrap(feat, w, h), wrap(range, w, h), lo, hi);
    store(q, out);
    return 0;
}
// convert_bearing<float, uchar> etc. are only used by the reference's unit
// tests (8-bit goldens): test_conversion_bearing.cpp:78-159.
int ref_convert_bearing_u8_f32(const float* feat, int w, int h, uint8_t* out) {
    const auto q = conversion::convert_bearing<float, uchar>(wrap(feat, w, h));
    store(q, out);
    return 0;
}
int ref_quantise_px_f32(int f, int pixel, const float* feat, int w, int h, void* out) {
    return quantise_px<float>(f, pixel, feat, w, h, out);
}
int ref_quantise_px_f64(int f, int pixel, const double* feat, int w, int h, void* out) {
    return quantise_px<double>(f, pixel, feat, w, h, out);
}
int ref_curvature_to_image_opt_f32(const float* feat, const float* range, int w, int h, int has_lo,
                                   double lo, int has_hi, double hi, uint16_t* out) {
    return curvature_opt<float>(feat, range, w, h, has_lo, lo, has_hi, hi, out);
}
int ref_curvature_to_image_opt_f64(const double* feat, const double* range, int w, int h,
                                   int has_lo, double lo, int has_hi, double hi, uint16_t* out) {
    return curvature_opt<double>(feat, range, w, h, has_lo, lo, has_hi, hi, out);
}
float ref_phi_f32(int w, int h, float fx, float fy, float cx, float cy, int u1,
                  int v1, int u2, int v2) {
    const pinhole<float> K(w, h, fx, fy, cx, cy);
    return camera_models::phi(K, math::pixel_coord<int>{u1, v1},
                              math::pixel_coord<int>{u2, v2});
}
void ref_pixel_to_sphere_f32(int w, int h, float fx, float fy, float cx, float cy,
                             int u, int v, float* xyz) {
    const pinhole<float> K(w, h, fx, fy, cx, cy);
    const auto           s = K.pixel_to_sphere(math::pixel_coord<int>{u, v});
    xyz[0] = s.Xs();
    xyz[1] = s.Ys();
    xyz[2] = s.Zs();
}
float ref_bearing_angle_f32(float b, float c, float cos_alpha) {
    return math::bearing_angle<float>(b, c, cos_alpha);
}
int ref_depth_scaling_u16(const uint16_t* depth, int w, int h, double scale,
                          double offset, uint16_t* out) {
    const auto r = conversion::depth_scaling(wrap<ushort>(depth, w, h), scale, offset);
    store(r, out);
    return 0;
}

// ---- Real = double: the precision the reference's golden PNGs were made in -
int ref_laserscan_f64(const uint16_t* depth, int w, int h, double fx, double fy,
                      double cx, double cy, double* out) {
    return laserscan_impl<double>(depth, pinhole<double>(w, h, fx, fy, cx, cy), out);
}
int ref_feature_f64(int f, const double* range, int w, int h, double fx,
                    double fy, double cx, double cy, double* out) {
    return feature_impl<double>(f, range, pinhole<double>(w, h, fx, fy, cx, cy), out);
}
int ref_feature_equi_f64(int f, const double* range, int w, int h,
                         double theta_min, double theta_max, double* out) {
    return feature_impl<double>(
        f, range, equirectangular<double>(w, h, {theta_min, theta_max}), out);
}
int ref_quantise_f64(int f, const double* feat, const double* range, int w,
                     int h, double lo, double hi, uint16_t* out) {
    const auto q = quantise<double>(f, wrap(feat, w, h), wrap(range, w, h), lo, hi);
    store(q, out);
    return 0;
}
int ref_convert_bearing_u8_f64(const double* feat, int w, int h, uint8_t* out) {
    const auto q = conversion::convert_bearing<double, uchar>(wrap(feat, w, h));
    store(q, out);
    return 0;
}

// ---- CLI-equivalent single frame and frame-parallel batch (CPU baseline) ---
// in_kind: 0 = pinhole-depth (orthographic u16), 1 = pinhole-range (u16).
// out_f32 / out_u16 may be NULL. Frames are dense, back to back.
int ref_cli_frame(int f, int in_kind, const uint16_t* depth, int w, int h,
                  float fx, float fy, float cx, float cy, double lo, double hi,
                  float* out_f32, uint16_t* out_u16) {
    cli_frame(f, in_kind, depth, pinhole<float>(w, h, fx, fy, cx, cy), lo, hi,
              out_f32, out_u16);
    return 0;
}
int ref_cli_frame_equi(int f, const uint16_t* depth, int w, int h,
                       float theta_min, float theta_max, double lo, double hi,
                       float* out_f32, uint16_t* out_u16) {
    cli_frame(f, 1, depth, equirectangular<float>(w, h, {theta_min, theta_max}),
              lo, hi, out_f32, out_u16);
    return 0;
}

// features: list of feature ids computed for every frame (bearing with four
// directions = four ids, as converter_bearing.h.inl does four full passes).
// outs_u16[k] receives n_frames images for features[k] (may be NULL entries).
// threads <= 0: tf::Executor default (hardware_concurrency), as the CLI does.
// n_distinct: number of frames actually present in `depth`; frame idx reads
// depth[idx % n_distinct] (lets the timed baseline run many frames from a small
// buffer). Outputs, when requested, must still hold n_frames images.
int ref_cli_batch_cyclic(const int* features, int n_features, int in_kind,
                         const uint16_t* depth, int n_distinct, int n_frames, int w,
                         int h, float fx, float fy, float cx, float cy, double lo,
                         double hi, uint16_t** outs_u16, int threads);

int ref_cli_batch(const int* features, int n_features, int in_kind,
                  const uint16_t* depth, int n_frames, int w, int h, float fx,
                  float fy, float cx, float cy, double lo, double hi,
                  uint16_t** outs_u16, int threads) {
    return ref_cli_batch_cyclic(features, n_features, in_kind, depth, n_frames,
                                n_frames, w, h, fx, fy, cx, cy, lo, hi, outs_u16,
                                threads);
}

int ref_cli_batch_cyclic(const int* features, int n_features, int in_kind,
                         const uint16_t* depth, int n_distinct, int n_frames, int w,
                         int h, float fx, float fy, float cx, float cy, double lo,
                         double hi, uint16_t** outs_u16, int threads) {
    const pinhole<float> K(w, h, fx, fy, cx, cy);
    const std::size_t    px = std::size_t(w) * std::size_t(h);
    tf::Executor executor(threads > 0 ? unsigned(threads)
                                      : std::thread::hardware_concurrency());
    tf::Taskflow flow;
    flow.parallel_for(0, n_frames, 1, [&](int idx) {
        const auto in16 = wrap<ushort>(depth + px * (idx % n_distinct), w, h);
        const math::image<float> range =
            in_kind == 0 ? conversion::depth_to_laserscan<float, ushort>(in16, K)
                         : math::convert<float>(in16);
        for (int k = 0; k < n_features; ++k) {
            const auto feat = run_feature<float>(features[k], range, K);
            const auto q    = quantise<float>(features[k], feat, range, lo, hi);
            if (outs_u16 != nullptr && outs_u16[k] != nullptr)
                store(q, outs_u16[k] + px * idx);
        }
    });
    executor.run(flow).wait();
    return int(executor.num_workers());
}
}
