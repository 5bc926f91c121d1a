// sens_loc::conversion::* -- the drop-in replacements of the reference's
// depth -> feature functions for Real = float (what depth2x instantiates) and
// Real = double (what the reference's unit tests instantiate,
// src/unittest/conversion/intrinsic.h:8-26), same names and argument meaning
// (src/include/sens_loc/conversion/*.h). Every function checks its sizes, hands
// the pixels to libdepthconv_b200.so (sm_100a kernels) and returns an owning
// image; nothing is computed on the host.
//
// Differences to the reference, all at the edges:
//   * precondition violations (image size != intrinsic size, empty image) throw
//     std::invalid_argument; the reference's GSL contracts terminate;
//   * a GPU failure throws sens_loc::gpu::error (the reference cannot fail);
//   * the par_* variants (pre-sized output image, border pixels left untouched)
//     need no tf::Taskflow: the three-argument form converts at once; the
//     four-argument form takes any flow type with emplace() (a tf::Taskflow) and
//     puts the conversion into it as one task, returned twice as the
//     (synchronisation, work) pair of the reference.
#ifndef DCB200_SENS_LOC_CONVERSION_H
#define DCB200_SENS_LOC_CONVERSION_H

#include "../gpu/context.h"
#include "../math/image.h"

#include <algorithm>
#include <optional>
#include <stdexcept>
#include <type_traits>
#include <utility>

namespace sens_loc::conversion {

/// conversion/util.h:18-27
enum class direction { horizontal, vertical, diagonal, antidiagonal };

namespace detail {
template <typename Intrinsic, typename T>
inline void expect_same_size(const math::image<T>& img, const Intrinsic& k) {
    if (img.empty() || img.w() != k.w() || img.h() != k.h())
        throw std::invalid_argument("conversion: image size does not match the intrinsic");
}
inline dc_job job(int feature, int in_kind, int out_kind, const void* in) {
    dc_job j{};
    j.feature = feature, j.in_kind = in_kind, j.out_kind = out_kind, j.n_frames = 1;
    j.in       = in;
    j.clamp_lo = -20.f, j.clamp_hi = 20.f;
    j.scale = 1., j.offset = 0.;
    return j;
}
template <typename Real>
struct kinds {
    static_assert(std::is_same_v<Real, float> || std::is_same_v<Real, double>,
                  "the B200 path carries Real = float and Real = double");
    static constexpr int in  = std::is_same_v<Real, float> ? DC_IN_F32_RANGE : DC_IN_F64_RANGE;
    static constexpr int out = std::is_same_v<Real, float> ? DC_OUT_F32 : DC_OUT_F64;
};
template <typename Real, typename Intrinsic>
inline math::image<Real> one_output(int feature, int slot, const math::image<Real>& range,
                                    const Intrinsic& k) {
    static_assert(std::is_same_v<typename Intrinsic::real_type, Real>,
                  "image and intrinsic must have the same Real");
    expect_same_size(range, k);
    gpu::context& g = gpu::context::mine();
    g.use(k);
    math::image<Real> out(range.w(), range.h());
    dc_job            j = job(feature, kinds<Real>::in, kinds<Real>::out, range.data());
    j.out[slot]         = out.data();
    g.convert_host(j);
    return out;
}
}  // namespace detail

/// depth_to_laserscan.h:70-93: orthographic depth -> Euclidean range.
template <typename Real = float, typename PixelType = ushort, typename Intrinsic>
inline math::image<Real> depth_to_laserscan(const math::image<PixelType>& depth_image,
                                            const Intrinsic&              intrinsic) {
    static_assert(std::is_same_v<PixelType, ushort>, "the B200 path carries PixelType = ushort");
    detail::expect_same_size(depth_image, intrinsic);
    gpu::context& g = gpu::context::mine();
    g.use(intrinsic);
    math::image<Real> out(depth_image.w(), depth_image.h());
    dc_job j = detail::job(DC_RANGE, DC_IN_U16_DEPTH, detail::kinds<Real>::out, depth_image.data());
    j.out[0] = out.data();
    g.convert_host(j);
    return out;
}

/// depth_to_bearing.h:221-248
template <direction Direction, typename Real = float, typename Intrinsic>
inline math::image<Real> depth_to_bearing(const math::image<Real>& depth_image,
                                          const Intrinsic&         intrinsic) {
    return detail::one_output<Real>(DC_BEARING, int(Direction), depth_image, intrinsic);
}
/// All requested directions in one pass over the image (the reference makes one
/// pass per direction, converter_bearing.h.inl:13-26). Slots not wanted stay empty.
template <typename Intrinsic>
inline void depth_to_bearing_all(const math::image<float>& depth_image, const Intrinsic& intrinsic,
                                 const bool (&want)[4], math::image<float> (&out)[4]) {
    detail::expect_same_size(depth_image, intrinsic);
    gpu::context& g = gpu::context::mine();
    g.use(intrinsic);
    dc_job j = detail::job(DC_BEARING, DC_IN_F32_RANGE, DC_OUT_F32, depth_image.data());
    for (int d = 0; d < 4; ++d)
        if (want[d]) {
            out[d]   = math::image<float>(depth_image.w(), depth_image.h());
            j.out[d] = out[d].data();
        }
    g.convert_host(j);
}

/// depth_to_flexion.h:147-167
template <typename Intrinsic, typename Real = typename Intrinsic::real_type>
inline math::image<Real> depth_to_flexion(const math::image<Real>& depth_image,
        const Intrinsic& intrinsic) {
    return detail::one_output<Real>(DC_FLEXION, 0, depth_image, intrinsic);
}
/// depth_to_curvature.h:166-184
template <typename Intrinsic, typename Real = typename Intrinsic::real_type>
inline math::image<Real> depth_to_gaussian_curvature(const math::image<Real>& depth_image,
        const Intrinsic& intrinsic) {
    return detail::one_output<Real>(DC_CURVATURE, 0, depth_image, intrinsic);
}
/// depth_to_curvature.h:188-206
template <typename Intrinsic, typename Real = typename Intrinsic::real_type>
inline math::image<Real> depth_to_mean_curvature(const math::image<Real>& depth_image,
        const Intrinsic& intrinsic) {
    return detail::one_output<Real>(DC_CURVATURE, 1, depth_image, intrinsic);
}
/// depth_to_max_curve.h:86-153
template <typename Intrinsic, typename Real = typename Intrinsic::real_type>
inline math::image<Real> depth_to_max_curve(const math::image<Real>& depth_image,
        const Intrinsic& intrinsic) {
    return detail::one_output<Real>(DC_MAX_CURVE, 0, depth_image, intrinsic);
}

// ---- par_* : results into a caller-owned image ----------------------------
namespace detail {
/// Copies the region [u0, u1) x [v0, v1) of `src` into `dst`; everything else of
/// `dst` keeps its content (the reference's par_ loops never touch the border).
template <typename Real>
inline void copy_region(const math::image<Real>& src, math::image<Real>& dst, int u0, int u1,
                        int v0, int v1) {
    if (u1 <= u0)
        return;
    for (int v = v0; v < v1; ++v)
        std::copy(src.data() + std::size_t(v) * src.w() + u0,
                  src.data() + std::size_t(v) * src.w() + u1,
                  dst.data() + std::size_t(v) * dst.w() + u0);
}
template <typename A, typename B>
inline void expect_same_image_size(const A& a, const B& b) {
    if (a.empty() || a.w() != b.w() || a.h() != b.h())
        throw std::invalid_argument("conversion: output image size does not match the input");
}
template <typename Flow, typename Work>
inline auto as_task(Flow& flow, Work&& work) {
    auto task = flow.emplace(std::forward<Work>(work));
    return std::make_pair(task, task);
}
}  // namespace detail

/// depth_to_bearing.h:254-279: writes the pixels direction `Direction` computes
/// (depth_to_bearing.h:135-178) into `out`; all others keep their value.
template <direction Direction, typename Real = float, typename Intrinsic>
inline void par_depth_to_bearing(const math::image<Real>& depth_image, const Intrinsic& intrinsic,
                                 math::image<Real>& out) {
    detail::expect_same_image_size(out, depth_image);
    const math::image<Real> res = depth_to_bearing<Direction, Real>(depth_image, intrinsic);
    const int w = res.w(), h = res.h();
    constexpr bool hor = Direction == direction::horizontal, ver = Direction == direction::vertical;
    constexpr bool anti = Direction == direction::antidiagonal;
    detail::copy_region(res, out, ver ? 0 : 1, w, (hor || anti) ? 0 : 1, anti ? h - 1 : h);
}
template <direction Direction, typename Real = float, typename Intrinsic, typename Flow>
inline auto par_depth_to_bearing(const math::image<Real>& depth_image, const Intrinsic& intrinsic,
                                 math::image<Real>& out, Flow& flow) {
    detail::expect_same_image_size(out, depth_image);
    return detail::as_task(flow, [&depth_image, &intrinsic, &out] {
        par_depth_to_bearing<Direction, Real>(depth_image, intrinsic, out);
    });
}
/// depth_to_flexion.h:171-191: interior pixels only.
template <typename Intrinsic, typename Real = typename Intrinsic::real_type>
inline void par_depth_to_flexion(const math::image<Real>& depth_image, const Intrinsic& intrinsic,
                                 math::image<Real>& out) {
    detail::expect_same_image_size(out, depth_image);
    const math::image<Real> res = depth_to_flexion(depth_image, intrinsic);
    detail::copy_region(res, out, 1, res.w() - 1, 1, res.h() - 1);
}
template <typename Intrinsic, typename Flow, typename Real = typename Intrinsic::real_type>
inline auto par_depth_to_flexion(const math::image<Real>& depth_image, const Intrinsic& intrinsic,
                                 math::image<Real>& out, Flow& flow) {
    detail::expect_same_image_size(out, depth_image);
    return detail::as_task(flow, [&depth_image, &intrinsic, &out] {
        par_depth_to_flexion(depth_image, intrinsic, out);
    });
}
/// depth_to_laserscan.h:100-121: every pixel.
template <typename Real = float, typename PixelType = ushort, typename Intrinsic>
inline void par_depth_to_laserscan(const math::image<PixelType>& depth_image,
                                   const Intrinsic& intrinsic, math::image<Real>& out) {
    detail::expect_same_image_size(out, depth_image);
    const math::image<Real> res = depth_to_laserscan<Real, PixelType>(depth_image, intrinsic);
    detail::copy_region(res, out, 0, res.w(), 0, res.h());
}
template <typename Real = float, typename PixelType = ushort, typename Intrinsic, typename Flow>
inline auto par_depth_to_laserscan(const math::image<PixelType>& depth_image,
                                   const Intrinsic& intrinsic, math::image<Real>& out, Flow& flow) {
    detail::expect_same_image_size(out, depth_image);
    return detail::as_task(flow, [&depth_image, &intrinsic, &out] {
        par_depth_to_laserscan<Real, PixelType>(depth_image, intrinsic, out);
    });
}

// ---- quantisers -----------------------------------------------------------
namespace detail {
template <typename PixelType>
constexpr int pixel_kind() {
    static_assert(std::is_same_v<PixelType, ushort> || std::is_same_v<PixelType, uchar> ||
                      std::is_same_v<PixelType, schar> || std::is_same_v<PixelType, short>,
                  "quantised images are 8 or 16 bit integers");
    return std::is_same_v<PixelType, ushort>  ? DC_PX_U16
           : std::is_same_v<PixelType, uchar> ? DC_PX_U8
           : std::is_same_v<PixelType, schar> ? DC_PX_I8
                                              : DC_PX_I16;
}
template <typename PixelType, typename Real>
inline math::image<PixelType> quantise(int quant, const math::image<Real>& img,
                                       const math::image<Real>* mask, std::optional<Real> lo,
                                       std::optional<Real> hi) {
    if (img.empty() || (mask != nullptr && (mask->w() != img.w() || mask->h() != img.h())))
        throw std::invalid_argument("quantiser: empty image or mask of another size");
    math::image<PixelType> out(img.w(), img.h());
    const double           dlo = lo ? double(*lo) : 0., dhi = hi ? double(*hi) : 0.;
    gpu::context::mine().quantise(quant, pixel_kind<PixelType>(), img.data(),
                                  mask != nullptr ? mask->data() : nullptr, img.pixels(),
                                  out.data(), lo ? &dlo : nullptr, hi ? &dhi : nullptr);
    return out;
}
}  // namespace detail

/// depth_to_bearing.h:282-301
template <typename Real = float, typename PixelType = ushort>
inline math::image<PixelType> convert_bearing(const math::image<Real>& bearing_image) {
    return detail::quantise<PixelType, Real>(DC_Q_BEARING, bearing_image, nullptr, Real(0), Real(0));
}
/// depth_to_flexion.h:194-215
template <typename PixelType = ushort, typename Real = float>
inline math::image<PixelType> convert_flexion(const math::image<Real>& flexion_image) {
    return detail::quantise<PixelType, Real>(DC_Q_FLEXION, flexion_image, nullptr, Real(0), Real(0));
}
/// depth_to_max_curve.h:156-176
template <typename PixelType = ushort, typename Real = float>
inline math::image<PixelType> convert_max_curve(const math::image<Real>& max_curve_image) {
    return detail::quantise<PixelType, Real>(DC_Q_MAX_CURVE, max_curve_image, nullptr, Real(0), Real(0));
}
/// depth_to_curvature.h:253-283. A clamp that is not given is the minimum /
/// maximum over the whole curvature image (:218-228; a reduction on the device).
/// 16 bit only: the reference's ushort intermediate re-types the result (:263-275).
template <typename PixelType = ushort, typename Real = float, typename MaskType = Real>
inline math::image<PixelType> curvature_to_image(const math::image<Real>&     curvature_img,
                                                 const math::image<MaskType>& depth_image_as_mask,
                                                 std::optional<Real> clamp_min = std::nullopt,
                                                 std::optional<Real> clamp_max = std::nullopt) {
    static_assert(std::is_same_v<PixelType, ushort>);
    // the mask is "saturate_cast<uchar>(depth) != 0" (:263-264): any arithmetic type
    const math::image<Real> mask = math::convert<Real>(depth_image_as_mask);
    return detail::quantise<PixelType, Real>(DC_Q_CURVATURE, curvature_img, &mask, clamp_min,
                                             clamp_max);
}
/// range image -> 16 bit, apps/depth2x/converter_laserscan.h.inl:9-10
inline math::image<ushort> convert_range(const math::image<float>& range_image) {
    return detail::quantise<ushort, float>(DC_Q_RANGE, range_image, nullptr, 0.f, 0.f);
}

/// depth_scaling.h:21-31: saturate(round(depth * scale + offset))
inline math::image<ushort> depth_scaling(const math::image<ushort>& depth_image, double scale,
                                         double offset) {
    if (depth_image.empty())
        throw std::invalid_argument("depth_scaling: empty image");
    gpu::context& g = gpu::context::mine();
    g.use_frame_size(depth_image.w(), depth_image.h());
    math::image<ushort> out(depth_image.w(), depth_image.h());
    dc_job j = detail::job(DC_SCALE, DC_IN_U16_RANGE, DC_OUT_U16, depth_image.data());
    j.out[0] = out.data();
    j.scale  = scale;
    j.offset = offset;
    g.convert_host(j);
    return out;
}

}  // namespace sens_loc::conversion
#endif
