// sens_loc::math::image<T>: the typed one-channel image container of the
// reference (src/include/sens_loc/math/image.h:43-128), without OpenCV: the
// pixels live in a reference-counted dense row-major buffer (copying an image
// shares the pixels, as copying a cv::Mat does). Same invariants: one channel,
// never empty once constructed from data.
#ifndef DCB200_SENS_LOC_MATH_IMAGE_H
#define DCB200_SENS_LOC_MATH_IMAGE_H

#include <cstddef>
#include <cmath>
#include <cstring>
#include <limits>
#include <memory>
#include <stdexcept>
#include <type_traits>

namespace sens_loc {
using uchar  = unsigned char;
using schar  = signed char;
using ushort = unsigned short;

namespace math {

/// Pixel coordinate (u = column, v = row), math/coordinate.h:99-246 reduced to
/// what image access needs.
template <typename Number = int>
class pixel_coord {
  public:
    constexpr pixel_coord(Number u, Number v) noexcept : _u(u), _v(v) {}
    [[nodiscard]] constexpr Number u() const noexcept { return _u; }
    [[nodiscard]] constexpr Number v() const noexcept { return _v; }

  private:
    Number _u, _v;
};

/// Unit ray in camera coordinates, math/coordinate.h sphere_coord.
template <typename Real = float>
class sphere_coord {
  public:
    constexpr sphere_coord(Real xs, Real ys, Real zs) noexcept : _x(xs), _y(ys), _z(zs) {}
    [[nodiscard]] constexpr Real Xs() const noexcept { return _x; }
    [[nodiscard]] constexpr Real Ys() const noexcept { return _y; }
    [[nodiscard]] constexpr Real Zs() const noexcept { return _z; }

  private:
    Real _x, _y, _z;
};

template <typename PixelType = ushort>
class image {
  public:
    static_assert(std::is_arithmetic_v<PixelType>);

    image() = default;
    /// A w x h image, zero-initialised (cv::Mat(h, w, type) followed by `= 0`).
    image(int w, int h) : _w(w), _h(h) {
        if (w <= 0 || h <= 0)
            throw std::invalid_argument("image: width and height must be positive");
        _px.reset(new PixelType[std::size_t(w) * std::size_t(h)]());
    }
    /// Copies w x h pixels from `src` (dense, row major).
    image(int w, int h, const PixelType* src) : image(w, h) {
        std::memcpy(_px.get(), src, bytes());
    }

    [[nodiscard]] int w() const noexcept { return _w; }
    [[nodiscard]] int h() const noexcept { return _h; }
    [[nodiscard]] bool empty() const noexcept { return _px == nullptr; }
    [[nodiscard]] std::size_t pixels() const noexcept { return std::size_t(_w) * std::size_t(_h); }
    [[nodiscard]] std::size_t bytes() const noexcept { return pixels() * sizeof(PixelType); }

    template <typename Number = int>
    [[nodiscard]] PixelType at(const pixel_coord<Number>& p) const noexcept {
        return _px[std::size_t(int(p.v())) * std::size_t(_w) + std::size_t(int(p.u()))];
    }
    template <typename Number = int>
    [[nodiscard]] PixelType& at(const pixel_coord<Number>& p) noexcept {
        return _px[std::size_t(int(p.v())) * std::size_t(_w) + std::size_t(int(p.u()))];
    }

    /// Dense row-major pixels (the reference hands out the cv::Mat here).
    [[nodiscard]] const PixelType* data() const noexcept { return _px.get(); }
    [[nodiscard]] PixelType*       data() noexcept { return _px.get(); }

  private:
    int                          _w = 0, _h = 0;
    std::shared_ptr<PixelType[]> _px;
};

/// math/image.h:117-128: cv::Mat::convertTo between pixel types. Integer -> real is
/// exact; real -> integer rounds half to even and saturates (cv::saturate_cast).
template <typename TargetType, typename PixelType>
inline image<TargetType> convert(const image<PixelType>& img) {
    static_assert(std::is_arithmetic_v<TargetType> && std::is_arithmetic_v<PixelType>);
    if constexpr (std::is_same_v<TargetType, PixelType>) {
        return img;
    } else {
        if (img.empty())
            return {};
        image<TargetType> out(img.w(), img.h());
        const PixelType*  src = img.data();
        TargetType*       dst = out.data();
        for (std::size_t i = 0, n = img.pixels(); i < n; ++i) {
            if constexpr (std::is_integral_v<TargetType> && std::is_floating_point_v<PixelType>) {
                const double r  = std::nearbyint(double(src[i]));  // round half to even
                const double lo = double(std::numeric_limits<TargetType>::min());
                const double hi = double(std::numeric_limits<TargetType>::max());
                dst[i] = !(r == r) ? TargetType(lo) : TargetType(r < lo ? lo : (r > hi ? hi : r));
            } else {
                dst[i] = TargetType(src[i]);
            }
        }
        return out;
    }
}

}  // namespace math
}  // namespace sens_loc
#endif
