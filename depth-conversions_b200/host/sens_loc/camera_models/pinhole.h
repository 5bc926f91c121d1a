// sens_loc::camera_models::pinhole<Real> and equirectangular<Real>: the
// intrinsic models of the reference, unchanged in meaning
// (src/include/sens_loc/camera_models/pinhole.h:26-171,
//  src/include/sens_loc/camera_models/equirectangular.h:43-181). The per-pixel
// ray arithmetic is restated for host-side use (tests, tools); the conversions
// themselves evaluate it once per camera inside libdepthconv_b200.so.
#ifndef DCB200_SENS_LOC_CAMERA_MODELS_H
#define DCB200_SENS_LOC_CAMERA_MODELS_H

#include "../math/image.h"

#include <cmath>
#include <stdexcept>
#include <type_traits>

namespace sens_loc::camera_models {

template <typename Real = float>
class pinhole {
  public:
    static_assert(std::is_floating_point_v<Real>);
    using real_type = Real;

    pinhole() = default;
    /// Preconditions as the reference's Expects (pinhole.h:37-50); a violation
    /// throws here instead of terminating.
    pinhole(int w, int h, Real fx, Real fy, Real cx, Real cy)
        : _w(w), _h(h), _fx(fx), _fy(fy), _cx(cx), _cy(cy) {
        if (!(w > 0 && h > 0 && fx > Real(0) && fy > Real(0) && cx > Real(0) && cy > Real(0)))
            throw std::invalid_argument("pinhole: w, h, fx, fy, cx, cy must be positive");
    }
    [[nodiscard]] int  w() const noexcept { return _w; }
    [[nodiscard]] int  h() const noexcept { return _h; }
    [[nodiscard]] Real fx() const noexcept { return _fx; }
    [[nodiscard]] Real fy() const noexcept { return _fy; }
    [[nodiscard]] Real cx() const noexcept { return _cx; }
    [[nodiscard]] Real cy() const noexcept { return _cy; }

    /// pinhole.h:139-171: ((u - cx) / fx, (v - cy) / fy) in Real, normalised in
    /// double, narrowed to Real.
    template <typename Number = int>
    [[nodiscard]] math::sphere_coord<Real>
    pixel_to_sphere(const math::pixel_coord<Number>& p) const noexcept {
        const Real   x   = (Real(p.u()) - _cx) / _fx;
        const Real   y   = (Real(p.v()) - _cy) / _fy;
        const double cse = 1. + double(x) * double(x) + double(y) * double(y);
        const double f   = 1. / std::sqrt(cse);
        return {Real(f * double(x)), Real(f * double(y)), Real(f)};
    }

  private:
    int  _w = 0, _h = 0;
    Real _fx = 0, _fy = 0, _cx = 0, _cy = 0;
};

template <typename Real = float>
class equirectangular {
  public:
    static_assert(std::is_floating_point_v<Real>);
    using real_type = Real;
    struct theta_range {
        Real min, max;
    };

    equirectangular() = default;
    equirectangular(int w, int h) : equirectangular(w, h, {Real(0), Real(3.141592653589793238462)}) {}
    /// equirectangular.h:59-72, 104-114
    equirectangular(int w, int h, theta_range t) : _w(w), _h(h), _t(t) {
        if (!(w > 0 && h > 0 && t.min >= Real(0) && t.max <= Real(3.141592653589793238462) &&
              t.min < t.max))
            throw std::invalid_argument(
                "equirectangular: w, h > 0 and 0 <= theta_min < theta_max <= pi");
    }
    [[nodiscard]] int  w() const noexcept { return _w; }
    [[nodiscard]] int  h() const noexcept { return _h; }
    [[nodiscard]] Real theta_min() const noexcept { return _t.min; }
    [[nodiscard]] Real theta_max() const noexcept { return _t.max; }

  private:
    int         _w = 0, _h = 0;
    theta_range _t{Real(0), Real(3.141592653589793238462)};
};

}  // namespace sens_loc::camera_models
#endif
