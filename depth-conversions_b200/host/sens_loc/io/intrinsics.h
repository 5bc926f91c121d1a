// Text intrinsic files, src/include/sens_loc/io/intrinsics.h:48-157: same
// grammar, same strictness (each pinhole line must be consumed exactly).
//   pinhole:          "w h" / "fx _ cx" / "_ fy cy" [/ anything]
//   equirectangular:  "w h" [/ "theta_min theta_max"]
#ifndef DCB200_SENS_LOC_IO_INTRINSICS_H
#define DCB200_SENS_LOC_IO_INTRINSICS_H

#include "../camera_models/pinhole.h"

#include <istream>
#include <optional>
#include <sstream>
#include <string>

namespace sens_loc::io {

template <typename Real, template <typename> typename Model>
class camera;

template <typename Real>
class camera<Real, camera_models::pinhole> {
  public:
    static std::optional<camera_models::pinhole<Real>> load_intrinsic(std::istream& in) noexcept {
        try {
            if (!in.good())
                return std::nullopt;
            int         w = 0, h = 0;
            Real        fx = 0, fy = 0, cx = 0, cy = 0, skip = 0;
            std::string line;
            auto        exhausted = [](std::istringstream& ss) {
                return !ss.fail() && ss.rdstate() == std::ios_base::eofbit;
            };
            std::getline(in, line);
            std::istringstream l1{line};
            l1 >> w >> h;
            if (!exhausted(l1) || w <= 0 || h <= 0)
                return std::nullopt;
            std::getline(in, line);
            std::istringstream l2{line};
            l2 >> fx >> skip >> cx;
            if (!exhausted(l2) || fx <= Real(0) || cx <= Real(0))
                return std::nullopt;
            std::getline(in, line);
            std::istringstream l3{line};
            l3 >> skip >> fy >> cy;
            if (!exhausted(l3) || fy <= Real(0) || cy <= Real(0))
                return std::nullopt;
            return camera_models::pinhole<Real>(w, h, fx, fy, cx, cy);
        } catch (...) {
            return std::nullopt;
        }
    }
};

template <typename Real>
class camera<Real, camera_models::equirectangular> {
  public:
    static std::optional<camera_models::equirectangular<Real>>
    load_intrinsic(std::istream& in) noexcept {
        try {
            if (!in.good())
                return std::nullopt;
            const Real  pi = Real(3.141592653589793238462);
            int         w = 0, h = 0;
            Real        tmin = Real(0), tmax = pi;
            std::string line;
            std::getline(in, line);
            std::istringstream l1{line};
            l1 >> w >> h;
            if (l1.fail() || w <= 0 || h <= 0)
                return std::nullopt;
            if (!in.eof()) {
                std::getline(in, line);
                std::istringstream l2{line};
                l2 >> tmin >> tmax;
                if (l2.fail() || tmin <= Real(0) || tmax <= Real(0) || tmin >= tmax || tmin >= pi ||
                    tmax >= pi)
                    return std::nullopt;
            }
            return camera_models::equirectangular<Real>(w, h, {tmin, tmax});
        } catch (...) {
            return std::nullopt;
        }
    }
};

}  // namespace sens_loc::io
#endif
