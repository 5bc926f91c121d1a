// util::average_pixel_error (src/lib/util/correctness_util.cpp:20-31): the L2 norm
// of the difference of two images divided by the pixel count -- the measure the
// reference's tests hold their golden images to.
#ifndef DCB200_SENS_LOC_UTIL_CORRECTNESS_UTIL_H
#define DCB200_SENS_LOC_UTIL_CORRECTNESS_UTIL_H

#include "../math/image.h"

#include <cmath>
#include <stdexcept>

namespace sens_loc::util {

template <typename PixelType>
inline double average_pixel_error(const math::image<PixelType>& i1,
                                  const math::image<PixelType>& i2) {
    if (i1.empty() || i1.w() != i2.w() || i1.h() != i2.h())
        throw std::invalid_argument("average_pixel_error: images of different size");
    double sum = 0.;
    for (std::size_t i = 0, n = i1.pixels(); i < n; ++i) {
        const double d = double(i1.data()[i]) - double(i2.data()[i]);
        sum += d * d;
    }
    return std::sqrt(sum) / double(i1.pixels());
}

}  // namespace sens_loc::util
#endif
