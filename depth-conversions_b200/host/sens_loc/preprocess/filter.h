// sens_loc::preprocess::median_blur -- drop-in for
// src/include/sens_loc/preprocess/filter.h:143-157 (cv::medianBlur, ksize 3 or 5)
// on the GPU. The other filters of that header (bilateral, guided, Gaussian) are
// OpenCV algorithms whose arithmetic is not specified by the reference: not
// provided ("parity unpinned").
#ifndef DCB200_SENS_LOC_PREPROCESS_FILTER_H
#define DCB200_SENS_LOC_PREPROCESS_FILTER_H

#include "../gpu/context.h"
#include "../math/image.h"

#include <stdexcept>
#include <type_traits>

namespace sens_loc::preprocess {

template <typename PixelType>
inline math::image<PixelType> median_blur(const math::image<PixelType>& input, int ksize) {
    static_assert(std::is_same_v<PixelType, ushort> || std::is_same_v<PixelType, float>,
                  "median_blur: 16 bit or float images (what cv::medianBlur takes for ksize 3, 5)");
    if (input.empty())
        throw std::invalid_argument("median_blur: empty image");
    if (ksize != 3 && ksize != 5)
        throw std::invalid_argument("median_blur: ksize must be 3 or 5");
    gpu::context& g = gpu::context::mine();
    g.use_frame_size(input.w(), input.h());
    math::image<PixelType> out(input.w(), input.h());
    const int kind = std::is_same_v<PixelType, ushort> ? DC_IN_U16_RANGE : DC_IN_F32_RANGE;
    const int rc   = dc_median_host(g.raw(), kind, input.data(), 1, ksize, out.data());
    if (rc != DC_OK)
        throw gpu::error(rc, g.last_error());
    return out;
}

}  // namespace sens_loc::preprocess
#endif
