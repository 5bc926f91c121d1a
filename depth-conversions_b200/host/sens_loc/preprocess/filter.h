// sens_loc::preprocess::{median_blur, gaussian_blur, bilateral_filter} -- drop-ins for
// src/include/sens_loc/preprocess/filter.h:49-82, 125-157 on the GPU. The median is bit
// exact against cv::medianBlur; the Gaussian blur and the bilateral filter restate
// OpenCV's published algorithms and agree with it within the tolerance stated in
// include/depthconv_b200.h (their arithmetic is not specified by the reference).
// guided_filter (cv::ximgproc, opencv_contrib) is not provided: nothing to pin it
// against exists in this image.
#ifndef DCB200_SENS_LOC_PREPROCESS_FILTER_H
#define DCB200_SENS_LOC_PREPROCESS_FILTER_H

#include "../gpu/context.h"
#include "../math/image.h"

#include <stdexcept>
#include <type_traits>
#include <variant>

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

// Stand-in for the cv::Size argument of the reference's gaussian_blur (no OpenCV here):
// `gaussian_blur(img, {5, 5}, 1.5)` reads the same.
struct kernel_size {
    int width = 0, height = 0;
};

namespace detail {
template <typename PixelType>
constexpr int elem_kind() {
    static_assert(std::is_same_v<PixelType, ushort> || std::is_same_v<PixelType, float>,
                  "preprocess filters: 16 bit or float images");
    return std::is_same_v<PixelType, ushort> ? DC_IN_U16_RANGE : DC_IN_F32_RANGE;
}
}  // namespace detail

// filter.h:125-136: cv::GaussianBlur(input, result, ksize, sigmaX, sigmaY)
template <typename PixelType>
inline math::image<PixelType> gaussian_blur(const math::image<PixelType>& input, kernel_size ksize,
                                            double sigmaX, double sigmaY = 0) {
    if (input.empty())
        throw std::invalid_argument("gaussian_blur: empty image");
    gpu::context& g = gpu::context::mine();
    g.use_frame_size(input.w(), input.h());
    math::image<PixelType> out(input.w(), input.h());
    const int rc = dc_gaussian_blur_host(g.raw(), detail::elem_kind<PixelType>(), input.data(), 1,
                                         ksize.width, ksize.height, sigmaX, sigmaY, out.data());
    if (rc != DC_OK)
        throw gpu::error(rc, g.last_error());
    return out;
}

// filter.h:49-82: `proximity` holds the pixel diameter (int) or sigma_space (double); the
// image is filtered as float and converted back to PixelType.
template <typename PixelType>
inline math::image<PixelType> bilateral_filter(const math::image<PixelType>& input, double sigma_color,
                                               std::variant<int, double> proximity) {
    if (input.empty())
        throw std::invalid_argument("bilateral_filter: empty image");
    const int    distance    = std::holds_alternative<int>(proximity) ? std::get<int>(proximity) : 0;
    const double sigma_space = std::holds_alternative<int>(proximity) ? 0. : std::get<double>(proximity);
    gpu::context& g = gpu::context::mine();
    g.use_frame_size(input.w(), input.h());
    math::image<PixelType> out(input.w(), input.h());
    const int rc = dc_bilateral_host(g.raw(), detail::elem_kind<PixelType>(), input.data(), 1, distance,
                                     sigma_color, sigma_space, out.data());
    if (rc != DC_OK)
        throw gpu::error(rc, g.last_error());
    return out;
}

}  // namespace sens_loc::preprocess
#endif
