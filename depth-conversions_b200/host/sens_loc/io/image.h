// io::load_image<PixelType> (src/include/sens_loc/io/image.h:15-31): read an image
// file, none if it cannot be read or its pixel type is not PixelType. The
// reference forwards to cv::imread; here the tool's own PNG codec (io/png.h)
// reads the one family depth2x moves: greyscale PNG, 8 or 16 bit.
#ifndef DCB200_SENS_LOC_IO_IMAGE_H
#define DCB200_SENS_LOC_IO_IMAGE_H

#include "../math/image.h"
#include "png.h"

#include <optional>
#include <string>
#include <vector>

namespace sens_loc::io {

/// Further arguments (cv::IMREAD_UNCHANGED in the reference's calls) are accepted
/// and ignored: the file is always read unchanged.
template <typename PixelType, typename... Ignored>
inline std::optional<math::image<PixelType>> load_image(const std::string& path, Ignored&&...) {
    static_assert(std::is_same_v<PixelType, ushort> || std::is_same_v<PixelType, uchar>,
                  "PNG holds 8 or 16 bit unsigned samples");
    std::vector<unsigned char> bytes, scratch;
    png::info                  hdr;
    if (png::read_file(path, bytes, hdr) != png::status::ok || hdr.color_type != 0 ||
        hdr.bit_depth != int(8 * sizeof(PixelType)))
        return std::nullopt;
    math::image<PixelType> img(hdr.w, hdr.h);
    if (png::decode_grey(bytes, hdr, img.data(), scratch) != png::status::ok)
        return std::nullopt;
    return img;
}

/// cv::imwrite(path, image.data()) of the reference's converters and tests.
template <typename PixelType>
inline bool write_image(const std::string& path, const math::image<PixelType>& img) {
    static_assert(std::is_same_v<PixelType, ushort> || std::is_same_v<PixelType, uchar>);
    std::vector<unsigned char> out, scratch;
    return !img.empty() &&
           png::encode_grey(img.data(), img.w(), img.h(), int(8 * sizeof(PixelType)), out, scratch) ==
               png::status::ok &&
           png::write_file(path, out) == png::status::ok;
}

}  // namespace sens_loc::io
#endif
