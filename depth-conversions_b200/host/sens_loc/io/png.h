// Minimal PNG codec on zlib for the one image format depth2x moves: greyscale,
// 8 or 16 bit, non-interlaced. Replaces cv::imread(IMREAD_UNCHANGED) /
// cv::imwrite on the tool's path (src/include/sens_loc/io/image.h:21-31,
// src/apps/depth2x/converter_*.h.inl): OpenCV's C++ libraries are not part of
// this build. 16 bit samples are big endian in the file and native in memory.
#ifndef DCB200_SENS_LOC_IO_PNG_H
#define DCB200_SENS_LOC_IO_PNG_H

#include <cstdint>
#include <string>
#include <vector>

namespace sens_loc::io::png {

struct info {
    int w = 0, h = 0;
    int bit_depth = 0;   // 1, 2, 4, 8, 16
    int color_type = 0;  // 0 grey, 2 RGB, 3 palette, 4 grey+alpha, 6 RGBA
    int interlace = 0;
};

enum class status { ok, io_error, not_png, corrupt, unsupported };

/// Reads the file and parses the header only.
status read_file(const std::string& path, std::vector<unsigned char>& bytes, info& hdr);

/// Decodes a greyscale image from `bytes` (a whole PNG file) into `dst`, which
/// must hold w*h samples of 1 (bit depth 8) or 2 bytes (bit depth 16).
/// `scratch` is reused between calls to avoid allocations.
status decode_grey(const std::vector<unsigned char>& bytes, const info& hdr, void* dst,
                   std::vector<unsigned char>& scratch);

/// How many successful decodes of this process went through io::fast_inflate and how
/// many through zlib's inflate (its fallback). For tests and diagnostics.
struct decode_counts {
    unsigned long fast, zlib;
};
decode_counts decode_statistics();

/// Encodes w*h greyscale samples (bit_depth 8 or 16) as PNG: filter "sub",
/// zlib level `level` with the RLE strategy (what OpenCV's imwrite uses).
status encode_grey(const void* src, int w, int h, int bit_depth, std::vector<unsigned char>& out,
                   std::vector<unsigned char>& scratch, int level = 1);

status write_file(const std::string& path, const std::vector<unsigned char>& bytes);

}  // namespace sens_loc::io::png
#endif
