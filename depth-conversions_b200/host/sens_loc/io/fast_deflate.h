// A small deflate ENCODER for PNG scanline data: run-length matches (distance 1,
// what zlib's Z_RLE strategy finds) + one dynamic Huffman code per 256 KB block,
// written through a 64-bit bit buffer. zlib level 1 / Z_RLE spends 37 ms per
// 1920x1080x16 bit frame in its per-symbol bit packing (65-110 MB/s), which makes
// the PNG writers the bottleneck of depth2x; this encoder produces streams of
// the same size class several times faster. Output is a complete zlib stream
// (RFC 1950/1951) that any inflate reads.
#ifndef DCB200_SENS_LOC_IO_FAST_DEFLATE_H
#define DCB200_SENS_LOC_IO_FAST_DEFLATE_H

#include <cstddef>
#include <cstdint>

namespace sens_loc::io {

/// Upper bound of the stream size for `n` input bytes.
std::size_t rle_deflate_bound(std::size_t n);

/// Compresses in[0..n) into out (capacity >= rle_deflate_bound(n)); returns the
/// number of bytes written.
std::size_t rle_deflate(const unsigned char* in, std::size_t n, unsigned char* out);

/// Whole-buffer inflate of a zlib stream (fast_inflate.cpp): decodes in[0..n_in)
/// into exactly n_out bytes. `in` must be followed by 32 readable bytes and `out`
/// by none. Returns true only when the stream decoded to exactly n_out bytes and
/// its Adler-32 trailer matches; on false the caller falls back to zlib.
bool fast_inflate(const unsigned char* in, std::size_t n_in, unsigned char* out, std::size_t n_out);

/// Adler-32 of p[0..n) continued from `adler` (1 for a fresh checksum); SSSE3
/// where available, zlib's otherwise (adler32.cpp).
std::uint32_t adler32_fast(std::uint32_t adler, const unsigned char* p, std::size_t n);

}  // namespace sens_loc::io
#endif
