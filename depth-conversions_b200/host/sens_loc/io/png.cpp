#include "png.h"

#include "fast_deflate.h"

#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <zlib.h>
#if defined(__SSE2__)
#include <emmintrin.h>
#endif

namespace sens_loc::io::png {
namespace {

std::atomic<unsigned long> g_fast{0}, g_zlib{0};

const unsigned char kSig[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};

std::uint32_t be32(const unsigned char* p) {
    return (std::uint32_t(p[0]) << 24) | (std::uint32_t(p[1]) << 16) | (std::uint32_t(p[2]) << 8) |
           std::uint32_t(p[3]);
}
void put32(std::vector<unsigned char>& v, std::uint32_t x) {
    v.push_back((unsigned char)(x >> 24));
    v.push_back((unsigned char)(x >> 16));
    v.push_back((unsigned char)(x >> 8));
    v.push_back((unsigned char)x);
}
void chunk(std::vector<unsigned char>& out, const char type[4], const unsigned char* data,
           std::size_t n) {
    put32(out, std::uint32_t(n));
    const std::size_t at = out.size();
    out.insert(out.end(), type, type + 4);
    if (n != 0)
        out.insert(out.end(), data, data + n);
    put32(out, std::uint32_t(crc32(0L, out.data() + at, uInt(n + 4))));
}
int paeth(int a, int b, int c) {
    const int p = a + b - c, pa = std::abs(p - a), pb = std::abs(p - b), pc = std::abs(p - c);
    return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
}


// Filter type 1 (sub) for 16 bit samples: every byte is the sum of the bytes two
// positions apart up to here. Writes the un-filtered big-endian row back to `cur`
// (the next row's filter may look at it) and the little-endian samples to `o`.
void unsub16(unsigned char* cur, unsigned char* o, std::size_t w) {
    std::size_t k = 0;
#if defined(__SSE2__)
    __m128i carry = _mm_setzero_si128();  // the previous pixel in all eight lanes
    for (; k + 8 <= w; k += 8) {
        __m128i v = _mm_loadu_si128(reinterpret_cast<const __m128i*>(cur + 2 * k));
        v         = _mm_add_epi8(v, _mm_slli_si128(v, 2));  // byte-wise prefix sums over 8 pixels
        v         = _mm_add_epi8(v, _mm_slli_si128(v, 4));
        v         = _mm_add_epi8(v, _mm_slli_si128(v, 8));
        v         = _mm_add_epi8(v, carry);
        _mm_storeu_si128(reinterpret_cast<__m128i*>(cur + 2 * k), v);
        _mm_storeu_si128(reinterpret_cast<__m128i*>(o + 2 * k),
                         _mm_or_si128(_mm_slli_epi16(v, 8), _mm_srli_epi16(v, 8)));
        carry = _mm_shuffle_epi32(_mm_shufflehi_epi16(v, 0xff), 0xff);
    }
    unsigned hi = k != 0 ? cur[2 * k - 2] : 0u, lo = k != 0 ? cur[2 * k - 1] : 0u;
#else
    unsigned hi = 0, lo = 0;
#endif
    for (; k < w; ++k) {
        hi             = (hi + cur[2 * k]) & 0xffu;
        lo             = (lo + cur[2 * k + 1]) & 0xffu;
        cur[2 * k]     = (unsigned char)hi;
        cur[2 * k + 1] = (unsigned char)lo;
        o[2 * k]       = (unsigned char)lo;
        o[2 * k + 1]   = (unsigned char)hi;
    }
}

}  // namespace

status read_file(const std::string& path, std::vector<unsigned char>& bytes, info& hdr) {
    std::FILE* f = std::fopen(path.c_str(), "rb");
    if (f == nullptr)
        return status::io_error;
    std::fseek(f, 0, SEEK_END);
    const long n = std::ftell(f);
    std::fseek(f, 0, SEEK_SET);
    if (n < 0) {
        std::fclose(f);
        return status::io_error;
    }
    bytes.resize(std::size_t(n));
    const std::size_t got = n != 0 ? std::fread(bytes.data(), 1, std::size_t(n), f) : 0;
    std::fclose(f);
    if (got != std::size_t(n))
        return status::io_error;
    if (n < 8 + 25 || std::memcmp(bytes.data(), kSig, 8) != 0)
        return status::not_png;
    const unsigned char* p = bytes.data() + 8;
    if (be32(p) != 13 || std::memcmp(p + 4, "IHDR", 4) != 0)
        return status::corrupt;
    hdr.w          = int(be32(p + 8));
    hdr.h          = int(be32(p + 12));
    hdr.bit_depth  = p[16];
    hdr.color_type = p[17];
    hdr.interlace  = p[20];
    if (hdr.w <= 0 || hdr.h <= 0 || p[18] != 0 || p[19] != 0)
        return status::corrupt;
    return status::ok;
}

status decode_grey(const std::vector<unsigned char>& bytes, const info& hdr, void* dst,
                   std::vector<unsigned char>& scratch) {
    if (hdr.color_type != 0 || hdr.interlace != 0 || (hdr.bit_depth != 8 && hdr.bit_depth != 16))
        return status::unsupported;
    const std::size_t bpp = std::size_t(hdr.bit_depth / 8), row = std::size_t(hdr.w) * bpp;
    const std::size_t raw_size = (row + 1) * std::size_t(hdr.h);
    if (raw_size > std::size_t(0xffffffffu))
        return status::unsupported;  // zlib's 32-bit counters (uInt avail_out)
    scratch.resize(raw_size);

    // Fast path: gather the IDAT payloads (libpng writes 8 KB chunks) and decode the
    // zlib stream in one go with the table-driven inflate of fast_inflate.cpp, which
    // also verifies the Adler-32 trailer. Anything it does not like -> zlib below.
    // DC_PNG_ZLIB=1 in the environment skips the fast path.
    static const bool zlib_only = std::getenv("DC_PNG_ZLIB") != nullptr;
    bool              complete  = false;
    if (!zlib_only) {
        static thread_local std::vector<unsigned char> idat;
        idat.clear();
        std::size_t pos = 8;
        while (pos + 12 <= bytes.size()) {
            const std::uint32_t  len  = be32(bytes.data() + pos);
            const unsigned char* type = bytes.data() + pos + 4;
            if (pos + 12 + len > bytes.size() || std::memcmp(type, "IEND", 4) == 0)
                break;
            if (std::memcmp(type, "IDAT", 4) == 0)
                idat.insert(idat.end(), type + 4, type + 4 + len);
            pos += 12 + len;
        }
        const std::size_t n = idat.size();
        idat.resize(n + 32, 0);  // fast_inflate reads whole words past the end
        complete = fast_inflate(idat.data(), n, scratch.data(), raw_size);
    }
    const bool via_zlib = !complete;
    if (via_zlib) {
        z_stream zs{};
        if (inflateInit(&zs) != Z_OK)
            return status::corrupt;
        zs.next_out  = scratch.data();
        zs.avail_out = uInt(raw_size);
        int         zr  = Z_OK;
        std::size_t pos = 8;
        bool        end = false;
        while (!end && pos + 12 <= bytes.size()) {
            const std::uint32_t  len  = be32(bytes.data() + pos);
            const unsigned char* type = bytes.data() + pos + 4;
            if (pos + 12 + len > bytes.size())
                break;
            if (std::memcmp(type, "IDAT", 4) == 0 && zr == Z_OK) {
                zs.next_in  = const_cast<unsigned char*>(type + 4);
                zs.avail_in = len;
                zr          = inflate(&zs, Z_NO_FLUSH);
            } else if (std::memcmp(type, "IEND", 4) == 0) {
                end = true;
            }
            pos += 12 + len;
        }
        // the whole stream, trailer included: Z_STREAM_END means zlib has verified the
        // Adler-32 (a cut or trailer-less stream that happens to fill the image is
        // refused, exactly as the fast path refuses it)
        complete = zs.total_out == raw_size && zr == Z_STREAM_END;
        inflateEnd(&zs);
    }
    if (!complete)
        return status::corrupt;
    (via_zlib ? g_zlib : g_fast).fetch_add(1, std::memory_order_relaxed);

    unsigned char*       out  = static_cast<unsigned char*>(dst);
    const unsigned char* prev = nullptr;
    for (int y = 0; y < hdr.h; ++y) {
        unsigned char* line = scratch.data() + std::size_t(y) * (row + 1);
        const int      ft   = line[0];
        unsigned char* cur  = line + 1;
        unsigned char* o = out + std::size_t(y) * row;
        if (ft == 1 && bpp == 2) {  // the common case, un-filtered and byte-swapped in one pass
            unsub16(cur, o, std::size_t(hdr.w));
            prev = cur;
            continue;
        }
        switch (ft) {
        case 0: break;
        case 1:
            for (std::size_t i = bpp; i < row; ++i)
                cur[i] = (unsigned char)(cur[i] + cur[i - bpp]);
            break;
        case 2:
            if (prev != nullptr)
                for (std::size_t i = 0; i < row; ++i)
                    cur[i] = (unsigned char)(cur[i] + prev[i]);
            break;
        case 3:
            for (std::size_t i = 0; i < row; ++i) {
                const int a = i >= bpp ? cur[i - bpp] : 0, b = prev != nullptr ? prev[i] : 0;
                cur[i]      = (unsigned char)(cur[i] + ((a + b) >> 1));
            }
            break;
        case 4:
            for (std::size_t i = 0; i < row; ++i) {
                const int a = i >= bpp ? cur[i - bpp] : 0, b = prev != nullptr ? prev[i] : 0;
                const int c = (i >= bpp && prev != nullptr) ? prev[i - bpp] : 0;
                cur[i]      = (unsigned char)(cur[i] + paeth(a, b, c));
            }
            break;
        default: return status::corrupt;
        }
        if (bpp == 2) {
            for (std::size_t k = 0; k < std::size_t(hdr.w); ++k) {  // big endian -> little endian
                unsigned short v;
                std::memcpy(&v, cur + 2 * k, 2);
                v = (unsigned short)((v >> 8) | (v << 8));
                std::memcpy(o + 2 * k, &v, 2);
            }
        } else {
            std::memcpy(o, cur, row);
        }
        prev = cur;
    }
    return status::ok;
}

decode_counts decode_statistics() { return {g_fast.load(), g_zlib.load()}; }

status encode_grey(const void* src, int w, int h, int bit_depth, std::vector<unsigned char>& out,
                   std::vector<unsigned char>& scratch, int level) {
    if (w <= 0 || h <= 0 || (bit_depth != 8 && bit_depth != 16))
        return status::unsupported;
    const std::size_t bpp = std::size_t(bit_depth / 8), row = std::size_t(w) * bpp;
    const std::size_t raw_size = (row + 1) * std::size_t(h);
    if (raw_size > std::size_t(0x7fffffffu))
        return status::unsupported;  // one IDAT chunk (31-bit length), zlib's 32-bit counters
    // DC_PNG_ZLIB=1 in the environment: zlib level-1 / Z_RLE (what cv::imwrite
    // produces) instead of the run-length + Huffman encoder of fast_deflate.cpp
    static const bool use_zlib = std::getenv("DC_PNG_ZLIB") != nullptr;
    const uLong       bound    = uLong(std::max<std::size_t>(compressBound(uLong(raw_size)),
                                                             rle_deflate_bound(raw_size)));
    scratch.resize(raw_size + bound);
    unsigned char*       raw = scratch.data();
    unsigned char*       zed = scratch.data() + raw_size;
    const unsigned char* in  = static_cast<const unsigned char*>(src);
    for (int y = 0; y < h; ++y) {  // filter type 1 (sub) on big-endian samples
        unsigned char*       line = raw + std::size_t(y) * (row + 1);
        const unsigned char* s    = in + std::size_t(y) * row;
        line[0]                   = 1;
        unsigned char* cur        = line + 1;
        if (bpp == 2) {
            // per pixel, no loop-carried state (vectorisable): the two bytes of the
            // big-endian sample minus the bytes of the previous sample
            const auto* px = reinterpret_cast<const unsigned short*>(s);
            unsigned    x0 = px[0];
            cur[0] = (unsigned char)(x0 >> 8), cur[1] = (unsigned char)x0;
#if defined(__GNUC__)
#pragma GCC ivdep
#endif
            for (std::size_t k = 1; k < std::size_t(w); ++k) {
                const unsigned x = px[k], p = px[k - 1];
                const unsigned hi = ((x >> 8) - (p >> 8)) & 0xffu, lo = (x - p) & 0xffu;
                const unsigned short o = (unsigned short)(hi | (lo << 8));
                std::memcpy(cur + 2 * k, &o, 2);
            }
        } else {
            unsigned p = 0;
            for (std::size_t i = 0; i < row; ++i) {
                cur[i] = (unsigned char)(s[i] - p);
                p      = s[i];
            }
        }
    }
    std::size_t zn = 0;
    if (!use_zlib) {
        zn = rle_deflate(raw, raw_size, zed);
    } else {
        z_stream zs{};
        if (deflateInit2(&zs, level, Z_DEFLATED, 15, 8, Z_RLE) != Z_OK)
            return status::corrupt;
        zs.next_in   = raw;
        zs.avail_in  = uInt(raw_size);
        zs.next_out  = zed;
        zs.avail_out = uInt(bound);
        const int zr = deflate(&zs, Z_FINISH);
        zn           = zs.total_out;
        deflateEnd(&zs);
        if (zr != Z_STREAM_END)
            return status::corrupt;
    }

    out.clear();
    out.reserve(zn + 64);
    out.insert(out.end(), kSig, kSig + 8);
    unsigned char ihdr[13];
    const std::uint32_t uw = std::uint32_t(w), uh = std::uint32_t(h);
    ihdr[0] = (unsigned char)(uw >> 24), ihdr[1] = (unsigned char)(uw >> 16);
    ihdr[2] = (unsigned char)(uw >> 8), ihdr[3] = (unsigned char)uw;
    ihdr[4] = (unsigned char)(uh >> 24), ihdr[5] = (unsigned char)(uh >> 16);
    ihdr[6] = (unsigned char)(uh >> 8), ihdr[7] = (unsigned char)uh;
    ihdr[8] = (unsigned char)bit_depth, ihdr[9] = 0, ihdr[10] = 0, ihdr[11] = 0, ihdr[12] = 0;
    chunk(out, "IHDR", ihdr, 13);
    chunk(out, "IDAT", zed, zn);
    chunk(out, "IEND", nullptr, 0);
    return status::ok;
}

status write_file(const std::string& path, const std::vector<unsigned char>& bytes) {
    std::FILE* f = std::fopen(path.c_str(), "wb");
    if (f == nullptr)
        return status::io_error;
    const std::size_t n = std::fwrite(bytes.data(), 1, bytes.size(), f);
    const int         c = std::fclose(f);
    return (n == bytes.size() && c == 0) ? status::ok : status::io_error;
}

}  // namespace sens_loc::io::png
