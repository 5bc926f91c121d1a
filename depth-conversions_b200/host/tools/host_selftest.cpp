// Drives the C++ host layer from C++ the way a reference user would:
//   host_selftest png <in.png> <out.png>
//       decode a 16 bit grey PNG with this build's codec and encode it again
//       (CPU only; tests compare both files with an independent decoder)
//   host_selftest inflatefuzz <seed> <iterations>
//       the PNG reader's inflate (io/fast_inflate.cpp) against streams zlib writes
//   host_selftest pngtime <in.png> <reps>
//       milliseconds per decode and per encode of that file on one thread
//   host_selftest api <intrinsic.txt> <depth.png> <outdir>
//       load_intrinsic + load PNG, then every sens_loc::conversion::* function
//       (Real = float) and quantiser; writes <outdir>/<name>.f32 (raw floats)
//       and <outdir>/<name>.png (what the quantisers return). Needs a B200.
#include "../sens_loc/conversion/depth_to_bearing.h"
#include "../sens_loc/io/intrinsics.h"
#include "../sens_loc/io/png.h"
#include "../sens_loc/preprocess/filter.h"

#include "../sens_loc/io/fast_deflate.h"

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <iostream>
#include <random>
#include <cstring>
#include <zlib.h>
#include <string>

using namespace sens_loc;
namespace png = sens_loc::io::png;

static bool load_u16(const std::string& path, math::image<ushort>& img) {
    std::vector<unsigned char> bytes, scratch;
    png::info                  hdr;
    if (png::read_file(path, bytes, hdr) != png::status::ok || hdr.color_type != 0 ||
        hdr.bit_depth != 16)
        return false;
    img = math::image<ushort>(hdr.w, hdr.h);
    return png::decode_grey(bytes, hdr, img.data(), scratch) == png::status::ok;
}
static bool save_u16(const std::string& path, const math::image<ushort>& img) {
    std::vector<unsigned char> out, scratch;
    return png::encode_grey(img.data(), img.w(), img.h(), 16, out, scratch) == png::status::ok &&
           png::write_file(path, out) == png::status::ok;
}
static bool save_f32(const std::string& path, const math::image<float>& img) {
    std::FILE* f = std::fopen(path.c_str(), "wb");
    if (f == nullptr)
        return false;
    const bool ok = std::fwrite(img.data(), sizeof(float), img.pixels(), f) == img.pixels();
    return std::fclose(f) == 0 && ok;
}

static bool save_f64(const std::string& path, const math::image<double>& img) {
    std::FILE* f = std::fopen(path.c_str(), "wb");
    if (f == nullptr)
        return false;
    const bool ok = std::fwrite(img.data(), sizeof(double), img.pixels(), f) == img.pixels();
    return std::fclose(f) == 0 && ok;
}

int main(int argc, char** argv) try {
    const std::string cmd = argc > 1 ? argv[1] : "";
    if (cmd == "png" && argc == 4) {
        math::image<ushort> img;
        if (!load_u16(argv[2], img))
            return 2;
        const auto dc = png::decode_statistics();
        std::printf("inflate %s\n", dc.fast != 0 ? "fast" : "zlib");
        return save_u16(argv[3], img) ? 0 : 3;
    }
    if (cmd == "inflatefuzz" && argc == 4) {
        // io::fast_inflate against zlib's deflate at every level / strategy / flush mode
        // on inputs from all-literal to all-match; then the same streams damaged
        // (rejected, or accepted exactly as zlib does; never a crash). Prints "ok <streams> <damaged accepted>".
        std::mt19937 rng(unsigned(std::atoi(argv[2])));
        const int    iters = std::atoi(argv[3]);
        long         streams = 0, accepted = 0;
        for (int it = 0; it < iters; ++it) {
            const std::size_t n = it < 8 ? std::size_t(it) : std::size_t(rng() % (it % 5 == 0 ? 600000 : 40000)) + 1;
            std::vector<unsigned char> src(n);
            const int                  kind = int(rng() % 6);
            for (std::size_t i = 0; i < n; ++i) {
                switch (kind) {
                case 0: src[i] = (unsigned char)rng(); break;                                  // incompressible
                case 1: src[i] = (unsigned char)(rng() % 4 == 0 ? rng() % 16 : 0); break;       // runs
                case 2: src[i] = i >= 3000 && rng() % 64 != 0 ? src[i - 3000] : (unsigned char)rng(); break;
                case 3: src[i] = (unsigned char)((i & 1) != 0 ? rng() % 32 + 100 : rng() % 3); break;  // depth-like
                case 4: src[i] = i >= 7 && rng() % 512 != 0 ? src[i - 1 - rng() % 7] : (unsigned char)rng(); break;
                default: src[i] = (unsigned char)(i >= 30000 && rng() % 256 != 0 ? src[i - 30000] : rng() % 200); break;
                }
            }
            const int level    = int(rng() % 10);
            const int strat[5] = {Z_DEFAULT_STRATEGY, Z_FILTERED, Z_HUFFMAN_ONLY, Z_RLE, Z_FIXED};
            z_stream  zs{};
            if (deflateInit2(&zs, level, Z_DEFLATED, 9 + int(rng() % 7), 1 + int(rng() % 9), strat[rng() % 5]) != Z_OK)
                return 4;
            std::vector<unsigned char> z(deflateBound(&zs, uLong(n)) + 4096 + n / 8);
            zs.next_out  = z.data();
            zs.avail_out = uInt(z.size());
            std::size_t fed = 0;
            const bool  chop = rng() % 3 == 0;  // several blocks through flushes
            while (fed < n) {
                const std::size_t step = chop ? std::min<std::size_t>(n - fed, rng() % 5000 + 1) : n - fed;
                zs.next_in             = src.data() + fed;
                zs.avail_in            = uInt(step);
                fed += step;
                const int fl = fed == n ? Z_FINISH : (rng() % 2 != 0 ? Z_FULL_FLUSH : Z_SYNC_FLUSH);
                if (deflate(&zs, fl) == Z_STREAM_ERROR)
                    return 4;
            }
            if (n == 0 && deflate(&zs, Z_FINISH) != Z_STREAM_END)
                return 4;
            const std::size_t zn = zs.total_out;
            deflateEnd(&zs);
            z.resize(zn + 32);
            std::vector<unsigned char> back(n + 1, 0xa5);
            if (!io::fast_inflate(z.data(), zn, back.data(), n) || (n != 0 && std::memcmp(back.data(), src.data(), n) != 0) ||
                back[n] != 0xa5) {
                std::printf("FAIL iter %d n %zu kind %d level %d\n", it, n, kind, level);
                return 5;
            }
            // wrong expected size must be refused
            if (n > 1 && io::fast_inflate(z.data(), zn, back.data(), n - 1)) {
                std::printf("FAIL short output accepted, iter %d\n", it);
                return 5;
            }
            ++streams;
            for (int d = 0; d < 6 && zn > 8; ++d) {  // damage: flipped bits, or a cut
                std::vector<unsigned char> bad(z.begin(), z.begin() + std::ptrdiff_t(zn));
                std::size_t                bn = zn;
                if (d % 3 == 2) {
                    bn = rng() % zn;
                } else {
                    const int flips = 1 + int(rng() % 3);
                    for (int f = 0; f < flips; ++f)
                        bad[2 + rng() % (zn - 2)] ^= (unsigned char)(1u << (rng() % 8));
                }
                bad.resize(bn);
                bad.resize(bn + 32, 0);
                std::fill(back.begin(), back.end(), 0xa5);
                if (io::fast_inflate(bad.data(), bn, back.data(), n)) {
                    // Adler-32 is weak against a few changed literals, so a damaged stream
                    // can be valid: then zlib must accept it too, with the same bytes
                    std::vector<unsigned char> zz(n + 64);
                    uLongf                     l = uLongf(n + 64);
                    const int zr = uncompress(zz.data(), &l, bad.data(), uLong(bn));
                    if (zr != Z_OK || l != n || std::memcmp(zz.data(), back.data(), n) != 0) {
                        std::printf("FAIL damaged stream accepted, zlib disagrees (%d, %lu of %zu), iter %d cut %d bn %zu zn %zu\n",
                                    zr, (unsigned long)l, n, it, int(d % 3 == 2), bn, zn);
                        return 5;
                    }
                    ++accepted;
                }
                if (back[n] != 0xa5) {
                    std::printf("FAIL wrote past the output, iter %d\n", it);
                    return 5;
                }
            }
        }
        std::printf("ok %ld %ld\n", streams, accepted);
        return 0;
    }
    if (cmd == "pngtime" && argc == 4) {  // ms per decode / encode, single thread
        std::vector<unsigned char> bytes, scratch, out;
        png::info                  hdr;
        if (png::read_file(argv[2], bytes, hdr) != png::status::ok || hdr.bit_depth != 16)
            return 2;
        math::image<ushort> img(hdr.w, hdr.h);
        const int           reps = std::atoi(argv[3]);
        auto                now  = [] { return std::chrono::steady_clock::now(); };
        auto                t0   = now();
        for (int i = 0; i < reps; ++i)
            if (png::decode_grey(bytes, hdr, img.data(), scratch) != png::status::ok)
                return 3;
        auto t1 = now();
        for (int i = 0; i < reps; ++i)
            if (png::encode_grey(img.data(), img.w(), img.h(), 16, out, scratch) != png::status::ok)
                return 3;
        auto t2 = now();
        std::printf("decode_ms %.3f encode_ms %.3f\n",
                    std::chrono::duration<double, std::milli>(t1 - t0).count() / reps,
                    std::chrono::duration<double, std::milli>(t2 - t1).count() / reps);
        return 0;
    }
    if (cmd == "api" && argc == 5) {
        std::ifstream in{argv[2]};
        const auto    k = io::camera<float, camera_models::pinhole>::load_intrinsic(in);
        if (!k)
            return 2;
        math::image<ushort> depth;
        if (!load_u16(argv[3], depth))
            return 3;
        const std::string dir = std::string(argv[4]) + "/";
        using namespace conversion;
        const math::image<float> range = depth_to_laserscan<float, ushort>(depth, *k);
        bool ok = save_f32(dir + "range.f32", range) && save_u16(dir + "range.png", convert_range(range));
        const auto bh = depth_to_bearing<direction::horizontal>(range, *k);
        const auto bv = depth_to_bearing<direction::vertical>(range, *k);
        const auto bd = depth_to_bearing<direction::diagonal>(range, *k);
        const auto ba = depth_to_bearing<direction::antidiagonal>(range, *k);
        ok = ok && save_f32(dir + "bearing_h.f32", bh) && save_f32(dir + "bearing_v.f32", bv) &&
             save_f32(dir + "bearing_d.f32", bd) && save_f32(dir + "bearing_a.f32", ba) &&
             save_u16(dir + "bearing_h.png", convert_bearing<float, ushort>(bh));
        const bool         want[4] = {true, false, true, false};
        math::image<float> all[4];
        depth_to_bearing_all(range, *k, want, all);
        ok = ok && save_f32(dir + "bearing_h_fused.f32", all[0]) && all[1].empty() &&
             save_f32(dir + "bearing_d_fused.f32", all[2]);
        const auto fl = depth_to_flexion(range, *k);
        ok = ok && save_f32(dir + "flexion.f32", fl) && save_u16(dir + "flexion.png", convert_flexion<ushort>(fl));
        const auto g = depth_to_gaussian_curvature(range, *k);
        const auto m = depth_to_mean_curvature(range, *k);
        ok = ok && save_f32(dir + "gauss.f32", g) && save_f32(dir + "mean.f32", m) &&
             save_u16(dir + "gauss.png", curvature_to_image<ushort>(g, range, {-20.f}, {20.f})) &&
             save_u16(dir + "mean.png", curvature_to_image<ushort>(m, range, {-20.f}, {20.f}));
        const auto mc = depth_to_max_curve(range, *k);
        ok = ok && save_f32(dir + "max_curve.f32", mc) &&
             save_u16(dir + "max_curve.png", convert_max_curve<ushort>(mc)) &&
             save_u16(dir + "scaled.png", depth_scaling(depth, 8., 3.)) &&
             save_u16(dir + "median5.png", preprocess::median_blur(depth, 5)) &&
             save_f32(dir + "median3.f32", preprocess::median_blur(range, 3)) &&
             save_f32(dir + "gaussblur.f32", preprocess::gaussian_blur(range, {5, 5}, 1.5)) &&
             save_u16(dir + "gaussblur.png", preprocess::gaussian_blur(depth, {3, 3}, 0.)) &&
             save_f32(dir + "bilateral_d.f32", preprocess::bilateral_filter(range, 30., 5)) &&
             save_f32(dir + "bilateral_s.f32", preprocess::bilateral_filter(range, 30., 2.0)) &&
             save_u16(dir + "bilateral.png", preprocess::bilateral_filter(depth, 30., 2.0));
        // preconditions: a size mismatch is rejected, not computed
        bool threw = false;
        try {
            (void)depth_to_flexion(math::image<float>(k->w() / 2, k->h()), *k);
        } catch (const std::invalid_argument&) {
            threw = true;
        }
        return ok && threw ? 0 : 4;
    }
    if (cmd == "api64" && argc == 5) {
        // the Real = double instantiations, as src/unittest/conversion/*.cpp call them
        std::ifstream in{argv[2]};
        const auto    k = io::camera<double, camera_models::pinhole>::load_intrinsic(in);
        if (!k)
            return 2;
        math::image<ushort> depth;
        if (!load_u16(argv[3], depth))
            return 3;
        const std::string dir = std::string(argv[4]) + "/";
        using namespace conversion;
        const math::image<double> range = depth_to_laserscan<double, ushort>(depth, *k);
        bool ok = save_f64(dir + "range.f64", range);
        const auto bh = depth_to_bearing<direction::horizontal, double>(range, *k);
        const auto ba = depth_to_bearing<direction::antidiagonal, double>(range, *k);
        ok = ok && save_f64(dir + "bearing_h.f64", bh) && save_f64(dir + "bearing_a.f64", ba) &&
             save_u16(dir + "bearing_h64.png", convert_bearing<double, ushort>(bh));
        const auto fl = depth_to_flexion(range, *k);
        ok = ok && save_f64(dir + "flexion.f64", fl) &&
             save_u16(dir + "flexion64.png", convert_flexion<ushort, double>(fl));
        const auto g = depth_to_gaussian_curvature(range, *k);
        const auto m = depth_to_mean_curvature(range, *k);
        ok = ok && save_f64(dir + "gauss.f64", g) && save_f64(dir + "mean.f64", m) &&
             save_u16(dir + "gauss64.png", curvature_to_image<ushort, double>(g, range, {-20.}, {20.}));
        const auto mc = depth_to_max_curve(range, *k);
        ok = ok && save_f64(dir + "max_curve.f64", mc) &&
             save_u16(dir + "max_curve64.png", convert_max_curve<ushort, double>(mc));
        return ok ? 0 : 4;
    }
    std::cerr << "usage: host_selftest png <in.png> <out.png> | api <intrinsic> <depth.png> <outdir>\n";
    return 64;
} catch (const std::exception& e) {
    std::cerr << "host_selftest: " << e.what() << "\n";
    return 70;
}
