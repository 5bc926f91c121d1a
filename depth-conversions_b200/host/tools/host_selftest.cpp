// Drives the C++ host layer from C++ the way a reference user would:
//   host_selftest png <in.png> <out.png>
//       decode a 16 bit grey PNG with this build's codec and encode it again
//       (CPU only; tests compare both files with an independent decoder)
//   host_selftest api <intrinsic.txt> <depth.png> <outdir>
//       load_intrinsic + load PNG, then every sens_loc::conversion::* function
//       (Real = float) and quantiser; writes <outdir>/<name>.f32 (raw floats)
//       and <outdir>/<name>.png (what the quantisers return). Needs a B200.
#include "../sens_loc/conversion/depth_to_bearing.h"
#include "../sens_loc/io/intrinsics.h"
#include "../sens_loc/io/png.h"
#include "../sens_loc/preprocess/filter.h"

#include <cstdio>
#include <fstream>
#include <iostream>
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
        return save_u16(argv[3], img) ? 0 : 3;
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
             save_f32(dir + "median3.f32", preprocess::median_blur(range, 3));
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
