// The image-level cases of the reference's doctest suites, on the drop-in headers:
//   src/unittest/conversion/test_conversion_bearing.cpp:78-195
//   src/unittest/conversion/test_conversion_flexion.cpp:14-48
//   src/unittest/conversion/test_conversion_curvature.cpp:14-119
//   src/unittest/conversion/test_conversion_max_curve.cpp:68-104
//   src/unittest/conversion/test_conversion_laser.cpp:16-68
//   src/unittest/conversion/test_conversion_scaling.cpp:13-39
// The bodies follow the reference line by line: same calls, same template
// arguments (Real = double and float, PixelType = uchar and ushort, par_* with a
// task flow, the unclamped curvature_to_image), same golden images, same
// thresholds. What differs is only where OpenCV shows: cv::imread flags are
// dropped, cv::imwrite(path, img.data()) is io::write_image(path, img), a zeroed
// cv::Mat is math::image<T>(w, h), and tf:: below is a twelve-line stand-in for
// cpp-taskflow (the build carries no third-party headers).
//
//   ref_unittests <dir with the PNGs of src/unittest/conversion/>
// exit code 0 = every case passed. Needs a B200 (there is no CPU path).
#include "../sens_loc/conversion/depth_scaling.h"
#include "../sens_loc/conversion/depth_to_bearing.h"
#include "../sens_loc/conversion/depth_to_curvature.h"
#include "../sens_loc/conversion/depth_to_flexion.h"
#include "../sens_loc/conversion/depth_to_laserscan.h"
#include "../sens_loc/conversion/depth_to_max_curve.h"
#include "../sens_loc/io/image.h"
#include "../sens_loc/util/correctness_util.h"

#include <cstdio>
#include <functional>
#include <iostream>
#include <string>
#include <vector>

// ---- stand-ins: doctest, taskflow -----------------------------------------
namespace {
struct test_case {
    const char*           name;
    std::function<void()> body;
};
std::vector<test_case>& registry() {
    static std::vector<test_case> r;
    return r;
}
struct registrar {
    registrar(const char* n, std::function<void()> b) { registry().push_back({n, std::move(b)}); }
};
struct requirement_failed {
    std::string what;
};
std::string g_dir;
std::string at(const std::string& rel) { return g_dir + "/" + rel.substr(rel.find('/') + 1); }
}  // namespace
#define DC_CAT2(a, b) a##b
#define DC_CAT(a, b) DC_CAT2(a, b)
#define TEST_CASE(name)                                                         \
    static void DC_CAT(test_fn_, __LINE__)();                                   \
    static registrar DC_CAT(test_reg_, __LINE__)(name, DC_CAT(test_fn_, __LINE__)); \
    static void DC_CAT(test_fn_, __LINE__)()
#define REQUIRE(cond)                                                                          \
    do {                                                                                       \
        if (!(cond))                                                                           \
            throw requirement_failed{std::string(#cond) + "  (line " + std::to_string(__LINE__) + ")"}; \
    } while (0)

namespace tf {
class Taskflow {
  public:
    int emplace(std::function<void()> work) {
        _work.push_back(std::move(work));
        return int(_work.size()) - 1;
    }
    std::vector<std::function<void()>> _work;
};
class Executor {
  public:
    struct done {
        void wait() {}
    };
    done run(Taskflow& flow) {
        for (auto& w : flow._work)
            w();
        return {};
    }
};
}  // namespace tf

// ---- src/unittest/conversion/intrinsic.h:8-31 -----------------------------
const sens_loc::camera_models::pinhole<double> p = {
    /*w=*/960,      /*h=*/540,     /*fx=*/519.226,
    /*fy=*/479.462, /*cx=*/522.23, /*cy=*/272.737,
};
const sens_loc::camera_models::pinhole<double> p_double = {
    /*w=*/960,      /*h=*/540,     /*fx=*/519.226,
    /*fy=*/479.462, /*cx=*/522.23, /*cy=*/272.737,
};
const sens_loc::camera_models::pinhole<float> p_float = {
    /*w=*/960,       /*h=*/540,      /*fx=*/519.226f,
    /*fy=*/479.462f, /*cx=*/522.23f, /*cy=*/272.737f,
};
constexpr double deg_to_rad(double d) { return d * 3.141592653589793238462 / 180.; }
const sens_loc::camera_models::equirectangular<double> e_double{
    /*width=*/1799,
    /*height=*/397,
    /*theta_range=*/{deg_to_rad(50.), deg_to_rad(130.)}};

using namespace sens_loc;
using namespace sens_loc::conversion;
using namespace sens_loc::math;
using namespace std;

// ---- test_conversion_bearing.cpp:78-195 -----------------------------------
TEST_CASE("Convert depth image to vertical bearing angle image") {
    auto depth_image = io::load_image<ushort>(at("conversion/data0-depth.png"));
    REQUIRE(depth_image);
    auto laser_double = depth_to_laserscan<double, ushort>(*depth_image, p_double);

    auto ref_vert = io::load_image<uchar>(at("conversion/bearing-vertical.png"));
    REQUIRE(ref_vert);

    auto vertical_bearing = depth_to_bearing<direction::vertical>(laser_double, p_double);
    auto converted        = convert_bearing<double, uchar>(vertical_bearing);
    io::write_image(at("conversion/test_vertical.png"), converted);

    REQUIRE(util::average_pixel_error(converted, *ref_vert) < 0.5);
}

TEST_CASE("Convert depth image to vertical bearing angle image in parallel") {
    auto depth_image = io::load_image<ushort>(at("conversion/data0-depth.png"));
    REQUIRE(depth_image);
    auto laser = depth_to_laserscan<double, ushort>(*depth_image, p_double);

    auto ref_vert = io::load_image<uchar>(at("conversion/bearing-vertical.png"));
    REQUIRE(ref_vert);

    math::image<double> out_img(laser.w(), laser.h());
    {
        tf::Taskflow flow;
        par_depth_to_bearing<direction::vertical>(laser, p_double, out_img, flow);
        tf::Executor().run(flow).wait();
    }
    auto converted = convert_bearing<double, uchar>(out_img);
    io::write_image(at("conversion/test_vertical_parallel.png"), converted);

    REQUIRE(util::average_pixel_error(converted, *ref_vert) < 0.5);
}

TEST_CASE("Convert depth image to diagonal bearing angle image") {
    auto depth_image = io::load_image<ushort>(at("conversion/data0-depth.png"));
    REQUIRE(depth_image);
    auto laser_double = depth_to_laserscan<double, ushort>(*depth_image, p_double);

    auto ref_diag = io::load_image<ushort>(at("conversion/bearing-diagonal.png"));
    REQUIRE(ref_diag);

    auto diagonal_bearing = depth_to_bearing<direction::diagonal>(laser_double, p_double);
    auto converted        = convert_bearing<double, ushort>(diagonal_bearing);
    io::write_image(at("conversion/test_diagonal.png"), converted);

    REQUIRE(util::average_pixel_error(converted, *ref_diag) < 1.0);
}

TEST_CASE("Convert depth image to antidiagonal bearing angle image") {
    auto depth_image = io::load_image<ushort>(at("conversion/data0-depth.png"));
    REQUIRE(depth_image);

    auto laser_float = depth_to_laserscan<float, ushort>(*depth_image, p_float);

    auto ref_anti = io::load_image<uchar>(at("conversion/bearing-antidiagonal.png"));
    REQUIRE(ref_anti);

    auto antidiagonal_bearing = depth_to_bearing<direction::antidiagonal>(laser_float, p_float);
    auto converted_bearing    = convert_bearing<float, uchar>(antidiagonal_bearing);
    io::write_image(at("conversion/test_antidiagonal.png"), converted_bearing);

    REQUIRE(util::average_pixel_error(converted_bearing, *ref_anti) < 0.5);
}

TEST_CASE("Convert depth image to horizontal bearing angle image") {
    auto depth_image = io::load_image<ushort>(at("conversion/data0-depth.png"));
    REQUIRE(depth_image);
    auto laser_float = depth_to_laserscan<float, ushort>(*depth_image, p_float);
    auto ref_hor     = io::load_image<ushort>(at("conversion/bearing-horizontal.png"));
    REQUIRE(ref_hor);

    auto hor_bear      = depth_to_bearing<direction::horizontal>(laser_float, p_float);
    auto converted_flt = convert_bearing(hor_bear);
    io::write_image(at("conversion/test_horizontal.png"), converted_flt);

    REQUIRE(util::average_pixel_error(converted_flt, *ref_hor) < 1.0);
}

TEST_CASE("convert laserscan to horizontal bearing angle") {
    auto depth_image = io::load_image<ushort>(at("conversion/laserscan-depth.png"));
    REQUIRE(depth_image);

    const auto laser_double = math::convert<double>(*depth_image);
    const auto bearing      = conversion::depth_to_bearing<direction::horizontal>(laser_double, e_double);
    const auto converted    = conversion::convert_bearing<double, ushort>(bearing);
    io::write_image(at("conversion/test_horizontal_bearing_laserscan.png"), converted);

    auto ref_image = io::load_image<ushort>(at("conversion/bearing-horizontal-laserscan-reference.png"));
    REQUIRE(ref_image);

    REQUIRE(util::average_pixel_error(*ref_image, converted) < 0.5);
}

// ---- test_conversion_flexion.cpp:14-48 ------------------------------------
TEST_CASE("flexion image pinhole") {
    auto depth_image = io::load_image<ushort>(at("conversion/data0-depth.png"));
    REQUIRE(depth_image);

    auto ref_image = io::load_image<ushort>(at("conversion/flexion-reference.png"));
    REQUIRE(ref_image);

    auto laser_double = conversion::depth_to_laserscan<double, ushort>(*depth_image, p);

    const auto triple    = conversion::depth_to_flexion(laser_double, p);
    const auto converted = conversion::convert_flexion<ushort>(triple);
    io::write_image(at("conversion/test_flexion.png"), converted);

    REQUIRE(util::average_pixel_error(*ref_image, converted) < 0.5);
}

TEST_CASE("flexion image equirectangular") {
    auto depth_image = io::load_image<ushort>(at("conversion/laserscan-depth.png"));
    REQUIRE(depth_image);

    const auto laser_double = math::convert<double>(*depth_image);
    const auto flexion      = conversion::depth_to_flexion(laser_double, e_double);
    const auto converted    = conversion::convert_flexion<ushort>(flexion);
    io::write_image(at("conversion/test_flexion_laserscan.png"), converted);

    auto ref_image = io::load_image<ushort>(at("conversion/flexion-laserscan-reference.png"));
    REQUIRE(ref_image);

    REQUIRE(util::average_pixel_error(*ref_image, converted) < 0.5);
}

// ---- test_conversion_curvature.cpp:14-119 ---------------------------------
TEST_CASE("gaussian curvature") {
    auto depth_image = io::load_image<ushort>(at("conversion/data0-depth.png"));
    REQUIRE(depth_image);

    auto laser_double = conversion::depth_to_laserscan<double, ushort>(*depth_image, p);

    const auto gauss     = conversion::depth_to_gaussian_curvature(laser_double, p);
    const auto converted = conversion::curvature_to_image(gauss, *depth_image, {-20.}, {20.});
    io::write_image(at("conversion/test_gauss.png"), converted);

    auto ref_image = io::load_image<ushort>(at("conversion/gauss-reference.png"));
    REQUIRE(ref_image);

    REQUIRE(util::average_pixel_error(converted, *ref_image) < 5.);
}

TEST_CASE("mean curvature, clamped and unclamped conversion") {
    auto depth_image = io::load_image<ushort>(at("conversion/data0-depth.png"));
    REQUIRE(depth_image);

    auto laser_double = conversion::depth_to_laserscan<double, ushort>(*depth_image, p);

    const auto mean = conversion::depth_to_mean_curvature(laser_double, p);
    {  // SUBCASE("clamped conversion")
        const auto converted = conversion::curvature_to_image(mean, *depth_image, {-20.}, {20.});
        io::write_image(at("conversion/test_mean.png"), converted);

        auto ref_image = io::load_image<ushort>(at("conversion/mean-reference.png"));
        REQUIRE(ref_image);

        REQUIRE(util::average_pixel_error(converted, *ref_image) < 5.);
    }
    {  // SUBCASE("unclamped conversion"): the reference only requires that it runs
        const auto converted = conversion::curvature_to_image<ushort>(mean, *depth_image);
        io::write_image(at("conversion/test_mean_unclamped.png"), converted);
        REQUIRE(converted.w() == mean.w());
        REQUIRE(converted.h() == mean.h());
    }
}

TEST_CASE("gaussian curvature equirectangular") {
    auto depth_image = io::load_image<ushort>(at("conversion/laserscan-depth.png"));
    REQUIRE(depth_image);

    const auto laser_double = math::convert<double>(*depth_image);
    const auto gauss        = conversion::depth_to_gaussian_curvature(laser_double, e_double);
    const auto converted    = conversion::curvature_to_image(gauss, *depth_image, {-20.}, {20.});
    io::write_image(at("conversion/test_gauss_laserscan.png"), converted);

    auto ref_image = io::load_image<ushort>(at("conversion/gauss-laserscan-reference.png"));
    REQUIRE(ref_image);

    REQUIRE(util::average_pixel_error(converted, *ref_image) < 2.5);
}

TEST_CASE("mean curvature equirectangular") {
    auto depth_image = io::load_image<ushort>(at("conversion/laserscan-depth.png"));
    REQUIRE(depth_image);

    const auto laser_double = math::convert<double>(*depth_image);
    const auto mean         = conversion::depth_to_mean_curvature(laser_double, e_double);
    {
        const auto converted = conversion::curvature_to_image(mean, *depth_image, {-20.}, {20.});
        io::write_image(at("conversion/test_mean_laserscan.png"), converted);

        auto ref_image = io::load_image<ushort>(at("conversion/mean-laserscan-reference.png"));
        REQUIRE(ref_image);

        REQUIRE(util::average_pixel_error(converted, *ref_image) < 2.5);
    }
    {
        const auto converted = conversion::curvature_to_image<ushort>(mean, *depth_image);
        REQUIRE(converted.w() == mean.w());
    }
}

// ---- test_conversion_max_curve.cpp:68-104 ---------------------------------
TEST_CASE("depth image to max curve") {
    auto depth_image = io::load_image<ushort>(at("conversion/data0-depth.png"));
    REQUIRE(depth_image);
    {  // SUBCASE("double accuracy")
        auto ref_double = io::load_image<ushort>(at("conversion/max-curve-double.png"));
        REQUIRE(ref_double);

        const auto laser_double = depth_to_laserscan<double, ushort>(*depth_image, p);
        const auto curve_double = depth_to_max_curve(laser_double, p);

        const auto curve_ushort = convert_max_curve<ushort>(curve_double);
        io::write_image(at("conversion/test_max_curve_double.png"), curve_ushort);

        REQUIRE(util::average_pixel_error(*ref_double, curve_ushort) < 0.5);
    }
}

TEST_CASE("laserscan to max curve") {
    auto depth_image = io::load_image<ushort>(at("conversion/laserscan-depth.png"));
    REQUIRE(depth_image);

    const auto laser_double = math::convert<double>(*depth_image);
    const auto curve_double = depth_to_max_curve(laser_double, e_double);
    const auto curve_ushort = convert_max_curve<ushort>(curve_double);
    io::write_image(at("conversion/test_max_curve_laserscan.png"), curve_ushort);

    auto ref_double = io::load_image<ushort>(at("conversion/max-curve-laserscan.png"));
    REQUIRE(ref_double);
    REQUIRE(util::average_pixel_error(*ref_double, curve_ushort) < 0.5);
}

// ---- test_conversion_laser.cpp:16-68 --------------------------------------
TEST_CASE("convert depth image to laser-scan image") {
    auto depth_image = io::load_image<ushort>(at("conversion/data0-depth.png"));
    auto ref_depth_laser_image = io::load_image<ushort>(at("conversion/data0-depth-laser.png"));
    REQUIRE(depth_image);
    REQUIRE(ref_depth_laser_image);

    auto laser_float  = depth_to_laserscan(*depth_image, p_float);
    auto laser_double = depth_to_laserscan<double>(*depth_image, p);

    const auto laser_float_16u  = math::convert<ushort>(laser_float);
    const auto laser_double_16u = math::convert<ushort>(laser_double);
    io::write_image(at("conversion/test_depth_to_laserscan_float.png"), laser_float_16u);

    REQUIRE(util::average_pixel_error(laser_float_16u, *ref_depth_laser_image) < 1.0);
    REQUIRE(util::average_pixel_error(laser_double_16u, laser_float_16u) < 1.0);
}

TEST_CASE("convert depth image to laser-scan image parallel") {
    auto depth_image = io::load_image<ushort>(at("conversion/data0-depth.png"));
    auto ref_depth_laser_image = io::load_image<ushort>(at("conversion/data0-depth-laser.png"));
    REQUIRE(depth_image);
    REQUIRE(ref_depth_laser_image);

    math::image<float> laser_img(depth_image->w(), depth_image->h());
    {
        tf::Taskflow flow;
        par_depth_to_laserscan<float>(*depth_image, p_float, laser_img, flow);
        tf::Executor().run(flow).wait();
    }
    const auto laser_float_16u = math::convert<ushort>(laser_img);
    REQUIRE(util::average_pixel_error(laser_float_16u, *ref_depth_laser_image) < 1.0);
}

// ---- test_conversion_scaling.cpp:13-39 ------------------------------------
TEST_CASE("scale a depth image") {
    auto depth_image = io::load_image<ushort>(at("conversion/data0-depth.png"));
    REQUIRE(depth_image);
    auto ref = io::load_image<ushort>(at("conversion/scale-up.png"));
    REQUIRE(ref);

    const auto scaled = depth_scaling(*depth_image, 8.0, 0.0);
    REQUIRE(util::average_pixel_error(scaled, *ref) < 0.5);
}

TEST_CASE("Constant offset") {
    auto depth_image = io::load_image<ushort>(at("conversion/data0-depth.png"));
    REQUIRE(depth_image);
    auto ref = io::load_image<ushort>(at("conversion/scale-offset.png"));
    REQUIRE(ref);

    const auto scaled = depth_scaling(*depth_image, 1.0, 10000.0);
    REQUIRE(util::average_pixel_error(scaled, *ref) < 0.5);
}

// ---- par_* leaves the border untouched (depth_to_flexion.h:171-191) -------
TEST_CASE("par_depth_to_flexion writes interior pixels only") {
    auto depth_image = io::load_image<ushort>(at("conversion/data0-depth.png"));
    REQUIRE(depth_image);
    auto laser = depth_to_laserscan<float, ushort>(*depth_image, p_float);

    math::image<float> out(laser.w(), laser.h());
    for (std::size_t i = 0; i < out.pixels(); ++i)
        out.data()[i] = -7.f;
    {
        tf::Taskflow flow;
        par_depth_to_flexion(laser, p_float, out, flow);
        tf::Executor().run(flow).wait();
    }
    const auto whole = depth_to_flexion(laser, p_float);
    for (int v = 0; v < out.h(); ++v)
        for (int u = 0; u < out.w(); ++u) {
            const bool  border = u == 0 || v == 0 || u == out.w() - 1 || v == out.h() - 1;
            const float got    = out.at(pixel_coord<int>(u, v));
            REQUIRE(border ? got == -7.f : got == whole.at(pixel_coord<int>(u, v)));
        }
}

int main(int argc, char** argv) {
    if (argc != 2) {
        std::cerr << "usage: ref_unittests <dir with the PNGs of src/unittest/conversion/>\n";
        return 64;
    }
    g_dir      = argv[1];
    int failed = 0;
    for (const auto& t : registry()) {
        try {
            t.body();
            std::printf("[ ok ] %s\n", t.name);
        } catch (const requirement_failed& f) {
            std::printf("[FAIL] %s: REQUIRE( %s )\n", t.name, f.what.c_str());
            ++failed;
        } catch (const std::exception& e) {
            std::printf("[FAIL] %s: exception: %s\n", t.name, e.what());
            ++failed;
        }
    }
    std::printf("%zu cases, %d failed\n", registry().size(), failed);
    return failed == 0 ? 0 : 1;
}
