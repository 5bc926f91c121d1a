// depth2x on B200: the reference's batch conversion tool
// (src/apps/depth2x/main.cpp:86-378) with the same flags, sub-commands,
// messages and exit codes; the conversions run in libdepthconv_b200.so.
// Exit 0 iff every index of [--start, --end] was converted (main.cpp:376).
#include "cli.h"
#include "console.h"
#include "pipeline.h"

#include "../../../../include/depthconv_b200.h"
#include "../../sens_loc/io/intrinsics.h"

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <iomanip>
#include <iostream>
#include <stdexcept>

using namespace sens_loc;
using namespace sens_loc::apps;

namespace {

int run(int argc, char** argv) {
    options opt;
    int     exit_code = 0;
    switch (parse_command_line(argc, argv, opt, exit_code)) {
    case parse_result::exit_ok:
    case parse_result::exit_error: return exit_code;
    case parse_result::run: break;
    }

    convert_spec spec;
    spec.input = opt.files.input;
    // apps/util/batch_converter.h:24-30 str_to_depth_type
    spec.in_kind = opt.input_type == "pinhole-range" ? DC_IN_U16_RANGE : DC_IN_U16_DEPTH;

    const bool is_scale = opt.subcommand == "scale";
    if (!is_scale) {
        // main.cpp:312-343: the intrinsic is mandatory for everything but `scale`
        std::ifstream cali{opt.calibration_file};
        bool          loaded = false;
        if (opt.camera_model == "pinhole") {
            if (auto k = io::camera<float, camera_models::pinhole>::load_intrinsic(cali)) {
                spec.camera = {1, k->w(), k->h(), {k->fx(), k->fy(), k->cx(), k->cy()}};
                loaded      = true;
            }
        } else {
            if (auto k = io::camera<float, camera_models::equirectangular>::load_intrinsic(cali)) {
                spec.camera = {2, k->w(), k->h(), {k->theta_min(), k->theta_max(), 0.f, 0.f}};
                loaded      = true;
            }
        }
        if (!loaded) {
            std::cerr << console::err() << "Could not load intrinsic calibration \""
                      << console::bold() << opt.calibration_file << console::reset() << "\"!\n";
            return 1;
        }
    }

    if (opt.subcommand == "bearing") {
        // converters.h:28-32
        if (opt.files.horizontal.empty() && opt.files.vertical.empty() &&
            opt.files.diagonal.empty() && opt.files.antidiagonal.empty())
            throw std::invalid_argument{"Missing output pattern for at least one bearing direction"};
        spec.feature   = DC_BEARING;
        spec.output[0] = opt.files.horizontal;
        spec.output[1] = opt.files.vertical;
        spec.output[2] = opt.files.diagonal;
        spec.output[3] = opt.files.antidiagonal;
    } else if (opt.subcommand == "range") {
        spec.feature   = DC_RANGE;
        spec.output[0] = opt.files.output;
    } else if (opt.subcommand == "flexion") {
        spec.feature   = DC_FLEXION;
        spec.output[0] = opt.files.output;
    } else if (opt.subcommand == "max-curve") {
        spec.feature   = DC_MAX_CURVE;
        spec.output[0] = opt.files.output;
    } else if (opt.subcommand == "gauss-curvature" || opt.subcommand == "mean-curvature") {
        spec.feature = DC_CURVATURE;
        spec.output[opt.subcommand == "gauss-curvature" ? 0 : 1] = opt.files.output;
        spec.clamp_lo = float(opt.lower_bound);
        spec.clamp_hi = float(opt.upper_bound);
        if (!(spec.clamp_lo < spec.clamp_hi))
            throw std::invalid_argument{"--lower-bound must be smaller than --upper-bound"};
    } else {  // scale: converter_scale.cpp:9-19, works on the raw u16 values
        spec.feature   = DC_SCALE;
        spec.in_kind   = DC_IN_U16_RANGE;
        spec.output[0] = opt.files.output;
        spec.scale     = opt.scale_factor;
        spec.offset    = opt.scale_delta;
    }

    if (opt.median != 0 && opt.median != 3 && opt.median != 5)
        throw std::invalid_argument{"--median-blur takes 3 or 5"};
    spec.median = opt.median;
    profile_mark("run_batch begin");
    const run_stats r =
        run_batch(spec, opt.start_idx, opt.end_idx, opt.devices, opt.batch_size, opt.threads);
    // parallel_processing.h:64-83
    std::cout << std::endl;
    std::cerr << console::info() << "Processing " << console::bold() << r.processed
              << console::reset() << " images took " << console::bold() << std::fixed
              << std::setprecision(2) << std::floor(r.seconds * 100.) / 100. << console::reset()
              << " seconds!\n";
    if (r.failed > 0)
        std::cerr << console::warn() << "Encountered " << console::bold() << r.failed
                  << console::reset() << " problematic files!\n";
    profile_mark("run_batch end");
    return r.failed == 0 ? 0 : 1;
}

}  // namespace

// apps/util/tool_macro.h:11-39 (MAIN_HEAD / MAIN_TAIL)
int main(int argc, char** argv) try {
    sens_loc::apps::profile_mark("main");
    const int rc = run(argc, argv);
    // every output file is written and closed: leave without the CUDA runtime's
    // exit handlers (0.3 s of context teardown that the kernel does anyway)
    std::cout.flush();
    std::cerr.flush();
    std::quick_exit(rc);
} catch (const std::exception& e) {
    std::cerr << console::err() << "Severe problem occured while system-setup.\n"
              << "Message:" << e.what() << "\n";
    return 1;
} catch (...) {
    std::cerr << console::err() << "Severe problem occured while system-setup.\n";
    return 1;
}
