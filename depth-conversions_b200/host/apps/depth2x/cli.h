// Command line of depth2x, re-implemented without CLI11: the flags, sub-commands,
// defaults and the "flags before or after the sub-command" behaviour of
// src/apps/depth2x/main.cpp:86-307 (app.fallthrough(), require_subcommand(1)).
#ifndef DCB200_DEPTH2X_CLI_H
#define DCB200_DEPTH2X_CLI_H

#include <string>
#include <vector>

namespace sens_loc::apps {

/// apps/util/batch_converter.h:52-63
struct file_patterns {
    std::string input;
    std::string output;
    std::string horizontal, vertical, diagonal, antidiagonal;
};

struct options {
    std::string   subcommand;  // bearing range mean-curvature gauss-curvature max-curve flexion scale
    std::string   calibration_file;
    std::string   camera_model = "pinhole";
    std::string   input_type   = "pinhole-depth";
    file_patterns files;
    int           start_idx = 0, end_idx = 0;
    double        upper_bound = 20., lower_bound = -20.;
    double        scale_factor = 1., scale_delta = 0.;
    // additive flags of this build (defaults keep the reference's behaviour)
    int devices    = 1;   // --devices N: shard the index range over N GPUs
    int batch_size = 0;   // --batch-size N: frames per GPU batch (0 = automatic)
    int threads    = 0;   // --threads N: PNG codec threads (0 = all host cores)
    int median     = 0;   // --median-blur K: depth_filter's median (K = 3, 5) fused in front
};

enum class parse_result { run, exit_ok, exit_error };

/// Parses argv. On exit_ok (--help, --version) / exit_error the message has been
/// printed and `exit_code` is what main() returns.
parse_result parse_command_line(int argc, char** argv, options& opt, int& exit_code);

/// fmt::format(pattern, idx) for the replacement fields depth2x patterns use:
/// "{}", "{:d}", "{:4d}", "{:04d}", "{0}", "{0:03d}", "{{" and "}}".
/// Returns false for a pattern fmt would reject.
bool format_index(const std::string& pattern, int idx, std::string& out);

}  // namespace sens_loc::apps
#endif
