#include "cli.h"

#include <cctype>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <map>
#include <sys/stat.h>

namespace sens_loc::apps {
namespace {

struct opt_def {
    const char* shortn;  // "-c" or ""
    const char* longn;   // "--calibration"
    const char* help;
    enum kind_t { string_v, int_v, real_v } kind;
    void* target;
    bool  required;
    const char* const* choices;  // nullptr-terminated, or nullptr
    bool  existing_file;
    bool  seen = false;
};

const char* const kModels[] = {"pinhole", "equirectangular", nullptr};
const char* const kTypes[]  = {"pinhole-depth", "pinhole-range", nullptr};

struct sub_def {
    const char*          name;
    const char*          help;
    std::vector<opt_def> opts;
};

bool file_exists(const std::string& p) {
    struct stat st{};
    return ::stat(p.c_str(), &st) == 0 && S_ISREG(st.st_mode);
}

void usage(const char* argv0, const std::vector<opt_def>& globals, const std::vector<sub_def>& subs,
           const sub_def* only) {
    std::cout << "Batch-conversion of depth images to various derived image-types.\n"
              << "Usage: " << argv0 << " [OPTIONS] SUBCOMMAND\n\nOptions:\n"
              << "  -h,--help                   Print this help message and exit\n"
              << "  -v,--version                Print version and exit\n";
    auto show = [](const opt_def& o) {
        std::string n = std::string(o.shortn[0] ? std::string(o.shortn) + "," : "") + o.longn;
        std::cout << "  " << n << std::string(n.size() < 28 ? 28 - n.size() : 1, ' ') << o.help
                  << (o.required ? " (REQUIRED)" : "") << "\n";
    };
    for (const auto& o : globals)
        show(o);
    if (only != nullptr) {
        std::cout << "\nSubcommand " << only->name << ": " << only->help << "\n";
        for (const auto& o : only->opts)
            show(o);
    } else {
        std::cout << "\nSubcommands:\n";
        for (const auto& s : subs)
            std::cout << "  " << s.name << std::string(28 - std::strlen(s.name), ' ') << s.help << "\n";
    }
    std::cout << "\nAn example invocation of the tool is:\n\n"
                 "depth2x bearing --calibration intrinsic.txt \\\n"
                 "                --input depth_{:04d}.png \\\n"
                 "                --start 0 \\\n"
                 "                --end 100 \\\n"
                 "                --horizontal horizontal_{:04d}.png\n\n"
                 "This will read 'depth_0000.png ...' and create 'horizontal_0000.png ...'\n"
                 "in the working directory\n";
}

bool assign(opt_def& o, const std::string& v, std::string& err) {
    if (o.choices != nullptr) {
        bool ok = false;
        for (const char* const* c = o.choices; *c != nullptr; ++c)
            ok = ok || v == *c;
        if (!ok) {
            err = std::string(o.longn) + ": " + v + " not in the allowed set";
            return false;
        }
    }
    if (o.existing_file && !file_exists(v)) {
        err = std::string(o.longn) + ": File does not exist: " + v;
        return false;
    }
    char* end = nullptr;
    switch (o.kind) {
    case opt_def::string_v: *static_cast<std::string*>(o.target) = v; break;
    case opt_def::int_v: {
        const long x = std::strtol(v.c_str(), &end, 10);
        if (v.empty() || *end != '\0') {
            err = std::string("Could not convert: ") + o.longn + " = " + v;
            return false;
        }
        *static_cast<int*>(o.target) = int(x);
        break;
    }
    case opt_def::real_v: {
        const double x = std::strtod(v.c_str(), &end);
        if (v.empty() || *end != '\0') {
            err = std::string("Could not convert: ") + o.longn + " = " + v;
            return false;
        }
        *static_cast<double*>(o.target) = x;
        break;
    }
    }
    o.seen = true;
    return true;
}

}  // namespace

parse_result parse_command_line(int argc, char** argv, options& opt, int& exit_code) {
    using K = opt_def;
    std::vector<opt_def> globals = {
        {"-c", "--calibration", "File that contains calibration parameters for the camera", K::string_v, &opt.calibration_file, false, nullptr, true},
        {"-m", "--model", "Camera model {pinhole, equirectangular}; must match the '--calibration' file", K::string_v, &opt.camera_model, false, kModels, false},
        {"-i", "--input", "Input pattern for image, e.g. \"depth-{}.png\"", K::string_v, &opt.files.input, true, nullptr, false},
        {"-t", "--type", "Type of input depth images {pinhole-depth, pinhole-range}", K::string_v, &opt.input_type, false, kTypes, false},
        {"-s", "--start", "Start index of batch, inclusive", K::int_v, &opt.start_idx, true, nullptr, false},
        {"-e", "--end", "End index of batch, inclusive", K::int_v, &opt.end_idx, true, nullptr, false},
        {"", "--devices", "Number of GPUs the index range is sharded over (default 1)", K::int_v, &opt.devices, false, nullptr, false},
        {"", "--batch-size", "Frames per GPU batch (default: automatic)", K::int_v, &opt.batch_size, false, nullptr, false},
        {"", "--threads", "PNG codec threads (default: all host cores)", K::int_v, &opt.threads, false, nullptr, false},
        {"", "--median-blur", "Median pre-filter of the depth image, window 3 or 5 (depth_filter median-blur -d K)", K::int_v, &opt.median, false, nullptr, false},
    };
    auto out_opt = [&](const char* help) {
        return opt_def{"-o", "--output", help, K::string_v, &opt.files.output, true, nullptr, false};
    };
    auto bounds = [&](std::vector<opt_def>& v) {
        v.push_back({"-u", "--upper-bound", "Upper bound that curvature values are clamped to (default 20)", K::real_v, &opt.upper_bound, false, nullptr, false});
        v.push_back({"-l", "--lower-bound", "Lower bound that curvature values are clamped to (default -20)", K::real_v, &opt.lower_bound, false, nullptr, false});
    };
    std::vector<sub_def> subs;
    subs.push_back({"bearing", "Convert depth images into bearing angle images",
                    {{"", "--horizontal", "Write the horizontal bearing angle image to this pattern", K::string_v, &opt.files.horizontal, false, nullptr, false},
                     {"", "--vertical", "Write the vertical bearing angle image to this pattern", K::string_v, &opt.files.vertical, false, nullptr, false},
                     {"", "--diagonal", "Write the diagonal bearing angle image to this pattern", K::string_v, &opt.files.diagonal, false, nullptr, false},
                     {"", "--anti-diagonal", "Write the anti-diagonal bearing angle image to this pattern", K::string_v, &opt.files.antidiagonal, false, nullptr, false}}});
    subs.push_back({"range", "Convert depth images into range images (laser-scan like)", {out_opt("Output pattern for the range images.")}});
    subs.push_back({"mean-curvature", "Convert depth images into mean-curvature images", {out_opt("Output pattern for the mean-curvature images.")}});
    bounds(subs.back().opts);
    subs.push_back({"gauss-curvature", "Convert depth images into gaussian-curvature images", {out_opt("Output pattern for the gaussian-curvature images.")}});
    bounds(subs.back().opts);
    subs.push_back({"max-curve", "Convert depth images into max-curve images", {out_opt("Output pattern for the max-curve images.")}});
    subs.push_back({"flexion", "Convert depth images into flexion images", {out_opt("Output pattern for the flexion images.")}});
    subs.push_back({"scale", "Scale depth images and add an optional offset.",
                    {out_opt("Output pattern for the scaled images."),
                     {"-f", "--factor", "Real number that is multipled to every depth value (default 1)", K::real_v, &opt.scale_factor, false, nullptr, false},
                     {"-d", "--delta", "Real number that is added to every depth value (default 0)", K::real_v, &opt.scale_delta, false, nullptr, false}}});

    sub_def* active = nullptr;
    auto     fail   = [&](const std::string& msg, int code) {
        std::cout << msg << "\nRun with --help for more information.\n";
        exit_code = code;
        return parse_result::exit_error;
    };
    auto find = [&](const std::string& name) -> opt_def* {
        if (active != nullptr)
            for (auto& o : active->opts)
                if (name == o.longn || (o.shortn[0] && name == o.shortn))
                    return &o;
        for (auto& o : globals)
            if (name == o.longn || (o.shortn[0] && name == o.shortn))
                return &o;
        return nullptr;
    };
    bool want_help = false;
    for (int i = 1; i < argc; ++i) {
        std::string a = argv[i];
        if (a == "-h" || a == "--help") {
            want_help = true;
            continue;
        }
        if (a == "-v" || a == "--version") {  // apps/util/version_printer.cpp:9-12
            std::cout << argv[0] << " v1.0.0\n";
            exit_code = 0;
            return parse_result::exit_ok;
        }
        if (a.size() > 1 && a[0] == '-' && !(a.size() > 1 && (std::isdigit((unsigned char)a[1]) || a[1] == '.'))) {
            std::string name = a, value;
            bool        has_value = false;
            const auto  eq        = a.find('=');
            if (a.rfind("--", 0) == 0 && eq != std::string::npos) {
                name = a.substr(0, eq), value = a.substr(eq + 1), has_value = true;
            } else if (a.rfind("--", 0) != 0 && a.size() > 2) {  // -s4
                name = a.substr(0, 2), value = a.substr(2), has_value = true;
            }
            opt_def* o = find(name);
            if (o == nullptr)
                return fail("The following argument was not expected: " + a, 109);
            if (!has_value) {
                if (i + 1 >= argc)
                    return fail(std::string(o->longn) + ": 1 required TEXT missing", 114);
                value = argv[++i];
            }
            std::string err;
            if (!assign(*o, value, err))
                return fail(err, 105);
            continue;
        }
        // positional: the sub-command (exactly one)
        sub_def* s = nullptr;
        for (auto& d : subs)
            if (a == d.name)
                s = &d;
        if (s == nullptr || active != nullptr)
            return fail("The following argument was not expected: " + a, 109);
        active         = s;
        opt.subcommand = a;
    }
    if (want_help) {
        usage(argv[0], globals, subs, active);
        exit_code = 0;
        return parse_result::exit_ok;
    }
    if (active == nullptr)
        return fail("A subcommand is required", 106);
    for (const auto& o : globals)
        if (o.required && !o.seen)
            return fail(std::string(o.longn) + " is required", 106);
    for (const auto& o : active->opts)
        if (o.required && !o.seen)
            return fail(std::string(o.longn) + " is required", 106);
    if (opt.devices < 1)
        return fail("--devices must be at least 1", 105);
    return parse_result::run;
}

bool format_index(const std::string& pattern, int idx, std::string& out) {
    out.clear();
    bool used_auto = false, used_manual = false;
    for (std::size_t i = 0; i < pattern.size(); ++i) {
        const char c = pattern[i];
        if (c == '{') {
            if (i + 1 < pattern.size() && pattern[i + 1] == '{') {
                out += '{';
                ++i;
                continue;
            }
            const std::size_t close = pattern.find('}', i);
            if (close == std::string::npos)
                return false;
            std::string spec = pattern.substr(i + 1, close - i - 1);
            std::string fmt_spec;
            const auto  colon = spec.find(':');
            std::string arg   = colon == std::string::npos ? spec : spec.substr(0, colon);
            if (colon != std::string::npos)
                fmt_spec = spec.substr(colon + 1);
            if (arg.empty())
                used_auto = true;
            else if (arg == "0")
                used_manual = true;
            else
                return false;  // only one argument exists
            if (used_auto && used_manual)
                return false;
            // [[fill]align][sign][0][width][type] restricted to integers
            bool        zero = false, left = false;
            std::size_t k = 0;
            char        fill = ' ';
            if (fmt_spec.size() >= 2 && (fmt_spec[1] == '<' || fmt_spec[1] == '>')) {
                fill = fmt_spec[0], left = fmt_spec[1] == '<', k = 2;
            } else if (!fmt_spec.empty() && (fmt_spec[0] == '<' || fmt_spec[0] == '>')) {
                left = fmt_spec[0] == '<', k = 1;
            }
            if (k < fmt_spec.size() && fmt_spec[k] == '0')
                zero = true, ++k;
            int width = 0;
            while (k < fmt_spec.size() && std::isdigit((unsigned char)fmt_spec[k]))
                width = width * 10 + (fmt_spec[k++] - '0');
            if (k < fmt_spec.size() && fmt_spec[k] == 'd')
                ++k;
            if (k != fmt_spec.size())
                return false;
            std::string num = std::to_string(idx);
            if (int(num.size()) < width) {
                const std::size_t pad = std::size_t(width) - num.size();
                if (zero && !left) {
                    const bool neg = idx < 0;
                    num = (neg ? "-" : "") + std::string(pad, '0') + num.substr(neg ? 1 : 0);
                } else if (left) {
                    num += std::string(pad, fill);
                } else {
                    num = std::string(pad, fill) + num;
                }
            }
            out += num;
            i = close;
        } else if (c == '}') {
            if (i + 1 < pattern.size() && pattern[i + 1] == '}') {
                out += '}';
                ++i;
                continue;
            }
            return false;
        } else {
            out += c;
        }
    }
    return true;
}

}  // namespace sens_loc::apps
