// RAII wrapper of the C ABI (include/depthconv_b200.h) for the C++ host layer.
// One context per (host thread, GPU), created on first use; the installed
// camera is remembered so that repeated conversions with the same intrinsic do
// not rebuild the per-camera tables. There is no CPU fallback: without a B200
// the first conversion throws.
#ifndef DCB200_SENS_LOC_GPU_CONTEXT_H
#define DCB200_SENS_LOC_GPU_CONTEXT_H

#include "../../../../include/depthconv_b200.h"
#include "../camera_models/pinhole.h"

#include <cstdint>
#include <cstdlib>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>

namespace sens_loc::gpu {

class error : public std::runtime_error {
  public:
    error(int code, const std::string& msg)
        : std::runtime_error("depthconv_b200 error " + std::to_string(code) + ": " + msg)
        , code(code) {}
    int code;
};

class context {
  public:
    explicit context(int device) {
        const int rc = dc_create(device, &_c);
        if (rc != DC_OK)
            throw error(rc, dc_last_error(nullptr));
    }
    ~context() { dc_destroy(_c); }
    context(const context&)            = delete;
    context& operator=(const context&) = delete;

    /// GPU used by the library functions: $DC_DEVICE, default 0.
    static int default_device() {
        const char* e = std::getenv("DC_DEVICE");
        return e != nullptr ? std::atoi(e) : 0;
    }
    /// The calling thread's context on `device`.
    static context& mine(int device = default_device()) {
        thread_local std::map<int, std::unique_ptr<context>> per_device;
        auto& slot = per_device[device];
        if (!slot)
            slot = std::make_unique<context>(device);
        return *slot;
    }

    void use(const camera_models::pinhole<float>& k) {
        const key want{1, k.w(), k.h(), {k.fx(), k.fy(), k.cx(), k.cy()}};
        if (!(want == _cam)) {
            check(dc_set_pinhole(_c, k.w(), k.h(), k.fx(), k.fy(), k.cx(), k.cy()));
            _cam = want;
        }
    }
    void use(const camera_models::equirectangular<float>& k) {
        const key want{2, k.w(), k.h(), {k.theta_min(), k.theta_max(), 0., 0.}};
        if (!(want == _cam)) {
            check(dc_set_equirectangular(_c, k.w(), k.h(), k.theta_min(), k.theta_max()));
            _cam = want;
        }
    }
    // Real = double cameras: kept in double on the library side (dc_set_*_f64)
    void use(const camera_models::pinhole<double>& k) {
        const key want{3, k.w(), k.h(), {k.fx(), k.fy(), k.cx(), k.cy()}};
        if (!(want == _cam)) {
            check(dc_set_pinhole_f64(_c, k.w(), k.h(), k.fx(), k.fy(), k.cx(), k.cy()));
            _cam = want;
        }
    }
    void use(const camera_models::equirectangular<double>& k) {
        const key want{4, k.w(), k.h(), {k.theta_min(), k.theta_max(), 0., 0.}};
        if (!(want == _cam)) {
            check(dc_set_equirectangular_f64(_c, k.w(), k.h(), k.theta_min(), k.theta_max()));
            _cam = want;
        }
    }
    void use_frame_size(int w, int h) {
        const key want{0, w, h, {0., 0., 0., 0.}};
        if (!(want == _cam)) {
            check(dc_set_frame_size(_c, w, h));
            _cam = want;
        }
    }

    void convert_host(const dc_job& job) { check(dc_convert_host(_c, &job)); }
    /// Status instead of an exception (the frame pipeline reports per index).
    int try_convert_host(const dc_job& job) noexcept { return dc_convert_host(_c, &job); }
    /// The general quantiser (dc_quantise_host_ex): `pixel` a dc_pixel, absent
    /// clamps (curvature only) are the image's own minimum / maximum.
    void quantise(int quant, int pixel, const float* feature, const float* mask, std::size_t n,
                  void* out, const double* lo, const double* hi) {
        check(dc_quantise_host_ex(_c, quant, pixel, DC_OUT_F32, feature, mask, n, out, lo, hi));
    }
    void quantise(int quant, int pixel, const double* feature, const double* mask, std::size_t n,
                  void* out, const double* lo, const double* hi) {
        check(dc_quantise_host_ex(_c, quant, pixel, DC_OUT_F64, feature, mask, n, out, lo, hi));
    }
    [[nodiscard]] const char* last_error() const noexcept { return dc_last_error(_c); }
    [[nodiscard]] dc_ctx*     raw() noexcept { return _c; }

  private:
    struct key {
        int    model = -1, w = 0, h = 0;
        double p[4] = {0, 0, 0, 0};
        bool  operator==(const key& o) const {
            return model == o.model && w == o.w && h == o.h && p[0] == o.p[0] && p[1] == o.p[1] &&
                   p[2] == o.p[2] && p[3] == o.p[3];
        }
    };
    void check(int rc) {
        if (rc != DC_OK)
            throw error(rc, dc_last_error(_c));
    }
    dc_ctx* _c = nullptr;
    key     _cam;
};

}  // namespace sens_loc::gpu
#endif
