// Real = double instantiation of the path: the
// precision the reference's unit tests and golden images were made in
// (src/unittest/conversion/intrinsic.h:8-26). Internal interface between
// dc_api.cu (context, staging, C ABI) and dc_f64.cu (tables + kernels).
#ifndef DC_F64_H
#define DC_F64_H

#include <cuda_runtime.h>

#include <cstddef>
#include <cstdint>
#include <string>

namespace dc64 {

struct camera {
    int    model = 0;  // 0 none, 1 pinhole, 2 equirectangular
    int    w = 0, h = 0;
    double p[4] = {0, 0, 0, 0};  // fx fy cx cy | theta_min theta_max
};

struct state;  // host tables (double, host libm) + their device copies

state* create();
void   destroy(state* s);
// Installs a camera; tables are rebuilt lazily by the next convert().
void set_camera(state* s, const camera& c);

// in_kind: 0 u16 depth, 1 u16 range, 3 f64 range (dc_in_kind). All pointers are
// device pointers. feature: dc_feature without DC_SCALE. out[k] as in dc_job.
// Returns cudaSuccess or the failing call's error (message in `err`).
cudaError_t convert(state* s, int feature, int in_kind, const void* d_in, int n_frames,
                    double* const d_out[4], cudaStream_t stream, std::string& err,
                    std::uint64_t& launches);

}  // namespace dc64
#endif
