// Internal interface between dc_api.cu (context, staging, C ABI) and the kernel
// translation units: one dc_k_<feature>.cu per feature, so that the ~170 kernel
// instantiations compile in parallel. Each function picks the kernel family for a
// job and enqueues it on `s`.
#ifndef DC_LAUNCH_H
#define DC_LAUNCH_H

#include "dc_kernels.cuh"

#include <cuda_runtime.h>

namespace dc {

struct launch_info {
    // what the caller established about the job
    bool vec4;  // w % 4 == 0, buffers aligned for 8-byte (u16) / 16-byte (f32) access
    bool even;  // w % 2 == 0, buffers aligned for 4-byte (u16) / 8-byte (f32) access
    bool fast;  // per-camera range proof (dc_tables.cpp fast_ok)
    // what the launch reports back
    const char* kernel = "";  // kernel family (dc_last_kernel)
    int         launches = 0;
};

// in_kind / out_kind: the dc:: IN_* / OUT_* enums of dc_kernels.cuh. For
// IN_F32_RANGE on a guard-free kernel p.guard / p.fallbacks must be set: the
// launch clears *p.guard first and enqueues the IEEE fallback kernel behind it.
cudaError_t launch_range(const kparams& p, int in_kind, int out_kind, int n, cudaStream_t s, launch_info& li);
cudaError_t launch_scale(const kparams& p, int n, cudaStream_t s, launch_info& li);
cudaError_t launch_bearing(const kparams& p, int in_kind, int out_kind, int n, cudaStream_t s, launch_info& li);
cudaError_t launch_flexion(const kparams& p, int in_kind, int out_kind, int n, cudaStream_t s, launch_info& li);
cudaError_t launch_curvature(const kparams& p, int in_kind, int out_kind, int n, cudaStream_t s, launch_info& li);
cudaError_t launch_max_curve(const kparams& p, int in_kind, int out_kind, int n, cudaStream_t s, launch_info& li);

}  // namespace dc
#endif
