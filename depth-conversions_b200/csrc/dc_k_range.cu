// Kernel instantiations of the two streaming jobs (dc_launch.h): range, scale.
#include "dc_launch.cuh"

namespace dc {
cudaError_t launch_range(const kparams& p, int in_kind, int out_kind, int n, cudaStream_t s, launch_info& li) {
    return ln::launch_feature<FEAT_RANGE>(p, in_kind, out_kind, n, s, li);
}
cudaError_t launch_scale(const kparams& p, int n, cudaStream_t s, launch_info& li) {
    li.fast = false;
    return ln::launch_px<FEAT_SCALE, IN_U16_RANGE, OUT_U16>(p, n, s, li);
}
}  // namespace dc
