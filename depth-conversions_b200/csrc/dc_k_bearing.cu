// Kernel instantiations of one feature (dc_launch.h): bearing.
#include "dc_launch.cuh"

namespace dc {
cudaError_t launch_bearing(const kparams& p, int in_kind, int out_kind, int n, cudaStream_t s, launch_info& li) {
    return ln::launch_feature<FEAT_BEARING>(p, in_kind, out_kind, n, s, li);
}
}  // namespace dc
