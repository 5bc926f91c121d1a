// The frame pipeline of depth2x: what src/apps/util/batch_converter.cpp:12-40 +
// src/apps/util/parallel_processing.h:23-91 do with one taskflow task per file
// (imread -> preprocess_depth -> process_file -> imwrite), re-hosted for a GPU:
//
//   reader   PNG files -> pinned batch buffer   (thread pool: inflate + unfilter)
//   gpu      ONE dc_convert_host per batch      (H2D, fused kernel, D2H, chunked
//                                                and double-buffered inside the ABI)
//   writer   pinned outputs -> PNG files        (thread pool: filter + deflate)
//
// Three batch buffers rotate between the stages, so decoding batch k+1, the GPU
// on batch k and encoding batch k-1 overlap. Per-index failure semantics are the
// reference's: a missing file, a file that is not 16 bit greyscale, or a size
// that does not match the intrinsic fails THAT index only.
#ifndef DCB200_DEPTH2X_PIPELINE_H
#define DCB200_DEPTH2X_PIPELINE_H

#include "cli.h"

#include <string>
#include <vector>

namespace sens_loc::apps {

struct camera_spec {
    int   model = 0;  // 0 none (scale), 1 pinhole, 2 equirectangular
    int   w = 0, h = 0;
    float p[4] = {0, 0, 0, 0};  // fx fy cx cy | theta_min theta_max
};

struct convert_spec {
    int         feature  = 0;  // dc_feature
    int         in_kind  = 0;  // dc_in_kind (u16 depth or u16 range)
    std::string input;         // input pattern
    std::string output[4];     // output pattern per dc_job out[] slot ("" = not wanted)
    float       clamp_lo = -20.f, clamp_hi = 20.f;
    double      scale = 1., offset = 0.;
    int         median = 0;  // 0, 3, 5: dc_job::median_ksize
    camera_spec camera;
};

struct run_stats {
    int    processed = 0, failed = 0;
    double seconds   = 0.;
};

/// Converts indices [start, end] (inclusive, swapped if reversed) on `devices`
/// GPUs, frames sharded by contiguous index blocks. Prints the per-index error
/// lines; returns the totals.
run_stats run_batch(const convert_spec& spec, int start, int end, int devices, int batch_size,
                    int threads);

/// DEPTH2X_PROFILE=1 in the environment: stage timestamps on stderr.
void profile_mark(const char* what, int value = -1);

}  // namespace sens_loc::apps
#endif
