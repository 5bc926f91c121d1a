// Per-camera constant tables (host side). Everything that depends on the
// intrinsic only -- unit rays, the orthographic->range factor, the ray angles
// phi() and their cosines -- is computed here once per dc_set_* call, with the
// host's libm, in exactly the operation order of the reference, and uploaded to
// the device. The per-frame kernels never call acosf/cosf on a ray angle.
#ifndef DC_TABLES_H
#define DC_TABLES_H

#include <cstdint>
#include <vector>

namespace dc {

enum class camera_model { none, pinhole, equirectangular };

struct camera {
    camera_model model = camera_model::none;
    int          w = 0, h = 0;
    // pinhole: fx fy cx cy. equirectangular: p[0] = theta_min, p[1] = theta_max.
    float p[4] = {0, 0, 0, 0};
};

// A per-pixel float table that collapses to one scalar when every entry of its
// valid region holds the same bit pattern (K4 in SURVEY.md: at >= 1080p every
// adjacent-pixel ray angle hits the clamp).
struct table {
    std::vector<float> v;        // w*h entries (empty until built)
    bool               uniform = false;
    float              value   = 0.f;  // the scalar when uniform
};

struct tables {
    camera cam;
    // conversion/util.h:44-50: sqrtf(float(double(x*x + y*y) + 1.)); all ones
    // for the equirectangular model (util.h:57-67, identity).
    table range_factor;
    // camera_models/pinhole.h:151-171 (or equirectangular.h:155-181): unit ray.
    std::vector<float> ray_x, ray_y, ray_z;
    // cos(phi(central, prior)) per bearing direction, index = direction
    // (0 horizontal (u-1,v), 1 vertical (u,v-1), 2 diagonal (u-1,v-1),
    //  3 antidiagonal (u-1,v+1)); 0 where the prior pixel is outside the image.
    table cos_prior[4];
    // phi over the 2-pixel spans curvature uses: d_phi (u-1,v)-(u+1,v),
    // d_theta (u,v-1)-(u,v+1), d_phi_theta (u-1,v-1)-(u+1,v+1); 1 at the border.
    table d_phi, d_theta, d_phi_theta;
    // True while every table built so far stays inside the value ranges for
    // which the kernels' guard-free exact division / square root are proven
    // (dc_fastmath.cuh): range factor in [1, 64], cos_prior in [0.5, 1 - 8e-7],
    // non-zero ray components >= 2^-40 in magnitude, curvature angles in
    // [1e-3, 3.2]. Otherwise the kernels with the IEEE intrinsics are used.
    bool fast_ok = true;
};

enum table_need : unsigned {
    NEED_RANGE_FACTOR = 1u,
    NEED_RAYS         = 2u,
    NEED_COS_PRIOR    = 4u,
    NEED_CURV_ANGLES  = 8u,
};

// Builds whatever of `need` is not built yet. Threaded over rows.
void build_tables(tables& t, unsigned need);

}  // namespace dc

#endif
