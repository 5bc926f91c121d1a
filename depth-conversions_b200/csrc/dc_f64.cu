// Real = double path (dc_f64.h). Same structure as the float path, without its
// tuning: per-camera tables are built on the host in double with the host's
// libm (bit identical to the reference running on the same machine), one
// thread converts one pixel of one frame, every operation in the reference's
// order (no FMA contraction: the library is compiled with -fmad=false). What is
// left to device libm is the final acos of math/triangles.h:58 and the pow of
// math/curvature.h:43 (<= 2 ulp against glibc).
//
// Eigen 3.3.7 reduces a fixed-size double 3-vector as (a0 + a1) + a2 (Packet2d
// first, Eigen/src/Core/Redux.h linear vectorised traversal), NOT as
// a0 + (a1 + a2) like the float vector: checked against the compiled reference.
#include "dc_f64.h"

#include <algorithm>
#include <cmath>
#include <functional>
#include <limits>
#include <thread>
#include <vector>

namespace dc64 {
namespace {

constexpr double kPi    = 3.141592653589793238462;  // math/constants.h:8-10
constexpr double kDelta = 10. * std::numeric_limits<double>::epsilon();

void parallel_rows(int h, const std::function<void(int, int)>& f) {
    unsigned nt = std::max(1u, std::thread::hardware_concurrency());
    nt          = std::min<unsigned>(nt, unsigned(std::max(1, h / 8)));
    const int                chunk = (h + int(nt) - 1) / int(nt);
    std::vector<std::thread> pool;
    for (unsigned i = 0; i < nt; ++i) {
        const int a = int(i) * chunk, b = std::min(h, a + chunk);
        if (a < b)
            pool.emplace_back([&f, a, b] { f(a, b); });
    }
    for (auto& t : pool)
        t.join();
}

struct host_tables {
    std::vector<double> factor, rx, ry, rz, cosp[4], ang[3];
};

// camera_models/pinhole.h:120-171, equirectangular.h:155-181
void build_rays(const camera& c, host_tables& t) {
    const std::size_t n = std::size_t(c.w) * c.h;
    t.rx.resize(n), t.ry.resize(n), t.rz.resize(n), t.factor.resize(n);
    if (c.model == 1) {
        const double fx = c.p[0], fy = c.p[1], cx = c.p[2], cy = c.p[3];
        parallel_rows(c.h, [&](int v0, int v1) {
            for (int v = v0; v < v1; ++v)
                for (int u = 0; u < c.w; ++u) {
                    const double x  = (double(u) - cx) / fx;
                    const double y  = (double(v) - cy) / fy;
                    const double xx = x * x, yy = y * y;
                    const double cse    = (1. + xx) + yy;
                    const double factor = 1. / std::sqrt(cse);
                    const std::size_t i = std::size_t(v) * c.w + u;
                    t.rx[i] = factor * x, t.ry[i] = factor * y, t.rz[i] = factor * 1.;
                    // conversion/util.h:44-50: sqrt(x*x + y*y + 1.)
                    t.factor[i] = std::sqrt((xx + yy) + 1.);
                }
        });
    } else {
        const double d_phi   = 2. * kPi / double(c.w);
        const double d_theta = (c.p[1] - c.p[0]) / double(c.h);
        parallel_rows(c.h, [&](int v0, int v1) {
            for (int v = v0; v < v1; ++v) {
                const double th = c.p[0] + double(v) * d_theta;
                const double st = std::sin(th), ct = std::cos(th);
                for (int u = 0; u < c.w; ++u) {
                    const double      ph = double(u) * d_phi - kPi;
                    const std::size_t i  = std::size_t(v) * c.w + u;
                    t.rx[i] = st * std::cos(ph), t.ry[i] = st * std::sin(ph), t.rz[i] = ct;
                    t.factor[i] = 1.;  // util.h:57-67: identity
                }
            }
        });
    }
}

// camera_models/utility.h:15-49
double phi(const camera& c, const host_tables& t, int u1, int v1, int u2, int v2) {
    const std::size_t i = std::size_t(v1) * c.w + u1, j = std::size_t(v2) * c.w + u2;
    const double      d = (t.rx[i] * t.rx[j] + t.ry[i] * t.ry[j]) + t.rz[i] * t.rz[j];
    return std::acos(std::clamp(d, -1. + kDelta, 1. - kDelta));
}

void build_angles(const camera& c, host_tables& t) {
    const std::size_t n     = std::size_t(c.w) * c.h;
    static const int  du[4] = {-1, 0, -1, -1}, dv[4] = {0, -1, -1, 1};
    for (auto& v : t.cosp)
        v.assign(n, 0.);
    for (auto& v : t.ang)
        v.assign(n, 1.);
    parallel_rows(c.h, [&](int v0, int v1) {
        for (int v = v0; v < v1; ++v)
            for (int u = 0; u < c.w; ++u) {
                const std::size_t i = std::size_t(v) * c.w + u;
                for (int d = 0; d < 4; ++d) {
                    const int pu = u + du[d], pv = v + dv[d];
                    if (pu >= 0 && pv >= 0 && pv < c.h)
                        t.cosp[d][i] = std::cos(phi(c, t, u, v, pu, pv));
                }
                if (u >= 1 && u < c.w - 1 && v >= 1 && v < c.h - 1) {
                    // depth_to_curvature.h:117-159
                    t.ang[0][i] = phi(c, t, u - 1, v, u + 1, v);
                    t.ang[1][i] = phi(c, t, u, v - 1, u, v + 1);
                    t.ang[2][i] = phi(c, t, u - 1, v - 1, u + 1, v + 1);
                }
            }
    });
}

}  // namespace

struct state {
    camera      cam;
    host_tables host;
    bool        built = false;
    double*     d_tab[11] = {nullptr};  // factor, rx, ry, rz, cosp[4], ang[3]
};

state* create() { return new state(); }
static void drop(state* s) {
    for (auto& p : s->d_tab) {
        if (p != nullptr)
            cudaFree(p);
        p = nullptr;
    }
    s->built = false;
}
void destroy(state* s) {
    if (s == nullptr)
        return;
    drop(s);
    delete s;
}
void set_camera(state* s, const camera& c) {
    drop(s);
    s->host = host_tables{};
    s->cam  = c;
}

static cudaError_t ensure(state* s) {
    if (s->built)
        return cudaSuccess;
    build_rays(s->cam, s->host);
    build_angles(s->cam, s->host);
    const std::vector<double>* src[11] = {&s->host.factor,  &s->host.rx,      &s->host.ry,
                                          &s->host.rz,      &s->host.cosp[0], &s->host.cosp[1],
                                          &s->host.cosp[2], &s->host.cosp[3], &s->host.ang[0],
                                          &s->host.ang[1],  &s->host.ang[2]};
    for (int k = 0; k < 11; ++k) {
        const std::size_t bytes = src[k]->size() * sizeof(double);
        cudaError_t       e     = cudaMalloc(reinterpret_cast<void**>(&s->d_tab[k]), bytes);
        if (e == cudaSuccess)
            e = cudaMemcpy(s->d_tab[k], src[k]->data(), bytes, cudaMemcpyHostToDevice);
        if (e != cudaSuccess)
            return e;
    }
    // pageable sources: the copies may still be in flight when cudaMemcpy returns,
    // and the kernels run on streams that do not wait for the legacy stream
    // (and the host tables are freed below)
    const cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess)
        return e;
    s->host  = host_tables{};  // the device copies are what is used from here on
    s->built = true;
    return cudaSuccess;
}

// ---------------------------------------------------------------------------
// device side
namespace {

struct kparams64 {
    const void* in;
    double*     out[4];
    int         w, h, in_kind, feature;
    const double *factor, *rx, *ry, *rz, *cosp[4], *ang[3];
};

struct v3 {
    double x, y, z;
};
__device__ __forceinline__ v3 sub3(v3 a, v3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
// Eigen 3.3.7, double: (a0 b0 + a1 b1) + a2 b2
__device__ __forceinline__ double dot3(v3 a, v3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
// Eigen/src/Core/Dot.h:124-134
__device__ __forceinline__ v3 normalized3(v3 a) {
    const double z = dot3(a, a);
    if (z > 0.) {
        const double s = sqrt(z);
        return {a.x / s, a.y / s, a.z / s};
    }
    return a;
}
// Eigen/src/Geometry/OrthoMethods.h:34-46
__device__ __forceinline__ v3 cross3(v3 a, v3 b) {
    return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}

// math/triangles.h:27-64
__device__ __forceinline__ double bearing_angle(double b, double c, double cos_alpha) {
    const double delta = 10. * 2.220446049250313e-16;
    const double nom   = b - c * cos_alpha;
    double       den   = sqrt(b * b + c * c - 2. * b * c * cos_alpha);
    den                = den == 0. ? den + delta : den;
    const double ratio = fmin(fmax(nom / den, -1. + delta), 1. - delta);
    return acos(ratio);
}

template <int IN>
__device__ __forceinline__ double range_at(const kparams64& p, const void* frame, int u, int v) {
    const size_t i = size_t(v) * p.w + u;
    if constexpr (IN == 3) {
        return static_cast<const double*>(frame)[i];
    } else {
        const unsigned short d = static_cast<const unsigned short*>(frame)[i];
        if constexpr (IN == 1)
            return double(d);  // math/image.h:117-128 convert<Real>
        // conversion/util.h:34-55
        return d == 0 ? 0. : double(d) * p.factor[i];
    }
}

template <int FEAT, int IN>
__global__ void __launch_bounds__(256) feature64_kernel(const kparams64 p) {
    const int u = blockIdx.x * 32 + threadIdx.x, v = blockIdx.y * 8 + threadIdx.y;
    if (u >= p.w || v >= p.h)
        return;
    const size_t fpx   = size_t(p.w) * p.h;
    const size_t eb    = IN == 3 ? 8 : 2;
    const void*  frame = static_cast<const char*>(p.in) + size_t(blockIdx.z) * fpx * eb;
    const size_t i     = size_t(v) * p.w + u;
    const size_t o     = size_t(blockIdx.z) * fpx + i;
    auto R = [&](int du, int dv) { return range_at<IN>(p, frame, u + du, v + dv); };
    const bool interior = u >= 1 && u < p.w - 1 && v >= 1 && v < p.h - 1;

    if constexpr (FEAT == 0) {  // DC_RANGE
        p.out[0][o] = R(0, 0);
    } else if constexpr (FEAT == 1) {  // DC_BEARING, depth_to_bearing.h:185-248
        const int    du[4] = {-1, 0, -1, -1}, dv[4] = {0, -1, -1, 1};
        const double b     = R(0, 0);
#pragma unroll
        for (int d = 0; d < 4; ++d) {
            if (p.out[d] == nullptr)
                continue;
            double    a  = 0.;
            const int pu = u + du[d], pv = v + dv[d];
            if (pu >= 0 && pv >= 0 && pv < p.h) {
                const double c = R(du[d], dv[d]);
                if (b != 0. && c != 0.)
                    a = bearing_angle(b, c, p.cosp[d][i]);
            }
            p.out[d][o] = a;
        }
    } else if constexpr (FEAT == 2) {  // DC_FLEXION, depth_to_flexion.h:79-167
        double f = 0.;
        if (interior) {
            auto P = [&](int du, int dv) {
                const size_t j = size_t(v + dv) * p.w + (u + du);
                const double d = R(du, dv);
                return v3{d * p.rx[j], d * p.ry[j], d * p.rz[j]};
            };
            const v3 n0 = normalized3(sub3(P(0, 1), P(0, -1)));
            const v3 n1 = normalized3(sub3(P(1, 0), P(-1, 0)));
            const v3 n2 = normalized3(sub3(P(-1, 1), P(1, -1)));
            const v3 n3 = normalized3(sub3(P(1, 1), P(-1, -1)));
            f           = fmin(fmax(fabs(dot3(cross3(n0, n1), cross3(n2, n3))), 0.), 1.);
        }
        p.out[0][o] = f;
    } else if constexpr (FEAT == 3) {  // DC_CURVATURE, depth_to_curvature.h:97-206
        double g = 0., m = 0.;
        if (interior) {
            const double d00 = R(-1, -1), d01 = R(0, -1), d02 = R(1, -1);
            const double d10 = R(-1, 0), d11 = R(0, 0), d12 = R(1, 0);
            const double d20 = R(-1, 1), d21 = R(0, 1), d22 = R(1, 1);
            const bool   any_zero = d00 == 0. || d01 == 0. || d02 == 0. || d10 == 0. || d11 == 0. ||
                                  d12 == 0. || d20 == 0. || d21 == 0. || d22 == 0.;
            if (!any_zero) {
                const double dphi = p.ang[0][i], dth = p.ang[1][i], dpt = p.ang[2][i];
                // math/derivatives.h:17-40, 66-93 (f_uv from the vertical triple: sic)
                const double f_u  = (d12 - d10) / (2. * dphi);
                const double f_v  = (d21 - d01) / (2. * dth);
                const double f_uu = (d12 + d10 - 2. * d11) / (dphi * dphi);
                const double f_vv = (d21 + d01 - 2. * d11) / (dth * dth);
                const double f_uv = (d21 + d01 - 2. * d11) / (dpt * dpt);
                // math/curvature.h:18-44
                g = (f_uu * f_vv - f_uv * f_uv) / (1. + f_u * f_u + f_v * f_v);
                m = ((1. + f_v * f_v) * f_uu - 2. * f_u * f_v * f_uv + (1. + f_u * f_u) * f_vv) /
                    pow(2. * sqrt(1. + f_u * f_u + f_v * f_v), 3.);
            }
        }
        if (p.out[0] != nullptr)
            p.out[0][o] = g;
        if (p.out[1] != nullptr)
            p.out[1][o] = m;
    } else if constexpr (FEAT == 4) {  // DC_MAX_CURVE, depth_to_max_curve.h:55-153
        double a = 0.;
        if (interior) {
            const double d0 = R(0, 0);
            auto form = [&](double dm, double dp, double c1, double c2) {
                if (dm == 0. || d0 == 0. || dp == 0.)
                    return 0.;
                return bearing_angle(d0, dm, c1) + bearing_angle(d0, dp, c2);
            };
            const size_t w  = size_t(p.w);
            const double ah = form(R(-1, 0), R(1, 0), p.cosp[0][i], p.cosp[0][i + 1]);
            const double av = form(R(0, -1), R(0, 1), p.cosp[1][i], p.cosp[1][i + w]);
            const double ad = form(R(-1, -1), R(1, 1), p.cosp[2][i], p.cosp[2][i + w + 1]);
            // :136-139: phi((u+1,v+1),(u,v)) with the depth of (u-1,v+1) (sic)
            const double aa = form(R(-1, 1), R(1, -1), p.cosp[2][i + w + 1], p.cosp[3][i - w + 1]);
            a               = fmax(ah, fmax(av, fmax(ad, aa)));
        }
        p.out[0][o] = a;
        if (p.out[1] != nullptr)
            p.out[1][o] = R(0, 0);
    }
}

template <int FEAT>
cudaError_t launch(const kparams64& p, int n_frames, cudaStream_t s) {
    const dim3 block(32, 8, 1);
    for (int f0 = 0; f0 < n_frames; f0 += 65535) {
        const int    nf  = std::min(65535, n_frames - f0);
        kparams64    q   = p;
        const size_t fpx = size_t(p.w) * p.h;
        q.in = static_cast<const char*>(p.in) + size_t(f0) * fpx * (p.in_kind == 3 ? 8 : 2);
        for (int k = 0; k < 4; ++k)
            q.out[k] = p.out[k] != nullptr ? p.out[k] + size_t(f0) * fpx : nullptr;
        const dim3 grid((p.w + 31) / 32, (p.h + 7) / 8, nf);
        switch (p.in_kind) {
        case 0: feature64_kernel<FEAT, 0><<<grid, block, 0, s>>>(q); break;
        case 1: feature64_kernel<FEAT, 1><<<grid, block, 0, s>>>(q); break;
        default: feature64_kernel<FEAT, 3><<<grid, block, 0, s>>>(q); break;
        }
        const cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess)
            return e;
    }
    return cudaSuccess;
}

}  // namespace

cudaError_t convert(state* s, int feature, int in_kind, const void* d_in, int n_frames,
                    double* const d_out[4], cudaStream_t stream, std::string& err,
                    std::uint64_t& launches) {
    if (s->cam.model == 0) {
        err = "no camera model installed";
        return cudaErrorInvalidValue;
    }
    cudaError_t e = ensure(s);
    if (e != cudaSuccess) {
        err = "uploading the double tables failed";
        return e;
    }
    kparams64 p{};
    p.in = d_in;
    for (int k = 0; k < 4; ++k)
        p.out[k] = d_out[k];
    p.w = s->cam.w, p.h = s->cam.h, p.in_kind = in_kind, p.feature = feature;
    p.factor = s->d_tab[0], p.rx = s->d_tab[1], p.ry = s->d_tab[2], p.rz = s->d_tab[3];
    for (int d = 0; d < 4; ++d)
        p.cosp[d] = s->d_tab[4 + d];
    for (int d = 0; d < 3; ++d)
        p.ang[d] = s->d_tab[8 + d];
    switch (feature) {
    case 0: e = launch<0>(p, n_frames, stream); break;
    case 1: e = launch<1>(p, n_frames, stream); break;
    case 2: e = launch<2>(p, n_frames, stream); break;
    case 3: e = launch<3>(p, n_frames, stream); break;
    case 4: e = launch<4>(p, n_frames, stream); break;
    default: err = "feature has no double instantiation"; return cudaErrorInvalidValue;
    }
    if (e != cudaSuccess)
        err = "launching the double kernel failed";
    else
        launches += std::uint64_t((n_frames + 65534) / 65535);
    return e;
}

}  // namespace dc64
