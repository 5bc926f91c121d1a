// Host-side per-camera precompute. Compile WITHOUT floating-point contraction
// (-ffp-contract=off, no -mfma, no -ffast-math): every operation below must
// round exactly as the reference's default -msse4.2 build does.
#include "dc_tables.h"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <thread>

namespace dc {
namespace {

constexpr float kPi    = float(3.141592653589793238462);  // math/constants.h:8-10
constexpr float kDelta = 10.0f * 1.1920928955078125e-07f;  // utility.h:39

template <typename F>
void parallel_rows(int h, F&& f) {
    unsigned nt = std::thread::hardware_concurrency();
    if (nt == 0)
        nt = 1;
    nt = std::min<unsigned>(nt, 32u);
    nt = std::min<unsigned>(nt, unsigned(std::max(1, h / 16)));
    if (nt <= 1) {
        f(0, h);
        return;
    }
    std::vector<std::thread> pool;
    const int                chunk = (h + int(nt) - 1) / int(nt);
    for (unsigned i = 0; i < nt; ++i) {
        const int a = int(i) * chunk, b = std::min(h, a + chunk);
        if (a < b)
            pool.emplace_back([&f, a, b] { f(a, b); });
    }
    for (auto& th : pool)
        th.join();
}

inline std::uint32_t bits(float x) {
    std::uint32_t u;
    std::memcpy(&u, &x, 4);
    return u;
}

// acosf / cosf(acosf) of a clamped ray dot product, memoised on the bit
// pattern: neighbouring rays give dot = 1 - n*2^-24 for a handful of n.
struct angle_memo {
    static constexpr int N = 1024;
    std::uint32_t key[N];
    float         phi[N];
    float         cphi[N];
    bool          used[N];
    angle_memo() { std::memset(used, 0, sizeof(used)); }
    void get(float clamped_dot, float& out_phi, float& out_cos) {
        const std::uint32_t b = bits(clamped_dot);
        const int           i = int((b ^ (b >> 10)) & (N - 1));
        if (!used[i] || key[i] != b) {
            used[i] = true;
            key[i]  = b;
            phi[i]  = std::acos(clamped_dot);  // utility.h:47  (acosf)
            cphi[i] = std::cos(phi[i]);        // depth_to_bearing.h:208 (cosf)
        }
        out_phi = phi[i];
        out_cos = cphi[i];
    }
};

struct ray_source {
    const float *x, *y, *z;
    int          w;
    // Eigen 3.3.7 fixed-size-3 dot: a0*b0 + (a1*b1 + a2*b2); then the clamp of
    // camera_models/utility.h:41-42.
    float clamped_dot(int u1, int v1, int u2, int v2) const {
        const std::size_t i = std::size_t(v1) * w + u1, j = std::size_t(v2) * w + u2;
        const float       p0 = x[i] * x[j];
        const float       p1 = y[i] * y[j];
        const float       p2 = z[i] * z[j];
        const float       t  = p1 + p2;
        const float       d  = p0 + t;
        return std::clamp(d, -1.0f + kDelta, 1.0f - kDelta);
    }
};

void build_rays(tables& t) {
    const camera&     c = t.cam;
    const std::size_t n = std::size_t(c.w) * c.h;
    t.ray_x.resize(n);
    t.ray_y.resize(n);
    t.ray_z.resize(n);
    if (c.model == camera_model::pinhole) {
        const float fx = c.p[0], fy = c.p[1], cx = c.p[2], cy = c.p[3];
        parallel_rows(c.h, [&](int v0, int v1) {
            for (int v = v0; v < v1; ++v)
                for (int u = 0; u < c.w; ++u) {
                    // pinhole.h:139 transform_to_image (float)
                    const float xu = float(u) - cx;
                    const float yv = float(v) - cy;
                    const float xf = xu / fx;
                    const float yf = yv / fy;
                    // pinhole.h:155-165 image_to_sphere (double, narrowed)
                    const double x = xf, y = yf;
                    const double xx     = x * x;
                    const double yy     = y * y;
                    const double cse0   = 1. + xx;
                    const double cse    = cse0 + yy;
                    const double factor = 1. / std::sqrt(cse);
                    const std::size_t i = std::size_t(v) * c.w + u;
                    t.ray_x[i]          = float(factor * x);
                    t.ray_y[i]          = float(factor * y);
                    t.ray_z[i]          = float(factor * 1.);
                }
        });
    } else {
        // equirectangular.h:16-21, 50-66: d_phi = 2*pi/w, d_theta = (max-min)/h
        const float two_pi  = 2.0f * kPi;
        const float d_phi   = two_pi / float(c.w);
        const float span    = c.p[1] - c.p[0];
        const float d_theta = span / float(c.h);
        const float tmin    = c.p[0];
        std::vector<float> sp(c.w), cp(c.w);
        for (int u = 0; u < c.w; ++u) {
            // equirectangular.h:165: phi = u * d_phi - pi
            const float m  = float(u) * d_phi;
            const float ph = m - kPi;
            sp[u]          = std::sin(ph);
            cp[u]          = std::cos(ph);
        }
        parallel_rows(c.h, [&](int v0, int v1) {
            for (int v = v0; v < v1; ++v) {
                // equirectangular.h:166: theta = theta_min + v * d_theta
                const float m  = float(v) * d_theta;
                const float th = tmin + m;
                const float st = std::sin(th), ct = std::cos(th);
                for (int u = 0; u < c.w; ++u) {
                    const std::size_t i = std::size_t(v) * c.w + u;
                    t.ray_x[i]          = st * cp[u];
                    t.ray_y[i]          = st * sp[u];
                    t.ray_z[i]          = ct;
                }
            }
        });
    }
}

// range proof for the guard-free arithmetic: all valid entries in [lo, hi]
bool all_within(const std::vector<float>& v, int w, int u0, int u1, int v0, int v1, float lo,
                float hi) {
    for (int y = v0; y < v1; ++y)
        for (int x = u0; x < u1; ++x) {
            const float f = v[std::size_t(y) * w + x];
            if (!(f >= lo && f <= hi))
                return false;
        }
    return true;
}

bool rays_in_fast_range(const tables& t) {
    const float tiny = 9.094947017729282e-13f;  // 2^-40
    for (const std::vector<float>* r : {&t.ray_x, &t.ray_y, &t.ray_z})
        for (float f : *r) {
            const float a = std::fabs(f);
            if (!(a == 0.0f || (a >= tiny && a <= 1.0f)))
                return false;
        }
    return true;
}

void finish_uniform(table& tb, int w, int h, int u0, int u1, int v0, int v1) {
    // uniform iff every entry of the valid region [u0,u1) x [v0,v1) is equal
    tb.uniform = false;
    if (u0 >= u1 || v0 >= v1)
        return;
    const std::uint32_t first = bits(tb.v[std::size_t(v0) * w + u0]);
    for (int v = v0; v < v1; ++v)
        for (int u = u0; u < u1; ++u)
            if (bits(tb.v[std::size_t(v) * w + u]) != first)
                return;
    (void)h;
    tb.uniform = true;
    std::memcpy(&tb.value, &first, 4);
}

void build_range_factor(tables& t) {
    const camera&     c = t.cam;
    const std::size_t n = std::size_t(c.w) * c.h;
    t.range_factor.v.assign(n, 1.0f);
    if (c.model == camera_model::pinhole) {
        const float fx = c.p[0], fy = c.p[1], cx = c.p[2], cy = c.p[3];
        parallel_rows(c.h, [&](int v0, int v1) {
            for (int v = v0; v < v1; ++v)
                for (int u = 0; u < c.w; ++u) {
                    const float xu = float(u) - cx;
                    const float yv = float(v) - cy;
                    const float x  = xu / fx;
                    const float y  = yv / fy;
                    const float xx = x * x;
                    const float yy = y * y;
                    const float s  = xx + yy;
                    // conversion/util.h:48: float sum + double literal 1.,
                    // narrowed on assignment to Real
                    const float z_helper = float(double(s) + 1.);
                    t.range_factor.v[std::size_t(v) * c.w + u] = std::sqrt(z_helper);
                }
        });
    }
    finish_uniform(t.range_factor, c.w, c.h, 0, c.w, 0, c.h);
    t.fast_ok = t.fast_ok && all_within(t.range_factor.v, c.w, 0, c.w, 0, c.h, 1.0f, 64.0f);
}

void build_cos_prior(tables& t) {
    const camera& c = t.cam;
    const int     w = c.w, h = c.h;
    // conversion/depth_to_bearing.h:96-163
    static const int du[4] = {-1, 0, -1, -1}, dv[4] = {0, -1, -1, 1};
    static const int xs[4] = {1, 0, 1, 1}, ys[4] = {0, 1, 1, 0}, ye[4] = {0, 0, 0, 1};
    const ray_source rs{t.ray_x.data(), t.ray_y.data(), t.ray_z.data(), w};
    for (int d = 0; d < 4; ++d) {
        table& tb = t.cos_prior[d];
        tb.v.assign(std::size_t(w) * h, 0.0f);
        parallel_rows(h, [&](int v0, int v1) {
            angle_memo memo;
            for (int v = std::max(v0, ys[d]); v < std::min(v1, h - ye[d]); ++v)
                for (int u = xs[d]; u < w; ++u) {
                    float ph, cph;
                    memo.get(rs.clamped_dot(u, v, u + du[d], v + dv[d]), ph, cph);
                    tb.v[std::size_t(v) * w + u] = cph;
                }
        });
        finish_uniform(tb, w, h, xs[d], w, ys[d], h - ye[d]);
        t.fast_ok = t.fast_ok &&
                    all_within(tb.v, w, xs[d], w, ys[d], h - ye[d], 0.5f, 1.0f - 8e-7f);
    }
}

void build_curv_angles(tables& t) {
    const camera&    c = t.cam;
    const int        w = c.w, h = c.h;
    const ray_source rs{t.ray_x.data(), t.ray_y.data(), t.ray_z.data(), w};
    t.d_phi.v.assign(std::size_t(w) * h, 1.0f);
    t.d_theta.v.assign(std::size_t(w) * h, 1.0f);
    t.d_phi_theta.v.assign(std::size_t(w) * h, 1.0f);
    parallel_rows(h, [&](int v0, int v1) {
        angle_memo memo;
        for (int v = std::max(v0, 1); v < std::min(v1, h - 1); ++v)
            for (int u = 1; u < w - 1; ++u) {
                // conversion/depth_to_curvature.h:126-128
                float             ph, cph;
                const std::size_t i = std::size_t(v) * w + u;
                memo.get(rs.clamped_dot(u - 1, v, u + 1, v), ph, cph);
                t.d_phi.v[i] = ph;
                memo.get(rs.clamped_dot(u, v - 1, u, v + 1), ph, cph);
                t.d_theta.v[i] = ph;
                memo.get(rs.clamped_dot(u - 1, v - 1, u + 1, v + 1), ph, cph);
                t.d_phi_theta.v[i] = ph;
            }
    });
    finish_uniform(t.d_phi, w, h, 1, w - 1, 1, h - 1);
    finish_uniform(t.d_theta, w, h, 1, w - 1, 1, h - 1);
    finish_uniform(t.d_phi_theta, w, h, 1, w - 1, 1, h - 1);
    for (const table* tb : {&t.d_phi, &t.d_theta, &t.d_phi_theta})
        t.fast_ok = t.fast_ok && all_within(tb->v, w, 1, w - 1, 1, h - 1, 1e-3f, 3.2f);
}

}  // namespace

void build_tables(tables& t, unsigned need) {
    if (t.cam.model == camera_model::none)
        return;
    const bool want_rays = (need & (NEED_RAYS | NEED_COS_PRIOR | NEED_CURV_ANGLES)) != 0;
    if (want_rays && t.ray_x.empty()) {
        build_rays(t);
        t.fast_ok = t.fast_ok && rays_in_fast_range(t);
    }
    if ((need & NEED_RANGE_FACTOR) && t.range_factor.v.empty())
        build_range_factor(t);
    if ((need & NEED_COS_PRIOR) && t.cos_prior[0].v.empty())
        build_cos_prior(t);
    if ((need & NEED_CURV_ANGLES) && t.d_phi.v.empty())
        build_curv_angles(t);
}

}  // namespace dc
