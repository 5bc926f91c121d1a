/* TEST INFRASTRUCTURE ONLY (oracle/). Never linked into, called from or
 * shipped with the product library. Only tests/, __graft_entry__.smoke() and
 * the cpu_baseline / --impl reference legs of bench.py may load this file's
 * shared object, and only as the checker.
 *
 * Plain-C restatement of the depth -> feature-image hot path of
 * JonasToth/depth-conversions for Real = float and the pinhole camera model
 * (the precision and model the depth2x CLI uses, src/apps/depth2x/main.cpp:29-31).
 * Every function cites the reference file:line it follows; paths are relative
 * to /root/reference/src/include/sens_loc/ unless stated otherwise.
 *
 * PARITY PIN: this restatement is asserted bit-identical to the reference's own
 * headers compiled in this container (oracle/_ref/liboracle_ref.so, recipe in
 * oracle/Makefile) by tests/test_oracle_vs_ref.py, and against the committed
 * golden vectors under tests/golden/ (generated from that compiled reference by
 * tests/golden/make_golden.py) by tests/test_oracle_golden.py. The compiled
 * reference itself reproduces the reference's checked-in golden PNGs
 * (src/unittest/conversion/ *.png) to <= 1 LSB on the double path.
 *
 * Build: gcc -O2 -msse4.2 -ffp-contract=off (no FMA contraction; the reference
 * default build is -msse4.2 without -ffast-math, cmake/configuration_functions.cmake:53-63)
 * so every float operation below is rounded individually, in the order written.
 * The third-party arithmetic on the path is Eigen 3.3.7 (vendored in the
 * reference tree under third_party/cpp-taskflow/3rd-party/eigen-3.3.7) for
 * fixed-size 3-vectors: dot/squaredNorm reduce as a0*b0 + (a1*b1 + a2*b2)
 * (Eigen/src/Core/Redux.h:91-105), normalized() divides by sqrt(squaredNorm)
 * iff squaredNorm > 0 (Eigen/src/Core/Dot.h:124-134), cross is the textbook
 * formula (Eigen/src/Geometry/OrthoMethods.h:34-46). OpenCV (pinned 4.3.0, not
 * vendored) contributes only convertTo's saturate_cast: round-half-to-even then
 * clamp.
 */
#include <math.h>
#include <stddef.h>
#include <stdint.h>
#include <string.h>

typedef struct {
    int   w, h;
    float fx, fy, cx, cy;
} dco_pinhole;

/* numeric_limits<float>::epsilon() * 10, camera_models/utility.h:39,
 * math/triangles.h:39 */
#define DCO_DELTA (10.0f * 1.1920928955078125e-07f)
/* math/constants.h:8-10, T(3.141592653589793238462) narrowed to float */
#define DCO_PI_F ((float)3.141592653589793238462)

/* ------------------------------------------------------------------------ */
/* camera_models/pinhole.h:120-140 transform_to_image: (u - cx)/fx, (v - cy)/fy
 * in float. */
static void image_coord(const dco_pinhole* k, int u, int v, float* x, float* y) {
    const float xu = (float)u - k->cx;
    const float yv = (float)v - k->cy;
    *x             = xu / k->fx;
    *y             = yv / k->fy;
}

/* camera_models/pinhole.h:142-171 pixel_to_sphere -> image_to_sphere: promote
 * to double, cse = 1 + x*x + y*y, factor = 1/sqrt(cse), narrow factor*{x,y,1}. */
void dco_pixel_to_sphere(const dco_pinhole* k, int u, int v, float s[3]) {
    float xf, yf;
    image_coord(k, u, v, &xf, &yf);
    const double x      = xf;
    const double y      = yf;
    const double z      = 1.;
    const double xx     = x * x;
    const double yy     = y * y;
    const double cse0   = 1. + xx;
    const double cse    = cse0 + yy;
    const double factor = 1. / sqrt(cse);
    s[0]                = (float)(factor * x);
    s[1]                = (float)(factor * y);
    s[2]                = (float)(factor * z);
}

/* Eigen 3.3.7 Redux.h:91-105 for size 3: c0 + (c1 + c2). */
static float dot3(const float a[3], const float b[3]) {
    const float p0 = a[0] * b[0];
    const float p1 = a[1] * b[1];
    const float p2 = a[2] * b[2];
    const float t  = p1 + p2;
    return p0 + t;
}

static float clampf(float v, float lo, float hi) {
    /* std::clamp: (v < lo) ? lo : (hi < v) ? hi : v  -- NaN passes through */
    return (v < lo) ? lo : ((hi < v) ? hi : v);
}

/* camera_models/utility.h:15-49 phi */
float dco_phi(const dco_pinhole* k, int u1, int v1, int u2, int v2) {
    float s1[3], s2[3];
    dco_pixel_to_sphere(k, u1, v1, s1);
    dco_pixel_to_sphere(k, u2, v2, s2);
    const float delta   = DCO_DELTA;
    const float cos_phi = clampf(dot3(s1, s2), -1.0f + delta, 1.0f - delta);
    return acosf(cos_phi);
}

/* math/triangles.h:27-64 bearing_angle */
float dco_bearing_angle(float b, float c, float cos_alpha) {
    const float delta = DCO_DELTA;
    const float t1    = c * cos_alpha;
    const float nom   = b - t1;
    const float bb    = b * b;
    const float cc    = c * c;
    const float sum   = bb + cc;
    const float b2    = 2.0f * b;
    const float b2c   = b2 * c;
    const float m     = b2c * cos_alpha;
    const float rad   = sum - m;
    const float d     = sqrtf(rad);
    const float den   = (d == 0.0f) ? d + delta : d;
    const float q     = nom / den;
    const float ratio = clampf(q, -1.0f + delta, 1.0f - delta);
    return acosf(ratio);
}

/* ------------------------------------------------------------------------ */
/* conversion/util.h:34-55 orthografic_to_euclidian +
 * conversion/depth_to_laserscan.h:54-93 */
void dco_laserscan(const uint16_t* depth, const dco_pinhole* k, float* out) {
    for (int v = 0; v < k->h; ++v)
        for (int u = 0; u < k->w; ++u) {
            const size_t   i = (size_t)v * k->w + u;
            const uint16_t d = depth[i];
            if (d == 0) {
                out[i] = 0.0f;
                continue;
            }
            float x, y;
            image_coord(k, u, v, &x, &y);
            const float xx = x * x;
            const float yy = y * y;
            const float s  = xx + yy;
            /* "+ 1." is a double literal: promoted, added in double, narrowed
             * on assignment to `const Real z_helper` (util.h:48). */
            const float z_helper = (float)((double)s + 1.);
            out[i]               = (float)d * sqrtf(z_helper);
        }
}

/* math/image.h:117-128 convert<float>(image<ushort>) -- the pinhole-range
 * input type (apps/util/batch_converter.h:160). */
void dco_u16_to_f32(const uint16_t* in, size_t n, float* out) {
    for (size_t i = 0; i < n; ++i)
        out[i] = (float)in[i];
}

/* conversion/depth_to_bearing.h:96-214. dir: 0 horizontal, 1 vertical,
 * 2 diagonal, 3 antidiagonal (conversion/util.h:18-27). */
void dco_bearing(int dir, const float* range, const dco_pinhole* k, float* out) {
    static const int du_t[4]      = {-1, 0, -1, -1}; /* get_du :96-104 */
    static const int dv_t[4]      = {0, -1, -1, 1};  /* get_dv :106-114 */
    static const int x_start_t[4] = {1, 0, 1, 1};    /* get_x_start :135-143 */
    static const int y_start_t[4] = {0, 1, 1, 0};    /* get_y_start :145-153 */
    static const int dy_end_t[4]  = {0, 0, 0, 1};    /* get_dy_end :155-163 */
    const int        w = k->w, h = k->h;
    memset(out, 0, sizeof(float) * (size_t)w * h); /* ba = Real(0.) :237 */
    for (int v = y_start_t[dir]; v < h - dy_end_t[dir]; ++v)
        for (int u = x_start_t[dir]; u < w; ++u) {
            const int   pu    = u + du_t[dir];
            const int   pv    = v + dv_t[dir];
            const float d_i   = range[(size_t)v * w + u];
            const float d_j   = range[(size_t)pv * w + pu];
            const float d_phi = dco_phi(k, u, v, pu, pv);
            const float angle = (d_i == 0.0f || d_j == 0.0f)
                                    ? 0.0f
                                    : dco_bearing_angle(d_i, d_j, cosf(d_phi));
            out[(size_t)v * w + u] = angle;
        }
}

/* conversion/depth_to_flexion.h:70-167 */
static void to_camera(const dco_pinhole* k, int u, int v, float d, float p[3]) {
    float s[3];
    dco_pixel_to_sphere(k, u, v, s);
    p[0] = d * s[0];
    p[1] = d * s[1];
    p[2] = d * s[2];
}
static void sub3(const float a[3], const float b[3], float r[3]) {
    r[0] = a[0] - b[0];
    r[1] = a[1] - b[1];
    r[2] = a[2] - b[2];
}
/* Eigen 3.3.7 Dot.h:124-134 */
static void normalized3(const float n[3], float r[3]) {
    const float z = dot3(n, n);
    if (z > 0.0f) {
        const float s = sqrtf(z);
        r[0]          = n[0] / s;
        r[1]          = n[1] / s;
        r[2]          = n[2] / s;
    } else {
        r[0] = n[0];
        r[1] = n[1];
        r[2] = n[2];
    }
}
/* Eigen 3.3.7 OrthoMethods.h:34-46 */
static void cross3(const float a[3], const float b[3], float r[3]) {
    const float t0 = a[1] * b[2], t1 = a[2] * b[1];
    const float t2 = a[2] * b[0], t3 = a[0] * b[2];
    const float t4 = a[0] * b[1], t5 = a[1] * b[0];
    r[0]           = t0 - t1;
    r[1]           = t2 - t3;
    r[2]           = t4 - t5;
}

void dco_flexion(const float* range, const dco_pinhole* k, float* out) {
    const int w = k->w, h = k->h;
    memset(out, 0, sizeof(float) * (size_t)w * h);
#define R(uu, vv) range[(size_t)(vv)*w + (uu)]
    for (int v = 1; v < h - 1; ++v)
        for (int u = 1; u < w - 1; ++u) {
            float p0[3], p1[3], p2[3], p3[3], p4[3], p5[3], p6[3], p7[3];
            to_camera(k, u, v - 1, R(u, v - 1), p0);
            to_camera(k, u, v + 1, R(u, v + 1), p1);
            to_camera(k, u - 1, v, R(u - 1, v), p2);
            to_camera(k, u + 1, v, R(u + 1, v), p3);
            to_camera(k, u + 1, v - 1, R(u + 1, v - 1), p4);
            to_camera(k, u - 1, v + 1, R(u - 1, v + 1), p5);
            to_camera(k, u - 1, v - 1, R(u - 1, v - 1), p6);
            to_camera(k, u + 1, v + 1, R(u + 1, v + 1), p7);
            float d0[3], d1[3], d2[3], d3[3];
            sub3(p1, p0, d0);
            sub3(p3, p2, d1);
            sub3(p5, p4, d2);
            sub3(p7, p6, d3);
            float n0[3], n1[3], n2[3], n3[3], c0[3], c1[3];
            normalized3(d0, n0);
            normalized3(d1, n1);
            normalized3(d2, n2);
            normalized3(d3, n3);
            cross3(n0, n1, c0);
            cross3(n2, n3, c1);
            const float f = clampf(fabsf(dot3(c0, c1)), 0.0f, 1.0f);
            out[(size_t)v * w + u] = f;
        }
#undef R
}

/* math/derivatives.h:17-40,66-93 and math/curvature.h:18-44;
 * conversion/depth_to_curvature.h:97-206. which: 0 gaussian, 1 mean. */
static void curvature(int which, const float* range, const dco_pinhole* k,
                      float* out) {
    const int w = k->w, h = k->h;
    memset(out, 0, sizeof(float) * (size_t)w * h);
#define R(uu, vv) range[(size_t)(vv)*w + (uu)]
    for (int v = 1; v < h - 1; ++v)
        for (int u = 1; u < w - 1; ++u) {
            const float d__1__1 = R(u - 1, v - 1), d__1__0 = R(u, v - 1),
                        d__1_1 = R(u + 1, v - 1);
            const float d__0__1 = R(u - 1, v), d__0__0 = R(u, v),
                        d__0_1 = R(u + 1, v);
            const float d_1__1 = R(u - 1, v + 1), d_1__0 = R(u, v + 1),
                        d_1_1 = R(u + 1, v + 1);
            if (d__1__1 == 0. || d__1__0 == 0. || d__1_1 == 0. || d__0__1 == 0. ||
                d__0__0 == 0. || d__0_1 == 0. || d_1__1 == 0. || d_1__0 == 0. ||
                d_1_1 == 0.) {
                out[(size_t)v * w + u] = 0.0f;
                continue;
            }
            const float d_phi       = dco_phi(k, u - 1, v, u + 1, v);
            const float d_theta     = dco_phi(k, u, v - 1, u, v + 1);
            const float d_phi_theta = dco_phi(k, u - 1, v - 1, u + 1, v + 1);
            /* first_derivative_central: (y_1 - y__1) / (2 * dx) */
            const float f_u = (d__0_1 - d__0__1) / (2.0f * d_phi);
            const float f_v = (d_1__0 - d__1__0) / (2.0f * d_theta);
            /* second_derivative_central: (y_1 + y__1 - 2*y_0) / (dx*dx) */
            const float two_d0 = 2.0f * d__0__0;
            const float f_uu   = ((d__0_1 + d__0__1) - two_d0) / (d_phi * d_phi);
            const float f_vv   = ((d_1__0 + d__1__0) - two_d0) / (d_theta * d_theta);
            /* sic: vertical neighbours with d_phi_theta, derivatives.h:89-90 */
            const float f_uv =
                ((d_1__0 + d__1__0) - two_d0) / (d_phi_theta * d_phi_theta);
            float res;
            if (which == 0) {
                /* (f_uu*f_vv - f_uv*f_uv) / (1 + f_u*f_u + f_v*f_v) */
                const float a  = f_uu * f_vv;
                const float b  = f_uv * f_uv;
                const float n  = a - b;
                const float e0 = f_u * f_u;
                const float e1 = f_v * f_v;
                const float dd = (1.0f + e0) + e1;
                res            = n / dd;
            } else {
                /* ((1+f_v^2) f_uu - 2 f_u f_v f_uv + (1+f_u^2) f_vv) /
                 *  pow(2 sqrt(1 + f_u^2 + f_v^2), 3) */
                const float fv2 = f_v * f_v;
                const float fu2 = f_u * f_u;
                const float t0  = (1.0f + fv2) * f_uu;
                const float t1  = ((2.0f * f_u) * f_v) * f_uv;
                const float t2  = (1.0f + fu2) * f_vv;
                const float num = (t0 - t1) + t2;
                const float rt  = sqrtf((1.0f + fu2) + fv2);
                const float den = powf(2.0f * rt, 3.0f);
                res             = num / den;
            }
            out[(size_t)v * w + u] = res;
        }
#undef R
}
void dco_gaussian_curvature(const float* range, const dco_pinhole* k, float* out) {
    curvature(0, range, k, out);
}
void dco_mean_curvature(const float* range, const dco_pinhole* k, float* out) {
    curvature(1, range, k, out);
}

/* conversion/depth_to_max_curve.h:55-82 angle_formula */
static float angle_formula(float d__1, float d__0, float d_1, float cos_alpha1,
                           float cos_alpha2) {
    if (d__1 == 0. || d__0 == 0. || d_1 == 0.)
        return 0.0f;
    const float a1 = dco_bearing_angle(d__0, d__1, cos_alpha1);
    const float a2 = dco_bearing_angle(d__0, d_1, cos_alpha2);
    return a1 + a2;
}
static float maxf(float a, float b) {
    /* std::max(a, b): (a < b) ? b : a */
    return (a < b) ? b : a;
}
/* conversion/depth_to_max_curve.h:86-153 */
void dco_max_curve(const float* range, const dco_pinhole* k, float* out) {
    const int w = k->w, h = k->h;
    memset(out, 0, sizeof(float) * (size_t)w * h);
#define R(uu, vv) range[(size_t)(vv)*w + (uu)]
    for (int v = 1; v < h - 1; ++v)
        for (int u = 1; u < w - 1; ++u) {
            const float d__1__1 = R(u - 1, v - 1), d__1__0 = R(u, v - 1),
                        d__1_1 = R(u + 1, v - 1);
            const float d__0__1 = R(u - 1, v), d__0__0 = R(u, v),
                        d__0_1 = R(u + 1, v);
            const float d_1__1 = R(u - 1, v + 1), d_1__0 = R(u, v + 1),
                        d_1_1 = R(u + 1, v + 1);
            const float phi_hor1 = dco_phi(k, u - 1, v, u, v);
            const float phi_hor2 = dco_phi(k, u, v, u + 1, v);
            const float a_hor    = angle_formula(d__0__1, d__0__0, d__0_1,
                                                 cosf(phi_hor1), cosf(phi_hor2));
            const float phi_ver1 = dco_phi(k, u, v - 1, u, v);
            const float phi_ver2 = dco_phi(k, u, v, u, v + 1);
            const float a_ver    = angle_formula(d__1__0, d__0__0, d_1__0,
                                                 cosf(phi_ver1), cosf(phi_ver2));
            const float phi_dia1 = dco_phi(k, u - 1, v - 1, u, v);
            const float phi_dia2 = dco_phi(k, u, v, u + 1, v + 1);
            const float a_dia    = angle_formula(d__1__1, d__0__0, d_1_1,
                                                 cosf(phi_dia1), cosf(phi_dia2));
            /* sic: the first antidiagonal angle is taken towards (u+1, v+1)
             * while the depth comes from (u-1, v+1); :136-139 */
            const float phi_ant1 = dco_phi(k, u + 1, v + 1, u, v);
            const float phi_ant2 = dco_phi(k, u, v, u + 1, v - 1);
            const float a_ant    = angle_formula(d_1__1, d__0__0, d__1_1,
                                                 cosf(phi_ant1), cosf(phi_ant2));
            out[(size_t)v * w + u] =
                maxf(a_hor, maxf(a_ver, maxf(a_dia, a_ant)));
        }
#undef R
}

/* ------------------------------------------------------------------------ */
/* Quantisers. cv::Mat::convertTo(dst, CV_16U, alpha, beta) on a 32F source is
 * saturate_cast<ushort>(src * (float)alpha + (float)beta): cvRound (round half
 * to even) then clamp to [0, 65535]. */
static long cv_round(float v) {
    /* OpenCV cvRound(float) is cvtss2si to a 32 bit int: NaN and anything
     * outside the int range become INT_MIN ("integer indefinite"). */
    if (!(fabsf(v) < 2147483648.0f))
        return -2147483647L - 1L;
    return lrintf(v);
}
static uint16_t sat_u16(float v) {
    const long r = cv_round(v);
    return (uint16_t)(r < 0 ? 0 : (r > 65535 ? 65535 : r));
}
static void convert_scale_u16(const float* in, size_t n, float alpha, float beta,
                              uint16_t* out) {
    for (size_t i = 0; i < n; ++i) {
        const float p = in[i] * alpha;
        const float s = p + beta;
        out[i]        = sat_u16(s);
    }
}
/* conversion/util.h:71-79 scaling_factor<float, ushort>(max_angle) */
static float scaling_factor_u16(float max_angle) {
    const float min = 0.0f, max = 65535.0f;
    return (max - min) / max_angle;
}
/* conversion/depth_to_bearing.h:282-301 */
void dco_convert_bearing_u16(const float* in, size_t n, uint16_t* out) {
    convert_scale_u16(in, n, scaling_factor_u16(DCO_PI_F), 0.0f, out);
}
/* conversion/depth_to_flexion.h:194-215 */
void dco_convert_flexion_u16(const float* in, size_t n, uint16_t* out) {
    convert_scale_u16(in, n, 65535.0f - 0.0f, 0.0f, out);
}
/* conversion/depth_to_max_curve.h:156-176: 2. * pi<float> evaluated in double,
 * narrowed to the float parameter of scaling_factor. */
void dco_convert_max_curve_u16(const float* in, size_t n, uint16_t* out) {
    const float max_angle = (float)(2. * (double)DCO_PI_F);
    convert_scale_u16(in, n, scaling_factor_u16(max_angle), 0.0f, out);
}
/* src/apps/depth2x/converter_laserscan.h.inl:9-10 convertTo(CV_16U), no scale */
void dco_convert_range_u16(const float* in, size_t n, uint16_t* out) {
    for (size_t i = 0; i < n; ++i)
        out[i] = sat_u16(in[i]);
}
/* conversion/depth_to_curvature.h:213-283 curvature_to_image<ushort> with both
 * clamps given (the only case the CLI produces, converter_curvature.h.inl:9-10)
 * + math/scaling.h:29-55 scale + gsl::narrow_cast (truncation), then masked by
 * range.convertTo(CV_8U) != 0. */
void dco_curvature_to_image_u16(const float* curv, const float* range, size_t n,
                                float lo, float hi, uint16_t* out) {
    const float tmin = 0.0f, tmax = 65535.0f;
    for (size_t i = 0; i < n; ++i) {
        const float a      = (tmax - tmin) * (curv[i] - lo);
        const float scaled = a / (hi - lo) + tmin;
        const float tc     = clampf(scaled, tmin, tmax);
        /* narrow_cast<ushort>(float): cvttss2si then truncate to 16 bit;
         * NaN -> 0x80000000 -> 0 */
        const uint16_t q = (uint16_t)(int)tc;
        /* mask: saturate_cast<uchar>(range) != 0 */
        const long     m  = cv_round(range[i]);
        const int      on = (m < 0 ? 0 : (m > 255 ? 255 : m)) != 0;
        out[i]            = on ? q : 0;
    }
}
/* conversion/depth_scaling.h:21-31: 16U -> 16U convertTo(alpha, beta) computes
 * in float (OpenCV cvt_32f for 16u sources). */
void dco_depth_scaling_u16(const uint16_t* in, size_t n, double scale,
                           double offset, uint16_t* out) {
    const float a = (float)scale, b = (float)offset;
    for (size_t i = 0; i < n; ++i) {
        const float p = (float)in[i] * a;
        const float s = p + b;
        out[i]        = sat_u16(s);
    }
}

/* src/lib/util/correctness_util.cpp:20-31 average_pixel_error = ||a-b||_2 / n */
double dco_average_pixel_error_u16(const uint16_t* a, const uint16_t* b, size_t n) {
    double s = 0.;
    for (size_t i = 0; i < n; ++i) {
        const double d = (double)a[i] - (double)b[i];
        s += d * d;
    }
    return sqrt(s) / (double)n;
}

/* ------------------------------------------------------------------------ */
/* Synthetic Kinect-like depth frames, integer-only so that C, numpy and the
 * device generator agree bit for bit (DESIGN.md "Synthetic input"). */
static uint64_t splitmix64(uint64_t x) {
    uint64_t z = x + 0x9E3779B97F4A7C15ull;
    z          = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z          = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
static int64_t tri(int64_t t, int64_t p) {
    const int64_t m = t % (2 * p);
    return m < p ? p - m : m - p;
}
void dco_synth_depth(uint64_t seed, int frame, int w, int h, uint16_t* out) {
    const uint64_t fkey = (uint64_t)frame * 0x9E3779B97F4A7C15ull;
    for (int v = 0; v < h; ++v)
        for (int u = 0; u < w; ++u) {
            const uint64_t hsh =
                splitmix64(seed ^ fkey ^ (uint64_t)((int64_t)v * w + u));
            int64_t d = 800 + 2 * tri((int64_t)u + 3 * (int64_t)frame, 400) +
                        3 * tri((int64_t)v + 5 * (int64_t)frame, 300) +
                        (int64_t)(hsh % 9) - 4;
            const uint64_t blk = splitmix64(
                seed ^ (fkey + 0xD1B54A32D192ED03ull) ^
                (uint64_t)(((int64_t)(v / 16)) * 4096 + (u / 16)));
            if (blk % 100 < 25 || (hsh >> 32) % 100 < 5)
                d = 0;
            out[(size_t)v * w + u] = (uint16_t)d;
        }
}
