// Device arithmetic of the depth -> feature hot path, written op for op in the
// rounding order of the reference (Real = float). Explicit __f*_rn intrinsics
// are never contracted into FMAs by nvcc, so the sequence below is what runs
// regardless of -fmad. Division and square root are the IEEE-correct
// (__fdiv_rn / __fsqrt_rn) forms: the reference build has neither FMA nor
// fast-math (cmake/configuration_functions.cmake:53-63), and the bearing
// formula cancels catastrophically, so "mathematically
// equivalent" shortcuts change the result at the 1e-2 level.
#ifndef DC_MATH_CUH
#define DC_MATH_CUH

#include "dc_fastmath.cuh"

#include <cuda_runtime.h>

namespace dc {

// 10 * numeric_limits<float>::epsilon(): camera_models/utility.h:39,
// math/triangles.h:39
#define DC_DELTA (10.0f * 1.1920928955078125e-07f)
#define DC_ONE_MINUS_DELTA (1.0f - DC_DELTA)
#define DC_MINUS_ONE_PLUS_DELTA (-1.0f + DC_DELTA)

__device__ __forceinline__ float mul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float add(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float sub(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float div(float a, float b) { return __fdiv_rn(a, b); }
__device__ __forceinline__ float sqrt_rn(float a) { return __fsqrt_rn(a); }

// std::clamp for in-contract (non-NaN) values. A NaN here means the
// reference's Ensures() would already have terminated the program.
__device__ __forceinline__ float clamp(float v, float lo, float hi) {
    return fminf(fmaxf(v, lo), hi);
}

struct vec3 {
    float x, y, z;
};

// Eigen 3.3.7 fixed-size-3 reduction order: c0 + (c1 + c2)  (Core/Redux.h:91-105)
__device__ __forceinline__ float dot3(const vec3& a, const vec3& b) {
    return add(mul(a.x, b.x), add(mul(a.y, b.y), mul(a.z, b.z)));
}
__device__ __forceinline__ vec3 sub3(const vec3& a, const vec3& b) {
    return {sub(a.x, b.x), sub(a.y, b.y), sub(a.z, b.z)};
}
// Eigen 3.3.7 Core/Dot.h:124-134: z = squaredNorm; z > 0 ? n / sqrt(z) : n
//
// FAST: the three divisions share one refined reciprocal, seeded from the
// rsqrt of the square-root step (tests/cuda/mathcheck.cu "normalise_shared_rcp":
// bit-identical to __fdiv_rn(x, __fsqrt_rn(z))). z == 0 happens only for the
// null vector, which must be returned unchanged: clamping z to 2^-100 makes the
// divisions yield 0 = the input without a branch. Needs z == 0 or z >= 2^-100
// and quotients that are 0 or normal: guaranteed per camera (dc_tables.cpp
// fast_ok) for u16 depths.
template <bool FAST>
__device__ __forceinline__ vec3 normalized3(const vec3& n) {
    const float z = dot3(n, n);
    if constexpr (FAST) {
        float       h;
        const float s = sqrt_rn_fast(fmaxf(z, 7.888609052210118e-31f), h);
        const float y = rcp_refined(s, add(h, h));
        return {div_rn_with_rcp(n.x, s, y), div_rn_with_rcp(n.y, s, y),
                div_rn_with_rcp(n.z, s, y)};
    } else {
        if (z > 0.0f) {
            const float s = sqrt_rn(z);
            return {div(n.x, s), div(n.y, s), div(n.z, s)};
        }
        return n;
    }
}
// Eigen 3.3.7 Geometry/OrthoMethods.h:34-46
__device__ __forceinline__ vec3 cross3(const vec3& a, const vec3& b) {
    return {sub(mul(a.y, b.z), mul(a.z, b.y)), sub(mul(a.z, b.x), mul(a.x, b.z)),
            sub(mul(a.x, b.y), mul(a.y, b.x))};
}

// math/triangles.h:27-64 bearing_angle(b, c, cos_alpha), with bb = b*b,
// cc = c*c and b2 = 2*b passed in (shared between directions / neighbours).
// Returns the clamped ratio's angle; `b2c` (= (2b)*c) is handed back so callers
// can test "either depth is zero" on it.
//
// FAST (b, c in [2^-10, 2^20], cos_alpha in (0, 1 - 8e-7]): the radicand is
// then strictly positive (b^2 + c^2 >= 2bc and the rounding errors, < 4e-7
// relative, are smaller than 1 - cos_alpha), so the reference's "d == 0 ?
// d + delta" never fires, and every operand of sqrt / div is normal.
template <bool FAST>
__device__ __forceinline__ float bearing_angle_pre(float b, float c, float bb, float cc,
                                                   float b2, float cos_alpha, float& b2c) {
    const float nom = sub(b, mul(c, cos_alpha));
    const float sum = add(bb, cc);
    b2c             = mul(b2, c);
    const float rad = sub(sum, mul(b2c, cos_alpha));
    float       q;
    if constexpr (FAST) {
        float       h;
        const float d = sqrt_rn_fast(fmaxf(rad, 7.888609052210118e-31f), h);
        q             = div_rn_with_rcp(nom, d, rcp_refined(d, add(h, h)));
    } else {
        const float d   = sqrt_rn(rad);
        const float den = (d == 0.0f) ? add(d, DC_DELTA) : d;
        q               = div(nom, den);
    }
    return acos_rel(clamp(q, DC_MINUS_ONE_PLUS_DELTA, DC_ONE_MINUS_DELTA));
}
template <bool FAST>
__device__ __forceinline__ float bearing_angle(float b, float c, float cos_alpha) {
    float b2c;
    return bearing_angle_pre<FAST>(b, c, mul(b, b), mul(c, c), mul(2.0f, b), cos_alpha, b2c);
}

// conversion/depth_to_flexion.h:105-135, given the eight neighbour points
// P = d * ray already multiplied out (to_camera, :70-76).
template <bool FAST>
__device__ __forceinline__ float flexion_from_points(const vec3& p0, const vec3& p1,
                                                     const vec3& p2, const vec3& p3,
                                                     const vec3& p4, const vec3& p5,
                                                     const vec3& p6, const vec3& p7) {
    const vec3  n0 = normalized3<FAST>(sub3(p1, p0));
    const vec3  n1 = normalized3<FAST>(sub3(p3, p2));
    const vec3  n2 = normalized3<FAST>(sub3(p5, p4));
    const vec3  n3 = normalized3<FAST>(sub3(p7, p6));
    const vec3  c0 = cross3(n0, n1);
    const vec3  c1 = cross3(n2, n3);
    return clamp(fabsf(dot3(c0, c1)), 0.0f, 1.0f);
}

struct derivs {
    float f_u, f_v, f_uu, f_vv, f_uv;
};
// math/derivatives.h:17-40, 66-93. left/right = (u-1,v)/(u+1,v), up/down =
// (u,v-1)/(u,v+1). f_uv uses the VERTICAL neighbours with d_phi_theta (sic).
// FAST: ray angles lie in [1.5e-3, pi] (checked per camera) and depths in
// [2^-10, 2^20], so each numerator is 0 or >= 2^-34 in magnitude and every
// quotient is 0 or normal: the guard-free division is exact.
template <bool FAST>
__device__ __forceinline__ float div_sel(float a, float b) {
    if constexpr (FAST)
        return div_rn_fast(a, b);
    else
        return div(a, b);
}
template <bool FAST>
__device__ __forceinline__ derivs derivatives(float left, float right, float up,
                                              float down, float centre, float d_phi,
                                              float d_theta, float d_phi_theta) {
    derivs      r;
    const float two_c = mul(2.0f, centre);
    const float sh    = sub(add(right, left), two_c);
    const float sv    = sub(add(down, up), two_c);
    r.f_u             = div_sel<FAST>(sub(right, left), mul(2.0f, d_phi));
    r.f_v             = div_sel<FAST>(sub(down, up), mul(2.0f, d_theta));
    r.f_uu            = div_sel<FAST>(sh, mul(d_phi, d_phi));
    if constexpr (FAST) {
        // f_vv and f_uv divide the same numerator
        r.f_vv = div_rn_fast(sv, mul(d_theta, d_theta));
        r.f_uv = div_rn_fast(sv, mul(d_phi_theta, d_phi_theta));
    } else {
        r.f_vv = div(sv, mul(d_theta, d_theta));
        r.f_uv = div(sv, mul(d_phi_theta, d_phi_theta));
    }
    return r;
}
// math/curvature.h:18-23
__device__ __forceinline__ float gaussian_curvature(const derivs& d) {
    const float n  = sub(mul(d.f_uu, d.f_vv), mul(d.f_uv, d.f_uv));
    const float dd = add(add(1.0f, mul(d.f_u, d.f_u)), mul(d.f_v, d.f_v));
    return div(n, dd);
}
// math/curvature.h:33-44. pow(x, 3.f): glibc's powf works in double and is
// correctly rounded in all but ~1e-4 of cases, so x*x*x evaluated in double
// (x*x is exact, the second product carries 53 bits) and narrowed once agrees
// with it far more often than CUDA's 4-ulp powf would, and is cheaper.
__device__ __forceinline__ float mean_curvature(const derivs& d) {
    const float fv2 = mul(d.f_v, d.f_v);
    const float fu2 = mul(d.f_u, d.f_u);
    const float t0  = mul(add(1.0f, fv2), d.f_uu);
    const float t1  = mul(mul(mul(2.0f, d.f_u), d.f_v), d.f_uv);
    const float t2  = mul(add(1.0f, fu2), d.f_vv);
    const float num = add(sub(t0, t1), t2);
    const float rt  = sqrt_rn(add(add(1.0f, fu2), fv2));
    const double x  = double(mul(2.0f, rt));
    const float den = __double2float_rn(__dmul_rn(__dmul_rn(x, x), x));
    return div(num, den);
}

// ---- quantisers (cv::Mat::convertTo semantics: round half to even, clamp) ---
__device__ __forceinline__ unsigned short sat_u16_rn(float v) {
    // cvRound + saturate_cast<ushort>. __float2int_rn saturates and maps NaN
    // to 0; x86 cvRound maps NaN/out-of-int-range to INT_MIN -> 0.
    int r = __float2int_rn(v);
    if (!(fabsf(v) < 2147483648.0f))
        r = 0;
    return (unsigned short)min(max(r, 0), 65535);
}
// convertTo(CV_16U, alpha, 0) on a float source
__device__ __forceinline__ unsigned short quant_scale(float v, float alpha) {
    return sat_u16_rn(mul(v, alpha));
}
// The same for features with a known bound (bearing in [0, pi), flexion in
// [0, 1], max-curve in [0, 2 pi)): 0 <= v*alpha < 2^22, so round-half-to-even
// is one FADD with 1.5 * 2^23 and the integer sits in the low mantissa bits --
// no F2I on the XU pipe. min() only guards the upper clamp of convertTo.
__device__ __forceinline__ unsigned short quant_scale_bounded(float v, float alpha) {
    const float x = fminf(mul(v, alpha), 65535.0f);
    return (unsigned short)(__float_as_uint(add(x, 12582912.0f)) & 0xffffu);
}
// math/scaling.h:29-55 scale({lo,hi},{0,65535}, v) then gsl::narrow_cast<ushort>
// (truncation), conversion/depth_to_curvature.h:239-243; masked by
// saturate_cast<uchar>(range) != 0 (:266-275).
__device__ __forceinline__ unsigned short quant_curvature(float v, float range,
                                                          float lo, float hi) {
    const float a      = mul(65535.0f, sub(v, lo));
    const float scaled = div(a, sub(hi, lo));  // "+ target_min" adds +0: no effect after truncation
    // std::clamp keeps NaN; (int)NaN is 0 on the device and INT_MIN -> ushort 0
    // on x86: same pixel.
    const float tc = (scaled < 0.0f) ? 0.0f : ((65535.0f < scaled) ? 65535.0f : scaled);
    const unsigned short q = (unsigned short)__float2int_rz(tc);
    const int m = __float2int_rn(range);  // round half to even, as cvRound
    const bool on = (fabsf(range) < 2147483648.0f) && (m > 0);
    return on ? q : (unsigned short)0;
}

}  // namespace dc
#endif
