// The frame-pair row-marching kernels (see dc_pair.cuh for the design).
#ifndef DC_PAIR_KERNELS_CUH
#define DC_PAIR_KERNELS_CUH

#include "dc_pair.cuh"

#include <type_traits>

namespace dc {
namespace pr {

constexpr int kPairThreads = 128;
// 1: the row loops that are instantiated several times with renamed registers
// (curvature, flexion, shared-edge max-curve) have ONE exit, at the bottom: a strip
// always runs whole groups of steps, steps past the strip's last row only suppress
// their stores. With an exit after every step ptxas treats the loop as divergent
// and keeps loop constants that could sit in uniform registers -- above all the
// -0.0 addend of every product (dc_pair.cuh mul2) -- in ordinary registers: an
// FFMA2 then reads three register operands instead of two (profiles/r2_ab_experiments.md).
#ifndef DC_SINGLE_EXIT
#define DC_SINGLE_EXIT 1
#endif

#ifndef DC_REFLECT_FSEL
#define DC_REFLECT_FSEL 1
#endif
// 1: the u16 bearing epilogue multiplies the acos polynomial's coefficients by
// 65535/pi and folds s * (Q P) into the rounding FFMA2 (16 FFMA2 per row less).
// Round 1 measured no gain with run-time scaled coefficients (register operands
// instead of immediates, 0.5723 vs 0.5704 ms); round 2 made the factor a compile-time
// constant (immediates again, dc_pair.cuh acos_parts_q2): still no gain (0.5215 vs
// 0.5213 ms, 78 instead of 82 packed operations per pixel pair; the -s select costs what
// the product saved). Off. What did pay is the constant itself: the +-Q / offset selects
// of the epilogue take immediates (0.5307 -> 0.5213 ms).
#ifndef DC_EDGE_RAD_FADD
#define DC_EDGE_RAD_FADD 1
#endif
#ifndef DC_BEARING_FOLD_Q
#define DC_BEARING_FOLD_Q 0
#endif
// Features whose three-stage row loop is instantiated three times with the
// window rows / stages renamed instead of rotated (bit mask over FEAT_*).
// Measured: curvature 0.443 -> 0.429 ms, bearing 0.605 -> 0.615 ms (kept rotating).
#ifndef DC_PAIR_RENAME_MASK
#define DC_PAIR_RENAME_MASK (1 << FEAT_CURVATURE)
#endif

// One window row: range of columns u0-1 .. u0+4 of both frames, plus (curvature
// only) the minimum over each 3-column neighbourhood for the "any of the nine
// depths is zero" test.
// 1: the squares of a bearing window row are computed once, when the row enters
// the window (5 FFMA2 per row step instead of 17), and carried with it (20 more
// live registers). profiles/r2_ab_experiments.md
#ifndef DC_BEARING_CARRY_SQ
#define DC_BEARING_CARRY_SQ 0
#endif
template <int FEAT>
struct wrow {
    f2 r[6];
    f2 m3[FEAT == FEAT_CURVATURE ? 4 : 1];
    f2 s[(FEAT == FEAT_BEARING && DC_BEARING_CARRY_SQ) ? 5 : 1];
};

template <int FEAT>
struct feat_traits {
    static constexpr bool right = FEAT != FEAT_BEARING;  // needs column u0+4
};

template <int FEAT, int IN>
__device__ __forceinline__ void commit(const rawrow& raw, wrow<FEAT>& dst, f32_window& win) {
    to_range<IN, feat_traits<FEAT>::right>(raw, dst.r, win);
    if constexpr (FEAT == FEAT_CURVATURE) {
#pragma unroll
        for (int i = 0; i < 4; ++i)
            dst.m3[i] = make_float2(fminf(fminf(dst.r[i].x, dst.r[i + 1].x), dst.r[i + 2].x),
                                    fminf(fminf(dst.r[i].y, dst.r[i + 1].y), dst.r[i + 2].y));
    }
    if constexpr (FEAT == FEAT_BEARING && DC_BEARING_CARRY_SQ) {
#pragma unroll
        for (int i = 0; i < 5; ++i)
            dst.s[i] = mul2(dst.r[i], dst.r[i]);
    }
}

struct pos {
    int    v, u0;
    bool   full;     // SPEC_HALF: columns u0+2, u0+3 exist (else always true)
    bool   live;     // the row belongs to the strip: stores happen (DC_SINGLE_EXIT)
    size_t oa, ob;   // element offset of (frame, v, u0) in frame A / B
    float  inside[4];  // 1 for pixels a 3x3 stencil computes, else 0 (border)
};

// ---- per-row camera tables, fetched one row ahead --------------------------
template <int FEAT>
struct rtabs {};
template <>
struct rtabs<FEAT_BEARING> {
    float4 c[4];
};
template <>
struct rtabs<FEAT_MAX_CURVE> {
    float cH[5], cV0[4], cV1[4], cD0[4], cD1[5], cA[5];
};
template <>
struct rtabs<FEAT_CURVATURE> {
    float4 phi, theta, pt;
};

__device__ __forceinline__ void put4(float* d, float4 v) {
    d[0] = v.x, d[1] = v.y, d[2] = v.z, d[3] = v.w;
}

template <int FEAT>
struct tab_slots {
    static constexpr int n = FEAT == FEAT_BEARING ? 4 : FEAT == FEAT_MAX_CURVE ? 7 : FEAT == FEAT_CURVATURE ? 3 : 0;
};

// Starts the copies of the camera tables that row `v` of the feature needs
// into the thread's table slots st[0], st[stride], ... `off` = v * w + u0.
template <int FEAT, int SPEC>
__device__ __forceinline__ void issue_tabs(const kparams& p, int v, int off, int u0, uint4* st,
                                           int stride, bool full) {
    const int w = p.w, h = p.h;
    if constexpr (FEAT == FEAT_BEARING) {
        const bool in = v < h;
        off           = in ? off : 0;
#pragma unroll
        for (int d = 0; d < 4; ++d)
            if (wanted<SPEC>(p.out[d]))
                issue_tab4<SPEC>(p.cosp[d], off, in, st + d * stride, full);
    } else if constexpr (FEAT == FEAT_MAX_CURVE) {
        // depth_to_max_curve.h:113-139: all eight phi() are angles between a
        // pixel and one of its bearing "prior" neighbours (see dc_kernels.cuh):
        //   hor1 = H(u,v)      hor2 = H(u+1,v)      ver1 = V(u,v)   ver2 = V(u,v+1)
        //   dia1 = D(u,v)      dia2 = D(u+1,v+1)
        //   ant1 = phi((u+1,v+1),(u,v)) = D(u+1,v+1) (sic, :136)   ant2 = A(u+1,v-1)
        const bool in = v >= 1 && v < h - 1, more = in && u0 + 4 < w;
        off           = in ? off : w;  // rows off-w .. off+w stay inside the table
        issue_tab4<SPEC>(p.cosp[0], off, in, st, full);
        issue_tab4<SPEC>(p.cosp[1], off, in, st + stride, full);
        issue_tab4<SPEC>(p.cosp[1], off + w, in, st + 2 * stride, full);
        issue_tab4<SPEC>(p.cosp[2], off, in, st + 3 * stride, full);
        issue_tab4<SPEC>(p.cosp[2], off + w, in, st + 4 * stride, full);
        issue_tab4<SPEC>(p.cosp[3], off - w, in, st + 5 * stride, full);
        char*     e  = reinterpret_cast<char*>(st + 6 * stride);
        const int o4 = more ? 4 : 0;
        issue_tab1<SPEC>(p.cosp[0], off + o4, more, e);
        issue_tab1<SPEC>(p.cosp[2], off + w + o4, more, e + 4);
        issue_tab1<SPEC>(p.cosp[3], off - w + o4, more, e + 8);
    } else if constexpr (FEAT == FEAT_CURVATURE) {
        const bool in = v >= 1 && v < h - 1;
        off           = in ? off : 0;
        issue_tab4<SPEC>(p.d_phi, off, in, st, full);
        issue_tab4<SPEC>(p.d_theta, off, in, st + stride, full);
        issue_tab4<SPEC>(p.d_phi_theta, off, in, st + 2 * stride, full);
    }
}

template <int FEAT, int SPEC>
__device__ __forceinline__ rtabs<FEAT> fetch_tabs(const kparams& p, const uint4* st, int stride) {
    rtabs<FEAT> t;
    if constexpr (FEAT == FEAT_BEARING) {
#pragma unroll
        for (int d = 0; d < 4; ++d)
            t.c[d] = wanted<SPEC>(p.out[d]) ? fetch_tab4<SPEC>(p.cosp[d], st + d * stride)
                                         : make_float4(.5f, .5f, .5f, .5f);
    } else if constexpr (FEAT == FEAT_MAX_CURVE) {
        put4(t.cH, fetch_tab4<SPEC>(p.cosp[0], st));
        put4(t.cV0, fetch_tab4<SPEC>(p.cosp[1], st + stride));
        put4(t.cV1, fetch_tab4<SPEC>(p.cosp[1], st + 2 * stride));
        put4(t.cD0, fetch_tab4<SPEC>(p.cosp[2], st + 3 * stride));
        put4(t.cD1, fetch_tab4<SPEC>(p.cosp[2], st + 4 * stride));
        put4(t.cA, fetch_tab4<SPEC>(p.cosp[3], st + 5 * stride));
        const float4 e = *reinterpret_cast<const float4*>(st + 6 * stride);
        t.cH[4]        = is_table<SPEC>(p.cosp[0]) ? e.x : p.cosp[0].value;
        t.cD1[4]       = is_table<SPEC>(p.cosp[2]) ? e.y : p.cosp[2].value;
        t.cA[4]        = is_table<SPEC>(p.cosp[3]) ? e.z : p.cosp[3].value;
    } else if constexpr (FEAT == FEAT_CURVATURE) {
        t.phi   = fetch_tab4<SPEC>(p.d_phi, st);
        t.theta = fetch_tab4<SPEC>(p.d_theta, st + stride);
        t.pt    = fetch_tab4<SPEC>(p.d_phi_theta, st + 2 * stride);
    }
    return t;
}

// ---- bearing: conversion/depth_to_bearing.h:185-214 ------------------------
template <int OUT, int SPEC>
__device__ __forceinline__ void bearing_row(const kparams& p, const wrow<FEAT_BEARING>& up,
                                            const wrow<FEAT_BEARING>& mid,
                                            const wrow<FEAT_BEARING>& dn,
                                            const rtabs<FEAT_BEARING>& tb, const pos& at) {
    f2 sm[5];  // squares of the centre row: b*b of every pixel, c*c for horizontal
#pragma unroll
    for (int i = 0; i < 5; ++i)
        sm[i] = DC_BEARING_CARRY_SQ ? mid.s[i] : mul2(mid.r[i], mid.r[i]);
#pragma unroll
    for (int d = 0; d < 4; ++d) {
        if (!wanted<SPEC>(p.out[d]))
            continue;
        const float cs[4] = {tb.c[d].x, tb.c[d].y, tb.c[d].z, tb.c[d].w};
        f2          o[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const f2 b = mid.r[i + 1], bb = sm[i + 1];
            // prior pixel: hor (u-1,v), ver (u,v-1), dia (u-1,v-1), anti (u-1,v+1)
            const f2   c  = d == 0 ? mid.r[i] : d == 1 ? up.r[i + 1] : d == 2 ? up.r[i] : dn.r[i];
            const f2   cc = d == 0 ? sm[i]
                            : !DC_BEARING_CARRY_SQ ? mul2(c, c)
                            : d == 1 ? up.s[i + 1] : d == 2 ? up.s[i] : dn.s[i];
            constexpr bool WITH_ACOS = OUT == OUT_U16 && !DC_BEARING_FOLD_Q;
            const bres k  = bearing_core2<WITH_ACOS, OUT == OUT_U16>(b, c, bb, cc, cs[i]);
            // depth_to_bearing.h:204-207: either depth 0 (or the prior pixel
            // outside the image, which reads as 0) -> 0
            const bool va = k.bc.x != 0.0f, vb = k.bc.y != 0.0f;
            if constexpr (OUT == OUT_U16) {
                // convert_bearing (:282-301): round(angle * 65535/pi). Reflection,
                // scale and the round-half-even magic constant in ONE FFMA2:
                // (pi - r) * Q + M = r * (-Q) + (pi*Q + M). Against scaling first
                // and rounding afterwards this moves a result by at most one LSB
                // (tolerance of the quantised images: +-1 LSB; acos itself is
                // only 1e-6 accurate against glibc).
                const float Q = kQBearing, M = 12582912.0f;
#if DC_BEARING_FOLD_Q
                // acos = s * P(a): the scale is inside the polynomial's (compile-time)
                // coefficients, the reflection's sign goes onto s, and the product
                // s * (Q P) happens inside the rounding FFMA2.
                const float C  = fmaf(DC_PI_F, Q, M);
                const bool  na = k.q.x < 0.0f, nb = k.q.y < 0.0f;
                f2          s, pk;
                acos_parts_q2(k.q, s, pk);
                const f2 ss  = make_float2(na ? -s.x : s.x, nb ? -s.y : s.y);
                const f2 off = make_float2(na ? C : M, nb ? C : M);
                const f2 x   = fma2(ss, pk, off);
#else
#if DC_REFLECT_FSEL
                const float C  = fmaf(DC_PI_F, Q, M);
                const bool  na = k.q.x < 0.0f, nb = k.q.y < 0.0f;
                const f2    sq  = make_float2(na ? -Q : Q, nb ? -Q : Q);
                const f2    off = make_float2(na ? C : M, nb ? C : M);
#else
                const f2 sq  = make_float2(sign_xor(Q, k.q.x), sign_xor(Q, k.q.y));
                const f2 off = make_float2(magic_refl(k.q.x), magic_refl(k.q.y));
#endif
                const f2    x   = fma2(k.r, sq, off);
#endif
#if DC_ABL & 2
                o[i] = fma2(k.r, B(Q), B(M));
                (void)va, (void)vb, (void)x;
#else
                o[i]            = make_float2(va ? x.x : M, vb ? x.y : M);
#endif
            } else {
                const f2 a = acos_signed2(k.q);
                o[i]       = make_float2(va ? a.x : 0.0f, vb ? a.y : 0.0f);
            }
        }
        if (at.live) {
            if constexpr (OUT == OUT_U16)
                store4_magic<SPEC == SPEC_HALF>(p.out[d], at.oa, at.ob, o, at.full);
            else
                store4<OUT, SPEC == SPEC_HALF>(p.out[d], at.oa, at.ob, o, at.full);
        }
    }
}

// ---- max-curve: conversion/depth_to_max_curve.h:55-153 ---------------------
template <int OUT, int SPEC>
__device__ __forceinline__ void max_curve_row(const kparams& p, const wrow<FEAT_MAX_CURVE>& up,
                                              const wrow<FEAT_MAX_CURVE>& mid,
                                              const wrow<FEAT_MAX_CURVE>& dn,
                                              const rtabs<FEAT_MAX_CURVE>& tb, const pos& at) {
    f2 su[6], sm[6], sd[6];
#pragma unroll
    for (int i = 0; i < 6; ++i) {
        su[i] = mul2(up.r[i], up.r[i]);
        sm[i] = mul2(mid.r[i], mid.r[i]);
        sd[i] = mul2(dn.r[i], dn.r[i]);
    }
    f2 o[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int c  = i + 1;
        const f2  d0 = mid.r[c], bb = sm[c];
        // angle_formula (:55-82): 0 if any of the three depths is 0, else the
        // sum of the two bearing angles
        auto pair = [&](f2 dm, f2 smm, f2 dp, f2 spp, float c1, float c2) {
            const bres k1 = bearing_core2<false, OUT == OUT_U16>(d0, dm, bb, smm, c1);
            const bres k2 = bearing_core2<false, OUT == OUT_U16>(d0, dp, bb, spp, c2);
            const f2   s  = add2(acos_signed2<OUT == OUT_U16>(k1.q), acos_signed2<OUT == OUT_U16>(k2.q));
            return make_float2((k1.bc.x != 0.0f && k2.bc.x != 0.0f) ? s.x : 0.0f,
                               (k1.bc.y != 0.0f && k2.bc.y != 0.0f) ? s.y : 0.0f);
        };
        const f2 ah = pair(mid.r[c - 1], sm[c - 1], mid.r[c + 1], sm[c + 1], tb.cH[i], tb.cH[i + 1]);
        const f2 av = pair(up.r[c], su[c], dn.r[c], sd[c], tb.cV0[i], tb.cV1[i]);
        const f2 ad = pair(up.r[c - 1], su[c - 1], dn.r[c + 1], sd[c + 1], tb.cD0[i], tb.cD1[i + 1]);
        const f2 aa = pair(dn.r[c - 1], sd[c - 1], up.r[c + 1], su[c + 1], tb.cD1[i + 1], tb.cA[i + 1]);
        f2       a  = make_float2(fmaxf(fmaxf(ah.x, av.x), fmaxf(ad.x, aa.x)),
                                  fmaxf(fmaxf(ah.y, av.y), fmaxf(ad.y, aa.y)));
        a           = mul2(a, B(at.inside[i]));  // border stays 0 (:101-102)
        o[i]        = OUT == OUT_U16 ? mul2(a, B(p.q_max_curve)) : a;
    }
    if (at.live)
        store4<OUT, SPEC == SPEC_HALF>(p.out[0], at.oa, at.ob, o, at.full);
    if (p.out[1] != nullptr && at.live) {
        // the fused `range` output (converter_laserscan.h.inl:9-10):
        // convertTo(CV_16U) = saturate(round half even); range >= 0
        f2 r[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            r[i] = mid.r[i + 1];
            if constexpr (OUT == OUT_U16)
                r[i] = make_float2(fminf(r[i].x, 65535.0f), fminf(r[i].y, 65535.0f));
        }
        store4<OUT, SPEC == SPEC_HALF>(p.out[1], at.oa, at.ob, r, at.full);
    }
}

// ---- curvature: depth_to_curvature.h:97-206, math/derivatives.h:66-93,
//      math/curvature.h:18-44 ------------------------------------------------
struct curv_consts {
    float q_div, q_rcp;  // hi - lo and its refined reciprocal (u16 output)
};

template <int OUT, int SPEC>
__device__ __forceinline__ void curvature_row(const kparams& p, const wrow<FEAT_CURVATURE>& up,
                                              const wrow<FEAT_CURVATURE>& mid,
                                              const wrow<FEAT_CURVATURE>& dn,
                                              const rtabs<FEAT_CURVATURE>& tb, const pos& at,
                                              const curv_consts& cc) {
    const float phi[4] = {tb.phi.x, tb.phi.y, tb.phi.z, tb.phi.w};
    const float the[4] = {tb.theta.x, tb.theta.y, tb.theta.z, tb.theta.w};
    const float pt[4]  = {tb.pt.x, tb.pt.y, tb.pt.z, tb.pt.w};
    f2          g[4], m[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int c = i + 1;
        // camera-only divisors, shared by the two frames: scalar
        float b_u, b_v, b_uu, b_vv, b_uv, y_u, y_v, y_uu, y_vv, y_uv;
        if constexpr (SPEC == SPEC_UNIFORM) {
            // host-computed (kparams::cv_b / cv_y): uniform operands. div2 rounds
            // correctly with any reciprocal as good as the refined MUFU seed; the
            // host's is the correctly rounded one.
            b_u = p.cv_b[0], b_v = p.cv_b[1], b_uu = p.cv_b[2], b_vv = p.cv_b[3], b_uv = p.cv_b[4];
            y_u = p.cv_y[0], y_v = p.cv_y[1], y_uu = p.cv_y[2], y_vv = p.cv_y[3], y_uv = p.cv_y[4];
        } else {
            b_u = mul(2.0f, phi[i]), b_v = mul(2.0f, the[i]);
            b_uu = mul(phi[i], phi[i]), b_vv = mul(the[i], the[i]);
            b_uv = mul(pt[i], pt[i]);
            y_u = rcp_refined(b_u, mufu_rcp(b_u)), y_v = rcp_refined(b_v, mufu_rcp(b_v));
            y_uu = rcp_refined(b_uu, mufu_rcp(b_uu));
            y_vv = rcp_refined(b_vv, mufu_rcp(b_vv));
            y_uv = rcp_refined(b_uv, mufu_rcp(b_uv));
        }
        const f2    left = mid.r[c - 1], right = mid.r[c + 1], upv = up.r[c], down = dn.r[c];
        const f2    two_c = add2(mid.r[c], mid.r[c]);
        const f2    sh = sub2(add2(right, left), two_c);
        const f2    sv = sub2(add2(down, upv), two_c);
        const f2    f_u  = div2(sub2(right, left), B(b_u), B(y_u));
        const f2    f_v  = div2(sub2(down, upv), B(b_v), B(y_v));
        const f2    f_uu = div2(sh, B(b_uu), B(y_uu));
        const f2    f_vv = div2(sv, B(b_vv), B(y_vv));
        const f2    f_uv = div2(sv, B(b_uv), B(y_uv));  // sic: derivatives.h:83-90
        const f2    fu2 = mul2(f_u, f_u), fv2 = mul2(f_v, f_v);
        const f2    dd = add2(add2(B(1.0f), fu2), fv2);
        // DIFF_STAR (:97-115): any of the nine depths zero -> exactly 0
        const float mna = fminf(fminf(up.m3[i].x, mid.m3[i].x), dn.m3[i].x);
        const float mnb = fminf(fminf(up.m3[i].y, mid.m3[i].y), dn.m3[i].y);
        const bool  oka = mna != 0.0f, okb = mnb != 0.0f;
        f2          kg = B(0.0f), km = B(0.0f);
        if (wanted<SPEC>(p.out[0])) {
            // (mul2x: f_uu * f_vv can be -0 and n's sign is the result's, see below)
            const f2 n = sub2(mul2x(f_uu, f_vv), mul2(f_uv, f_uv));
            kg         = div2_fast(n, dd);
            // n can be -0 (f_uu < 0, f_vv = f_uv = +0); the residual step
            // would turn the quotient into +0. dd > 0: the zero keeps n's sign.
            kg = make_float2(n.x == 0.0f ? n.x : kg.x, n.y == 0.0f ? n.y : kg.y);
            kg         = make_float2(oka ? kg.x : 0.0f, okb ? kg.y : 0.0f);
        }
        if (wanted<SPEC>(p.out[1])) {
            const f2 t0  = mul2(add2(B(1.0f), fv2), f_uu);
            const f2 t1  = mul2(mul2(add2(f_u, f_u), f_v), f_uv);
            const f2 t2  = mul2(add2(B(1.0f), fu2), f_vv);
            const f2 num = add2(sub2(t0, t1), t2);
            f2       y;
            const f2 rt = sqrt2(dd, y);
            const f2 x  = add2(rt, rt);
            // pow(x, 3.f): x*x*x with an exact-product correction, rounded once
            // (glibc powf is correctly rounded in all but ~1e-4 of cases)
            const f2 x2h = mul2(x, x);
            const f2 x2l = fma2(x, x, neg2(x2h));
            const f2 ph  = mul2(x2h, x);
            const f2 pe  = fma2(x2h, x, neg2(ph));
            const f2 den = add2(ph, fma2(x2l, x, pe));
            km           = div2_fast(num, den);
            km           = make_float2(oka ? km.x : 0.0f, okb ? km.y : 0.0f);
        }
        if constexpr (OUT == OUT_U16) {
            // math/scaling.h:29-55 scale({lo,hi},{0,65535},v), truncated
            // (depth_to_curvature.h:239-243), zero where the depth is invalid
            // (:266-275). FADD.RZ with 2^23 truncates for 0 <= x < 2^23.
            const bool ca = mid.r[c].x != 0.0f, cb = mid.r[c].y != 0.0f;
            auto       q  = [&](f2 v) {
                const f2 a  = mul2(B(65535.0f), sub2(v, B(p.lo)));
                f2       s  = div2(a, B(cc.q_div), B(cc.q_rcp));
                s           = make_float2(fminf(fmaxf(s.x, 0.0f), 65535.0f),
                                          fminf(fmaxf(s.y, 0.0f), 65535.0f));
                const f2 t  = __fadd2_rz(s, B(8388608.0f));
                // store4 adds 1.5*2^23 and keeps the low 16 bits: hand it the
                // integer as an exactly representable float
                const f2 iv = sub2(t, B(8388608.0f));
                return make_float2(ca ? iv.x : 0.0f, cb ? iv.y : 0.0f);
            };
            g[i] = q(kg);
            m[i] = q(km);
        } else {
            g[i] = kg;
            m[i] = km;
        }
    }
    if (wanted<SPEC>(p.out[0]) && at.live)
        store4<OUT, SPEC == SPEC_HALF>(p.out[0], at.oa, at.ob, g, at.full);
    if (wanted<SPEC>(p.out[1]) && at.live)
        store4<OUT, SPEC == SPEC_HALF>(p.out[1], at.oa, at.ob, m, at.full);
}

// ---------------------------------------------------------------------------
#ifndef DC_MC_UNIFORM_BLOCKS
#define DC_MC_UNIFORM_BLOCKS 4
#endif
template <int FEAT, int SPEC = SPEC_GENERIC>
constexpr int min_blocks() {
#ifdef DC_MIN_BLOCKS
    return DC_MIN_BLOCKS;
#else
    if (FEAT == FEAT_MAX_CURVE && SPEC == SPEC_UNIFORM)
        return DC_MC_UNIFORM_BLOCKS;
    // curvature wants 168 registers (no spills); bearing fits 128 but measures 3 %
    // faster with 168 (0.605 -> 0.586 ms per 256 frames of 960x540)
    return FEAT == FEAT_MAX_CURVE ? 4 : 3;
#endif
}

// staging slots per stage: SPEC_UNIFORM reads no table
template <int FEAT, int SPEC>
__host__ __device__ constexpr int stage_slots() {
    return kRawSlots + (SPEC == SPEC_UNIFORM ? 0 : tab_slots<FEAT>::n);
}
// Staging stages: three (the copies of a row are issued two row steps before
// the row is consumed: curvature 0.478 -> 0.443 ms) except for max-curve, whose
// 11 slots per stage would not fit four CTAs per SM three times and which is
// register bound (the third stage pointer costs it spills).
template <int FEAT, int SPEC>
__host__ __device__ constexpr int stages() {
#ifdef DC_PAIR_STAGES
    return DC_PAIR_STAGES;
#else
    return FEAT == FEAT_MAX_CURVE ? 2 : 3;  // (max-curve measured faster with two)
#endif
}
template <int FEAT, int SPEC>
constexpr size_t pair_smem_bytes() {
    return size_t(stages<FEAT, SPEC>()) * stage_slots<FEAT, SPEC>() * kPairThreads * sizeof(uint4);
}

template <int FEAT, int IN, int OUT, int SPEC>
__global__ void __launch_bounds__(kPairThreads, min_blocks<FEAT, SPEC>())
    pair_kernel(const __grid_constant__ kparams p, int n_frames, int rows_per_strip) {
    const int       groups_x = (p.w + 3) >> 2;
    const int       strips   = (p.h + rows_per_strip - 1) / rows_per_strip;
    const long long gid      = (long long)blockIdx.x * kPairThreads + threadIdx.x;
    const int       gx       = int(gid % groups_x);
    const long long t        = gid / groups_x;
    // Frame pair varies faster than the strip: the CTAs resident at any moment
    // cover all frames of a few strips, so a table row is fetched from HBM once
    // per launch and then hit in L2 by every frame pair (with the strip fastest,
    // the 1 GB of streamed output evicted the 8-35 MB tables between pairs).
    const int       pairs    = (n_frames + 1) >> 1;
    const long long fp       = t % pairs;
    const int       strip    = int(t / pairs);
    if (strip >= strips)
        return;
    const long long fa = 2 * fp, fb = min(fa + 1, (long long)n_frames - 1);
    const int  u0 = gx * 4, w = p.w, h = p.h;
    const int  v0 = strip * rows_per_strip, v1 = min(v0 + rows_per_strip, h);
    constexpr bool R = feat_traits<FEAT>::right;
    constexpr bool HALF = SPEC == SPEC_HALF;
    constexpr int  FC = IN != IN_U16_DEPTH ? FAC_NONE
                        : (SPEC == SPEC_TABLES || SPEC == SPEC_UNIFORM) ? FAC_TABLE : FAC_RUNTIME;
    lane L;
    if constexpr (IN == IN_F32_RANGE) {
        L.Af  = static_cast<const float*>(p.in) + size_t(fa) * p.frame_px + u0;
        L.Bff = static_cast<const float*>(p.in) + size_t(fb) * p.frame_px + u0;
        L.A = L.Bf = nullptr;
    } else {
        L.A   = static_cast<const unsigned short*>(p.in) + size_t(fa) * p.frame_px + u0;
        L.Bf  = static_cast<const unsigned short*>(p.in) + size_t(fb) * p.frame_px + u0;
        L.Af = L.Bff = nullptr;
    }
    L.fac = (IN == IN_U16_DEPTH && p.range_factor.p != nullptr) ? p.range_factor.p + u0 : nullptr;
    L.hl  = u0 > 0;
    L.hr  = R && (u0 + 4 < w);
    L.full = !HALF || u0 + 4 <= w;
    // u16: the 4-byte word that holds column u0-1 starts at u0-2; f32: the pixel itself
    L.offl = L.hl ? (IN == IN_F32_RANGE ? -1 : -2) : 0;
    L.offr = L.hr ? 4 : 0;
    f32_window win;
    const float ufac = IN == IN_U16_DEPTH ? p.range_factor.value : 1.0f;

    curv_consts cc{1.f, 1.f};
    if constexpr (FEAT == FEAT_CURVATURE && OUT == OUT_U16) {
        cc.q_div = sub(p.hi, p.lo);
        cc.q_rcp = rcp_refined(cc.q_div, mufu_rcp(cc.q_div));
    }

    // ---- staging slots of this thread (dc_pair.cuh "Staging") ----
    extern __shared__ uint4 smem[];
    constexpr int NSLOT = stage_slots<FEAT, SPEC>();
    constexpr int T     = kPairThreads;
    constexpr int STAGES = stages<FEAT, SPEC>();
    uint4* const  stg0  = smem + threadIdx.x;
    uint4* const  stg1  = smem + NSLOT * T + threadIdx.x;
    uint4* const  stg2  = smem + 2 * NSLOT * T + threadIdx.x;  // (STAGES == 3 only)

    // A stage holds input row r together with the tables of row r-1 (what the
    // step that consumes row r as its "down" row needs). `ro` = r * w.
    const int  ro_last = (h - 1) * w;
    auto issue_stage = [&](int r, int ro, uint4* st) {
        const bool in = r < h;
        issue_raw<R, FC, IN, HALF>(L, in ? ro : ro_last, in, st, T);
        issue_tabs<FEAT, SPEC>(p, r - 1, ro - w + u0, u0, st + kRawSlots * T, T, L.full);
    };
    // the three window rows; they rotate by renaming in the three-stage loop
    wrow<FEAT> w0, w1, w2;
    issue_raw<R, FC, IN, HALF>(L, max(v0 - 1, 0) * w, v0 >= 1, stg0, T);
    issue_raw<R, FC, IN, HALF>(L, v0 * w, true, stg1, T);
    cp_async_commit();
    if constexpr (STAGES == 3) {
        issue_stage(v0 + 1, (v0 + 1) * w, stg2);
        cp_async_commit();
        cp_async_wait_but_one();
    } else {
        cp_async_wait_all();
    }
    commit<FEAT, IN>(fetch_raw<R, FC, IN>(L, ufac, stg0, T), w0, win);
    commit<FEAT, IN>(fetch_raw<R, FC, IN>(L, ufac, stg1, T), w1, win);
    if constexpr (STAGES == 3)
        issue_stage(v0 + 2, (v0 + 2) * w, stg0);
    else
        issue_stage(v0 + 1, (v0 + 1) * w, stg0);
    cp_async_commit();

    pos at;
    at.u0 = u0;
    at.full = L.full;
    at.live = true;
    at.oa = size_t(fa) * p.frame_px + size_t(v0) * w + u0;
    at.ob = size_t(fb) * p.frame_px + size_t(v0) * w + u0;
    float col_in[4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
        col_in[i] = (u0 + i >= 1 && u0 + i < w - 1) ? 1.0f : 0.0f;
    int  v       = v0;
    auto do_row = [&](const wrow<FEAT>& up, const wrow<FEAT>& mid, const wrow<FEAT>& dn,
                      const rtabs<FEAT>& tb) {
        at.v               = v;
        const float row_in = (v >= 1 && v < h - 1) ? 1.0f : 0.0f;
#pragma unroll
        for (int i = 0; i < 4; ++i)
            at.inside[i] = col_in[i] * row_in;
        if constexpr (FEAT == FEAT_BEARING)
            bearing_row<OUT, SPEC>(p, up, mid, dn, tb, at);
        else if constexpr (FEAT == FEAT_MAX_CURVE)
            max_curve_row<OUT, SPEC>(p, up, mid, dn, tb, at);
        else if constexpr (FEAT == FEAT_CURVATURE)
            curvature_row<OUT, SPEC>(p, up, mid, dn, tb, at, cc);
        at.oa += w;
        at.ob += w;
    };
    if constexpr (STAGES == 3) {
        // Step k = (v - v0) % 3 consumes row v+1 (and the tables of row v) from
        // stage SC and refills stage SF, consumed one step earlier, with row v+3.
        // Window rows and stages rotate with the same period, so the loop body is
        // instantiated three times with the names permuted (no register moves)
        // where that measured faster, see DC_PAIR_RENAME_MASK.
        int  ro   = (v0 + 3) * w;  // row offset of row v + 3
        auto step = [&](const wrow<FEAT>& up, const wrow<FEAT>& mid, wrow<FEAT>& dn, uint4* sc,
                        uint4* sf) {
            cp_async_wait_but_one();
            commit<FEAT, IN>(fetch_raw<R, FC, IN>(L, ufac, sc, T), dn, win);
            const rtabs<FEAT> tb = fetch_tabs<FEAT, SPEC>(p, sc + kRawSlots * T, T);
            issue_stage(v + 3, ro, sf);
            cp_async_commit();
            ro += w;
            do_row(up, mid, dn, tb);
            return ++v < v1;
        };
        if constexpr (((DC_PAIR_RENAME_MASK) >> FEAT) & 1) {
#if DC_SINGLE_EXIT
            do {
                at.live = v < v1;
                step(w0, w1, w2, stg2, stg1);
                at.live = v < v1;
                step(w1, w2, w0, stg0, stg2);
                at.live = v < v1;
                step(w2, w0, w1, stg1, stg0);
            } while (v < v1);
#else
            for (;;) {
                if (!step(w0, w1, w2, stg2, stg1))
                    break;
                if (!step(w1, w2, w0, stg0, stg2))
                    break;
                if (!step(w2, w0, w1, stg1, stg0))
                    break;
            }
#endif
        } else {
            uint4 *sc = stg2, *sf = stg1, *sn = stg0;
#pragma unroll 1
            for (;;) {
                const bool more = step(w0, w1, w2, sc, sf);
                uint4* const consumed = sc;
                sc = sn, sn = sf, sf = consumed;
                w0 = w1, w1 = w2;
                if (!more)
                    break;
            }
        }
    } else {
        int cur  = 0;
        int roff = (v0 + 1) * w;  // row offset of row v + 1
#pragma unroll 1
        for (; v < v1; ++v) {
            uint4* const s_now = cur ? stg1 : stg0;  // row v+1 and the tables of row v
            uint4* const s_nxt = cur ? stg0 : stg1;
            cur ^= 1;
            cp_async_wait_all();
            commit<FEAT, IN>(fetch_raw<R, FC, IN>(L, ufac, s_now, T), w2, win);
            const rtabs<FEAT> tb = fetch_tabs<FEAT, SPEC>(p, s_now + kRawSlots * T, T);
            const bool in2 = v + 2 < h;
            issue_raw<R, FC, IN, HALF>(L, in2 ? roff + w : roff - w, in2, s_nxt, T);
            issue_tabs<FEAT, SPEC>(p, v + 1, roff + u0, u0, s_nxt + kRawSlots * T, T, L.full);
            cp_async_commit();
            roff += w;
            do_row(w0, w1, w2, tb);
            w0 = w1;
            w1 = w2;
        }
    }
    cp_async_wait_all();
    if constexpr (IN == IN_F32_RANGE)
        win.raise(p.guard);
}

// ===========================================================================
// Max-curve with shared edges (conversion/depth_to_max_curve.h:55-153) for
// cameras whose four neighbour-angle tables are ONE scalar (the clamp of
// camera_models/utility.h:39-45: Kinect-like intrinsics from 1080p upward).
//
// Every pixel pair (X, Y) of the 8-neighbourhood graph is looked along from both
// ends: bearing_angle(X, Y, c) at pixel X and bearing_angle(Y, X, c) at pixel Y
// (math/triangles.h:27-64). With the same cosine c at both ends
//     rad = (XX + YY) - 2 (X Y) c,   den = sqrt(rad),   1 / den
// are bit-identical for the two calls (sums and products commute), so they are
// computed once per edge: 9 of the 22 FP32 operations of a bearing angle. Only
// nom = X - Y c, the quotient and the acos are per end. A step of the row march
// owns the edges inside row v (horizontal) and from row v down to row v+1
// (vertical, diagonal, anti-diagonal); the angles at the lower ends are carried
// to the next step, where they are the "prior" angles of row v+1. The strip's
// first step runs once without output to produce the carried angles.
// An invalid edge end (either depth zero, depth_to_max_curve.h:60-61) is carried
// as a large negative number: it loses every max, and max(., 0) restores the 0.
struct wrow_sq {
    f2 r[6];  // range of columns u0-1 .. u0+4, both frames
    f2 s[6];  // its square
    f2 c[6];  // and range * cos: with ONE cosine the subtrahend of nom = X - Y c is a
              // property of pixel Y, shared by the (up to eight) edge ends that look at Y
};
struct edge {
    f2 d, y;  // sqrt(rad), refined reciprocal of it
    f2 bc;    // X * Y: zero iff either depth is zero
};
template <bool RAD_FADD, bool LSB_DIV = false>
__device__ __forceinline__ edge edge_shared(f2 x, f2 y, f2 xx, f2 yy, float cu) {
    edge     e;
    const f2 sum = add2(xx, yy);
    e.bc         = mul2(x, y);
    f2 rad;
    if constexpr (RAD_FADD && DC_EDGE_RAD_FADD) {
        // 2 RN(bc cu) == RN(bc (2 cu)) exactly: the doubling moves into the scalar
        // (loop invariant with ONE cosine) and rad = sum - t2 is an FADD2 with two
        // register operands (2.07 clocks) instead of an FFMA2 with an immediate
        // multiplicand (3.7 clocks): 0.975 -> 0.968 ms at 4096x2160. With per-pixel
        // cosines the extra scalar add per edge makes it slower (1.056 -> 1.074 ms).
        const f2 t2 = mul2(e.bc, B(cu + cu));
        rad         = sub2(sum, t2);
    } else {
        const f2 t = mul2(e.bc, B(cu));
        rad        = fma2(B(-2.0f), t, sum);
    }
    f2       seed;
    e.d = sqrt2(rad, seed);
    e.y = (LSB_DIV && DC_U16_SKIP_REFINE) ? seed : rcp_refine2(e.d, seed);  // (dc_pair.cuh bearing_core2)
    return e;
}
#define DC_MC_INVALID (-1.0e30f)
// bearing angle at pixel x towards pixel y over the shared edge e; yc = y * cos
template <bool LSB = false>
__device__ __forceinline__ f2 edge_end(f2 x, f2 yc, const edge& e) {
    const f2 nom = sub2(x, yc);
    const f2 q   = div2(nom, e.d, e.y);
    const f2 a   = acos_signed2<LSB>(q);
    return make_float2(e.bc.x != 0.0f ? a.x : DC_MC_INVALID, e.bc.y != 0.0f ? a.y : DC_MC_INVALID);
}

// Register bound: with 128 registers (4 CTAs per SM) ptxas spills 130 bytes and
// the kernel is SLOWER than pair_kernel<MAX_CURVE> (1.25 vs 1.21 ms on 16 frames
// of 4096x2160), with 168 it is on par, with 255 (2 CTAs per SM, no spills)
// 1.06 ms: the 19 edge chains of a step give the scheduler enough independent
// work from two warps per sub-partition.
#ifndef DC_MCE_BLOCKS
#define DC_MCE_BLOCKS 2
#endif
constexpr int kMceStages = 3;
// 1: the twelve carried lower-end angles of a step live in shared memory (12
// conflict-free 8-byte words per thread) instead of registers, so that the kernel
// fits 168 registers / three CTAs per SM. profiles/r2_ab_experiments.md
#ifndef DC_MCE_CARRY_SMEM
#define DC_MCE_CARRY_SMEM 0
#endif
#ifndef DC_MCE_SINGLE
#define DC_MCE_SINGLE 1
#endif
// SPEC_TABLES: per-pixel cosine tables (960x540 Kinect). Per stage, next to the
// raw slots of input row r:
//   slot 4: V(r, u0..u0+3)   slot 5: D(r, u0..u0+3)   slot 6: H(r-1, u0..u0+3)
//   slot 7: A(r-1, u0..u0+3) slot 8: [D(r, u0+4) | H(r-1, u0+4) | A(r-1, u0+4) | -]
// (X(u, v) = cos(phi((u, v), prior pixel of direction X)), kparams::cosp). The
// step with rows v (mid) and v+1 = r (down) uses H and A of row v, V and D of row v+1.
template <int SPEC>
__host__ __device__ constexpr int mce_slots() {
    return kRawSlots + (SPEC == SPEC_TABLES ? 5 : 0);
}
template <int SPEC>
constexpr size_t mce_smem_bytes() {
    return size_t(kMceStages) * mce_slots<SPEC>() * kPairThreads * sizeof(uint4) +
           (DC_MCE_CARRY_SMEM ? size_t(12) * kPairThreads * sizeof(f2) : 0);
}
// cosines of one step; SPEC_UNIFORM: every entry is the one scalar
struct mce_cos {
    float h[5];  // horizontal edge j = (u0-1+j, v) - (u0+j, v):        H(u0+j, v)
    float v[4];  // vertical edge of column u0+i:                       V(u0+i, v+1)
    float d[5];  // diagonal edge (u0-1+c, v) - (u0+c, v+1):            D(u0+c, v+1)
    float a[4];  // anti-diagonal, seen from (u0+i, v+1) ("second"):    A(u0+i+1, v)
};

template <int IN, int OUT, int SPEC>
__global__ void __launch_bounds__(kPairThreads, DC_MCE_BLOCKS)
    maxcurve_edge_kernel(const __grid_constant__ kparams p, int n_frames, int rows_per_strip) {
    static_assert(SPEC == SPEC_UNIFORM || SPEC == SPEC_TABLES, "");
    constexpr bool TAB = SPEC == SPEC_TABLES;
    const int       groups_x = p.w >> 2;
    const int       strips   = (p.h + rows_per_strip - 1) / rows_per_strip;
    const long long gid      = (long long)blockIdx.x * kPairThreads + threadIdx.x;
    const int       gx       = int(gid % groups_x);
    const long long t        = gid / groups_x;
    const int       pairs    = (n_frames + 1) >> 1;  // pair fastest, see pair_kernel
    const long long fp       = t % pairs;
    const int       strip    = int(t / pairs);
    if (strip >= strips)
        return;
    const long long fa = 2 * fp, fb = min(fa + 1, (long long)n_frames - 1);
    const int  u0 = gx * 4, w = p.w, h = p.h;
    const int  v0 = strip * rows_per_strip, v1 = min(v0 + rows_per_strip, h);
    constexpr int FC = IN != IN_U16_DEPTH ? FAC_NONE : FAC_TABLE;
    lane L;
    if constexpr (IN == IN_F32_RANGE) {
        L.Af  = static_cast<const float*>(p.in) + size_t(fa) * p.frame_px + u0;
        L.Bff = static_cast<const float*>(p.in) + size_t(fb) * p.frame_px + u0;
        L.A = L.Bf = nullptr;
    } else {
        L.A   = static_cast<const unsigned short*>(p.in) + size_t(fa) * p.frame_px + u0;
        L.Bf  = static_cast<const unsigned short*>(p.in) + size_t(fb) * p.frame_px + u0;
        L.Af = L.Bff = nullptr;
    }
    L.fac = IN == IN_U16_DEPTH ? p.range_factor.p + u0 : nullptr;
    L.hl  = u0 > 0;
    L.hr  = u0 + 4 < w;
    L.full = true;
    L.offl = L.hl ? (IN == IN_F32_RANGE ? -1 : -2) : 0;
    L.offr = L.hr ? 4 : 0;
    f32_window win;
    const float cu = p.cosp[0].value;  // SPEC_UNIFORM: == cosp[1..3].value (checked by the launch code)

    // Three staging stages (dc_pair.cuh "Staging"): `sc` is consumed by this
    // step, `sn` is in flight for the next one, `sf` was consumed one step ago
    // and is refilled now: two rows of lead, no barrier.
    extern __shared__ uint4 smem[];
    constexpr int T  = kPairThreads;
    constexpr int NS = mce_slots<SPEC>();
    uint4*        sc = smem + threadIdx.x;
    uint4*        sn = smem + NS * T + threadIdx.x;
    uint4*        sf = smem + 2 * NS * T + threadIdx.x;
    const int     ro_last = (h - 1) * w;
    int           rn = v0 + 1, ro = (v0 + 1) * w;  // next row to fetch and its offset
    auto issue = [&](int r, int roff, uint4* st) {
        const bool in = r >= 0 && r < h;
        issue_raw<true, FC, IN>(L, in ? roff : ro_last, in, st, T);
        if constexpr (TAB) {
            // tables of row r (V, D) and of row r-1 (H, A); entries of pixels whose
            // prior neighbour lies outside the image are never used unmasked
            const bool in1 = r >= 1 && r <= h, more = u0 + 4 < w;
            const int  o0 = (in ? roff : ro_last) + u0, o1 = (in1 ? roff - w : 0) + u0;
            const int  o4 = more ? 4 : 0;
            cp_async<16>(st + 4 * T, p.cosp[1].p + o0, in);
            cp_async<16>(st + 5 * T, p.cosp[2].p + o0, in);
            cp_async<16>(st + 6 * T, p.cosp[0].p + o1, in1);
            cp_async<16>(st + 7 * T, p.cosp[3].p + o1, in1);
            char* e = reinterpret_cast<char*>(st + 8 * T);
            cp_async<4>(e, p.cosp[2].p + o0 + o4, in && more);
            cp_async<4>(e + 4, p.cosp[0].p + o1 + o4, in1 && more);
            cp_async<4>(e + 8, p.cosp[3].p + o1 + o4, in1 && more);
        }
        cp_async_commit();
    };
    auto consume = [&](wrow_sq& dst, mce_cos& cs) {
        cp_async_wait_but_one();
        to_range<IN, true>(fetch_raw<true, FC, IN>(L, 1.0f, sc, T), dst.r, win);
#pragma unroll
        for (int i = 0; i < 6; ++i) {
            dst.s[i] = mul2(dst.r[i], dst.r[i]);
            if constexpr (!TAB)
                dst.c[i] = mul2(dst.r[i], B(cu));
        }
        if constexpr (TAB) {
            const float4 tv = *reinterpret_cast<const float4*>(sc + 4 * T);
            const float4 td = *reinterpret_cast<const float4*>(sc + 5 * T);
            const float4 th = *reinterpret_cast<const float4*>(sc + 6 * T);
            const float4 ta = *reinterpret_cast<const float4*>(sc + 7 * T);
            const float4 te = *reinterpret_cast<const float4*>(sc + 8 * T);
            cs.v[0] = tv.x, cs.v[1] = tv.y, cs.v[2] = tv.z, cs.v[3] = tv.w;
            cs.d[0] = td.x, cs.d[1] = td.y, cs.d[2] = td.z, cs.d[3] = td.w, cs.d[4] = te.x;
            cs.h[0] = th.x, cs.h[1] = th.y, cs.h[2] = th.z, cs.h[3] = th.w, cs.h[4] = te.y;
            cs.a[0] = ta.y, cs.a[1] = ta.z, cs.a[2] = ta.w, cs.a[3] = te.z;
        }
        issue(rn, ro, sf);
        ++rn;
        ro += w;
        uint4* const done = sc;
        sc = sn, sn = sf, sf = done;
    };
    issue(v0 - 1, (v0 - 1) * w, sc);
    issue(v0, v0 * w, sn);

    // carried: angles at row v towards row v-1 (ver 0..3, dia 4..7, anti-second 8..11)
#if DC_MCE_CARRY_SMEM
    f2* const carry = reinterpret_cast<f2*>(smem + kMceStages * NS * T) + threadIdx.x;
    auto cget = [&](int k) { return carry[k * T]; };
    auto cset = [&](int k, f2 x) { carry[k * T] = x; };
#else
    f2   cr[12];
    auto cget = [&](int k) { return cr[k]; };
    auto cset = [&](int k, f2 x) { cr[k] = x; };
#endif
    // One step: edges of row `mid` and from `mid` down to `dn`. STORE = false is
    // the strip prologue (mid = row v0-1): only the carried angles are produced.
    pos at;
    at.u0 = u0;
    at.oa = size_t(fa) * p.frame_px + size_t(v0) * w + u0;
    at.ob = size_t(fb) * p.frame_px + size_t(v0) * w + u0;
    float col_in[4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
        col_in[i] = (u0 + i >= 1 && u0 + i < w - 1) ? 1.0f : 0.0f;
    int  v    = v0;
    auto step = [&](const wrow_sq& mid, const wrow_sq& dn, const mce_cos& cs, auto store_tag) {
        constexpr bool STORE = decltype(store_tag)::value;
        // y * cos for the end that looks at y: shared per pixel with one cosine
        auto yc = [&](const wrow_sq& row, int i, float c) {
            if constexpr (TAB)
                return mul2(row.r[i], B(c));
            else
                return row.c[i];
        };
        auto cs_ = [&](float c) { return TAB ? c : cu; };
        // Direction by direction with a running maximum, the carried angles
        // replaced in place: few values are live at a time (the kernel is
        // register bound: 128 registers spill 600 bytes, 255 none).
        f2 m[4];
        if constexpr (STORE) {
            // horizontal edges (j, v) - (j+1, v), j = u0-1 .. u0+3
            f2 hb = B(0.0f);
#pragma unroll
            for (int j = 0; j < 5; ++j) {
                const float c = cs_(cs.h[j]);
                const edge  e = edge_shared<!TAB, OUT == OUT_U16>(mid.r[j], mid.r[j + 1], mid.s[j], mid.s[j + 1], c);
                if (j >= 1)  // pixel j looks right; angle_formula (:55-82): back + forward
                    m[j - 1] = add2(hb, edge_end<OUT == OUT_U16>(mid.r[j], yc(mid, j + 1, c), e));
                if (j <= 3)  // pixel j+1 looks left
                    hb = edge_end<OUT == OUT_U16>(mid.r[j + 1], yc(mid, j, c), e);
            }
        }
        auto take = [&](int i, f2 back, f2 fwd) {
            const f2 a = add2(back, fwd);
            m[i]       = make_float2(fmaxf(m[i].x, a.x), fmaxf(m[i].y, a.y));
        };
        // vertical edges (c, v) - (c, v+1), columns u0 .. u0+3
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int   c  = i + 1;
            const float co = cs_(cs.v[i]);
            const edge  e  = edge_shared<!TAB, OUT == OUT_U16>(mid.r[c], dn.r[c], mid.s[c], dn.s[c], co);
            if constexpr (STORE)
                take(i, cget(i), edge_end<OUT == OUT_U16>(mid.r[c], yc(dn, c, co), e));
            cset(i, edge_end<OUT == OUT_U16>(dn.r[c], yc(mid, c, co), e));
        }
        // diagonal edges (c, v) - (c+1, v+1), c = u0-1 .. u0+3: pixel c looks
        // down-right (forward), pixel c+1 of the next row up-left (carried)
        {
            f2 pending = B(0.0f);  // the new bd[c-1], stored once the old one is used
#pragma unroll
            for (int c = 0; c < 5; ++c) {
                const float co = cs_(cs.d[c]);
                const edge  e  = edge_shared<!TAB, OUT == OUT_U16>(mid.r[c], dn.r[c + 1], mid.s[c], dn.s[c + 1], co);
                if (c >= 1) {
                    if constexpr (STORE)
                        take(c - 1, cget(4 + c - 1), edge_end<OUT == OUT_U16>(mid.r[c], yc(dn, c + 1, co), e));
                    cset(4 + c - 1, pending);
                }
                if (c <= 3)
                    pending = edge_end<OUT == OUT_U16>(dn.r[c + 1], yc(mid, c, co), e);
            }
        }
        // anti-diagonal edges (c, v) - (c-1, v+1), c = u0 .. u0+4: pixel c looks
        // down-left ("first": with the cosine of ITS diagonal successor, sic :136),
        // pixel c-1 of the next row up-right ("second", A(c, v)). Two cosines with
        // per-pixel tables: nothing to share there.
#pragma unroll
        for (int c = 1; c < 6; ++c) {
            if constexpr (!TAB) {
                const edge e = edge_shared<!TAB, OUT == OUT_U16>(mid.r[c], dn.r[c - 1], mid.s[c], dn.s[c - 1], cu);
                if constexpr (STORE)
                    if (c <= 4)
                        take(c - 1, cget(8 + c - 1), edge_end<OUT == OUT_U16>(mid.r[c], dn.c[c - 1], e));
                if (c >= 2)
                    cset(8 + c - 2, edge_end<OUT == OUT_U16>(dn.r[c - 1], mid.c[c], e));
            } else {
                if constexpr (STORE)
                    if (c <= 4) {
                        const float co = cs.d[c];  // D(u + 1, v + 1), u = u0 + c - 1
                        const edge  e  = edge_shared<!TAB, OUT == OUT_U16>(mid.r[c], dn.r[c - 1], mid.s[c], dn.s[c - 1], co);
                        take(c - 1, cget(8 + c - 1), edge_end<OUT == OUT_U16>(mid.r[c], mul2(dn.r[c - 1], B(co)), e));
                    }
                if (c >= 2) {
                    const float co = cs.a[c - 2];  // A(u0 + c - 1, v)
                    const edge  e  = edge_shared<!TAB, OUT == OUT_U16>(dn.r[c - 1], mid.r[c], dn.s[c - 1], mid.s[c], co);
                    cset(8 + c - 2, edge_end<OUT == OUT_U16>(dn.r[c - 1], mul2(mid.r[c], B(co)), e));
                }
            }
        }
        if constexpr (STORE) {
            const bool  live   = !DC_SINGLE_EXIT || v < v1;
            const float row_in = (v >= 1 && v < h - 1) ? 1.0f : 0.0f;
            f2          o[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                // maximum over the directions (:141-149); invalid ones lose to 0
                f2 a = make_float2(fmaxf(m[i].x, 0.0f), fmaxf(m[i].y, 0.0f));
                a    = mul2(a, B(col_in[i] * row_in));  // border stays 0 (:101-102)
                o[i] = OUT == OUT_U16 ? mul2(a, B(p.q_max_curve)) : a;
            }
            if (live)
                store4<OUT>(p.out[0], at.oa, at.ob, o);
            if (p.out[1] != nullptr && live) {
                // the fused `range` output (converter_laserscan.h.inl:9-10)
                f2 r[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    r[i] = mid.r[i + 1];
                    if constexpr (OUT == OUT_U16)
                        r[i] = make_float2(fminf(r[i].x, 65535.0f), fminf(r[i].y, 65535.0f));
                }
                store4<OUT>(p.out[1], at.oa, at.ob, r);
            }
            at.oa += w;
            at.ob += w;
        }
    };

    // rows alternate between wa and wb (period 2: the loop body is instantiated
    // twice with the names swapped)
    wrow_sq wa, wb;
    mce_cos cs;
    consume(wa, cs);  // row v0-1
    consume(wb, cs);  // row v0 (+ the tables of the step rows v0-1 / v0)
    step(wa, wb, cs, std::false_type{});
#if DC_MCE_SINGLE
    // one instantiation of the step (half the code: the kernel shows instruction-
    // fetch stalls), the window row handed on by 36 register moves
#pragma unroll 1
    for (;;) {
        wa = wb;
        consume(wb, cs);  // row v+1
        step(wa, wb, cs, std::true_type{});
        if (++v >= v1)
            break;
    }
#elif DC_SINGLE_EXIT
    do {
        consume(wa, cs);  // row v+1
        step(wb, wa, cs, std::true_type{});
        ++v;
        consume(wb, cs);
        step(wa, wb, cs, std::true_type{});
        ++v;
    } while (v < v1);
#else
    for (;;) {
        consume(wa, cs);  // row v+1
        step(wb, wa, cs, std::true_type{});
        if (++v >= v1)
            break;
        consume(wb, cs);
        step(wa, wb, cs, std::true_type{});
        if (++v >= v1)
            break;
    }
#endif
    cp_async_wait_all();
    if constexpr (IN == IN_F32_RANGE)
        win.raise(p.guard);
}

// ===========================================================================
// Flexion: conversion/depth_to_flexion.h:79-167. Same frame-pair marching, but a
// thread owns 2 columns: the window holds the camera-frame points P = d * ray
// (3 rows x 4 columns x 3 components x 2 frames = 72 registers), computed once
// per position. Rows rotate by renaming (the step is instantiated three times).
struct v3 {
    f2 x, y, z;
};
__device__ __forceinline__ v3 sub3(const v3& a, const v3& b) {
    return {sub2(a.x, b.x), sub2(a.y, b.y), sub2(a.z, b.z)};
}
// Eigen 3.3.7 Core/Redux.h:91-105: c0 + (c1 + c2)
__device__ __forceinline__ f2 dot3(const v3& a, const v3& b) {
    return add2(mul2(a.x, b.x), add2(mul2(a.y, b.y), mul2(a.z, b.z)));
}
// Eigen 3.3.7 Core/Dot.h:124-134 normalized(): n / sqrt(squaredNorm), the null
// vector unchanged. Guard-free as in dc_math.cuh normalized3<true>: z clamped to
// 2^-100 makes the three quotients of the null vector 0 = the input.
__device__ __forceinline__ v3 normalized3(const v3& n) {
    f2 z = dot3(n, n);
    z    = make_float2(fmaxf(z.x, 7.888609052210118e-31f), fmaxf(z.y, 7.888609052210118e-31f));
    f2       y;
    const f2 s = sqrt2(z, y);
    y          = rcp_refine2(s, y);
    return {div2(n.x, s, y), div2(n.y, s, y), div2(n.z, s, y)};
}
// Eigen 3.3.7 Geometry/OrthoMethods.h:34-46
__device__ __forceinline__ v3 cross3(const v3& a, const v3& b) {
    return {sub2(mul2(a.y, b.z), mul2(a.z, b.y)), sub2(mul2(a.z, b.x), mul2(a.x, b.z)),
            sub2(mul2(a.x, b.y), mul2(a.y, b.x))};
}

// ---- flexion for 16-bit output, within the +-1 LSB contract (BASELINE.json north_star:
// "8/16-bit quantised output images within +-1 LSB") instead of bit for bit --------------
// One LSB of convert_flexion<ushort> is 1 / 65535 = 1.5e-5 of a value in [0, 1]. The
// exact path spends 11 of the 22 packed operations of every normalisation on a correctly
// rounded sqrt and three correctly rounded divisions; what the 16-bit image needs is the
// value to 1e-5. Differences of points, dot and cross products keep the reference's
// operations and order; only the division by the norm changes:
//   mode 1: n / |n| as n * rsqrt(n.n)            (MUFU.RSQ: 2^-22 relative)
//   mode 2: the four divisions moved behind the products,
//             dot(v0 x v1, v2 x v3) * rsqrt(|v0|^2 |v1|^2) * rsqrt(|v2|^2 |v3|^2)
//           (range values lie in [2^-10, 2^22]: the products stay inside [2e-22, 4e28]).
// Error against the exact path: each cross-product component carries <= 2e-7 of
// |a||b|, the dot product of two of them <= 1.5e-6, the two scale factors <= 6e-7:
// together < 3e-6 = 0.2 LSB, so the rounded images differ by at most 1 (where the exact
// value sits within 0.2 LSB of a rounding boundary). A null vector (both points at the
// origin) gives 0 as in the reference: 0 * rsqrt(2^-100) = 0.
// DC_EXACT_U16=1 in the environment selects the bit-exact kernel for 16-bit output too.
#ifndef DC_FLEX_LSB_MODE
#define DC_FLEX_LSB_MODE 1
#endif
__device__ __forceinline__ f2 clamp_tiny2(f2 z) {
    return make_float2(fmaxf(z.x, 7.888609052210118e-31f), fmaxf(z.y, 7.888609052210118e-31f));
}
__device__ __forceinline__ v3 normalized3_lsb(const v3& n) {
    const f2 y = rsq2(clamp_tiny2(dot3(n, n)));
    return {mul2(n.x, y), mul2(n.y, y), mul2(n.z, y)};
}
// |dot(n0 x n1, n2 x n3)| before the clamp to [0, 1], from the four unnormalised vectors
template <bool LSB>
__device__ __forceinline__ f2 flexion_value(const v3& v0, const v3& v1, const v3& v2, const v3& v3_) {
    if constexpr (!LSB) {
        return dot3(cross3(normalized3(v0), normalized3(v1)), cross3(normalized3(v2), normalized3(v3_)));
    } else if constexpr (DC_FLEX_LSB_MODE == 1) {
        return dot3(cross3(normalized3_lsb(v0), normalized3_lsb(v1)),
                    cross3(normalized3_lsb(v2), normalized3_lsb(v3_)));
    } else {
        const f2 s01 = clamp_tiny2(mul2(dot3(v0, v0), dot3(v1, v1)));
        const f2 s23 = clamp_tiny2(mul2(dot3(v2, v2), dot3(v3_, v3_)));
        const f2 x   = dot3(cross3(v0, v1), cross3(v2, v3_));
        return mul2(mul2(x, rsq2(s01)), rsq2(s23));
    }
}

constexpr int kFlexSlots = 6;
// slot 0: [A: 2 px | B: 2 px | A word with u0-1 (high half) | B word with u0-1]
// slot 1: [A word with u0+2 (low half) | B word | factor(u0-1) | factor(u0+2)]
// slot 2: [factor(u0), factor(u0+1) | ray_x(u0-1) | ray_x(u0+2)]
// slot 3: [ray_x(u0), ray_x(u0+1) | ray_y(u0), ray_y(u0+1)]
// slot 4: [ray_z(u0), ray_z(u0+1) | ray_y(u0-1) | ray_y(u0+2)]
// slot 5: [ray_z(u0-1) | ray_z(u0+2) | - | -]
// f32 range input (no factor): slot 0: [A(u0), A(u0+1) | B(u0), B(u0+1)],
// slot 1: [A(u0-1) | B(u0-1) | A(u0+2) | B(u0+2)]; the ray slots as above.
struct flane {
    const unsigned short *A, *Bf;
    const float *         Af, *Bff;  // f32 input
    const float *         fac, *rx, *ry, *rz;  // tables at column u0 (fac may be nullptr)
    bool                  hl, hr;
};

template <int IN>
__device__ __forceinline__ void flex_issue(const flane& L, int roff, bool in, uint4* st, int T) {
    const bool l = in && L.hl, r = in && L.hr;
    const int  ol = L.hl ? -2 : 0, orr = L.hr ? 2 : 0, ol1 = L.hl ? -1 : 0;
    char *     s0 = reinterpret_cast<char*>(st), *s1 = reinterpret_cast<char*>(st + T),
         *s2 = reinterpret_cast<char*>(st + 2 * T), *s3 = reinterpret_cast<char*>(st + 3 * T),
         *s4 = reinterpret_cast<char*>(st + 4 * T), *s5 = reinterpret_cast<char*>(st + 5 * T);
    if constexpr (IN == IN_F32_RANGE) {
        const float *a = L.Af + roff, *b = L.Bff + roff;
        cp_async<8>(s0, a, in);
        cp_async<8>(s0 + 8, b, in);
        cp_async<4>(s1, a + ol1, l);
        cp_async<4>(s1 + 4, b + ol1, l);
        cp_async<4>(s1 + 8, a + orr, r);
        cp_async<4>(s1 + 12, b + orr, r);
    } else {
        const unsigned short *a = L.A + roff, *b = L.Bf + roff;
        cp_async<4>(s0, a, in);
        cp_async<4>(s0 + 4, b, in);
        cp_async<4>(s0 + 8, a + ol, l);
        cp_async<4>(s0 + 12, b + ol, l);
        cp_async<4>(s1, a + orr, r);
        cp_async<4>(s1 + 4, b + orr, r);
    }
    if (IN == IN_U16_DEPTH && L.fac != nullptr) {
        const float* f = L.fac + roff;
        cp_async<4>(s1 + 8, f + ol1, l);
        cp_async<4>(s1 + 12, f + orr, r);
        cp_async<8>(s2, f, in);
    }
    const float *x = L.rx + roff, *y = L.ry + roff, *z = L.rz + roff;
    cp_async<4>(s2 + 8, x + ol1, l);
    cp_async<4>(s2 + 12, x + orr, r);
    cp_async<8>(s3, x, in);
    cp_async<8>(s3 + 8, y, in);
    cp_async<8>(s4, z, in);
    cp_async<4>(s4 + 8, y + ol1, l);
    cp_async<4>(s4 + 12, y + orr, r);
    cp_async<4>(s5, z + ol1, l);
    cp_async<4>(s5 + 4, z + orr, r);
}

// to_camera (:70-76): P = d * ray for columns u0-1 .. u0+2 of both frames
template <int IN>
__device__ __forceinline__ void flex_points(const flane& L, float ufac, const uint4* st, int T,
                                            v3 (&P)[4], f32_window& win) {
    const uint4  q0 = st[0], q1 = st[T];
    const float4 q2 = *reinterpret_cast<const float4*>(st + 2 * T);
    const float4 q3 = *reinterpret_cast<const float4*>(st + 3 * T);
    const float4 q4 = *reinterpret_cast<const float4*>(st + 4 * T);
    const float4 q5 = *reinterpret_cast<const float4*>(st + 5 * T);
    f2           d[4];
    if constexpr (IN == IN_F32_RANGE) {
        d[0] = make_float2(__uint_as_float(q1.x), __uint_as_float(q1.y));
        d[1] = make_float2(__uint_as_float(q0.x), __uint_as_float(q0.z));
        d[2] = make_float2(__uint_as_float(q0.y), __uint_as_float(q0.w));
        d[3] = make_float2(__uint_as_float(q1.z), __uint_as_float(q1.w));
#pragma unroll
        for (int i = 0; i < 4; ++i)
            win.see(d[i].x, d[i].y);
    } else {
        d[0] = make_float2(magic_hi(q0.z), magic_hi(q0.w));
        d[1] = make_float2(magic_lo(q0.x), magic_lo(q0.y));
        d[2] = make_float2(magic_hi(q0.x), magic_hi(q0.y));
        d[3] = make_float2(magic_lo(q1.x), magic_lo(q1.y));
    }
    const bool  tab   = IN == IN_U16_DEPTH && L.fac != nullptr;
    const float fc[4] = {tab ? __uint_as_float(q1.z) : ufac, tab ? q2.x : ufac, tab ? q2.y : ufac,
                         tab ? __uint_as_float(q1.w) : ufac};
    const float rx[4] = {q2.z, q3.x, q3.y, q2.w};
    const float ry[4] = {q4.z, q3.z, q3.w, q4.w};
    const float rz[4] = {q5.x, q4.x, q4.y, q5.y};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        f2 r = IN == IN_F32_RANGE ? d[i] : add2(d[i], B(-8388608.0f));
        if constexpr (IN == IN_U16_DEPTH)
            r = mul2(r, B(fc[i]));
        P[i] = {mul2(r, B(rx[i])), mul2(r, B(ry[i])), mul2(r, B(rz[i]))};
    }
}

// ---- conflict-free staging layout of the 16-bit kernel (DC_FLEX_STAGE_WORDS=1) ----------
// The slot layout above puts the 16 bytes of a thread's slot side by side: a 4- or 8-byte
// cp.async of a warp then lands 16 bytes apart, four lanes per bank (ncu, round 2: 60.6 M
// of the 79 M shared-memory wavefronts of the kernel's LDGSTS were bank conflicts, and the
// L1TEX data stage was the busiest unit of the SM at 62 %; with the 16-bit kernel's shorter
// arithmetic it became the bound). Here every staged item is its own array over the CTA:
//   4-byte item j of thread t:  word  [j * T + t]                      (j < kFlexWords)
//   8-byte item p of thread t:  dword [kFlexWords * T / 2 + p * T + t] (p < kFlexPairs)
// consecutive lanes hit consecutive banks for writes and read-backs alike. Same footprint
// (96 bytes per thread and stage); 18 LDS.32 / LDS.64 per row step instead of 6 LDS.128.
// u16 input: words 0 A(u0,u0+1) 1 B(..) 2 A-left word 3 B-left 4 A-right 5 B-right
//            6 fac(u0-1) 7 fac(u0+2); pair 0 fac(u0,u0+1)
// f32 input: words 0 A(u0-1) 1 B(u0-1) 2 A(u0+2) 3 B(u0+2); pairs 0 A(u0,u0+1) 4 B(u0,u0+1)
// both:      words 8 rx(u0-1) 9 rx(u0+2) 10 ry(u0-1) 11 ry(u0+2) 12 rz(u0-1) 13 rz(u0+2);
//            pairs 1 rx(u0,u0+1) 2 ry(..) 3 rz(..)
#ifndef DC_FLEX_STAGE_WORDS
#define DC_FLEX_STAGE_WORDS 1
#endif
constexpr int kFlexWords = 14, kFlexPairs = 5;
static_assert(kFlexWords * 4 + kFlexPairs * 8 == kFlexSlots * 16, "same footprint as the slot layout");
struct flex_stage {
    char* base;  // of the stage, for this thread: word j at base + j * 4 T, pair p at pairs + p * 8 T
    char* pairs;
    __device__ __forceinline__ char* word(int j, int T) const { return base + j * 4 * T; }
    __device__ __forceinline__ char* pair(int p, int T) const { return pairs + p * 8 * T; }
};
__device__ __forceinline__ flex_stage flex_stage_of(uint4* smem, int stage, int T, int t) {
    char* b = reinterpret_cast<char*>(smem) + size_t(stage) * kFlexSlots * 16 * T;
    return {b + 4 * t, b + kFlexWords * 4 * T + 8 * t};
}

template <int IN>
__device__ __forceinline__ void flex_issue(const flane& L, int roff, bool in, const flex_stage& st, int T) {
    const bool l = in && L.hl, r = in && L.hr;
    const int  ol = L.hl ? -2 : 0, orr = L.hr ? 2 : 0, ol1 = L.hl ? -1 : 0;
    if constexpr (IN == IN_F32_RANGE) {
        const float *a = L.Af + roff, *b = L.Bff + roff;
        cp_async<8>(st.pair(0, T), a, in);
        cp_async<8>(st.pair(4, T), b, in);
        cp_async<4>(st.word(0, T), a + ol1, l);
        cp_async<4>(st.word(1, T), b + ol1, l);
        cp_async<4>(st.word(2, T), a + orr, r);
        cp_async<4>(st.word(3, T), b + orr, r);
    } else {
        const unsigned short *a = L.A + roff, *b = L.Bf + roff;
        cp_async<4>(st.word(0, T), a, in);
        cp_async<4>(st.word(1, T), b, in);
        cp_async<4>(st.word(2, T), a + ol, l);
        cp_async<4>(st.word(3, T), b + ol, l);
        cp_async<4>(st.word(4, T), a + orr, r);
        cp_async<4>(st.word(5, T), b + orr, r);
    }
    if (IN == IN_U16_DEPTH && L.fac != nullptr) {
        const float* f = L.fac + roff;
        cp_async<4>(st.word(6, T), f + ol1, l);
        cp_async<4>(st.word(7, T), f + orr, r);
        cp_async<8>(st.pair(0, T), f, in);
    }
    const float *x = L.rx + roff, *y = L.ry + roff, *z = L.rz + roff;
    cp_async<4>(st.word(8, T), x + ol1, l);
    cp_async<4>(st.word(9, T), x + orr, r);
    cp_async<4>(st.word(10, T), y + ol1, l);
    cp_async<4>(st.word(11, T), y + orr, r);
    cp_async<4>(st.word(12, T), z + ol1, l);
    cp_async<4>(st.word(13, T), z + orr, r);
    cp_async<8>(st.pair(1, T), x, in);
    cp_async<8>(st.pair(2, T), y, in);
    cp_async<8>(st.pair(3, T), z, in);
}

template <int IN>
__device__ __forceinline__ void flex_points(const flane& L, float ufac, const flex_stage& st, int T,
                                            v3 (&P)[4], f32_window& win) {
    auto W  = [&](int j) { return *reinterpret_cast<const unsigned*>(st.word(j, T)); };
    auto Wf = [&](int j) { return *reinterpret_cast<const float*>(st.word(j, T)); };
    auto Pf = [&](int p) { return *reinterpret_cast<const float2*>(st.pair(p, T)); };
    f2 d[4];
    if constexpr (IN == IN_F32_RANGE) {
        const float2 a = Pf(0), b = Pf(4);
        d[0] = make_float2(Wf(0), Wf(1));
        d[1] = make_float2(a.x, b.x);
        d[2] = make_float2(a.y, b.y);
        d[3] = make_float2(Wf(2), Wf(3));
#pragma unroll
        for (int i = 0; i < 4; ++i)
            win.see(d[i].x, d[i].y);
    } else {
        const unsigned a = W(0), b = W(1);
        d[0] = make_float2(magic_hi(W(2)), magic_hi(W(3)));
        d[1] = make_float2(magic_lo(a), magic_lo(b));
        d[2] = make_float2(magic_hi(a), magic_hi(b));
        d[3] = make_float2(magic_lo(W(4)), magic_lo(W(5)));
    }
    const bool tab = IN == IN_U16_DEPTH && L.fac != nullptr;
    float      fc[4] = {ufac, ufac, ufac, ufac};
    if (tab) {
        const float2 f = Pf(0);
        fc[0] = Wf(6), fc[1] = f.x, fc[2] = f.y, fc[3] = Wf(7);
    }
    const float2 x = Pf(1), y = Pf(2), z = Pf(3);
    const float  rx[4] = {Wf(8), x.x, x.y, Wf(9)};
    const float  ry[4] = {Wf(10), y.x, y.y, Wf(11)};
    const float  rz[4] = {Wf(12), z.x, z.y, Wf(13)};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        f2 r = IN == IN_F32_RANGE ? d[i] : add2(d[i], B(-8388608.0f));
        if constexpr (IN == IN_U16_DEPTH)
            r = mul2(r, B(fc[i]));
        P[i] = {mul2(r, B(rx[i])), mul2(r, B(ry[i])), mul2(r, B(rz[i]))};
    }
}

#ifndef DC_FLEX_BLOCKS
#define DC_FLEX_BLOCKS 3
#endif
template <int IN, int OUT, bool LSB = false>
__global__ void __launch_bounds__(kPairThreads, DC_FLEX_BLOCKS)
    flexion_pair_kernel(const __grid_constant__ kparams p, int n_frames, int rows_per_strip) {
    const int       groups_x = p.w >> 1;
    const int       strips   = (p.h + rows_per_strip - 1) / rows_per_strip;
    const long long gid      = (long long)blockIdx.x * kPairThreads + threadIdx.x;
    const int       gx       = int(gid % groups_x);
    const long long t        = gid / groups_x;
    // Frame pair varies faster than the strip: the CTAs resident at any moment
    // cover all frames of a few strips, so a table row is fetched from HBM once
    // per launch and then hit in L2 by every frame pair (with the strip fastest,
    // the 1 GB of streamed output evicted the 8-35 MB tables between pairs).
    const int       pairs    = (n_frames + 1) >> 1;
    const long long fp       = t % pairs;
    const int       strip    = int(t / pairs);
    if (strip >= strips)
        return;
    const long long fa = 2 * fp, fb = min(fa + 1, (long long)n_frames - 1);
    const int u0 = gx * 2, w = p.w, h = p.h;
    const int v0 = strip * rows_per_strip, v1 = min(v0 + rows_per_strip, h);
    flane     L;
    if constexpr (IN == IN_F32_RANGE) {
        L.Af  = static_cast<const float*>(p.in) + size_t(fa) * p.frame_px + u0;
        L.Bff = static_cast<const float*>(p.in) + size_t(fb) * p.frame_px + u0;
        L.A = L.Bf = nullptr;
    } else {
        L.A   = static_cast<const unsigned short*>(p.in) + size_t(fa) * p.frame_px + u0;
        L.Bf  = static_cast<const unsigned short*>(p.in) + size_t(fb) * p.frame_px + u0;
        L.Af = L.Bff = nullptr;
    }
    f32_window win;
    L.fac = (IN == IN_U16_DEPTH && p.range_factor.p != nullptr) ? p.range_factor.p + u0 : nullptr;
    L.rx  = p.ray_x + u0;
    L.ry  = p.ray_y + u0;
    L.rz  = p.ray_z + u0;
    L.hl  = u0 > 0;
    L.hr  = u0 + 2 < w;
    const float ufac = IN == IN_U16_DEPTH ? p.range_factor.value : 1.0f;

    // Three staging stages: the copies of a row are issued THREE steps before the
    // row is consumed (two full row steps in flight). With one step of lead the
    // kernel (3 warps per sub-partition) spent 23 % of its stall samples waiting
    // for cp.async groups (ncu: long scoreboard at the wait).
    extern __shared__ uint4 smem[];
    constexpr int T    = kPairThreads;
    // the word layout for the 16-bit kernel (L1TEX bound: 0.537 -> 0.495 ms), the slot layout
    // for the exact ones (register-read / FMA bound: the twelve extra LDS cost them 1-4 %)
    auto stage = [&](int g) {
        if constexpr (LSB && DC_FLEX_STAGE_WORDS)
            return flex_stage_of(smem, g, T, threadIdx.x);
        else
            return smem + g * kFlexSlots * T + threadIdx.x;
    };
    const auto stg0 = stage(0), stg1 = stage(1), stg2 = stage(2);

    v3 w0[4], w1[4], w2[4];
    flex_issue<IN>(L, max(v0 - 1, 0) * w, v0 >= 1, stg0, T);
    flex_issue<IN>(L, v0 * w, true, stg1, T);
    cp_async_commit();
    flex_issue<IN>(L, min(v0 + 1, h - 1) * w, v0 + 1 < h, stg2, T);
    cp_async_commit();
    cp_async_wait_but_one();
    flex_points<IN>(L, ufac, stg0, T, w0, win);
    flex_points<IN>(L, ufac, stg1, T, w1, win);
    flex_issue<IN>(L, min(v0 + 2, h - 1) * w, v0 + 2 < h, stg0, T);
    cp_async_commit();

    using TO         = typename out_elem<OUT>::type;
    TO* const   out  = static_cast<TO*>(p.out[0]);
    size_t      oa   = size_t(fa) * p.frame_px + size_t(v0) * w + u0;
    size_t      ob   = size_t(fb) * p.frame_px + size_t(v0) * w + u0;
    const bool  cin0 = u0 >= 1, cin1 = u0 + 1 < w - 1;
    int         v = v0;
    // step k = (v - v0) % 3 consumes row v+1 from stage SC and refills stage SN
    // (consumed in the previous step) with row v+3
#define DC_FLEX_STEP(UP, MID, DN, SC, SN)                                                      \
    {                                                                                          \
        cp_async_wait_but_one();                                                               \
        flex_points<IN>(L, ufac, SC, T, DN, win);                                                   \
        const bool in3 = v + 3 < h;                                                            \
        flex_issue<IN>(L, (in3 ? v + 3 : h - 1) * w, in3, SN, T);                                  \
        cp_async_commit();                                                                     \
        const bool rin = v >= 1 && v < h - 1;                                                  \
        f2         f[2];                                                                       \
        _Pragma("unroll") for (int i = 0; i < 2; ++i) {                                        \
            const int c  = i + 1;                                                              \
            /* :105-127: (u,v-1)->(u,v+1), (u-1,v)->(u+1,v), (u+1,v-1)->(u-1,v+1),        */   \
            /*           (u-1,v-1)->(u+1,v+1)                                              */   \
            const f2 x  = flexion_value<LSB>(sub3(DN[c], UP[c]), sub3(MID[c + 1], MID[c - 1]),  \
                                             sub3(DN[c - 1], UP[c + 1]), sub3(DN[c + 1], UP[c - 1])); \
            const bool ok = rin && (i == 0 ? cin0 : cin1);                                     \
            f[i] = make_float2(ok ? fminf(fabsf(x.x), 1.0f) : 0.0f,                            \
                               ok ? fminf(fabsf(x.y), 1.0f) : 0.0f);                           \
        }                                                                                      \
        if (DC_SINGLE_EXIT && v >= v1) {                                                       \
            /* a step past the strip (whole groups of three run): no store */                  \
        } else if constexpr (OUT == OUT_U16) {                                                 \
            const f2 s0 = add2(mul2(f[0], B(65535.0f)), B(12582912.0f));                       \
            const f2 s1 = add2(mul2(f[1], B(65535.0f)), B(12582912.0f));                       \
            *reinterpret_cast<unsigned*>(out + oa) =                                           \
                __byte_perm(__float_as_uint(s0.x), __float_as_uint(s1.x), 0x5410);             \
            *reinterpret_cast<unsigned*>(out + ob) =                                           \
                __byte_perm(__float_as_uint(s0.y), __float_as_uint(s1.y), 0x5410);             \
        } else {                                                                               \
            *reinterpret_cast<float2*>(out + oa) = make_float2(f[0].x, f[1].x);                \
            *reinterpret_cast<float2*>(out + ob) = make_float2(f[0].y, f[1].y);                \
        }                                                                                      \
        oa += w;                                                                               \
        ob += w;                                                                               \
        ++v;                                                                                   \
        if (!DC_SINGLE_EXIT && v >= v1)                                                        \
            break;                                                                             \
    }
    do {
        DC_FLEX_STEP(w0, w1, w2, stg2, stg1)
        DC_FLEX_STEP(w1, w2, w0, stg0, stg2)
        DC_FLEX_STEP(w2, w0, w1, stg1, stg0)
    } while (v < v1);
#undef DC_FLEX_STEP
    cp_async_wait_all();
    if constexpr (IN == IN_F32_RANGE)
        win.raise(p.guard);
}

// ===========================================================================
// Streaming kernel for the two HBM-bound jobs: DC_RANGE (depth_to_laserscan.h:
// 54-93 + the u16 conversion of converter_laserscan.h.inl:9-10) and DC_SCALE
// (depth_scaling.h:21-31). A thread owns 8 consecutive pixels of the frame
// (one 16-byte load), keeps their range factors in registers and walks over the
// frames of the batch, four frames in flight, so the factor table is read once
// per thread instead of once per frame.
template <int FEAT, int IN, int OUT>
__global__ void __launch_bounds__(256)
    stream_kernel(const __grid_constant__ kparams p, int n_frames, int groups_per_frame) {
    const int g = blockIdx.x * 256 + threadIdx.x;
    if (g >= groups_per_frame)
        return;
    f2 fac[4];
    if constexpr (FEAT == FEAT_RANGE && IN == IN_U16_DEPTH) {
        if (p.range_factor.p != nullptr) {
            const float4 a = __ldg(reinterpret_cast<const float4*>(p.range_factor.p) + 2 * g);
            const float4 b = __ldg(reinterpret_cast<const float4*>(p.range_factor.p) + 2 * g + 1);
            fac[0] = make_float2(a.x, a.y), fac[1] = make_float2(a.z, a.w);
            fac[2] = make_float2(b.x, b.y), fac[3] = make_float2(b.z, b.w);
        } else {
            fac[0] = fac[1] = fac[2] = fac[3] = B(p.range_factor.value);
        }
    }
    const uint4* in = static_cast<const uint4*>(p.in) + g;
    const size_t fstride = size_t(groups_per_frame);  // uint4 per frame
    constexpr int K = 4;
    for (int f0 = blockIdx.y; f0 < n_frames; f0 += K * gridDim.y) {
        uint4 raw[K];
#pragma unroll
        for (int k = 0; k < K; ++k) {
            const int f = f0 + k * gridDim.y;
            if (f < n_frames)
                raw[k] = __ldcs(in + size_t(f) * fstride);
        }
#pragma unroll
        for (int k = 0; k < K; ++k) {
            const int f = f0 + k * gridDim.y;
            if (f >= n_frames)
                break;
            const unsigned wd[4] = {raw[k].x, raw[k].y, raw[k].z, raw[k].w};
            f2             v[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                v[i] = add2(make_float2(magic_lo(wd[i]), magic_hi(wd[i])), B(-8388608.0f));
                if constexpr (FEAT == FEAT_RANGE && IN == IN_U16_DEPTH)
                    v[i] = mul2(v[i], fac[i]);
                if constexpr (FEAT == FEAT_SCALE)
                    v[i] = add2(mul2(v[i], B(p.s_scale)), B(p.s_offset));
            }
            const size_t o = size_t(f) * fstride + g;
            if constexpr (OUT == OUT_U16) {
                unsigned q[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    if constexpr (FEAT == FEAT_SCALE) {
                        // any sign / magnitude: cvRound + saturate_cast<ushort>
                        q[i] = unsigned(sat_u16_rn(v[i].x)) | (unsigned(sat_u16_rn(v[i].y)) << 16);
                    } else {
                        // 0 <= range: clamp above, round half to even by the magic add
                        const f2 c = make_float2(fminf(v[i].x, 65535.0f), fminf(v[i].y, 65535.0f));
                        const f2 m = add2(c, B(12582912.0f));
                        q[i] = __byte_perm(__float_as_uint(m.x), __float_as_uint(m.y), 0x5410);
                    }
                }
                __stcs(static_cast<uint4*>(p.out[0]) + o, make_uint4(q[0], q[1], q[2], q[3]));
            } else {
                float4* dst = static_cast<float4*>(p.out[0]) + 2 * o;
                __stcs(dst, make_float4(v[0].x, v[0].y, v[1].x, v[1].y));
                __stcs(dst + 1, make_float4(v[2].x, v[2].y, v[3].x, v[3].y));
            }
        }
    }
}

}  // namespace pr
}  // namespace dc
#endif
