// Frame-pair, row-marching kernels for sm_100a: the fast path for u16 input.
//
// The exact (reference-order, IEEE-rounded) arithmetic makes every stencil
// kernel FP32-pipe bound on B200, not HBM bound. These
// kernels therefore minimise FP32-pipe cycles and issue slots per pixel:
//
//   * A thread converts the SAME pixels of TWO frames at once and keeps the two
//     values of a pixel in one 64-bit register pair, so that all arithmetic
//     runs as packed FFMA2 / FMUL2 / FADD2 (fma.rn.f32x2: one issue slot for
//     two IEEE-rounded results). Per-camera table values are identical for the
//     two frames and enter as broadcast scalar operands; vector loads and
//     stores stay per frame and fully coalesced.
//   * A thread owns 4 columns and marches down a strip of rows with the 3-row
//     window (range, squares) in registers: each input row is loaded, unpacked
//     and converted to range once instead of three times, and the column halo
//     is two extra scalar loads per row (L1 hits) instead of a shuffle.
//   * Compares, selects, clamps, exponent tricks and u16 packing are scalar
//     ALU-pipe instructions, which issue beside the FMA pipe.
//
// Results are bit-identical to the one-frame kernels of dc_kernels.cuh (same
// operation order; the packed instructions round each half like the scalar
// ones). Preconditions, checked by the host: u16 input, width % 4 == 0,
// 8-byte aligned frames, and the per-camera range proof `fast_ok`.
#ifndef DC_PAIR_CUH
#define DC_PAIR_CUH

#include "dc_kernels.cuh"

namespace dc {
namespace pr {

typedef float2 f2;

__device__ __forceinline__ f2 B(float s) { return make_float2(s, s); }
__device__ __forceinline__ f2 neg2(f2 a) { return make_float2(-a.x, -a.y); }
// ptxas 12.9 contracts mul.rn.f32x2 + add.rn.f32x2 into FFMA2 even under
// --fmad=false (the scalar .rn forms are never contracted), which breaks the
// reference's rounding order. A product is therefore written as an FMA with a
// zero addend, which cannot be fused into a following add:
//   mul2(a, b)  = fma(a, b, +0.0): the addend is the zero register (FFMA2 .., RZ),
//                 no third register read. RN(a*b + +0) == RN(a*b) for every a, b
//                 EXCEPT that a product of -0 comes out as +0. No result of the
//                 kernels depends on the sign of an intermediate zero (squares,
//                 products of ranges >= 0, terms of sums and of |.|), except the
//                 one product named below.
//   mul2x(a, b) = fma(a, b, -0.0) with the -0.0 read from constant memory (opaque to
//                 the compiler, which would fold a literal -0.0 away and then
//                 contract): exact for every a, b including signed zeros. Costs a
//                 third operand: a uniform register at best, an ordinary register in
//                 the unrolled row loops, where ptxas reloads the constant per
//                 iteration (one more register-file read per product; the register
//                 file's two reads per clock are what bounds these kernels,
//                 profiles/r2_pipebench_rf_b200.txt). Used where the sign of a zero product
//                 reaches an output: Gaussian curvature's f_uu * f_vv.
// DC_MUL_NEGZERO=1 restores the -0.0 form everywhere (A/B switch).
#ifndef DC_MUL_NEGZERO
#define DC_MUL_NEGZERO 0
#endif
static __constant__ float k_neg_zero = -0.0f;  // (one copy per translation unit)
__device__ __forceinline__ f2 mul2x(f2 a, f2 b) { return __ffma2_rn(a, b, make_float2(k_neg_zero, k_neg_zero)); }
__device__ __forceinline__ f2 mul2(f2 a, f2 b) {
#if DC_MUL_NEGZERO
    return mul2x(a, b);
#else
    return __ffma2_rn(a, b, make_float2(0.0f, 0.0f));
#endif
}
__device__ __forceinline__ f2 add2(f2 a, f2 b) { return __fadd2_rn(a, b); }
__device__ __forceinline__ f2 sub2(f2 a, f2 b) { return __fadd2_rn(a, neg2(b)); }
__device__ __forceinline__ f2 fma2(f2 a, f2 b, f2 c) { return __ffma2_rn(a, b, c); }
__device__ __forceinline__ f2 rsq2(f2 x) { return make_float2(mufu_rsq(x.x), mufu_rsq(x.y)); }
__device__ __forceinline__ f2 rcp2(f2 x) { return make_float2(mufu_rcp(x.x), mufu_rcp(x.y)); }
// x * 0.5 for normal x: exponent decrement on the ALU pipe
__device__ __forceinline__ float halve(float x) {
    return __uint_as_float(__float_as_uint(x) - 0x00800000u);
}
// Measured (profiles/r1_ab_experiments.md): one FMUL2 by 0.5 instead of the two
// integer adds makes bearing 4 % and curvature 2 % SLOWER although it removes an
// instruction: the row loops are bound by FMA-pipe slots, ALU-pipe work is hidden.
#ifndef DC_HALVE_ALU
#define DC_HALVE_ALU 1
#endif
__device__ __forceinline__ f2 halve2(f2 x) {
#if DC_HALVE_ALU
    return make_float2(halve(x.x), halve(x.y));
#else
    return mul2(x, B(0.5f));
#endif
}

#ifndef DC_ACOS_MUFU_SQRT
#define DC_ACOS_MUFU_SQRT 1
#endif
#ifndef DC_SQRT_MUFU_SEED
#define DC_SQRT_MUFU_SEED 0
#endif
#ifndef DC_APPROX_RATIO
#define DC_APPROX_RATIO 0
#endif
__device__ __forceinline__ float mufu_sqrt(float x) {
#if DC_ABL & 1
    return x;
#else
    float y;
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
#endif
}
// RN(sqrt(x)) per half (dc_fastmath.cuh sqrt_rn_fast), y = the rsqrt seed.
__device__ __forceinline__ f2 sqrt2(f2 x, f2& y) {
    y          = rsq2(x);
#if DC_SQRT_MUFU_SEED
    // first estimate from MUFU.SQRT instead of x * rsqrt(x): an XU-pipe slot per
    // value instead of an FMA-pipe slot per pair (exactness: tests/cuda/mathcheck.cu)
    f2       g = make_float2(mufu_sqrt(x.x), mufu_sqrt(x.y));
#else
    f2       g = mul2(x, y);
#endif
    const f2 h = halve2(y);
    const f2 r = fma2(neg2(g), g, x);
    return fma2(r, h, g);
}
// one Newton step on a reciprocal seed (dc_fastmath.cuh rcp_refined)
__device__ __forceinline__ f2 rcp_refine2(f2 b, f2 seed) {
    return fma2(seed, fma2(neg2(b), seed, B(1.0f)), seed);
}
// RN(a / b) given the refined reciprocal y of b (dc_fastmath.cuh div_rn_with_rcp)
__device__ __forceinline__ f2 div2(f2 a, f2 b, f2 y) {
    const f2 q = mul2(a, y);
    const f2 r = fma2(neg2(b), q, a);
    return fma2(y, r, q);
}
__device__ __forceinline__ f2 div2_fast(f2 a, f2 b) { return div2(a, b, rcp_refine2(b, rcp2(b))); }

// acos(min(|q|, 1 - delta)), relative error < 1.2e-6: sqrt(1 - a) * P5(a).
// The caller reflects for q < 0 (acos(q) = pi - acos(|q|)); together that is
// acos(clamp(q, -1 + delta, 1 - delta)) of math/triangles.h:55-58. A NaN q
// (only on lanes that are masked out afterwards) comes out finite.
// P(a) of acos(a) = sqrt(1 - a) P(a). P5: the float path (relative error < 1.2e-6).
// P4 (16-bit output only, DC_U16_ACOS_DEG=4): fitted for the ABSOLUTE error of the angle,
// max 5.0e-6 rad in float evaluation = 0.10 LSB of convert_bearing<ushort> (pi / 65535); near
// the clamp, where the unrefined quotient may cost 0.8 LSB (bearing_core2), its error is
// 3.6e-5 of an angle of 1.5e-3 rad: nothing. One FFMA2 less per angle pair.
#ifndef DC_U16_ACOS_DEG
#define DC_U16_ACOS_DEG 4
#endif
template <bool LSB = false>
__device__ __forceinline__ f2 acos_poly2(f2 a) {
    if constexpr (LSB && DC_U16_ACOS_DEG == 4) {
        f2 p = fma2(B(0.009730911813676357f), a, B(-0.03761417418718338f));
        p    = fma2(p, a, B(0.08563580363988876f));
        p    = fma2(p, a, B(-0.21428003907203674f));
        return fma2(p, a, B(1.570791482925415f));
    } else {
        f2 p = fma2(B(-0.004307856317609549f), a, B(0.0192753653973341f));
        p    = fma2(p, a, B(-0.044893890619277954f));
        p    = fma2(p, a, B(0.08785323798656464f));
        p    = fma2(p, a, B(-0.2145095020532608f));
        return fma2(p, a, B(1.5707951784133911f));
    }
}
template <bool LSB = false>
__device__ __forceinline__ f2 acos_abs2(f2 q) {
    const f2 a = make_float2(fminf(fabsf(q.x), DC_ONE_MINUS_DELTA),
                             fminf(fabsf(q.y), DC_ONE_MINUS_DELTA));
#if DC_ABL & 4
    return a;
#endif
    const f2 t = sub2(B(1.0f), a);
#if DC_ACOS_MUFU_SQRT
    // MUFU.SQRT (max. relative error 2^-23) instead of t * MUFU.RSQ(t): one
    // FMA-pipe slot less per angle pair, the same two XU-pipe slots
    const f2 s = make_float2(mufu_sqrt(t.x), mufu_sqrt(t.y));
#else
    const f2 s = mul2(t, rsq2(t));
#endif
    const f2 p = acos_poly2<LSB>(a);
    return mul2(s, p);
}
// The two factors separately, the polynomial with its coefficients pre-multiplied
// by a scale K (k[i] = K * coefficient i, lowest first): acos(min(|q|, 1 - delta)) * K
// = s * pk. The u16 bearing epilogue folds s * pk into its final FFMA2 and saves
// the product (one FMA-pipe slot per angle pair).
struct acos_scaled {
    float k[6];
};
__device__ __forceinline__ acos_scaled acos_scale(float K) {
    acos_scaled c;
    c.k[0] = __fmul_rn(K, 1.5707951784133911f);
    c.k[1] = __fmul_rn(K, -0.2145095020532608f);
    c.k[2] = __fmul_rn(K, 0.08785323798656464f);
    c.k[3] = __fmul_rn(K, -0.044893890619277954f);
    c.k[4] = __fmul_rn(K, 0.0192753653973341f);
    c.k[5] = __fmul_rn(K, -0.004307856317609549f);
    return c;
}
__device__ __forceinline__ void acos_parts2(f2 q, const acos_scaled& c, f2& s, f2& pk) {
    const f2 a = make_float2(fminf(fabsf(q.x), DC_ONE_MINUS_DELTA),
                             fminf(fabsf(q.y), DC_ONE_MINUS_DELTA));
    const f2 t = sub2(B(1.0f), a);
    s          = make_float2(mufu_sqrt(t.x), mufu_sqrt(t.y));
    pk         = fma2(B(c.k[5]), a, B(c.k[4]));
    pk         = fma2(pk, a, B(c.k[3]));
    pk         = fma2(pk, a, B(c.k[2]));
    pk         = fma2(pk, a, B(c.k[1]));
    pk         = fma2(pk, a, B(c.k[0]));
}
#define DC_PI_F 3.14159274101257324f
// convert_bearing<float, ushort>'s factor, conversion/util.h:71-79: 65535 / pi<float>. A
// compile-time constant, so that the polynomial's coefficients times the factor are
// immediates (dc_api.cu checks that kparams::q_bearing is this very float).
constexpr float kQBearing = 65535.0f / DC_PI_F;
// acos(min(|q|, 1 - delta)) * kQBearing = s * pk for the 16-bit bearing epilogue, which
// multiplies s into the rounding FFMA2 (DC_BEARING_FOLD_Q): one FFMA2 less per angle pair.
__device__ __forceinline__ void acos_parts_q2(f2 q, f2& s, f2& pk) {
    const f2 a = make_float2(fminf(fabsf(q.x), DC_ONE_MINUS_DELTA),
                             fminf(fabsf(q.y), DC_ONE_MINUS_DELTA));
    const f2 t = sub2(B(1.0f), a);
    s          = make_float2(mufu_sqrt(t.x), mufu_sqrt(t.y));
    constexpr float K = kQBearing;
#if DC_U16_ACOS_DEG == 4
    pk = fma2(B(K * 0.009730911813676357f), a, B(K * -0.03761417418718338f));
    pk = fma2(pk, a, B(K * 0.08563580363988876f));
    pk = fma2(pk, a, B(K * -0.21428003907203674f));
    pk = fma2(pk, a, B(K * 1.570791482925415f));
#else
    pk = fma2(B(K * -0.004307856317609549f), a, B(K * 0.0192753653973341f));
    pk = fma2(pk, a, B(K * -0.044893890619277954f));
    pk = fma2(pk, a, B(K * 0.08785323798656464f));
    pk = fma2(pk, a, B(K * -0.2145095020532608f));
    pk = fma2(pk, a, B(K * 1.5707951784133911f));
#endif
}
// acos(clamp(q, -1 + delta, 1 - delta)), what the kernels call.
// DC_ACOS_FOLD_REFLECT=1 puts the reflection inside the last FFMA2 (q < 0:
// pi - s p = fma(-s, p, pi), else fma(s, p, 0)): one packed instruction less per
// angle pair, but an FFMA2 with three distinct register operands costs 3.1 clocks
// against 2 x 2.07 with an immediate each, and measured SLOWER (max-curve 960x540
// 1.054 -> 1.122 ms, profiles/r1_ab_experiments.md section 8). Off.
#ifndef DC_ACOS_FOLD_REFLECT
#define DC_ACOS_FOLD_REFLECT 0
#endif
template <bool LSB = false>
__device__ __forceinline__ f2 acos_signed2(f2 q) {
    const f2 a = make_float2(fminf(fabsf(q.x), DC_ONE_MINUS_DELTA),
                             fminf(fabsf(q.y), DC_ONE_MINUS_DELTA));
    const f2 t = sub2(B(1.0f), a);
    const f2 s = make_float2(mufu_sqrt(t.x), mufu_sqrt(t.y));
    const f2 p = acos_poly2<LSB>(a);
#if DC_ACOS_FOLD_REFLECT
    const bool na = q.x < 0.0f, nb = q.y < 0.0f;
    const f2   ss  = make_float2(na ? -s.x : s.x, nb ? -s.y : s.y);
    const f2   off = make_float2(na ? DC_PI_F : 0.0f, nb ? DC_PI_F : 0.0f);
    return fma2(ss, p, off);
#else
    const f2 r = mul2(s, p);
    const f2 m = sub2(B(DC_PI_F), r);
    return make_float2(q.x < 0.0f ? m.x : r.x, q.y < 0.0f ? m.y : r.y);
#endif
}
// u16 reflection without compares (bearing epilogue; DC_REFLECT_FSEL=0, measured
// 1.5 % slower than the compare + select form, see halve2): for the ratio q
//   sign_xor(Q, q)  = q < 0 ? -Q : Q             (one LOP3)
//   magic_refl(q)   = q < 0 ? M + 65535 : M      (one PRMT: the low two bytes
//                     replicate q's sign bit, M = 1.5 * 2^23 = 0x4B400000)
// so that (q < 0 ? pi - r : r) * Q + M = r * sign_xor(Q, q) + magic_refl(q) with
// pi * Q = 65535. A ratio of -0 cannot occur (RN(b - x) is +0 when b == x).
__device__ __forceinline__ float sign_xor(float v, float q) {
    return __uint_as_float(__float_as_uint(v) ^ (__float_as_uint(q) & 0x80000000u));
}
__device__ __forceinline__ float magic_refl(float q) {
    unsigned d;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(__float_as_uint(q)), "r"(0x4B400000u), "r"(0x76BBu));
    return __uint_as_float(d);
}
__device__ __forceinline__ f2 reflect2(f2 r, f2 q) {
    const f2 m = sub2(B(DC_PI_F), r);
    return make_float2(q.x < 0.0f ? m.x : r.x, q.y < 0.0f ? m.y : r.y);
}

// math/triangles.h:27-64 for both frames: b, c the two ranges, bb / cc their
// squares, cosine of the (camera-only) ray angle broadcast.
//   nom = b - c*cos;  rad = (bb + cc) - ((2b)*c)*cos;  q = nom / sqrt(rad)
// (2b)*c = 2*RN(b*c) and (..)*cos = 2*RN(RN(b*c)*cos) exactly (scaling by two
// commutes with rounding), so rad = RN(sum - 2t) is one FFMA2 with t = RN(bc*cos).
// bc = b*c is handed back for the "either depth is zero" test ((2b)*c == 0).
struct bres {
    f2 r;   // acos(min(|q|, 1 - delta))
    f2 q;   // the ratio: its sign selects the reflection
    f2 bc;  // b * c
};
// 16-bit output (LSB_DIV): the quotient is taken with the raw MUFU.RSQ seed of the radicand
// as the reciprocal of its (still correctly rounded) square root, without the Newton step
// that makes the division provably correctly rounded. The residual step squares the seed's
// 2^-22 error: the quotient is the correctly rounded one except where nom / d lies within
// 2e-6 ulp of a rounding boundary (about one division in 10^5..10^6), there it is the
// neighbouring float. One ulp of a ratio at the clamp |q| = 1 - 10 eps, where acos is
// steepest, moves the angle by 3.9e-5 rad = 0.8 LSB of convert_bearing<ushort> (pi / 65535)
// and 0.4 LSB of convert_max_curve: inside the +-1 LSB contract of the quantised images.
// Float output keeps the refined reciprocal (bit-exact ratio).
#ifndef DC_U16_SKIP_REFINE
#define DC_U16_SKIP_REFINE 1
#endif
template <bool WITH_ACOS = true, bool LSB_DIV = false>
__device__ __forceinline__ bres bearing_core2(f2 b, f2 c, f2 bb, f2 cc, float cos_alpha) {
    bres     o;
    const f2 ca  = B(cos_alpha);
    const f2 nom = sub2(b, mul2(c, ca));
    const f2 sum = add2(bb, cc);
    o.bc         = mul2(b, c);
    const f2 t   = mul2(o.bc, ca);
    const f2 rad = fma2(B(-2.0f), t, sum);
#if DC_APPROX_RATIO
    // EXPERIMENT ONLY (profiles/r1_ab_experiments.md section 7): the ratio as
    // nom * rsqrt(rad), 1 FFMA2 instead of 8. Not the reference's
    // RN(nom / RN(sqrt(rad))): fails the 1e-5 contract where acos is steep.
    o.q = mul2(nom, rsq2(rad));
#else
    f2       y;
    const f2 d = sqrt2(rad, y);
    o.q        = div2(nom, d, (LSB_DIV && DC_U16_SKIP_REFINE) ? y : rcp_refine2(d, y));
#endif
    if constexpr (WITH_ACOS)
        o.r = acos_abs2<LSB_DIV>(o.q);
    return o;
}

// ---------------------------------------------------------------------------
// Staging. Everything a thread needs for a later row (input pixels, range
// factors, per-row camera tables) is fetched with cp.async into shared-memory
// slots that belong to that thread alone, one row step ahead of its use:
// the copies are in flight without holding registers (a register prefetch made
// ptxas spill the freshly loaded values, stalling on them at once), the compiler
// cannot sink them towards their use, and no barrier is needed because a thread
// only reads back its own slots after cp.async.wait_group.
// Slot s of stage g of thread t is the uint4 at smem[(g * NSLOT + s) * T + t]:
// consecutive threads hit consecutive 16-byte words (conflict-free LDS.128).
template <int BYTES>
__device__ __forceinline__ void cp_async(void* smem_dst, const void* gsrc, bool pred) {
    const unsigned d = static_cast<unsigned>(__cvta_generic_to_shared(smem_dst));
    const int      n = pred ? BYTES : 0;  // 0: nothing is read, the slot is zero-filled
    asm volatile("cp.async.ca.shared.global [%0], [%1], %2, %3;" ::"r"(d), "l"(gsrc), "n"(BYTES),
                 "r"(n)
                 : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }
// all groups but the most recently committed one are complete
__device__ __forceinline__ void cp_async_wait_but_one() { asm volatile("cp.async.wait_group 1;" ::: "memory"); }

// SPEC: what the launch code has already established about a job, so that the
// row loop carries no pointer tests. At run time those tests cost ~60
// instructions per row and, worse, split the loop body into one basic block per
// output (ptxas does not schedule across the uniform branches: 4 independent
// dependency chains in flight instead of 16).
//   SPEC_GENERIC  nothing known: every table / output pointer is tested
//   SPEC_TABLES   every output of the feature is wanted, every camera table of
//                 the feature is a per-pixel array (and, for depth input, so is
//                 the range factor)
//   SPEC_UNIFORM  every output is wanted, every angle table is one scalar (the
//                 clamp of camera_models/utility.h:39-45, 1080p and larger);
//                 the range factor of depth input is still an array
//   SPEC_HALF     SPEC_GENERIC for widths with w % 4 == 2 (rows are only 4-byte
//                 (u16) / 8-byte (f32) aligned): every 4-pixel access is made as
//                 two halves, the second one masked in the last column group
enum { SPEC_GENERIC = 0, SPEC_TABLES = 1, SPEC_UNIFORM = 2, SPEC_HALF = 3 };
template <int SPEC>
__device__ __forceinline__ bool is_table(const tab_ref& t) {
    if constexpr (SPEC == SPEC_TABLES)
        return true;
    else if constexpr (SPEC == SPEC_UNIFORM)
        return false;
    else
        return t.p != nullptr;
}
template <int SPEC>
__device__ __forceinline__ bool wanted(const void* out) {
    if constexpr (SPEC == SPEC_TABLES || SPEC == SPEC_UNIFORM)
        return true;
    else
        return out != nullptr;
}
// range factor: FAC_RUNTIME tests lane::fac, FAC_TABLE / FAC_NONE know
enum { FAC_RUNTIME = 0, FAC_TABLE = 1, FAC_NONE = 2 };
template <int FAC>
__device__ __forceinline__ bool has_fac(const float* fac) {
    if constexpr (FAC == FAC_TABLE)
        return true;
    else if constexpr (FAC == FAC_NONE)
        return false;
    else
        return fac != nullptr;
}

constexpr int kRawSlots = 4;
// u16 input:
// slot 0: 4 px of frame A | 4 px of frame B        (u16 x 8)
// slot 1: the u16 pairs holding column u0-1 of A, of B, column u0+4 of A, of B
// slot 2: range factor of the 4 columns            (f32 x 4)
// slot 3: range factor of columns u0-1, u0+4       (f32 x 2)
// f32 range input (no factor):
// slot 0: 4 px of frame A   slot 1: 4 px of frame B
// slot 2: [A(u0-1) | B(u0-1) | A(u0+4) | B(u0+4)]

// one input row of the two frames: columns u0-1 .. u0+4
struct rawrow {
    uint2    a, b;            // u16: 4 pixels of frame A / B
    unsigned la, lb, ra, rb;  // u16: words holding the halo pixels (0 outside the image):
                              // left pixel in the HIGH half, right pixel in the LOW half
    float4   f;               // range factor of the 4 columns
    float    fl, fr;          // and of the halo columns
    float4   fa, fb;          // f32: 4 pixels of frame A / B
    float    hla, hlb, hra, hrb;  // f32: halo pixels
};

// Per-thread constants of the marching loop: every stream as a pointer to the
// thread's first column, so that a row is reached by adding one 32-bit row
// offset; the halo predicates / offsets do not change from row to row.
struct lane {
    const unsigned short *A, *Bf;  // frames A and B at column u0 (u16 input)
    const float *         Af, *Bff;  // the same for f32 input
    const float*          fac;     // range factor table at column u0 (or nullptr)
    int                   offl, offr;  // element offset of the word holding u0-1 / u0+4
    bool                  hl, hr;      // halo column inside the image
    bool                  full;        // SPEC_HALF: columns u0+2, u0+3 exist
};

// Starts the copies of input row `roff / w` (roff = row * w, row clamped into
// the image; `in` false = outside: zero-fill) into raw slots st[0..3*stride].
template <bool RIGHT, int FAC = FAC_RUNTIME, int IN = IN_U16_DEPTH, bool HALF = false>
__device__ __forceinline__ void issue_raw(const lane& L, int roff, bool in, uint4* st,
                                          int stride) {
    const bool l = in && L.hl, rt = in && L.hr;
    const bool in2 = HALF ? in && L.full : in;  // second half of a 4-pixel access
    char*      s0 = reinterpret_cast<char*>(st);
    char*      s1 = reinterpret_cast<char*>(st + stride);
    if constexpr (IN == IN_F32_RANGE) {
        char*        s2 = reinterpret_cast<char*>(st + 2 * stride);
        const float* a  = L.Af + roff;
        const float* b  = L.Bff + roff;
        if constexpr (HALF) {
            cp_async<8>(s0, a, in);
            cp_async<8>(s0 + 8, a + (L.full ? 2 : 0), in2);
            cp_async<8>(s1, b, in);
            cp_async<8>(s1 + 8, b + (L.full ? 2 : 0), in2);
        } else {
            cp_async<16>(s0, a, in);
            cp_async<16>(s1, b, in);
        }
        cp_async<4>(s2, a + L.offl, l);
        cp_async<4>(s2 + 4, b + L.offl, l);
        if constexpr (RIGHT) {
            cp_async<4>(s2 + 8, a + L.offr, rt);
            cp_async<4>(s2 + 12, b + L.offr, rt);
        }
        return;
    } else {
        const unsigned short* a = L.A + roff;
        const unsigned short* b = L.Bf + roff;
        if constexpr (HALF) {
            cp_async<4>(s0, a, in);
            cp_async<4>(s0 + 4, a + (L.full ? 2 : 0), in2);
            cp_async<4>(s0 + 8, b, in);
            cp_async<4>(s0 + 12, b + (L.full ? 2 : 0), in2);
        } else {
            cp_async<8>(s0, a, in);
            cp_async<8>(s0 + 8, b, in);
        }
        cp_async<4>(s1, a + L.offl, l);
        cp_async<4>(s1 + 4, b + L.offl, l);
        if constexpr (RIGHT) {
            cp_async<4>(s1 + 8, a + L.offr, rt);
            cp_async<4>(s1 + 12, b + L.offr, rt);
        }
        if (has_fac<FAC>(L.fac)) {
            char*        s2 = reinterpret_cast<char*>(st + 2 * stride);
            char*        s3 = reinterpret_cast<char*>(st + 3 * stride);
            const float* f  = L.fac + roff;
            if constexpr (HALF) {
                cp_async<8>(s2, f, in);
                cp_async<8>(s2 + 8, f + (L.full ? 2 : 0), in2);
            } else {
                cp_async<16>(s2, f, in);
            }
            cp_async<4>(s3, f + (L.hl ? -1 : 0), l);
            if constexpr (RIGHT)
                cp_async<4>(s3 + 4, f + L.offr, rt);
        }
    }
}

template <bool RIGHT, int FAC = FAC_RUNTIME, int IN = IN_U16_DEPTH>
__device__ __forceinline__ rawrow fetch_raw(const lane& L, float uniform_factor, const uint4* st,
                                            int stride) {
    rawrow x;
    if constexpr (IN == IN_F32_RANGE) {
        x.fa            = *reinterpret_cast<const float4*>(st);
        x.fb            = *reinterpret_cast<const float4*>(st + stride);
        const float4 hl = *reinterpret_cast<const float4*>(st + 2 * stride);
        x.hla = hl.x, x.hlb = hl.y;
        x.hra = RIGHT ? hl.z : 0.0f, x.hrb = RIGHT ? hl.w : 0.0f;
        return x;
    } else {
        const uint4 ab = st[0], hl = st[stride];
        x.a            = make_uint2(ab.x, ab.y);
        x.b            = make_uint2(ab.z, ab.w);
        x.la = hl.x, x.lb = hl.y;
        x.ra = RIGHT ? hl.z : 0u, x.rb = RIGHT ? hl.w : 0u;
        if (has_fac<FAC>(L.fac)) {
            x.f             = *reinterpret_cast<const float4*>(st + 2 * stride);
            const float4 fe = *reinterpret_cast<const float4*>(st + 3 * stride);
            x.fl = fe.x, x.fr = fe.y;
        } else {
            const float v = uniform_factor;
            x.f           = make_float4(v, v, v, v);
            x.fl = x.fr = v;
        }
        return x;
    }
}

// f32 range input: running bounds of the values a thread has loaded, as unsigned
// bit patterns. hi = max(bits) catches negative values, NaN and infinities too
// (their patterns exceed every admissible one); lo = min(bits - 1) skips the
// zeros (0 - 1 wraps to the largest pattern). Raised once, after the strip.
struct f32_window {
    unsigned hi = 0u, lo = 0xffffffffu;
    __device__ __forceinline__ void see(float a, float b) {
        const unsigned x = __float_as_uint(a), y = __float_as_uint(b);
        hi = max(hi, max(x, y));
        lo = min(lo, min(x - 1u, y - 1u));
    }
    __device__ __forceinline__ void raise(int* guard) const {
        if (hi > kF32HiBits || lo < kF32LoBits - 1u)
            *guard = 1;
    }
};

// 2^23 + v as a float has v in its low mantissa bits
__device__ __forceinline__ float magic_lo(unsigned w) {
    return __uint_as_float(__byte_perm(w, 0x4b000000u, 0x7410));
}
__device__ __forceinline__ float magic_hi(unsigned w) {
    return __uint_as_float(__byte_perm(w, 0x4b000000u, 0x7432));
}

// raw -> range pairs rng[0..5] = columns u0-1 .. u0+4 (conversion/util.h:41-51:
// float(d) * sqrtf(z); d == 0 gives 0). rng[5] only when RIGHT.
template <int IN, bool RIGHT>
__device__ __forceinline__ void to_range(const rawrow& x, f2 (&rng)[6], f32_window& win) {
    if constexpr (IN == IN_F32_RANGE) {
        rng[0] = make_float2(x.hla, x.hlb);
        rng[1] = make_float2(x.fa.x, x.fb.x);
        rng[2] = make_float2(x.fa.y, x.fb.y);
        rng[3] = make_float2(x.fa.z, x.fb.z);
        rng[4] = make_float2(x.fa.w, x.fb.w);
        rng[5] = RIGHT ? make_float2(x.hra, x.hrb) : B(0.0f);
#pragma unroll
        for (int i = 0; i < (RIGHT ? 6 : 5); ++i)
            win.see(rng[i].x, rng[i].y);
    } else {
        const f2 m = B(-8388608.0f);
        f2       d[6];
        d[0] = make_float2(magic_hi(x.la), magic_hi(x.lb));
        d[1] = make_float2(magic_lo(x.a.x), magic_lo(x.b.x));
        d[2] = make_float2(magic_hi(x.a.x), magic_hi(x.b.x));
        d[3] = make_float2(magic_lo(x.a.y), magic_lo(x.b.y));
        d[4] = make_float2(magic_hi(x.a.y), magic_hi(x.b.y));
        d[5] = make_float2(magic_lo(x.ra), magic_lo(x.rb));
        const float fac[6] = {x.fl, x.f.x, x.f.y, x.f.z, x.f.w, x.fr};
#pragma unroll
        for (int i = 0; i < (RIGHT ? 6 : 5); ++i) {
            rng[i] = add2(d[i], m);
            if constexpr (IN == IN_U16_DEPTH)
                rng[i] = mul2(rng[i], B(fac[i]));
        }
        if constexpr (!RIGHT)
            rng[5] = B(0.0f);
    }
}

__device__ __forceinline__ unsigned short q16(float magic_sum) {
    return (unsigned short)(__float_as_uint(magic_sum) & 0xffffu);
}
__device__ __forceinline__ uint2 pack4(float a0, float a1, float a2, float a3) {
    return make_uint2(__byte_perm(__float_as_uint(a0), __float_as_uint(a1), 0x5410),
                      __byte_perm(__float_as_uint(a2), __float_as_uint(a3), 0x5410));
}

// Stores 4 u16 pixels per frame from floats that already hold 1.5*2^23 + value
// (the integer sits in the low mantissa bits). HALF: as two 4-byte stores, the
// second one only where columns u0+2, u0+3 exist (`full`).
template <bool HALF = false>
__device__ __forceinline__ void store4_magic(void* base, size_t oa, size_t ob, const f2 (&s)[4],
                                             bool full = true) {
    unsigned short* o  = static_cast<unsigned short*>(base);
    const uint2     pa = pack4(s[0].x, s[1].x, s[2].x, s[3].x);
    const uint2     pb = pack4(s[0].y, s[1].y, s[2].y, s[3].y);
    if constexpr (HALF) {
        *reinterpret_cast<unsigned*>(o + oa) = pa.x;
        *reinterpret_cast<unsigned*>(o + ob) = pb.x;
        if (full) {
            *reinterpret_cast<unsigned*>(o + oa + 2) = pa.y;
            *reinterpret_cast<unsigned*>(o + ob + 2) = pb.y;
        }
    } else {
        *reinterpret_cast<uint2*>(o + oa) = pa;
        *reinterpret_cast<uint2*>(o + ob) = pb;
    }
}

// Stores 4 pixels of frame A (.x halves) and frame B (.y halves).
// OUT_U16: v already holds value * scale; round half to even by the magic add.
template <int OUT, bool HALF = false>
__device__ __forceinline__ void store4(void* base, size_t oa, size_t ob, const f2 (&v)[4],
                                       bool full = true) {
    if constexpr (OUT == OUT_U16) {
        f2 s[4];
#pragma unroll
        for (int i = 0; i < 4; ++i)
            s[i] = add2(v[i], B(12582912.0f));
        store4_magic<HALF>(base, oa, ob, s, full);
    } else {
        float* o = static_cast<float*>(base);
        if constexpr (HALF) {
            *reinterpret_cast<float2*>(o + oa) = make_float2(v[0].x, v[1].x);
            *reinterpret_cast<float2*>(o + ob) = make_float2(v[0].y, v[1].y);
            if (full) {
                *reinterpret_cast<float2*>(o + oa + 2) = make_float2(v[2].x, v[3].x);
                *reinterpret_cast<float2*>(o + ob + 2) = make_float2(v[2].y, v[3].y);
            }
        } else {
            *reinterpret_cast<float4*>(o + oa) = make_float4(v[0].x, v[1].x, v[2].x, v[3].x);
            *reinterpret_cast<float4*>(o + ob) = make_float4(v[0].y, v[1].y, v[2].y, v[3].y);
        }
    }
}

// per-row camera table values: 4 columns -> one slot; uniform tables are
// passed by value and need no copy. `off` = row * w + u0 (clamped into the table).
template <int SPEC = SPEC_GENERIC>
__device__ __forceinline__ void issue_tab4(const tab_ref& t, int off, bool in, uint4* slot,
                                           bool full = true) {
    if (is_table<SPEC>(t)) {
        if constexpr (SPEC == SPEC_HALF) {
            char* d = reinterpret_cast<char*>(slot);
            cp_async<8>(d, t.p + off, in);
            cp_async<8>(d + 8, t.p + off + (full ? 2 : 0), in && full);
        } else {
            cp_async<16>(slot, t.p + off, in);
        }
    }
}
template <int SPEC = SPEC_GENERIC>
__device__ __forceinline__ float4 fetch_tab4(const tab_ref& t, const uint4* slot) {
    if (!is_table<SPEC>(t))
        return make_float4(t.value, t.value, t.value, t.value);
    return *reinterpret_cast<const float4*>(slot);
}
template <int SPEC = SPEC_GENERIC>
__device__ __forceinline__ void issue_tab1(const tab_ref& t, int off, bool in, void* dst) {
    if (is_table<SPEC>(t))
        cp_async<4>(dst, t.p + off, in);
}

}  // namespace pr
}  // namespace dc
#endif
