// Kernels that live in ONE translation unit (dc_api.cu): the stand-alone
// quantisers and the synthetic depth generator.
#ifndef DC_MISC_KERNELS_CUH
#define DC_MISC_KERNELS_CUH

#include "dc_kernels.cuh"

#include <climits>

namespace dc {

// Stand-alone quantisers of the library API (dc_quantise_host*): a float or
// double feature image -> an 8 / 16 bit image.
//   convert_bearing / convert_flexion / convert_max_curve<Real, PixelType>
//   (depth_to_bearing.h:282-301, depth_to_flexion.h:194-215, depth_to_max_curve.h:
//   156-176): cv::Mat::convertTo(type, scale, offset) with scaling_factor<Real,
//   PixelType>(max) = ((PixelMax - PixelMin) / max, PixelMin) (conversion/util.h:
//   71-79), i.e. saturate_cast<PixelType>(cvRound(v * scale + offset)): round half
//   to even, clamp; NaN and values beyond int become INT_MIN -> the type's minimum.
//   curvature_to_image (depth_to_curvature.h:213-283): math::scale({lo, hi},
//   {0, 65535}, v) truncated, zero where saturate_cast<uchar>(range) == 0; always
//   a 16 bit image (the intermediate reals_to_image<Real> is ushort and copyTo
//   re-types its destination).
enum : int { Q_RANGE = 0, Q_BEARING = 1, Q_FLEXION = 2, Q_CURVATURE = 3, Q_MAX_CURVE = 4 };

template <typename Px>
struct px_limits;
template <>
struct px_limits<unsigned short> { static constexpr int lo = 0, hi = 65535; };
template <>
struct px_limits<unsigned char> { static constexpr int lo = 0, hi = 255; };
template <>
struct px_limits<short> { static constexpr int lo = -32768, hi = 32767; };
template <>
struct px_limits<signed char> { static constexpr int lo = -128, hi = 127; };

__device__ __forceinline__ int cv_round(float v) {
    return (fabsf(v) < 2147483648.0f) ? __float2int_rn(v) : INT_MIN;  // NaN fails the test
}
__device__ __forceinline__ int cv_round(double v) {
    return (fabs(v) < 2147483648.0) ? __double2int_rn(v) : INT_MIN;
}
__device__ __forceinline__ float  scale_offset(float v, float a, float b) { return add(mul(v, a), b); }
__device__ __forceinline__ double scale_offset(double v, double a, double b) { return __dadd_rn(__dmul_rn(v, a), b); }

template <typename Real, typename Px>
__global__ void __launch_bounds__(256)
    quantise_kernel(int quant, const Real* __restrict__ in, const Real* __restrict__ mask, size_t n,
                    Px* __restrict__ out, Real alpha, Real offset, Real lo, Real hi) {
    for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
         i += size_t(gridDim.x) * blockDim.x) {
        const Real v = in[i];
        int        r;
        if (quant == Q_CURVATURE) {
            // math/scaling.h:29-55 with target {0, 65535}; "+ target_min" adds 0
            const Real scaled = (Real(65535) * (v - lo)) / (hi - lo);
            const Real tc     = scaled < Real(0) ? Real(0) : (Real(65535) < scaled ? Real(65535) : scaled);
            // gsl::narrow_cast<ushort>: truncation; (int)NaN is INT_MIN on x86 -> ushort 0
            r = (tc == tc) ? int(tc) : 0;
            const int m = cv_round(mask[i]);  // convertTo(CV_8U) != 0
            r           = m > 0 ? r : 0;
        } else {
            r = cv_round(quant == Q_RANGE ? v : scale_offset(v, alpha, offset));
            r = min(max(r, px_limits<Px>::lo), px_limits<Px>::hi);
        }
        out[i] = Px(r);
    }
}

// min / max of an image (curvature_to_image without clamps: std::minmax_element
// over every pixel, depth_to_curvature.h:218-228). NaN pixels are skipped.
// res[0] = min, res[1] = max, initialised to +inf / -inf by the caller.
template <typename Real>
struct ordered;
template <>
struct ordered<float> {
    using U = unsigned;
    __device__ static U enc(float f) {
        const U u = __float_as_uint(f);
        return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
    }
};
template <>
struct ordered<double> {
    using U = unsigned long long;
    __device__ static U enc(double f) {
        const U u = (U)__double_as_longlong(f);
        return (u >> 63) ? ~u : (u | 0x8000000000000000ull);
    }
};
// res holds the ORDER-PRESERVING unsigned encodings (decoded on the host)
template <typename Real>
__global__ void __launch_bounds__(256)
    minmax_kernel(const Real* __restrict__ in, size_t n, typename ordered<Real>::U* res) {
    using U = typename ordered<Real>::U;
    U lo = ~U(0), hi = 0;
    for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
         i += size_t(gridDim.x) * blockDim.x) {
        const Real v = in[i];
        if (v == v) {
            const U e = ordered<Real>::enc(v);
            lo        = e < lo ? e : lo;
            hi        = e > hi ? e : hi;
        }
    }
    for (int o = 16; o > 0; o >>= 1) {
        const U a = __shfl_xor_sync(0xffffffffu, lo, o), b = __shfl_xor_sync(0xffffffffu, hi, o);
        lo = a < lo ? a : lo;
        hi = b > hi ? b : hi;
    }
    if ((threadIdx.x & 31) == 0) {
        atomicMin(res, lo);
        atomicMax(res + 1, hi);
    }
}

// Synthetic depth generator (bench / test tooling), bit-identical to
// oracle/dc_oracle.c dco_synth_depth.
__device__ __forceinline__ unsigned long long splitmix64(unsigned long long x) {
    unsigned long long z = x + 0x9E3779B97F4A7C15ull;
    z                    = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z                    = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
__device__ __forceinline__ long long tri(long long t, long long p) {
    const long long m = t % (2 * p);
    return m < p ? p - m : m - p;
}
__global__ void __launch_bounds__(256)
    synth_depth_kernel(unsigned long long seed, int frame0, int w, int h,
                       unsigned short* out) {
    const int u = blockIdx.x * blockDim.x + threadIdx.x;
    const int v = blockIdx.y * blockDim.y + threadIdx.y;
    if (u >= w || v >= h)
        return;
    const int                frame = frame0 + int(blockIdx.z);
    const unsigned long long fkey  = (unsigned long long)frame * 0x9E3779B97F4A7C15ull;
    const unsigned long long hsh =
        splitmix64(seed ^ fkey ^ (unsigned long long)((long long)v * w + u));
    long long d = 800 + 2 * tri((long long)u + 3ll * frame, 400) +
                  3 * tri((long long)v + 5ll * frame, 300) + (long long)(hsh % 9) - 4;
    const unsigned long long blk =
        splitmix64(seed ^ (fkey + 0xD1B54A32D192ED03ull) ^
                   (unsigned long long)(((long long)(v / 16)) * 4096 + (u / 16)));
    if (blk % 100 < 25 || (hsh >> 32) % 100 < 5)
        d = 0;
    out[size_t(blockIdx.z) * size_t(w) * h + size_t(v) * w + u] = (unsigned short)d;
}

}  // namespace dc
#endif
