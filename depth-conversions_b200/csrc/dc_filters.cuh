// Gaussian blur and bilateral filter of sens_loc::preprocess (row f4 of SURVEY.md
// section 8, after the median of dc_median.cuh):
//   preprocess::gaussian_blur    src/include/sens_loc/preprocess/filter.h:125-136  (cv::GaussianBlur)
//   preprocess::bilateral_filter src/include/sens_loc/preprocess/filter.h:49-82    (cv::bilateralFilter,
//                                apps/depth_filter `bilateral`, filter_functor.h:24-50)
// The arithmetic is OpenCV's (4.3.0 pinned by the reference, not vendored): what is
// restated here is its published algorithm, checked against the cv2 of this image
// within a stated tolerance (oracle/filters_oracle.py, tests/test_gpu_filters.py);
// unlike the median these filters are NOT bit-exact against OpenCV (its SIMD
// accumulation order and FMA use are build dependent).
//
// Both kernels work on a shared-memory tile with the BORDER_REFLECT_101 halo
// (cv::BORDER_DEFAULT), compute in float without FMA contraction (-fmad=false)
// and store float, or u16 by cv::saturate_cast (round half to even, clamp):
// math::convert<PixelType> of math/image.h:117-128.
#ifndef DC_FILTERS_CUH
#define DC_FILTERS_CUH

#include <cuda_runtime.h>

#include <cmath>
#include <cstdint>

namespace dc {
namespace flt {

constexpr int kMaxRadius = 15;  // kernel sizes up to 31 x 31

// cv::borderInterpolate(i, n, BORDER_REFLECT_101)
__host__ __device__ inline int reflect101(int i, int n) {
    if (n == 1)
        return 0;
    while (i < 0 || i >= n)
        i = i < 0 ? -i : 2 * (n - 1) - i;
    return i;
}

template <typename T>
__device__ __forceinline__ float load_px(const T* p) {
    return float(*p);
}
template <typename T>
__device__ __forceinline__ void store_px(T* p, float v);
template <>
__device__ __forceinline__ void store_px<float>(float* p, float v) {
    *p = v;
}
template <>
__device__ __forceinline__ void store_px<unsigned short>(unsigned short* p, float v) {
    // saturate_cast<ushort>(float): cvRound (half to even), clamped; NaN -> 0
    const float c = fminf(fmaxf(v, 0.0f), 65535.0f);
    *p            = (unsigned short)__float2uint_rn(c);
}

// ---------------------------------------------------------------------------
// Gaussian blur. Separable: rows, then columns, each
//   k[c] x[c] + sum_i k[c + i] (x[c + i] + x[c - i])        (i = 1 .. radius)
// accumulated in float in that order (cv::SymmRowSmallFilter / SymmColumnFilter
// add the two symmetric taps before multiplying).
struct gauss_params {
    int   w, h;
    int   rx, ry;                 // radii: kernel sizes 2 r + 1
    float kx[kMaxRadius + 1];     // kx[i] = coefficient at distance i from the centre
    float ky[kMaxRadius + 1];
};

// cv::getGaussianKernel(n, sigma, CV_32F), centre-out: k[i] for distance i.
// sigma <= 0: the fixed tables for n <= 9, else sigma = 0.3 ((n - 1) 0.5 - 1) + 0.8.
inline void gaussian_kernel_host(int n, double sigma, float* centre_out) {
    static const float t1[] = {1.f};
    static const float t3[] = {0.5f, 0.25f};
    static const float t5[] = {0.375f, 0.25f, 0.0625f};
    static const float t7[] = {0.28125f, 0.21875f, 0.109375f, 0.03125f};
    static const float t9[] = {60.f / 256, 51.f / 256, 30.f / 256, 13.f / 256, 4.f / 256};
    const int          r    = n / 2;
    if (sigma <= 0 && n <= 9) {
        const float* t = n == 1 ? t1 : n == 3 ? t3 : n == 5 ? t5 : n == 7 ? t7 : t9;
        for (int i = 0; i <= r; ++i)
            centre_out[i] = t[i];
        return;
    }
    if (sigma <= 0)
        sigma = 0.3 * ((n - 1) * 0.5 - 1) + 0.8;
    const double scale = -0.5 / (sigma * sigma);
    double       e[kMaxRadius + 1], sum = 0;
    // sum over all n taps in index order (left to right), as the reference loop does
    for (int i = 0; i < n; ++i) {
        const double x = i - (n - 1) * 0.5;
        const double t = std::exp(scale * x * x);
        if (i >= r)
            e[i - r] = t;
        sum += t;
    }
    for (int i = 0; i <= r; ++i)
        centre_out[i] = float(e[i] / sum);
}

constexpr int kGaussTW = 64, kGaussTH = 32, kGaussThreads = 256;

inline size_t gauss_smem_bytes(int rx, int ry) {
    return size_t((kGaussTH + 2 * ry) * (kGaussTW + 2 * rx) + (kGaussTH + 2 * ry) * kGaussTW) *
           sizeof(float);
}

// 256 threads as 64 columns x 4 row groups: a thread walks its column (no index division);
// a warp covers 32 consecutive columns of one row (conflict-free shared memory, coalesced).
template <typename T>
__global__ void __launch_bounds__(kGaussThreads)
    gaussian_blur_kernel(const T* __restrict__ in, T* __restrict__ out,
                         const __grid_constant__ gauss_params p) {
    extern __shared__ float sm[];
    const int rx = p.rx, ry = p.ry, w = p.w, h = p.h;
    const int pw = kGaussTW + 2 * rx, ph = kGaussTH + 2 * ry;
    float*    tile = sm;            // ph x pw input pixels
    float*    rows = sm + ph * pw;  // ph x TW row-filtered values
    const int u0 = blockIdx.x * kGaussTW, v0 = blockIdx.y * kGaussTH;
    const int tx = threadIdx.x & (kGaussTW - 1), ty = threadIdx.x / kGaussTW;  // 64 x 4
    constexpr int TY = kGaussThreads / kGaussTW;
    const T*  src = in + size_t(blockIdx.z) * size_t(w) * h;
    T*        dst = out + size_t(blockIdx.z) * size_t(w) * h;
    // a tile whose halo lies inside the image needs no border arithmetic
    const bool inner = u0 - rx >= 0 && u0 + kGaussTW + rx <= w && v0 - ry >= 0 && v0 + kGaussTH + ry <= h;
    for (int y = ty; y < ph; y += TY) {
        const int v   = inner ? v0 - ry + y : reflect101(v0 - ry + y, h);
        const T*  row = src + size_t(v) * w;
        for (int x = tx; x < pw; x += kGaussTW) {
            const int u      = inner ? u0 - rx + x : reflect101(u0 - rx + x, w);
            tile[y * pw + x] = load_px(row + u);
        }
    }
    __syncthreads();
    for (int y = ty; y < ph; y += TY) {
        const float* c = tile + y * pw + tx + rx;
        float        s = __fmul_rn(c[0], p.kx[0]);
        for (int k = 1; k <= rx; ++k)
            s = __fadd_rn(s, __fmul_rn(p.kx[k], __fadd_rn(c[k], c[-k])));
        rows[y * kGaussTW + tx] = s;
    }
    __syncthreads();
    const int u = u0 + tx;
    if (u >= w)
        return;
    for (int y = ty; y < kGaussTH; y += TY) {
        const int v = v0 + y;
        if (v >= h)
            break;
        const float* c = rows + (y + ry) * kGaussTW + tx;
        float        s = __fmul_rn(c[0], p.ky[0]);
        for (int k = 1; k <= ry; ++k)
            s = __fadd_rn(s, __fmul_rn(p.ky[k], __fadd_rn(c[k * kGaussTW], c[-k * kGaussTW])));
        store_px(dst + size_t(v) * w + u, s);
    }
}

template <typename T>
inline cudaError_t launch_gaussian(const T* in, T* out, int n_frames, const gauss_params& p,
                                   cudaStream_t s) {
    const size_t smem = gauss_smem_bytes(p.rx, p.ry);
    cudaError_t  e    = cudaFuncSetAttribute(gaussian_blur_kernel<T>,
                                             cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
    if (e != cudaSuccess)
        return e;
    for (int f0 = 0; f0 < n_frames; f0 += 65535) {
        const int    nf = n_frames - f0 < 65535 ? n_frames - f0 : 65535;
        const dim3   grid((p.w + kGaussTW - 1) / kGaussTW, (p.h + kGaussTH - 1) / kGaussTH, nf);
        const size_t off = size_t(f0) * size_t(p.w) * p.h;
        gaussian_blur_kernel<T><<<grid, kGaussThreads, smem, s>>>(in + off, out + off, p);
        e = cudaGetLastError();
        if (e != cudaSuccess)
            return e;
    }
    return cudaSuccess;
}

// ---------------------------------------------------------------------------
// Bilateral filter, CV_32FC1 algorithm of OpenCV (bilateralFilter_32f):
//   radius = d <= 0 ? round(1.5 sigma_space) : d / 2, at least 1; sigma <= 0 -> 1
//   colour weight of a neighbour = linear interpolation in a table of 4096 + 2
//     entries lut[i] = exp((i / scale)^2 * (-0.5 / sigma_color^2)), scale =
//     4096 / (max - min) over the frame, at |neighbour - centre| * scale
//   spatial weight = exp(r^2 * (-0.5 / sigma_space^2)) for r = distance <= radius
//   out = (sum w v + centre) / (sum w + 1); a frame with max - min < FLT_EPSILON
//   is copied.
constexpr int kLutBins = 1 << 12;
// per-frame scratch: [0] min (ordered-int), [1] max (ordered-int), [2] scale_index,
// [3] flat flag, [4 ..] the table
constexpr int kBilatScratchWords = 4 + kLutBins + 2;

__device__ __forceinline__ unsigned ordered_bits(float f) {
    const unsigned u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float from_ordered(unsigned o) {
    return __uint_as_float((o & 0x80000000u) ? (o & 0x7fffffffu) : ~o);
}

__global__ void bilateral_init_kernel(unsigned* scratch) {
    unsigned* s = scratch + size_t(blockIdx.x) * kBilatScratchWords;
    if (threadIdx.x == 0) {
        s[0] = 0xffffffffu;
        s[1] = 0u;
    }
}

template <typename T>
__global__ void __launch_bounds__(256)
    bilateral_minmax_kernel(const T* __restrict__ in, size_t frame_px, unsigned* scratch) {
    const T* src = in + size_t(blockIdx.y) * frame_px;
    unsigned lo = 0xffffffffu, hi = 0u;
    for (size_t i = size_t(blockIdx.x) * 256 + threadIdx.x; i < frame_px; i += size_t(gridDim.x) * 256) {
        const unsigned o = ordered_bits(float(src[i]));
        lo               = min(lo, o);
        hi               = max(hi, o);
    }
    for (int d = 16; d > 0; d >>= 1) {
        lo = min(lo, __shfl_xor_sync(0xffffffffu, lo, d));
        hi = max(hi, __shfl_xor_sync(0xffffffffu, hi, d));
    }
    if ((threadIdx.x & 31) == 0) {
        unsigned* s = scratch + size_t(blockIdx.y) * kBilatScratchWords;
        atomicMin(s + 0, lo);
        atomicMax(s + 1, hi);
    }
}

__global__ void __launch_bounds__(256)
    bilateral_lut_kernel(unsigned* scratch, double gauss_color_coeff) {
    unsigned*    s  = scratch + size_t(blockIdx.y) * kBilatScratchWords;
    const double mn = double(from_ordered(s[0])), mx = double(from_ordered(s[1]));
    // float len = (float)(max - min); float scale_index = kExpNumBins / len;
    const float len   = float(mx - mn);
    const float scale = float(kLutBins) / len;
    const bool  flat  = fabs(mn - mx) < 1.1920928955078125e-07;  // FLT_EPSILON
    const int   i     = blockIdx.x * 256 + threadIdx.x;
    if (i < kLutBins + 2) {
        const double val = double(i) / double(scale);
        s[4 + i]         = flat ? 0u : __float_as_uint(float(exp(val * val * gauss_color_coeff)));
    }
    if (i == 0) {
        s[2] = __float_as_uint(scale);
        s[3] = flat ? 1u : 0u;
    }
}

constexpr int kBilatTile = 32, kBilatThreads = 256;

inline size_t bilateral_smem_bytes(int radius) {
    const int side = kBilatTile + 2 * radius, win = 2 * radius + 1;
    return size_t(side * side + kLutBins + 2 + win * win) * sizeof(float);
}

template <typename T>
__global__ void __launch_bounds__(kBilatThreads)
    bilateral_kernel(const T* __restrict__ in, T* __restrict__ out, int w, int h, int radius,
                     double gauss_space_coeff, const unsigned* __restrict__ scratch) {
    extern __shared__ float sm[];
    const int side = kBilatTile + 2 * radius, win = 2 * radius + 1;
    float*    tile = sm;                           // side x side pixels
    float*    lut  = sm + side * side;             // kLutBins + 2
    float*    spw  = lut + kLutBins + 2;           // win x win spatial weights (0 = not used)
    const unsigned* fs = scratch + size_t(blockIdx.z) * kBilatScratchWords;
    const T*  src = in + size_t(blockIdx.z) * size_t(w) * h;
    T*        dst = out + size_t(blockIdx.z) * size_t(w) * h;
    const int u0 = blockIdx.x * kBilatTile, v0 = blockIdx.y * kBilatTile;
    const bool flat = fs[3] != 0u;
    if (flat) {  // |max - min| < FLT_EPSILON: src.copyTo(dst)
        for (int i = threadIdx.x; i < kBilatTile * kBilatTile; i += kBilatThreads) {
            const int u = u0 + i % kBilatTile, v = v0 + i / kBilatTile;
            if (u < w && v < h)
                dst[size_t(v) * w + u] = src[size_t(v) * w + u];
        }
        return;
    }
    const float scale = __uint_as_float(fs[2]);
    for (int i = threadIdx.x; i < side * side; i += kBilatThreads) {
        const int y = i / side, x = i - y * side;
        tile[i] = load_px(src + size_t(reflect101(v0 - radius + y, h)) * w + reflect101(u0 - radius + x, w));
    }
    for (int i = threadIdx.x; i < kLutBins + 2; i += kBilatThreads)
        lut[i] = __uint_as_float(fs[4 + i]);
    for (int i = threadIdx.x; i < win * win; i += kBilatThreads) {
        const int    dy = i / win - radius, dx = i % win - radius;
        const double r  = sqrt(double(dy) * dy + double(dx) * dx);
        spw[i] = (r > radius || (dy == 0 && dx == 0)) ? 0.0f : float(exp(r * r * gauss_space_coeff));
    }
    __syncthreads();
    for (int i = threadIdx.x; i < kBilatTile * kBilatTile; i += kBilatThreads) {
        const int y = i / kBilatTile, x = i - y * kBilatTile;
        const int u = u0 + x, v = v0 + y;
        if (u >= w || v >= h)
            continue;
        const float* c    = tile + (y + radius) * side + x + radius;
        const float  val0 = c[0];
        float        sum = 0.0f, wsum = 0.0f;
        for (int dy = -radius; dy <= radius; ++dy)
            for (int dx = -radius; dx <= radius; ++dx) {
                const float sw = spw[(dy + radius) * win + dx + radius];
                if (sw == 0.0f)  // outside the circle, or the centre
                    continue;
                const float val   = c[dy * side + dx];
                float       alpha = __fmul_rn(fabsf(__fsub_rn(val, val0)), scale);
                // (<= kLutBins for finite input; the bound keeps an infinite pixel inside the table)
                const int   idx   = min(int(floorf(alpha)), kLutBins);
                alpha             = __fsub_rn(alpha, float(idx));
                const float e0 = lut[idx], e1 = lut[idx + 1];
                const float wgt =
                    __fmul_rn(sw, __fadd_rn(e0, __fmul_rn(alpha, __fsub_rn(e1, e0))));
                sum  = __fadd_rn(sum, __fmul_rn(val, wgt));
                wsum = __fadd_rn(wsum, wgt);
            }
        store_px(dst + size_t(v) * w + u, __fdiv_rn(__fadd_rn(sum, val0), __fadd_rn(wsum, 1.0f)));
    }
}

// `scratch`: n_frames * kBilatScratchWords words of device memory.
template <typename T>
inline cudaError_t launch_bilateral(const T* in, T* out, int w, int h, int n_frames, int radius,
                                    double sigma_color, double sigma_space, unsigned* scratch,
                                    cudaStream_t s) {
    const double gc = -0.5 / (sigma_color * sigma_color), gs = -0.5 / (sigma_space * sigma_space);
    const size_t smem = bilateral_smem_bytes(radius);
    cudaError_t  e    = cudaFuncSetAttribute(bilateral_kernel<T>,
                                             cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
    if (e != cudaSuccess)
        return e;
    const size_t px = size_t(w) * h;
    for (int f0 = 0; f0 < n_frames; f0 += 65535) {
        const int    nf  = n_frames - f0 < 65535 ? n_frames - f0 : 65535;
        const size_t off = size_t(f0) * px;
        unsigned*    sc  = scratch + size_t(f0) * kBilatScratchWords;
        bilateral_init_kernel<<<nf, 32, 0, s>>>(sc);
        const unsigned bx = unsigned(px / 256 / 8 + 1 < 1184 ? px / 256 / 8 + 1 : 1184);
        bilateral_minmax_kernel<T><<<dim3(bx, nf), 256, 0, s>>>(in + off, px, sc);
        bilateral_lut_kernel<<<dim3((kLutBins + 2 + 255) / 256, nf), 256, 0, s>>>(sc, gc);
        const dim3 grid((w + kBilatTile - 1) / kBilatTile, (h + kBilatTile - 1) / kBilatTile, nf);
        bilateral_kernel<T><<<grid, kBilatThreads, smem, s>>>(in + off, out + off, w, h, radius, gs, sc);
        e = cudaGetLastError();
        if (e != cudaSuccess)
            return e;
    }
    return cudaSuccess;
}

}  // namespace flt
}  // namespace dc
#endif
