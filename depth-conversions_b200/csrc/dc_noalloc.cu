// dc_curvature_to_image_noalloc: curvature_to_image without clamps (the minimum /
// maximum of the image are its bounds, conversion/depth_to_curvature.h:218-228) for
// callers that cannot allocate: code under CUDA graph capture, serving loops,
// harnesses that forbid cudaMalloc. The entry point has NO counterpart in the
// reference. It is a variant
// of the min/max reduction + quantiser of dc_quantise_host_ex (dc_api.cu) with
//   - DEVICE buffers from the caller (float curvature image, float range image,
//     16 bit result), nothing staged;
//   - the reduction's scratch (per-CTA partial minima / maxima, the folded result,
//     a ticket counter and a status word) in ONE file-scope __device__ array of
//     fixed size: every GPU that loads the library holds its 33 KB until exit;
//   - no cudaMalloc, cudaFree, cudaHostAlloc or allocating library routine.
// The array is shared by every call on a GPU and keeps the last call's contents:
// the words that must start at zero are cleared on the caller's stream at the head
// of each call, and two calls that overlap on one GPU (different streams, or two
// host threads) RACE on it - callers serialise.
#include "../../include/depthconv_b200.h"

#include <cuda_runtime.h>

#include <algorithm>
#include <climits>
#include <cstdio>
#include <mutex>

namespace {

constexpr unsigned kBlockThreads = 256;
constexpr unsigned kBlockPixels  = 4096;  // pixels one CTA of the reduction folds
constexpr unsigned kMaxBlocks    = unsigned(DC_NOALLOC_MAX_PIXELS / kBlockPixels);
static_assert(DC_NOALLOC_MAX_PIXELS % kBlockPixels == 0, "whole CTAs at the maximum size");

// scratch layout, 32-bit words
//   [W_MIN] [W_MAX]  order-preserving encodings (enc_ordered) of the image's
//                    minimum / maximum, written by the last CTA of the reduction
//   [W_TICKET]       CTAs of the reduction that have published their partial:
//                    MUST be zero when the reduction starts
//   [W_STATUS]       1 = the bounds fail lower < upper (constant or all-NaN image):
//                    MUST be zero when the quantiser starts
//   [kHeaderWords + 2 b], [.. + 1]   partial minimum / maximum of CTA b
enum : unsigned { W_MIN = 0, W_MAX = 1, W_TICKET = 2, W_STATUS = 3, kHeaderWords = 64 };
constexpr size_t kScratchWords = kHeaderWords + 2 * size_t(kMaxBlocks);

// In HOST code this name is a host-side stand-in, not a device address: the
// address a kernel can use comes from cudaGetSymbolAddress (scratch_of below).
static __device__ __align__(256) unsigned g_noalloc_scratch[kScratchWords];

// dc_misc_kernels.cuh (one translation unit only: dc_api.cu) holds the kernels this
// file follows; the three helpers are restated here.
// order-preserving unsigned encoding of a float (dc::ordered<float>::enc) and back
__device__ __forceinline__ unsigned enc_ordered(float f) {
    const unsigned u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
// cvRound as cv::Mat::convertTo uses it (dc::cv_round): NaN / beyond int -> INT_MIN
__device__ __forceinline__ int cv_round(float v) {
    return (fabsf(v) < 2147483648.0f) ? __float2int_rn(v) : INT_MIN;
}
__device__ __forceinline__ float dec_ordered(unsigned e) {
    return __uint_as_float((e & 0x80000000u) ? (e & 0x7fffffffu) : ~e);
}

// CTA-wide fold of (lo, hi); the result is valid in thread 0
__device__ __forceinline__ void block_fold(unsigned& lo, unsigned& hi) {
    __shared__ unsigned s_lo[kBlockThreads / 32], s_hi[kBlockThreads / 32];
    for (int o = 16; o > 0; o >>= 1) {
        const unsigned a = __shfl_xor_sync(0xffffffffu, lo, o), b = __shfl_xor_sync(0xffffffffu, hi, o);
        lo = a < lo ? a : lo;
        hi = b > hi ? b : hi;
    }
    if ((threadIdx.x & 31) == 0) {
        s_lo[threadIdx.x >> 5] = lo;
        s_hi[threadIdx.x >> 5] = hi;
    }
    __syncthreads();
    if (threadIdx.x == 0)
        for (unsigned k = 1; k < kBlockThreads / 32; ++k) {
            lo = s_lo[k] < lo ? s_lo[k] : lo;
            hi = s_hi[k] > hi ? s_hi[k] : hi;
        }
    __syncthreads();  // s_lo / s_hi are reused by the second fold of the last CTA
}

// Stage 1: CTA b folds pixels [b * kBlockPixels, (b + 1) * kBlockPixels) into its
// partial; the CTA that draws the last ticket folds the partials (NaN pixels are
// skipped, as dc::minmax_kernel does) and re-arms the ticket counter.
__global__ void __launch_bounds__(kBlockThreads)
    minmax_partials_kernel(const float* __restrict__ in, size_t n, unsigned* scratch) {
    unsigned     lo = ~0u, hi = 0u;
    const size_t first = size_t(blockIdx.x) * kBlockPixels;
    const size_t last  = first + kBlockPixels < n ? first + kBlockPixels : n;
    for (size_t i = first + threadIdx.x; i < last; i += kBlockThreads) {
        const float v = in[i];
        if (v == v) {
            const unsigned e = enc_ordered(v);
            lo               = e < lo ? e : lo;
            hi               = e > hi ? e : hi;
        }
    }
    block_fold(lo, hi);
    __shared__ bool is_last;
    if (threadIdx.x == 0) {
        scratch[kHeaderWords + 2 * blockIdx.x]     = lo;
        scratch[kHeaderWords + 2 * blockIdx.x + 1] = hi;
        __threadfence();
        is_last = atomicAdd(&scratch[W_TICKET], 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (!is_last)
        return;
    __threadfence();
    lo = ~0u, hi = 0u;
    for (unsigned b = threadIdx.x; b < gridDim.x; b += kBlockThreads) {
        const unsigned a = __ldcg(&scratch[kHeaderWords + 2 * b]), c = __ldcg(&scratch[kHeaderWords + 2 * b + 1]);
        lo = a < lo ? a : lo;
        hi = c > hi ? c : hi;
    }
    block_fold(lo, hi);
    if (threadIdx.x == 0) {
        scratch[W_MIN]    = lo;
        scratch[W_MAX]    = hi;
        scratch[W_TICKET] = 0u;
    }
}

// Stage 2: dc::quantise_kernel's Q_CURVATURE arm (math::scale of math/scaling.h:
// 29-55 onto {0, 65535}, truncated, blanked where the range image rounds to 0)
// with the bounds read from the scratch instead of passed by value.
__global__ void __launch_bounds__(kBlockThreads)
    curvature_image_kernel(const float* __restrict__ in, const float* __restrict__ mask, size_t n,
                           unsigned short* __restrict__ out, unsigned* scratch) {
    const float lo = dec_ordered(scratch[W_MIN]), hi = dec_ordered(scratch[W_MAX]);
    if (!(lo < hi)) {  // Expects(source_range.min < source_range.max), math/scaling.h:36
        if (blockIdx.x == 0 && threadIdx.x == 0)
            scratch[W_STATUS] = 1u;
        return;
    }
    for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
         i += size_t(gridDim.x) * blockDim.x) {
        const float v      = in[i];
        const float scaled = (65535.0f * (v - lo)) / (hi - lo);
        const float tc     = scaled < 0.0f ? 0.0f : (65535.0f < scaled ? 65535.0f : scaled);
        int         r      = (tc == tc) ? int(tc) : 0;
        const int   m      = cv_round(mask[i]);
        r                  = m > 0 ? r : 0;
        out[i]             = static_cast<unsigned short>(r);
    }
}

// Device address of the scratch, looked up once per device with that device current.
constexpr int kMaxDevices = 64;
std::mutex    g_lookup;
unsigned*     g_scratch_of[kMaxDevices] = {};

cudaError_t scratch_of(int device, unsigned** ptr) {
    std::lock_guard<std::mutex> lock(g_lookup);
    if (g_scratch_of[device] == nullptr) {
        unsigned*         p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_noalloc_scratch);
        if (e != cudaSuccess)
            return e;
        g_scratch_of[device] = p;
    }
    *ptr = g_scratch_of[device];
    return cudaSuccess;
}

int report(char* err, size_t err_len, int code, const char* what, cudaError_t e = cudaSuccess) {
    if (err != nullptr && err_len > 0) {
        if (e != cudaSuccess)
            std::snprintf(err, err_len, "dc_curvature_to_image_noalloc: %s: %s", what, cudaGetErrorString(e));
        else
            std::snprintf(err, err_len, "dc_curvature_to_image_noalloc: %s", what);
    }
    return code;
}

}  // namespace

extern "C" int dc_curvature_to_image_noalloc(int device, const float* curvature_device,
                                             const float* range_device, size_t n_px,
                                             uint16_t* out_device, void* cuda_stream, char* err,
                                             size_t err_len) {
    if (err != nullptr && err_len > 0)
        err[0] = '\0';
    if (n_px > size_t(DC_NOALLOC_MAX_PIXELS)) {
        char msg[160];
        std::snprintf(msg, sizeof msg, "%zu pixels exceed DC_NOALLOC_MAX_PIXELS = %zu (the fixed scratch)", n_px,
                      size_t(DC_NOALLOC_MAX_PIXELS));
        return report(err, err_len, DC_ERR_INVALID, msg);
    }
    if (n_px == 0)
        return DC_OK;
    if (curvature_device == nullptr || range_device == nullptr || out_device == nullptr)
        return report(err, err_len, DC_ERR_INVALID, "buffers are NULL");
    int         n_dev = 0;
    cudaError_t e     = cudaGetDeviceCount(&n_dev);
    if (e != cudaSuccess || device < 0 || device >= n_dev || device >= kMaxDevices)
        return report(err, err_len, DC_ERR_NO_DEVICE, "no such CUDA device", e);
    if ((e = cudaSetDevice(device)) != cudaSuccess)
        return report(err, err_len, DC_ERR_CUDA, "cudaSetDevice", e);
    unsigned* scratch = nullptr;
    if ((e = scratch_of(device, &scratch)) != cudaSuccess)
        return report(err, err_len, DC_ERR_CUDA, "cudaGetSymbolAddress(scratch)", e);
    cudaStream_t s = static_cast<cudaStream_t>(cuda_stream);
    // the ticket counter and the status word start at zero, whatever the last call left
    if ((e = cudaMemsetAsync(scratch, 0, kHeaderWords * sizeof(unsigned), s)) != cudaSuccess)
        return report(err, err_len, DC_ERR_CUDA, "clearing the scratch header", e);
    const unsigned blocks = unsigned((n_px + kBlockPixels - 1) / kBlockPixels);  // <= kMaxBlocks
    minmax_partials_kernel<<<blocks, kBlockThreads, 0, s>>>(curvature_device, n_px, scratch);
    if ((e = cudaGetLastError()) != cudaSuccess)
        return report(err, err_len, DC_ERR_CUDA, "min/max kernel launch", e);
    const unsigned qblocks = unsigned(std::min<size_t>((n_px + kBlockThreads - 1) / kBlockThreads, 148 * 16));
    curvature_image_kernel<<<qblocks, kBlockThreads, 0, s>>>(curvature_device, range_device, n_px, out_device,
                                                             scratch);
    if ((e = cudaGetLastError()) != cudaSuccess)
        return report(err, err_len, DC_ERR_CUDA, "quantiser kernel launch", e);
    // synchronous, like dc_quantise_host_ex: the status word tells whether the
    // image had lower < upper (4 bytes onto the stack; nothing is allocated)
    unsigned status = 0;
    if ((e = cudaMemcpyAsync(&status, scratch + W_STATUS, sizeof status, cudaMemcpyDeviceToHost, s)) != cudaSuccess)
        return report(err, err_len, DC_ERR_CUDA, "status read", e);
    if ((e = cudaStreamSynchronize(s)) != cudaSuccess)
        return report(err, err_len, DC_ERR_CUDA, "kernels", e);
    if (status != 0)
        return report(err, err_len, DC_ERR_INVALID,
                      "the image's minimum is not below its maximum (constant or all-NaN image); "
                      "nothing was written");
    return DC_OK;
}
