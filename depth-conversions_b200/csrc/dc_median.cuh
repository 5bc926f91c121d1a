// Median pre-filter: preprocess::median_blur,
// src/include/sens_loc/preprocess/filter.h:143-157, i.e. cv::medianBlur(src, dst,
// ksize) with ksize 3 or 5 on a single-channel 16U or 32F image. cv::medianBlur
// replicates the border (BORDER_REPLICATE) and the median of an odd window is
// unique, so the result is defined without reference to OpenCV's algorithm;
// tests compare with cv2.medianBlur itself.
//
// A thread owns two adjacent columns and marches down a strip of rows with the
// last K rows of its K+1 columns in registers. u16 pixels travel as u16x2
// pairs (one pair per window column: (px[u+c], px[u+1+c])), so that one
// VIMNMX.U16x2 orders the window entry of both output pixels. The 3x3 median is
// found by "forgetful selection" (keep n/2+2 candidates, drop their minimum and
// maximum, take in the next element, until one is left: correct by construction),
// the 5x5 median by a 99-exchange selection network verified exhaustively.
#ifndef DC_MEDIAN_CUH
#define DC_MEDIAN_CUH

#include <cuda_runtime.h>

namespace dc {
namespace med {

struct u16x2 {
    unsigned v;
};
__device__ __forceinline__ void cex(u16x2& a, u16x2& b) {  // a <- min, b <- max (per half)
    const unsigned lo = __vminu2(a.v, b.v), hi = __vmaxu2(a.v, b.v);
    a.v = lo, b.v = hi;
}
__device__ __forceinline__ void cex(float& a, float& b) {
    const float lo = fminf(a, b), hi = fmaxf(a, b);
    a = lo, b = hi;
}

// After the call v[0] holds the minimum and v[N-1] the maximum of v[0..N-1].
template <int N, typename T>
__device__ __forceinline__ void min_max_to_ends(T (&v)[N]) {
#pragma unroll
    for (int i = 0; i < N / 2; ++i)
        cex(v[i], v[N - 1 - i]);  // minima in the lower half, maxima in the upper half
    constexpr int lo_end = (N + 1) / 2;  // the middle element (N odd) may be either
#pragma unroll
    for (int i = 1; i < lo_end; ++i)
        cex(v[0], v[i]);
#pragma unroll
    for (int i = N / 2; i < N - 1; ++i)
        cex(v[i], v[N - 1]);
}

// Median of the n = K*K values w[0..n-1] (destroys w).
template <int NTOT, int LIVE, typename T>
struct select {
    // cand[0..LIVE-1] are the live candidates, `next` the index of the next
    // element of w to take in
    static __device__ __forceinline__ T run(T (&cand)[LIVE], const T* w, int next) {
        min_max_to_ends<LIVE>(cand);
        if constexpr (LIVE == 3) {
            (void)w;
            (void)next;
            return cand[1];
        } else {
            // drop cand[0] (min) and cand[LIVE-1] (max), append w[next]
            T nxt[LIVE - 1];
#pragma unroll
            for (int i = 0; i < LIVE - 2; ++i)
                nxt[i] = cand[i + 1];
            nxt[LIVE - 2] = w[next];
            return select<NTOT, LIVE - 1, T>::run(nxt, w, next + 1);
        }
    }
};

// Median of 25 by a 99-exchange selection network (the classic opt_med25
// sequence). tests/test_median_network.py parses this table and checks it
// exhaustively with the zero-one principle (all 2^25 inputs). Exchanges whose
// minimum or maximum is never read again lose that half to dead-code elimination.
// MEDIAN25_NETWORK_BEGIN
__device__ constexpr unsigned char kNet25[99][2] = {
    {0, 1},   {3, 4},   {2, 4},   {2, 3},   {6, 7},   {5, 7},   {5, 6},   {9, 10},  {8, 10},
    {8, 9},   {12, 13}, {11, 13}, {11, 12}, {15, 16}, {14, 16}, {14, 15}, {18, 19}, {17, 19},
    {17, 18}, {21, 22}, {20, 22}, {20, 21}, {23, 24}, {2, 5},   {3, 6},   {0, 6},   {0, 3},
    {4, 7},   {1, 7},   {1, 4},   {11, 14}, {8, 14},  {8, 11},  {12, 15}, {9, 15},  {9, 12},
    {13, 16}, {10, 16}, {10, 13}, {20, 23}, {17, 23}, {17, 20}, {21, 24}, {18, 24}, {18, 21},
    {19, 22}, {8, 17},  {9, 18},  {0, 18},  {0, 9},   {10, 19}, {1, 19},  {1, 10},  {11, 20},
    {2, 20},  {2, 11},  {12, 21}, {3, 21},  {3, 12},  {13, 22}, {4, 22},  {4, 13},  {14, 23},
    {5, 23},  {5, 14},  {15, 24}, {6, 24},  {6, 15},  {7, 16},  {7, 19},  {13, 21}, {15, 23},
    {7, 13},  {7, 15},  {1, 9},   {3, 11},  {5, 17},  {11, 17}, {9, 17},  {4, 10},  {6, 12},
    {7, 14},  {4, 6},   {4, 7},   {12, 14}, {10, 14}, {6, 7},   {10, 12}, {6, 10},  {6, 17},
    {12, 17}, {7, 17},  {7, 10},  {12, 18}, {7, 12},  {10, 18}, {12, 20}, {10, 20}, {10, 12}};
// MEDIAN25_NETWORK_END  (result: element 12)

template <int K, typename T>
__device__ __forceinline__ T median_of(const T (&w)[K * K]) {
    if constexpr (K == 5) {
        T v[25];
#pragma unroll
        for (int i = 0; i < 25; ++i)
            v[i] = w[i];
#pragma unroll
        for (int i = 0; i < 99; ++i)
            cex(v[kNet25[i][0]], v[kNet25[i][1]]);
        return v[12];
    } else {
        constexpr int n = K * K, live = n / 2 + 2;  // 9 -> 6
        T             cand[live];
#pragma unroll
        for (int i = 0; i < live; ++i)
            cand[i] = w[i];
        return select<n, live, T>::run(cand, w, live);
    }
}

// ---- u16 -------------------------------------------------------------------
// grid: x = column pairs, y = row strips, z = frames
template <int K>
__global__ void __launch_bounds__(128)
    median_u16_kernel(const unsigned short* __restrict__ in, unsigned short* __restrict__ out, int w,
                      int h, int rows_per_strip) {
    constexpr int R  = K / 2;
    const int     u  = 2 * (blockIdx.x * blockDim.x + threadIdx.x);
    const int     v0 = blockIdx.y * rows_per_strip;
    if (u >= w || v0 >= h)
        return;
    const size_t          fpx = size_t(w) * h;
    const unsigned short* src = in + size_t(blockIdx.z) * fpx;
    unsigned short*       dst = out + size_t(blockIdx.z) * fpx;
    int                   col[K + 1];  // clamped column of window column c (BORDER_REPLICATE)
#pragma unroll
    for (int c = 0; c < K + 1; ++c)
        col[c] = min(max(u - R + c, 0), w - 1);
    // pairs of one input row: p[c] = (px[u-R+c], px[u-R+c+1]), c = 0 .. K-1
    auto load_row = [&](int v, u16x2 (&p)[K]) {
        const unsigned short* row = src + size_t(min(max(v, 0), h - 1)) * w;
        unsigned              px[K + 1];
#pragma unroll
        for (int c = 0; c < K + 1; ++c)
            px[c] = __ldg(row + col[c]);
#pragma unroll
        for (int c = 0; c < K; ++c)
            p[c].v = px[c] | (px[c + 1] << 16);
    };
    u16x2 win[K][K];  // win[r] = row v - R + r
#pragma unroll
    for (int r = 1; r < K; ++r)
        load_row(v0 - R + r - 1, win[r]);
    const int v1 = min(v0 + rows_per_strip, h);
    for (int v = v0; v < v1; ++v) {
#pragma unroll
        for (int r = 0; r < K - 1; ++r)
#pragma unroll
            for (int c = 0; c < K; ++c)
                win[r][c] = win[r + 1][c];
        load_row(v + R, win[K - 1]);
        u16x2 flat[K * K];
#pragma unroll
        for (int r = 0; r < K; ++r)
#pragma unroll
            for (int c = 0; c < K; ++c)
                flat[r * K + c] = win[r][c];
        const unsigned m = median_of<K>(flat).v;
        unsigned short* o = dst + size_t(v) * w + u;
        if (u + 1 < w && (reinterpret_cast<uintptr_t>(o) & 3u) == 0)
            *reinterpret_cast<unsigned*>(o) = m;
        else {
            o[0] = (unsigned short)(m & 0xffffu);
            if (u + 1 < w)
                o[1] = (unsigned short)(m >> 16);
        }
    }
}

// ---- f32 (the library API on math::image<float>) ----------------------------
template <int K>
__global__ void __launch_bounds__(128)
    median_f32_kernel(const float* __restrict__ in, float* __restrict__ out, int w, int h,
                      int rows_per_strip) {
    constexpr int R  = K / 2;
    const int     u  = blockIdx.x * blockDim.x + threadIdx.x;
    const int     v0 = blockIdx.y * rows_per_strip;
    if (u >= w || v0 >= h)
        return;
    const size_t fpx = size_t(w) * h;
    const float* src = in + size_t(blockIdx.z) * fpx;
    float*       dst = out + size_t(blockIdx.z) * fpx;
    int          col[K];
#pragma unroll
    for (int c = 0; c < K; ++c)
        col[c] = min(max(u - R + c, 0), w - 1);
    auto load_row = [&](int v, float (&p)[K]) {
        const float* row = src + size_t(min(max(v, 0), h - 1)) * w;
#pragma unroll
        for (int c = 0; c < K; ++c)
            p[c] = __ldg(row + col[c]);
    };
    float win[K][K];
#pragma unroll
    for (int r = 1; r < K; ++r)
        load_row(v0 - R + r - 1, win[r]);
    const int v1 = min(v0 + rows_per_strip, h);
    for (int v = v0; v < v1; ++v) {
#pragma unroll
        for (int r = 0; r < K - 1; ++r)
#pragma unroll
            for (int c = 0; c < K; ++c)
                win[r][c] = win[r + 1][c];
        load_row(v + R, win[K - 1]);
        float flat[K * K];
#pragma unroll
        for (int r = 0; r < K; ++r)
#pragma unroll
            for (int c = 0; c < K; ++c)
                flat[r * K + c] = win[r][c];
        dst[size_t(v) * w + u] = median_of<K>(flat);
    }
}

constexpr int kMedianRows = 32;

template <typename T>
inline cudaError_t launch_median(const T* in, T* out, int w, int h, int n_frames, int ksize,
                                 cudaStream_t s) {
    constexpr bool u16   = sizeof(T) == 2;
    const int      cols  = u16 ? (w + 1) / 2 : w;
    const dim3     block(128, 1, 1);
    for (int f0 = 0; f0 < n_frames; f0 += 65535) {
        const int    nf  = n_frames - f0 < 65535 ? n_frames - f0 : 65535;
        const dim3   grid((cols + 127) / 128, (h + kMedianRows - 1) / kMedianRows, nf);
        const size_t off = size_t(f0) * size_t(w) * h;
        if constexpr (u16) {
            if (ksize == 3)
                median_u16_kernel<3><<<grid, block, 0, s>>>(in + off, out + off, w, h, kMedianRows);
            else
                median_u16_kernel<5><<<grid, block, 0, s>>>(in + off, out + off, w, h, kMedianRows);
        } else {
            if (ksize == 3)
                median_f32_kernel<3><<<grid, block, 0, s>>>(in + off, out + off, w, h, kMedianRows);
            else
                median_f32_kernel<5><<<grid, block, 0, s>>>(in + off, out + off, w, h, kMedianRows);
        }
        const cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess)
            return e;
    }
    return cudaSuccess;
}

}  // namespace med
}  // namespace dc
#endif
