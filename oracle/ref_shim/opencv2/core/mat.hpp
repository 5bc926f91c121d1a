// TEST INFRASTRUCTURE ONLY (oracle/): a minimal stand-in for the handful of
// cv::Mat operations that the reference's hot-path headers use, so that the
// reference's OWN headers under /root/reference/src/include can be compiled in
// a container that has no OpenCV C++ development files.
//
// This is not OpenCV and not a copy of it. It restates the documented
// behaviour of the operations the reference calls on the path:
//   * Mat(rows, cols, type), shallow copy with shared storage, at<T>(r, c),
//     rows/cols/data/type()/channels()/empty(), operator=(scalar) (fill),
//     begin<T>()/end<T>() (dense, continuous storage => raw pointers).
//   * convertTo(dst, rtype, alpha, beta): dst = saturate_cast<D>(src*alpha+beta)
//     computed in float for sources up to 32F and in double for 64F sources,
//     where saturate_cast to an integer type rounds half-to-even (cvRound /
//     lrint) and clamps to the destination range.
//     Reference call sites: conversion/depth_to_bearing.h:292,
//     depth_to_flexion.h:206, depth_to_max_curve.h:167,
//     depth_to_curvature.h:267, depth_scaling.h:28, math/image.h:126.
//   * copyTo(dst, mask): copy elements where the 8-bit mask is non-zero
//     (depth_to_curvature.h:275).
// Single channel only: the reference's math::image<T> enforces one channel.
#ifndef ORACLE_REF_SHIM_OPENCV2_CORE_MAT_HPP
#define ORACLE_REF_SHIM_OPENCV2_CORE_MAT_HPP

#include <cmath>
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <limits>
#include <memory>
#include <stdexcept>
#include <vector>

typedef unsigned char  uchar;
typedef signed char    schar;
typedef unsigned short ushort;

#define CV_8U 0
#define CV_8S 1
#define CV_16U 2
#define CV_16S 3
#define CV_32S 4
#define CV_32F 5
#define CV_64F 6

namespace cv {

template <typename T>
using MatConstIterator_ = const T*;
template <typename T>
using MatIterator_ = T*;

namespace shim_detail {
inline std::size_t elem_size(int type) {
    switch (type) {
    case CV_8U:
    case CV_8S: return 1;
    case CV_16U:
    case CV_16S: return 2;
    case CV_32S:
    case CV_32F: return 4;
    case CV_64F: return 8;
    }
    throw std::invalid_argument("cv shim: unsupported type");
}

// Round half to even, then clamp to the destination integer range.
template <typename D, typename W>
inline D saturate_round(W v) {
    if constexpr (std::is_floating_point_v<D>) {
        return static_cast<D>(v);
    } else {
        // cvRound: lrint with the default rounding mode is round-half-to-even;
        // it converts to a 32 bit int (cvtss2si / cvtsd2si), so NaN and values
        // outside the int range become INT_MIN, which then clamps to the
        // minimum of the destination type.
        const long r = (std::fabs(v) < W(2147483648.0)) ? std::lrint(v)
                                                        : (-2147483647L - 1L);
        const long lo = static_cast<long>(std::numeric_limits<D>::min());
        const long hi = static_cast<long>(std::numeric_limits<D>::max());
        return static_cast<D>(r < lo ? lo : (r > hi ? hi : r));
    }
}

template <typename S, typename D, typename W>
inline void convert_loop(const void* src, void* dst, std::size_t n, double alpha,
                         double beta, bool scaled) {
    const S* s = static_cast<const S*>(src);
    D*       d = static_cast<D*>(dst);
    const W  a = static_cast<W>(alpha);
    const W  b = static_cast<W>(beta);
    for (std::size_t i = 0; i < n; ++i) {
        if (scaled) {
            const W prod = static_cast<W>(s[i]) * a;
            const W sum  = prod + b;
            d[i]         = saturate_round<D, W>(sum);
        } else {
            if constexpr (std::is_floating_point_v<S>)
                d[i] = saturate_round<D, S>(s[i]);
            else if constexpr (std::is_floating_point_v<D>)
                d[i] = static_cast<D>(s[i]);
            else {
                const long v  = static_cast<long>(s[i]);
                const long lo = static_cast<long>(std::numeric_limits<D>::min());
                const long hi = static_cast<long>(std::numeric_limits<D>::max());
                d[i]          = static_cast<D>(v < lo ? lo : (v > hi ? hi : v));
            }
        }
    }
}

template <typename S, typename W>
inline void convert_from(const void* src, void* dst, int dtype, std::size_t n,
                         double alpha, double beta, bool scaled) {
    switch (dtype) {
    case CV_8U: convert_loop<S, uchar, W>(src, dst, n, alpha, beta, scaled); return;
    case CV_8S: convert_loop<S, schar, W>(src, dst, n, alpha, beta, scaled); return;
    case CV_16U: convert_loop<S, ushort, W>(src, dst, n, alpha, beta, scaled); return;
    case CV_16S: convert_loop<S, short, W>(src, dst, n, alpha, beta, scaled); return;
    case CV_32S: convert_loop<S, int, W>(src, dst, n, alpha, beta, scaled); return;
    case CV_32F: convert_loop<S, float, W>(src, dst, n, alpha, beta, scaled); return;
    case CV_64F: convert_loop<S, double, W>(src, dst, n, alpha, beta, scaled); return;
    }
    throw std::invalid_argument("cv shim: unsupported destination type");
}
}  // namespace shim_detail

class Mat {
  public:
    Mat() = default;
    Mat(int r, int c, int t) { create(r, c, t); }
    // Wrap caller-owned memory (no ownership taken).
    Mat(int r, int c, int t, void* external)
        : rows(r)
        , cols(c)
        , data(static_cast<uchar*>(external))
        , _type(t) {}

    void create(int r, int c, int t) {
        if (r == rows && c == cols && t == _type && data != nullptr)
            return;
        rows  = r;
        cols  = c;
        _type = t;
        const std::size_t bytes =
            static_cast<std::size_t>(r) * static_cast<std::size_t>(c) *
            shim_detail::elem_size(t);
        _store = std::shared_ptr<uchar[]>(new uchar[bytes == 0 ? 1 : bytes]);
        data   = _store.get();
    }

    [[nodiscard]] int  type() const { return _type; }
    [[nodiscard]] int  channels() const { return 1; }
    [[nodiscard]] bool empty() const {
        return data == nullptr || rows == 0 || cols == 0;
    }
    [[nodiscard]] std::size_t total() const {
        return static_cast<std::size_t>(rows) * static_cast<std::size_t>(cols);
    }

    template <typename T>
    T& at(int r, int c) {
        return reinterpret_cast<T*>(data)[static_cast<std::size_t>(r) * cols + c];
    }
    template <typename T>
    const T& at(int r, int c) const {
        return reinterpret_cast<const T*>(
            data)[static_cast<std::size_t>(r) * cols + c];
    }

    template <typename T>
    T* begin() {
        return reinterpret_cast<T*>(data);
    }
    template <typename T>
    T* end() {
        return reinterpret_cast<T*>(data) + total();
    }
    template <typename T>
    const T* begin() const {
        return reinterpret_cast<const T*>(data);
    }
    template <typename T>
    const T* end() const {
        return reinterpret_cast<const T*>(data) + total();
    }

    // Fill with a scalar (cv::Mat::operator=(const Scalar&)).
    Mat& operator=(double v) {
        const std::size_t n = total();
        switch (_type) {
        case CV_8U: fill<uchar>(v, n); break;
        case CV_8S: fill<schar>(v, n); break;
        case CV_16U: fill<ushort>(v, n); break;
        case CV_16S: fill<short>(v, n); break;
        case CV_32S: fill<int>(v, n); break;
        case CV_32F: fill<float>(v, n); break;
        case CV_64F: fill<double>(v, n); break;
        }
        return *this;
    }

    void convertTo(Mat& dst, int rtype, double alpha = 1., double beta = 0.) const {
        const bool scaled = !(alpha == 1. && beta == 0.);
        Mat        out;
        out.create(rows, cols, rtype);
        const std::size_t n = total();
        using namespace shim_detail;
        switch (_type) {
        case CV_8U: convert_from<uchar, float>(data, out.data, rtype, n, alpha, beta, scaled); break;
        case CV_8S: convert_from<schar, float>(data, out.data, rtype, n, alpha, beta, scaled); break;
        case CV_16U: convert_from<ushort, float>(data, out.data, rtype, n, alpha, beta, scaled); break;
        case CV_16S: convert_from<short, float>(data, out.data, rtype, n, alpha, beta, scaled); break;
        case CV_32S: convert_from<int, double>(data, out.data, rtype, n, alpha, beta, scaled); break;
        case CV_32F: convert_from<float, float>(data, out.data, rtype, n, alpha, beta, scaled); break;
        case CV_64F: convert_from<double, double>(data, out.data, rtype, n, alpha, beta, scaled); break;
        default: throw std::invalid_argument("cv shim: unsupported source type");
        }
        dst = out;
    }

    void copyTo(Mat& dst, const Mat& mask) const {
        if (dst.rows != rows || dst.cols != cols || dst._type != _type ||
            dst.data == nullptr) {
            dst.create(rows, cols, _type);
            dst = 0.;
        }
        const std::size_t es = shim_detail::elem_size(_type);
        const std::size_t n  = total();
        for (std::size_t i = 0; i < n; ++i)
            if (mask.data[i] != 0)
                std::memcpy(dst.data + i * es, data + i * es, es);
    }

    int    rows = 0;
    int    cols = 0;
    uchar* data = nullptr;

  private:
    template <typename T>
    void fill(double v, std::size_t n) {
        T* p = reinterpret_cast<T*>(data);
        for (std::size_t i = 0; i < n; ++i)
            p[i] = static_cast<T>(v);
    }

    int                      _type = -1;
    std::shared_ptr<uchar[]> _store;
};

}  // namespace cv

#endif
