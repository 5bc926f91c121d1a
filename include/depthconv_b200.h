/* depthconv_b200 -- C ABI of the B200 (sm_100a) implementation of the
 * depth-conversions hot path: depth map -> feature image stencil filters.
 *
 * This header is the drop-in boundary. Each entry point names the reference
 * interface it replaces; paths are relative to the reference's src/ directory
 * (JonasToth/depth-conversions). Plain pointers and sizes only: no C++ types,
 * no torch types, no exceptions cross this boundary. INTEGRATION.md shows the
 * binding a reference maintainer would add on top of it.
 *
 * Conventions
 *   - Images are single channel, dense, row major: pixel (u, v) is data[v*w+u]
 *     (include/sens_loc/math/image.h:95-108). Frames of a batch lie back to
 *     back: frame f starts at element f*w*h.
 *   - Every function returns a dc_status (0 = ok, < 0 = error) unless stated
 *     otherwise; dc_last_error() gives the message for the last failure of a
 *     context. CUDA errors are sticky per context.
 *   - A context is bound to one GPU and is not re-entrant; use one context per
 *     (host thread, GPU). Distinct contexts are independent.
 *   - There is no CPU fallback: every conversion runs in the CUDA kernels of
 *     this library or fails with an error.
 */
#ifndef DEPTHCONV_B200_H
#define DEPTHCONV_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct dc_ctx dc_ctx;

typedef enum dc_status {
    DC_OK              = 0,
    DC_ERR_INVALID     = -1, /* bad argument (null pointer, unknown enum, ...) */
    DC_ERR_NO_DEVICE   = -2, /* no usable CUDA device / device index out of range */
    DC_ERR_CUDA        = -3, /* a CUDA call or kernel failed (sticky) */
    DC_ERR_NO_CAMERA   = -4, /* conversion requested before dc_set_* */
    DC_ERR_UNSUPPORTED = -5, /* combination not provided by the reference either */
    DC_ERR_NOMEM       = -6
} dc_status;

/* What to compute. One job = one fused kernel launch per chunk of frames. */
typedef enum dc_feature {
    /* conversion::depth_to_laserscan<float, ushort>
     *   include/sens_loc/conversion/depth_to_laserscan.h:70-93
     * and the `depth2x range` converter, apps/depth2x/converter_laserscan.h.inl:1-15.
     *   out[0] = range image */
    DC_RANGE = 0,
    /* conversion::depth_to_bearing<direction>
     *   include/sens_loc/conversion/depth_to_bearing.h:221-248
     * and `depth2x bearing`, apps/depth2x/converter_bearing.h.inl:1-29 (which
     * makes up to four full passes; here all requested directions come out of
     * one pass over the input).
     *   out[0] horizontal, out[1] vertical, out[2] diagonal, out[3] antidiagonal;
     *   NULL = direction not wanted, at least one must be given */
    DC_BEARING = 1,
    /* conversion::depth_to_flexion
     *   include/sens_loc/conversion/depth_to_flexion.h:147-167
     * and `depth2x flexion`, apps/depth2x/converter_flexion.h.inl:1-12.
     *   out[0] = flexion image */
    DC_FLEXION = 2,
    /* conversion::depth_to_gaussian_curvature / depth_to_mean_curvature
     *   include/sens_loc/conversion/depth_to_curvature.h:166-206
     * and `depth2x gauss-curvature|mean-curvature`,
     * apps/depth2x/converter_curvature.h.inl:1-30.
     *   out[0] = Gaussian curvature, out[1] = mean curvature (either may be NULL) */
    DC_CURVATURE = 3,
    /* conversion::depth_to_max_curve
     *   include/sens_loc/conversion/depth_to_max_curve.h:86-153
     * and `depth2x max-curve`, apps/depth2x/converter_max_curve.h.inl:1-13.
     *   out[0] = max-curve image; out[1] (optional) = the range image of the
     *   same pass, i.e. `depth2x range` fused in (BASELINE config 4) */
    DC_MAX_CURVE = 4,
    /* conversion::depth_scaling
     *   include/sens_loc/conversion/depth_scaling.h:21-31
     * and `depth2x scale`, apps/depth2x/converter_scale.cpp:9-19.
     * Needs no camera. in_kind must be DC_IN_U16_RANGE, out_kind DC_OUT_U16.
     *   out[0] = saturate(round(depth * scale + offset)) */
    DC_SCALE = 5
} dc_feature;

/* What the input pixels are (apps/util/batch_converter.h:24-30 depth_type and
 * the element type of math::image). */
typedef enum dc_in_kind {
    DC_IN_U16_DEPTH = 0, /* --type pinhole-depth: orthographic u16, converted to
                            range first (batch_converter.h:157-159) */
    DC_IN_U16_RANGE = 1, /* --type pinhole-range: Euclidean u16, widened to
                            float (batch_converter.h:160) */
    DC_IN_F32_RANGE = 2, /* math::image<float> range image: the argument type of
                            the conversion::depth_to_* library functions */
    DC_IN_F64_RANGE = 3  /* math::image<double> range image: the Real = double
                            instantiation (src/unittest/conversion/intrinsic.h:8-26);
                            requires DC_OUT_F64 */
} dc_in_kind;

/* What the output pixels are. */
typedef enum dc_out_kind {
    DC_OUT_F32 = 0, /* the float feature image the library functions return */
    DC_OUT_U16 = 1, /* the 16 bit image depth2x writes: convert_bearing<float,
                       ushort> (depth_to_bearing.h:282-301), convert_flexion
                       <ushort> (depth_to_flexion.h:194-215), convert_max_curve
                       <ushort> (depth_to_max_curve.h:156-176), curvature_to_image
                       <ushort> with both clamps (depth_to_curvature.h:253-283),
                       convertTo(CV_16U) for range (converter_laserscan.h.inl:9-10).
                       Accuracy: within +-1 of the reference's 16 bit image (range
                       and the curvatures: equal). The fused kernels use that
                       margin: bearing / max-curve fold scale and rounding into
                       one operation and take the ratio with an unrefined
                       reciprocal (<= 0.8 LSB at the acos clamp), flexion divides
                       by the norms with a 2^-22 reciprocal square root (< 0.2 LSB).
                       DC_EXACT_U16=1 in the environment selects the bit-exact
                       flexion kernel; quantising a DC_OUT_F32 image with
                       dc_quantise_host reproduces the reference's two-step result. */
    DC_OUT_F64 = 2  /* Real = double: every operation of the path in double, as the
                       reference's templates instantiated with double compute it
                       (camera from dc_set_*_f64, or the float parameters widened).
                       Input DC_IN_U16_DEPTH / DC_IN_U16_RANGE / DC_IN_F64_RANGE.
                       Not for DC_SCALE. */
} dc_out_kind;

#define DC_MAX_OUTPUTS 4

typedef struct dc_job {
    int32_t     feature;  /* dc_feature */
    int32_t     in_kind;  /* dc_in_kind */
    int32_t     out_kind; /* dc_out_kind */
    int32_t     n_frames; /* frames in `in` and in every non-NULL out[] */
    const void* in;       /* n_frames * w * h elements of the in_kind type */
    void*       out[DC_MAX_OUTPUTS]; /* per-feature slots, see dc_feature */
    float       clamp_lo; /* curvature, DC_OUT_U16: --lower-bound (default -20) */
    float       clamp_hi; /* curvature, DC_OUT_U16: --upper-bound (default +20) */
    double      scale;    /* DC_SCALE: --factor */
    double      offset;   /* DC_SCALE: --delta */
    int32_t     median_ksize; /* 0 = none; 3 or 5: preprocess::median_blur
                                 (preprocess/filter.h:143-157) applied to the u16
                                 input frames on the device before the conversion
                                 (what `depth_filter median-blur -d K` followed by
                                 depth2x does, scripts/analysis/transformation_step.py:80-96,
                                 without the PNG round trip). u16 input kinds only. */
} dc_job;

/* Library / device discovery. */
const char* dc_version(void);
int         dc_device_count(void); /* >= 0, or a negative dc_status */

/* Create / destroy a context on GPU `device`. Owns all device memory, streams
 * and pinned staging buffers it uses. */
int  dc_create(int device, dc_ctx** ctx);
void dc_destroy(dc_ctx* ctx);

/* Message of the last error on this context ("" if none). Never NULL. With a
 * NULL ctx: the message of the last failed dc_create on this thread. */
const char* dc_last_error(const dc_ctx* ctx);

/* Install the camera model; replaces the construction of
 * camera_models::pinhole<float>(w, h, fx, fy, cx, cy)
 *   include/sens_loc/camera_models/pinhole.h:26-50
 * and hoists everything that depends on the camera only out of the per-frame
 * path: unit rays (pinhole.h:142-171), the orthographic->range factor
 * (conversion/util.h:41-51) and the ray angles phi() and their cosines
 * (camera_models/utility.h:15-49) are computed once, with the host's libm, and
 * kept in device tables. Same preconditions as the reference constructor
 * (w, h > 0; fx, fy, cx, cy > 0): violated -> DC_ERR_INVALID. */
int dc_set_pinhole(dc_ctx* ctx, int w, int h, float fx, float fy, float cx, float cy);

/* camera_models::equirectangular<float>(w, h, {theta_min, theta_max})
 *   include/sens_loc/camera_models/equirectangular.h:43-153
 * (`depth2x -m equirectangular`). For this model the range conversion is the
 * identity (conversion/util.h:57-67). */
int dc_set_equirectangular(dc_ctx* ctx, int w, int h, float theta_min, float theta_max);

/* The same two constructors for Real = double (camera_models::pinhole<double>,
 * equirectangular<double>): the parameters are kept in double for DC_OUT_F64
 * jobs; float jobs on this context see them narrowed to float. */
int dc_set_pinhole_f64(dc_ctx* ctx, int w, int h, double fx, double fy, double cx, double cy);
int dc_set_equirectangular_f64(dc_ctx* ctx, int w, int h, double theta_min, double theta_max);

/* Frame size only, no camera model: all that DC_SCALE needs (`depth2x scale`
 * takes no --calibration; apps/depth2x/main.cpp:338-352). Clears any camera. */
int dc_set_frame_size(dc_ctx* ctx, int w, int h);

/* Size of the installed camera / frame size (0 if none). */
int dc_width(const dc_ctx* ctx);
int dc_height(const dc_ctx* ctx);

/* Run a job whose `in` and `out[]` pointers are DEVICE memory on this
 * context's GPU. Asynchronous: enqueued on `cuda_stream` (a cudaStream_t /
 * CUstream, NULL = the context's own stream); returns after the launch.
 * Buffers must stay valid until the stream reaches this point. */
int dc_convert_device(dc_ctx* ctx, const dc_job* job, void* cuda_stream);

/* Run a job whose `in` and `out[]` pointers are HOST memory. Synchronous: on
 * return all outputs are written. Frames are processed in chunks through a
 * double-buffered H2D -> kernel -> D2H pipeline on the context's streams.
 * Pinned memory (dc_host_alloc, cudaHostAlloc, torch pin_memory) is copied
 * straight by DMA; pageable memory is staged through the context's pinned ring.
 * This is the call the conversion::depth_to_* drop-in wrappers and the depth2x
 * frame pipeline make. */
int dc_convert_host(dc_ctx* ctx, const dc_job* job);

/* The stand-alone quantisers of the library API: float feature image -> the
 * 16 bit image depth2x writes. (The depth2x path never calls this: DC_OUT_U16
 * fuses the same arithmetic into the feature kernel.) */
typedef enum dc_quant {
    DC_Q_RANGE     = 0, /* cv::Mat::convertTo(CV_16U): apps/depth2x/converter_laserscan.h.inl:9-10 */
    DC_Q_BEARING   = 1, /* convert_bearing<float, ushort>: conversion/depth_to_bearing.h:282-301 */
    DC_Q_FLEXION   = 2, /* convert_flexion<ushort>: conversion/depth_to_flexion.h:194-215 */
    DC_Q_CURVATURE = 3, /* curvature_to_image<ushort> with both clamps:
                           conversion/depth_to_curvature.h:253-283; `mask` = the
                           range image whose zeros blank the result */
    DC_Q_MAX_CURVE = 4  /* convert_max_curve<ushort>: conversion/depth_to_max_curve.h:156-176 */
} dc_quant;

/* HOST buffers, synchronous. `mask` is read for DC_Q_CURVATURE only (n_px
 * floats), clamp_lo < clamp_hi likewise. Needs no camera. 16 bit output, both
 * clamps: what depth2x uses (converter_curvature.h.inl:9-10). */
int dc_quantise_host(dc_ctx* ctx, int quant, const float* feature, const float* mask,
                     size_t n_px, uint16_t* out, float clamp_lo, float clamp_hi);

/* Pixel type of a quantised image: the PixelType template argument of
 * convert_bearing / convert_flexion / convert_max_curve (scale and offset from
 * detail::scaling_factor<Real, PixelType>, conversion/util.h:71-79; the reference's
 * unit tests compare 8 bit images, src/unittest/conversion/
 * test_conversion_bearing.cpp:78-159). */
typedef enum dc_pixel { DC_PX_U16 = 0, DC_PX_U8 = 1, DC_PX_I8 = 2, DC_PX_I16 = 3 } dc_pixel;

/* The general quantiser. real_kind: DC_OUT_F32 (feature / mask are float) or
 * DC_OUT_F64 (double). `out` holds n_px pixels of `pixel` type. clamp_lo / clamp_hi
 * are read for DC_Q_CURVATURE only and may each be NULL = std::nullopt of
 * curvature_to_image (conversion/depth_to_curvature.h:85-92): a missing bound is the
 * minimum / maximum over the whole curvature image (std::minmax_element, :218-228),
 * found by a reduction on the device. DC_Q_CURVATURE and DC_Q_RANGE produce 16 bit
 * images only (DC_ERR_UNSUPPORTED otherwise); DC_ERR_INVALID unless lower < upper. */
int dc_quantise_host_ex(dc_ctx* ctx, int quant, int pixel, int real_kind, const void* feature,
                        const void* mask, size_t n_px, void* out, const double* clamp_lo,
                        const double* clamp_hi);

/* curvature_to_image<ushort> WITHOUT clamps (the bounds are the image's own minimum
 * and maximum, as dc_quantise_host_ex with both clamps NULL) for callers that must
 * not allocate: code under CUDA graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc. This entry point has NO counterpart in the reference.
 * It needs no dc_ctx and calls no
 * cudaMalloc / cudaFree / cudaHostAlloc or allocating library routine: the caller
 * passes DEVICE buffers on GPU `device` (n_px floats of curvature, n_px floats of
 * range, n_px 16 bit pixels of result), and the reduction's scratch is a fixed
 * file-scope __device__ array in the library (33 KB, held by every GPU that loads
 * the library until exit), which bounds the image: n_px > DC_NOALLOC_MAX_PIXELS is
 * rejected with DC_ERR_INVALID. Enqueued on `cuda_stream` (NULL = the default
 * stream), then waits for it; makes `device` the calling thread's current device.
 * DC_ERR_INVALID with nothing written unless minimum < maximum. On failure a
 * message is copied into err[0 .. err_len) when err is not NULL.
 * ALL calls of one process on one GPU share the scratch, which keeps the last
 * call's contents (each call clears the words that must start at zero): calls that
 * overlap on the GPU (two streams, two host threads) race on it; serialise them. */
#define DC_NOALLOC_MAX_PIXELS 16777216u /* 4096 x 4096 */
int dc_curvature_to_image_noalloc(int device, const float* curvature_device,
                                  const float* range_device, size_t n_px, uint16_t* out_device,
                                  void* cuda_stream, char* err, size_t err_len);

/* The same for Real = double images (convert_bearing<double, ushort> etc.). */
int dc_quantise_host_f64(dc_ctx* ctx, int quant, const double* feature, const double* mask,
                         size_t n_px, uint16_t* out, double clamp_lo, double clamp_hi);

/* preprocess::median_blur<PixelType>(image, ksize) = cv::medianBlur, ksize 3 or 5,
 *   include/sens_loc/preprocess/filter.h:143-157 (apps/depth_filter `median-blur`)
 * on n_frames frames of the installed frame size. elem_kind: DC_IN_U16_RANGE
 * (u16 pixels) or DC_IN_F32_RANGE (float pixels); `out` has the type of `in`
 * and must not alias it. HOST buffers, synchronous / DEVICE buffers, enqueued. */
int dc_median_host(dc_ctx* ctx, int elem_kind, const void* in, int n_frames, int ksize, void* out);
int dc_median_device(dc_ctx* ctx, int elem_kind, const void* in, int n_frames, int ksize,
                     void* out, void* cuda_stream);

/* preprocess::gaussian_blur<PixelType>(image, ksize, sigmaX, sigmaY) = cv::GaussianBlur
 *   include/sens_loc/preprocess/filter.h:125-136
 * on n_frames frames of the installed frame size, BORDER_REFLECT_101. ksize_w / ksize_h:
 * positive and odd (up to 31), or <= 0 to derive the size from sigma as OpenCV does;
 * sigma_y <= 0 means sigma_x; sigma <= 0 means the kernel OpenCV derives from the size.
 * elem_kind DC_IN_U16_RANGE or DC_IN_F32_RANGE as for the median; `out` has the type of
 * `in` and must not alias it. The arithmetic is OpenCV's (not vendored by the reference):
 * float results agree with cv2 to 1e-6 relative, u16 results to +-1. */
int dc_gaussian_blur_host(dc_ctx* ctx, int elem_kind, const void* in, int n_frames, int ksize_w,
                          int ksize_h, double sigma_x, double sigma_y, void* out);
int dc_gaussian_blur_device(dc_ctx* ctx, int elem_kind, const void* in, int n_frames, int ksize_w,
                            int ksize_h, double sigma_x, double sigma_y, void* out,
                            void* cuda_stream);

/* preprocess::bilateral_filter<PixelType>(image, sigma_color, proximity) = cv::bilateralFilter
 * on the image converted to float
 *   include/sens_loc/preprocess/filter.h:49-82, apps/depth_filter/filter_functor.h:24-50
 * `proximity` as the reference's variant: d > 0 is the pixel diameter (sigma_space then
 * only shapes the spatial weights), d <= 0 derives the radius from sigma_space. sigma <= 0
 * means 1 (OpenCV 4.3.0). Radius up to 15. u16 frames are converted to float, filtered and
 * converted back with saturate_cast (math::convert, math/image.h:117-128). The input must
 * hold no NaN. Agrees with cv2 to 2e-5 of the frame's value range. */
int dc_bilateral_host(dc_ctx* ctx, int elem_kind, const void* in, int n_frames, int d,
                      double sigma_color, double sigma_space, void* out);
int dc_bilateral_device(dc_ctx* ctx, int elem_kind, const void* in, int n_frames, int d,
                        double sigma_color, double sigma_space, void* out, void* cuda_stream);

/* Block until all work this context enqueued on its own streams has finished. */
int dc_synchronize(dc_ctx* ctx);

/* Pinned host memory for zero-staging dc_convert_host calls. */
void* dc_host_alloc(size_t bytes);
void  dc_host_free(void* p);

/* Synthetic Kinect-like orthographic depth frames (DESIGN.md "Synthetic
 * input"), generated on the device: frames [frame0, frame0 + n_frames) of the
 * stream identified by `seed`, w x h taken from the installed camera.
 * Benchmark / test tooling; bit-identical to oracle/dc_oracle.c dco_synth_depth. */
int dc_synth_depth_device(dc_ctx* ctx, uint64_t seed, int frame0, int n_frames,
                          uint16_t* out_device, void* cuda_stream);

/* Number of kernels of this library launched on this context since creation
 * (for the benchmark's gpu_launches claim). */
uint64_t dc_kernel_launches(const dc_ctx* ctx);

/* Name of the kernel family the most recent dc_convert_* of this context launched
 * for its (last) chunk, e.g. "dc::pr::maxcurve_edge_kernel"; "" before the first
 * conversion. The string is static. Test / benchmark introspection: which kernel
 * serves a job depends on the camera (per-pixel or scalar angle tables), the
 * input type, the frame width and buffer alignment. */
const char* dc_last_kernel(const dc_ctx* ctx);

/* DC_IN_F32_RANGE (what sens_loc::conversion::depth_to_*(math::image<float>) sends,
 * depth_to_flexion.h:31-33, depth_to_bearing.h:44-49) runs on the same guard-free
 * frame-pair kernels as u16 input while every non-zero value lies in [2^-10, 2^22]
 * (the window the exact division / square root sequences are proven for). The
 * kernels check that on the values they load; a batch that holds anything else
 * (negative, NaN, infinite, out of the window) is converted again, inside the same
 * call, by the kernels with the guarded IEEE intrinsics. Returns how many batches of
 * this context needed that (synchronises the device). Results are identical either
 * way; this is introspection for tests and benchmarks. */
uint64_t dc_f32_fallbacks(dc_ctx* ctx);

#ifdef __cplusplus
}
#endif
#endif /* DEPTHCONV_B200_H */
