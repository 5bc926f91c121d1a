// C ABI of libdepthconv_b200.so: context, per-camera device tables, kernel
// dispatch and the host-buffer pipeline. See include/depthconv_b200.h.
#include "../../include/depthconv_b200.h"
#include "dc_launch.h"
#include "dc_filters.cuh"
#include "dc_median.cuh"
#include "dc_misc_kernels.cuh"
#include "dc_f64.h"
#include "dc_tables.h"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <map>
#include <string>
#include <vector>

#define DC_VERSION_STRING "depthconv_b200 0.1.0 (sm_100a)"

namespace {

thread_local std::string g_create_error;

constexpr int    kSlots            = 3;
constexpr size_t kChunkDefaultBytes = size_t(48) << 20;  // in + out bytes per chunk
// (tuning knob: DC_CHUNK_MB in the environment, read once)
const size_t kChunkTargetBytes = [] {
    const char* e = std::getenv("DC_CHUNK_MB");
    const long  n = e != nullptr ? std::atol(e) : 0;
    return n > 0 ? size_t(n) << 20 : kChunkDefaultBytes;
}();

struct dev_buf {
    void*  p   = nullptr;
    size_t cap = 0;
};

struct slot {
    cudaStream_t stream = nullptr;
    cudaEvent_t  done   = nullptr;
    dev_buf      d_in, d_out[DC_MAX_OUTPUTS];
    dev_buf      h_in, h_out[DC_MAX_OUTPUTS];  // pinned staging for pageable callers
    bool         busy = false;
    // pending copy-out to pageable user memory
    void*  user_out[DC_MAX_OUTPUTS] = {nullptr, nullptr, nullptr, nullptr};
    size_t out_bytes                = 0;
};

struct dev_table {
    float* p = nullptr;
};

}  // namespace

struct dc_ctx {
    int          device = 0;
    cudaStream_t stream = nullptr;  // default stream for dc_convert_device(NULL)
    std::string  err;
    bool         sticky = false;
    dc::tables   tabs;
    unsigned     uploaded = 0;  // table_need bits present on the device
    dev_table    d_range_factor, d_ray[3], d_cos[4], d_curv[3];
    slot         slots[kSlots];
    uint64_t     launches = 0;
    const char*  last_kernel = "";  // kernel family of the most recent launch
    // Real = double (dc_f64.h): camera parameters in double, tables on demand
    dc64::camera cam64;
    dc64::state* f64 = nullptr;
    // filtered copy of the input frames (dc_job::median_ksize), one per stream:
    // the chunks of dc_convert_host run concurrently on the slots' streams
    std::map<cudaStream_t, dev_buf> d_median;
    // per-frame scratch of the bilateral filter (min / max, table), one per stream
    std::map<cudaStream_t, dev_buf> d_bilateral;
    // f32 range input on the guard-free kernels (dc_kernels.cuh kparams::guard):
    // a ring of flags, one per launch in flight, and the count of batches the IEEE
    // kernel had to redo (dc_f32_fallbacks)
    static constexpr int kGuards = 256;
    int*                 d_guard     = nullptr;
    unsigned long long*  d_fallbacks = nullptr;
    unsigned             next_guard  = 0;
};

namespace {

int fail(dc_ctx* c, int code, const std::string& msg) {
    if (c != nullptr)
        c->err = msg;
    else
        g_create_error = msg;
    return code;
}

int cuda_fail(dc_ctx* c, cudaError_t e, const char* what) {
    if (c != nullptr)
        c->sticky = true;
    return fail(c, DC_ERR_CUDA,
                std::string(what) + ": " + cudaGetErrorName(e) + " (" +
                    cudaGetErrorString(e) + ")");
}

#define DC_CUDA(ctx, call)                                 \
    do {                                                   \
        const cudaError_t e__ = (call);                    \
        if (e__ != cudaSuccess)                            \
            return cuda_fail((ctx), e__, #call);           \
    } while (0)

void free_dev(dev_buf& b) {
    if (b.p != nullptr)
        cudaFree(b.p);
    b = dev_buf{};
}
void free_pinned(dev_buf& b) {
    if (b.p != nullptr)
        cudaFreeHost(b.p);
    b = dev_buf{};
}
cudaError_t reserve_dev(dev_buf& b, size_t bytes) {
    if (b.cap >= bytes)
        return cudaSuccess;
    free_dev(b);
    const cudaError_t e = cudaMalloc(&b.p, bytes);
    if (e == cudaSuccess)
        b.cap = bytes;
    return e;
}
cudaError_t reserve_pinned(dev_buf& b, size_t bytes) {
    if (b.cap >= bytes)
        return cudaSuccess;
    free_pinned(b);
    const cudaError_t e = cudaHostAlloc(&b.p, bytes, cudaHostAllocDefault);
    if (e == cudaSuccess)
        b.cap = bytes;
    return e;
}

void drop_device_tables(dc_ctx* c) {
    auto drop = [](dev_table& t) {
        if (t.p != nullptr)
            cudaFree(t.p);
        t.p = nullptr;
    };
    drop(c->d_range_factor);
    for (auto& t : c->d_ray)
        drop(t);
    for (auto& t : c->d_cos)
        drop(t);
    for (auto& t : c->d_curv)
        drop(t);
    c->uploaded = 0;
}

int upload(dc_ctx* c, dev_table& dst, const std::vector<float>& src) {
    const size_t bytes = src.size() * sizeof(float);
    if (dst.p == nullptr)
        DC_CUDA(c, cudaMalloc(reinterpret_cast<void**>(&dst.p), bytes));
    DC_CUDA(c, cudaMemcpy(dst.p, src.data(), bytes, cudaMemcpyHostToDevice));
    return DC_OK;
}
int upload(dc_ctx* c, dev_table& dst, const dc::table& src) {
    if (src.uniform)
        return DC_OK;  // passed by value
    return upload(c, dst, src.v);
}

int ensure_tables(dc_ctx* c, unsigned need) {
    const unsigned missing = need & ~c->uploaded;
    if (missing == 0)
        return DC_OK;
    dc::build_tables(c->tabs, missing);
    int rc = DC_OK;
    if ((missing & dc::NEED_RANGE_FACTOR) != 0)
        rc = upload(c, c->d_range_factor, c->tabs.range_factor);
    if (rc == DC_OK && (missing & dc::NEED_RAYS) != 0) {
        rc = upload(c, c->d_ray[0], c->tabs.ray_x);
        if (rc == DC_OK)
            rc = upload(c, c->d_ray[1], c->tabs.ray_y);
        if (rc == DC_OK)
            rc = upload(c, c->d_ray[2], c->tabs.ray_z);
    }
    if (rc == DC_OK && (missing & dc::NEED_COS_PRIOR) != 0)
        for (int d = 0; d < 4 && rc == DC_OK; ++d)
            rc = upload(c, c->d_cos[d], c->tabs.cos_prior[d]);
    if (rc == DC_OK && (missing & dc::NEED_CURV_ANGLES) != 0) {
        rc = upload(c, c->d_curv[0], c->tabs.d_phi);
        if (rc == DC_OK)
            rc = upload(c, c->d_curv[1], c->tabs.d_theta);
        if (rc == DC_OK)
            rc = upload(c, c->d_curv[2], c->tabs.d_phi_theta);
    }
    // The sources are pageable: cudaMemcpy may return once they are staged, before
    // the DMA has landed, and the kernels run on non-blocking streams that are not
    // ordered after the legacy stream. Wait for the copies before anything launches.
    if (rc == DC_OK)
        DC_CUDA(c, cudaDeviceSynchronize());
    if (rc == DC_OK)
        c->uploaded |= missing;
    return rc;
}

dc::tab_ref ref_of(const dc::table& t, const dev_table& d) {
    return t.uniform ? dc::tab_ref{nullptr, t.value} : dc::tab_ref{d.p, 0.0f};
}

unsigned tables_needed(int feature, int in_kind) {
    unsigned need = 0;
    if (in_kind == DC_IN_U16_DEPTH)
        need |= dc::NEED_RANGE_FACTOR;
    switch (feature) {
    case DC_BEARING:
    case DC_MAX_CURVE: need |= dc::NEED_COS_PRIOR; break;
    case DC_FLEXION: need |= dc::NEED_RAYS; break;
    case DC_CURVATURE: need |= dc::NEED_CURV_ANGLES; break;
    default: break;
    }
    return need;
}

size_t in_elem_bytes(int in_kind) {
    return in_kind == DC_IN_F64_RANGE ? 8 : in_kind == DC_IN_F32_RANGE ? 4 : 2;
}
size_t out_elem_bytes(int out_kind) {
    return out_kind == DC_OUT_F64 ? 8 : out_kind == DC_OUT_F32 ? 4 : 2;
}

int validate_job(dc_ctx* c, const dc_job* j) {
    if (j == nullptr)
        return fail(c, DC_ERR_INVALID, "job is NULL");
    if (j->feature < DC_RANGE || j->feature > DC_SCALE)
        return fail(c, DC_ERR_INVALID, "unknown feature");
    if (j->in_kind < DC_IN_U16_DEPTH || j->in_kind > DC_IN_F64_RANGE)
        return fail(c, DC_ERR_INVALID, "unknown in_kind");
    if (j->out_kind < DC_OUT_F32 || j->out_kind > DC_OUT_F64)
        return fail(c, DC_ERR_INVALID, "unknown out_kind");
    // Real = double: double out from u16 or double in; nothing mixed with float
    if (j->in_kind == DC_IN_F64_RANGE && j->out_kind != DC_OUT_F64)
        return fail(c, DC_ERR_UNSUPPORTED, "DC_IN_F64_RANGE needs DC_OUT_F64");
    if (j->out_kind == DC_OUT_F64 && j->in_kind == DC_IN_F32_RANGE)
        return fail(c, DC_ERR_UNSUPPORTED, "double output from a float range image is not defined");
    if (j->out_kind == DC_OUT_F64 && j->feature == DC_SCALE)
        return fail(c, DC_ERR_UNSUPPORTED, "scale is u16 -> u16 (depth_scaling.h:21-31)");
    if (j->median_ksize != 0) {
        if (j->median_ksize != 3 && j->median_ksize != 5)
            return fail(c, DC_ERR_INVALID, "median_ksize must be 0, 3 or 5 (filter_functor.h:57-60)");
        if (j->in_kind != DC_IN_U16_DEPTH && j->in_kind != DC_IN_U16_RANGE)
            return fail(c, DC_ERR_UNSUPPORTED, "the fused median pre-filter takes u16 input frames");
    }
    if (j->n_frames < 0)
        return fail(c, DC_ERR_INVALID, "n_frames < 0");
    if (j->n_frames > 0 && j->in == nullptr)
        return fail(c, DC_ERR_INVALID, "input pointer is NULL");
    bool any = false;
    for (int k = 0; k < DC_MAX_OUTPUTS; ++k)
        any = any || j->out[k] != nullptr;
    if (!any)
        return fail(c, DC_ERR_INVALID, "no output requested");
    switch (j->feature) {
    case DC_RANGE:
    case DC_FLEXION:
        if (j->out[0] == nullptr || j->out[1] || j->out[2] || j->out[3])
            return fail(c, DC_ERR_INVALID, "feature has exactly one output: out[0]");
        break;
    case DC_CURVATURE:
        if (j->out[2] || j->out[3])
            return fail(c, DC_ERR_INVALID, "curvature outputs are out[0] (gauss), out[1] (mean)");
        if (j->out_kind == DC_OUT_U16 && !(j->clamp_lo < j->clamp_hi))
            return fail(c, DC_ERR_INVALID, "curvature clamp needs lower bound < upper bound");
        break;
    case DC_MAX_CURVE:
        if (j->out[0] == nullptr || j->out[2] || j->out[3])
            return fail(c, DC_ERR_INVALID, "max-curve outputs are out[0], optional out[1] (range)");
        break;
    case DC_SCALE:
        if (j->out[0] == nullptr || j->out[1] || j->out[2] || j->out[3])
            return fail(c, DC_ERR_INVALID, "scale has exactly one output: out[0]");
        if (j->in_kind != DC_IN_U16_RANGE || j->out_kind != DC_OUT_U16)
            return fail(c, DC_ERR_UNSUPPORTED, "scale is u16 -> u16 (depth_scaling.h:21-31)");
        break;
    default: break;
    }
    if (j->feature != DC_SCALE && c->tabs.cam.model == dc::camera_model::none)
        return fail(c, DC_ERR_NO_CAMERA, "no camera model installed (dc_set_pinhole)");
    return DC_OK;
}

bool aligned(const void* p, size_t a) { return (reinterpret_cast<uintptr_t>(p) % a) == 0; }

// Enqueue the fused kernel(s) for `job` (device pointers) on stream `s`.
int launch_job(dc_ctx* c, const dc_job* j, cudaStream_t s) {
    if (j->n_frames == 0)
        return DC_OK;
    dc::kparams p{};
    int         w = 0, h = 0;
    w = c->tabs.cam.w;
    h = c->tabs.cam.h;
    if (w <= 0 || h <= 0)
        return fail(c, DC_ERR_NO_CAMERA, "frame size unknown: dc_set_pinhole / dc_set_frame_size");
    dc_job filtered;
    if (j->median_ksize != 0) {
        const size_t bytes = size_t(j->n_frames) * size_t(w) * h * sizeof(uint16_t);
        dev_buf&     tmp   = c->d_median[s];
        if (tmp.cap < bytes) {
            // (a growing buffer is freed only after everything that reads it ran)
            DC_CUDA(c, cudaStreamSynchronize(s));
            DC_CUDA(c, reserve_dev(tmp, bytes));
        }
        const cudaError_t e = dc::med::launch_median(
            static_cast<const unsigned short*>(j->in), static_cast<unsigned short*>(tmp.p), w, h,
            j->n_frames, j->median_ksize, s);
        if (e != cudaSuccess)
            return cuda_fail(c, e, "median kernel launch");
        c->launches += uint64_t((j->n_frames + 65534) / 65535);
        filtered              = *j;
        filtered.in           = tmp.p;
        filtered.median_ksize = 0;
        j                     = &filtered;
    }
    if (j->out_kind == DC_OUT_F64) {
        if (j->feature == DC_SCALE)
            return fail(c, DC_ERR_UNSUPPORTED, "scale is u16 -> u16");
        if (c->f64 == nullptr) {
            c->f64 = dc64::create();
            dc64::set_camera(c->f64, c->cam64);
        }
        double* outs[4];
        for (int k = 0; k < 4; ++k)
            outs[k] = static_cast<double*>(j->out[k]);
        std::string       msg;
        const cudaError_t e = dc64::convert(c->f64, j->feature, j->in_kind, j->in, j->n_frames, outs,
                                            s, msg, c->launches);
        if (e != cudaSuccess)
            return cuda_fail(c, e, msg.c_str());
        c->last_kernel = "dc64 (Real = double, one thread per pixel)";
        return DC_OK;
    }
    int in_kind = j->in_kind;
    if (j->feature != DC_SCALE) {
        // equirectangular: the range conversion is the identity (util.h:57-67)
        if (in_kind == DC_IN_U16_DEPTH && c->tabs.cam.model == dc::camera_model::equirectangular)
            in_kind = DC_IN_U16_RANGE;
        const int rc = ensure_tables(c, tables_needed(j->feature, in_kind));
        if (rc != DC_OK)
            return rc;
    }
    p.in = j->in;
    for (int k = 0; k < 4; ++k)
        p.out[k] = j->out[k];
    p.w            = w;
    p.h            = h;
    p.frame_px     = (long long)w * h;
    p.range_factor = ref_of(c->tabs.range_factor, c->d_range_factor);
    p.ray_x        = c->d_ray[0].p;
    p.ray_y        = c->d_ray[1].p;
    p.ray_z        = c->d_ray[2].p;
    for (int d = 0; d < 4; ++d)
        p.cosp[d] = ref_of(c->tabs.cos_prior[d], c->d_cos[d]);
    p.d_phi       = ref_of(c->tabs.d_phi, c->d_curv[0]);
    p.d_theta     = ref_of(c->tabs.d_theta, c->d_curv[1]);
    p.d_phi_theta = ref_of(c->tabs.d_phi_theta, c->d_curv[2]);
    if (c->tabs.d_phi.uniform && c->tabs.d_theta.uniform && c->tabs.d_phi_theta.uniform) {
        // math/derivatives.h:66-93: the divisors of the five central differences
        // (float arithmetic, no contraction: -ffp-contract=off)
        const volatile float phi = c->tabs.d_phi.value, the = c->tabs.d_theta.value,
                             pt = c->tabs.d_phi_theta.value;
        const float b[5] = {2.0f * phi, 2.0f * the, phi * phi, the * the, pt * pt};
        for (int k = 0; k < 5; ++k) {
            p.cv_b[k] = b[k];
            p.cv_y[k] = 1.0f / b[k];
        }
    }
    p.lo          = j->clamp_lo;
    p.hi          = j->clamp_hi;
    // conversion/util.h:71-79 scaling_factor<float, ushort>(max_angle) =
    // (65535 - 0) / max_angle; pi<float> and float(2. * pi<float>)
    const float pi_f  = float(3.141592653589793238462);
    p.q_bearing       = 65535.0f / pi_f;
    p.q_max_curve     = 65535.0f / float(2. * double(pi_f));
    p.s_scale         = float(j->scale);
    p.s_offset        = float(j->offset);

    // vector paths: rows and frames must keep 8-byte (u16) / 16-byte (f32) alignment
    // for a width that is a multiple of four, half of that for any other even width
    const size_t in_al  = in_kind == DC_IN_F32_RANGE ? 16 : 8;
    const size_t out_al = j->out_kind == DC_OUT_F32 ? 16 : 8;
    dc::launch_info li{};
    li.vec4 = (w % 4) == 0 && aligned(j->in, in_al);
    li.even = (w % 2) == 0 && aligned(j->in, in_al / 2);
    for (int k = 0; k < 4; ++k)
        if (j->out[k] != nullptr) {
            li.vec4 = li.vec4 && aligned(j->out[k], out_al);
            li.even = li.even && aligned(j->out[k], out_al / 2);
        }
    // guard-free exact arithmetic needs the per-camera range proof (dc_tables)
    li.fast = c->tabs.fast_ok;
    if (in_kind == DC_IN_F32_RANGE) {
        if (c->d_guard == nullptr) {
            DC_CUDA(c, cudaMalloc(reinterpret_cast<void**>(&c->d_guard), dc_ctx::kGuards * sizeof(int)));
            DC_CUDA(c, cudaMalloc(reinterpret_cast<void**>(&c->d_fallbacks), sizeof(unsigned long long)));
            DC_CUDA(c, cudaMemset(c->d_fallbacks, 0, sizeof(unsigned long long)));
            DC_CUDA(c, cudaDeviceSynchronize());
        }
        p.guard     = c->d_guard + (c->next_guard++ % dc_ctx::kGuards);
        p.fallbacks = c->d_fallbacks;
    }
    const int   ik = in_kind == DC_IN_U16_DEPTH ? dc::IN_U16_DEPTH
                     : in_kind == DC_IN_U16_RANGE ? dc::IN_U16_RANGE : dc::IN_F32_RANGE;
    const int   ok = j->out_kind == DC_OUT_F32 ? dc::OUT_F32 : dc::OUT_U16;
    cudaError_t e  = cudaSuccess;
    switch (j->feature) {
    case DC_RANGE: e = dc::launch_range(p, ik, ok, j->n_frames, s, li); break;
    case DC_BEARING: e = dc::launch_bearing(p, ik, ok, j->n_frames, s, li); break;
    case DC_FLEXION: e = dc::launch_flexion(p, ik, ok, j->n_frames, s, li); break;
    case DC_CURVATURE: e = dc::launch_curvature(p, ik, ok, j->n_frames, s, li); break;
    case DC_MAX_CURVE: e = dc::launch_max_curve(p, ik, ok, j->n_frames, s, li); break;
    case DC_SCALE: e = dc::launch_scale(p, j->n_frames, s, li); break;
    default: return fail(c, DC_ERR_INVALID, "unknown feature");
    }
    if (e != cudaSuccess)
        return cuda_fail(c, e, "kernel launch");
    c->launches += uint64_t(li.launches);
    c->last_kernel = li.kernel;
    return DC_OK;
}

bool is_pinned(const void* p) {
    cudaPointerAttributes a{};
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return a.type == cudaMemoryTypeHost;
}

int finish_slot(dc_ctx* c, slot& s) {
    if (!s.busy)
        return DC_OK;
    DC_CUDA(c, cudaEventSynchronize(s.done));
    for (int k = 0; k < DC_MAX_OUTPUTS; ++k)
        if (s.user_out[k] != nullptr) {
            std::memcpy(s.user_out[k], s.h_out[k].p, s.out_bytes);
            s.user_out[k] = nullptr;
        }
    s.busy = false;
    return DC_OK;
}

}  // namespace

namespace {
template <typename Real, typename Px>
cudaError_t run_quantise(int quant, const void* in, const void* mask, size_t n, void* out,
                         double max_value, double lo, double hi, cudaStream_t s) {
    // conversion/util.h:71-79 scaling_factor<Real, PixelType>(max): both factors
    // computed in Real, as the reference does
    const Real pmin = Real(dc::px_limits<Px>::lo), pmax = Real(dc::px_limits<Px>::hi);
    const Real alpha = (pmax - pmin) / Real(max_value), offset = pmin;
    const unsigned blocks = unsigned(std::min<size_t>((n + 255) / 256, 148 * 16));
    dc::quantise_kernel<Real, Px><<<blocks, 256, 0, s>>>(
        quant, static_cast<const Real*>(in), static_cast<const Real*>(mask), n,
        static_cast<Px*>(out), alpha, offset, Real(lo), Real(hi));
    return cudaGetLastError();
}
template <typename Real>
cudaError_t run_quantise_px(int quant, int pixel, const void* in, const void* mask, size_t n,
                            void* out, double max_value, double lo, double hi, cudaStream_t s) {
    switch (pixel) {
    case DC_PX_U8: return run_quantise<Real, unsigned char>(quant, in, mask, n, out, max_value, lo, hi, s);
    case DC_PX_I8: return run_quantise<Real, signed char>(quant, in, mask, n, out, max_value, lo, hi, s);
    case DC_PX_I16: return run_quantise<Real, short>(quant, in, mask, n, out, max_value, lo, hi, s);
    default: return run_quantise<Real, unsigned short>(quant, in, mask, n, out, max_value, lo, hi, s);
    }
}
// decode of dc::ordered<Real>::enc
float dec_ordered(unsigned e) {
    const unsigned u = (e & 0x80000000u) ? (e & 0x7fffffffu) : ~e;
    float f;
    std::memcpy(&f, &u, sizeof f);
    return f;
}
double dec_ordered(unsigned long long e) {
    const unsigned long long u = (e >> 63) ? (e & 0x7fffffffffffffffull) : ~e;
    double f;
    std::memcpy(&f, &u, sizeof f);
    return f;
}
}  // namespace

// ===========================================================================
extern "C" {

const char* dc_version(void) { return DC_VERSION_STRING; }

int dc_device_count(void) {
    int               n = 0;
    const cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) {
        cudaGetLastError();
        g_create_error = std::string("cudaGetDeviceCount: ") + cudaGetErrorString(e);
        return DC_ERR_NO_DEVICE;
    }
    return n;
}

int dc_create(int device, dc_ctx** out) {
    if (out == nullptr)
        return fail(nullptr, DC_ERR_INVALID, "ctx out pointer is NULL");
    *out        = nullptr;
    const int n = dc_device_count();
    if (n <= 0)
        return fail(nullptr, DC_ERR_NO_DEVICE,
                    "no CUDA device available (this library has no CPU fallback): " +
                        g_create_error);
    if (device < 0 || device >= n)
        return fail(nullptr, DC_ERR_NO_DEVICE, "device index out of range");
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess)
        return cuda_fail(nullptr, e, "cudaSetDevice");
    cudaDeviceProp prop{};
    e = cudaGetDeviceProperties(&prop, device);
    if (e != cudaSuccess)
        return cuda_fail(nullptr, e, "cudaGetDeviceProperties");
    if (prop.major != 10)
        return fail(nullptr, DC_ERR_NO_DEVICE,
                    std::string("device '") + prop.name +
                        "' is not sm_100 (this library carries sm_100a code only)");
    dc_ctx* c = new (std::nothrow) dc_ctx();
    if (c == nullptr)
        return fail(nullptr, DC_ERR_NOMEM, "out of host memory");
    c->device = device;
    e         = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking);
    for (int i = 0; i < kSlots && e == cudaSuccess; ++i) {
        e = cudaStreamCreateWithFlags(&c->slots[i].stream, cudaStreamNonBlocking);
        if (e == cudaSuccess)
            e = cudaEventCreateWithFlags(&c->slots[i].done, cudaEventDisableTiming);
    }
    if (e != cudaSuccess) {
        const int rc = cuda_fail(nullptr, e, "stream/event creation");
        dc_destroy(c);
        return rc;
    }
    *out = c;
    return DC_OK;
}

void dc_destroy(dc_ctx* c) {
    if (c == nullptr)
        return;
    cudaSetDevice(c->device);
    cudaDeviceSynchronize();
    drop_device_tables(c);
    dc64::destroy(c->f64);
    for (auto& kv : c->d_bilateral)
        if (kv.second.p != nullptr)
            cudaFree(kv.second.p);
    for (auto& kv : c->d_median)
        free_dev(kv.second);
    if (c->d_guard != nullptr)
        cudaFree(c->d_guard);
    if (c->d_fallbacks != nullptr)
        cudaFree(c->d_fallbacks);
    for (auto& s : c->slots) {
        free_dev(s.d_in);
        free_pinned(s.h_in);
        for (int k = 0; k < DC_MAX_OUTPUTS; ++k) {
            free_dev(s.d_out[k]);
            free_pinned(s.h_out[k]);
        }
        if (s.done != nullptr)
            cudaEventDestroy(s.done);
        if (s.stream != nullptr)
            cudaStreamDestroy(s.stream);
    }
    if (c->stream != nullptr)
        cudaStreamDestroy(c->stream);
    delete c;
}

const char* dc_last_error(const dc_ctx* c) {
    return c != nullptr ? c->err.c_str() : g_create_error.c_str();
}

static int set_camera(dc_ctx* c, const dc::camera& cam, const double* precise = nullptr) {
    DC_CUDA(c, cudaSetDevice(c->device));
    DC_CUDA(c, cudaDeviceSynchronize());
    drop_device_tables(c);
    c->tabs     = dc::tables{};
    c->tabs.cam = cam;
    c->cam64.model = cam.model == dc::camera_model::pinhole ? 1
                     : cam.model == dc::camera_model::equirectangular ? 2 : 0;
    c->cam64.w = cam.w, c->cam64.h = cam.h;
    for (int k = 0; k < 4; ++k)
        c->cam64.p[k] = precise != nullptr ? precise[k] : double(cam.p[k]);
    if (c->f64 != nullptr)
        dc64::set_camera(c->f64, c->cam64);
    return DC_OK;
}

int dc_set_pinhole(dc_ctx* c, int w, int h, float fx, float fy, float cx, float cy) {
    if (c == nullptr)
        return DC_ERR_INVALID;
    if (c->sticky)
        return DC_ERR_CUDA;
    // preconditions of pinhole<Real>::pinhole, camera_models/pinhole.h:37-50
    if (!(w > 0 && h > 0 && fx > 0.f && fy > 0.f && cx > 0.f && cy > 0.f))
        return fail(c, DC_ERR_INVALID, "pinhole needs w, h > 0 and fx, fy, cx, cy > 0");
    dc::camera cam;
    cam.model = dc::camera_model::pinhole;
    cam.w     = w;
    cam.h     = h;
    cam.p[0] = fx, cam.p[1] = fy, cam.p[2] = cx, cam.p[3] = cy;
    return set_camera(c, cam);
}

int dc_set_equirectangular(dc_ctx* c, int w, int h, float theta_min, float theta_max) {
    if (c == nullptr)
        return DC_ERR_INVALID;
    if (c->sticky)
        return DC_ERR_CUDA;
    // camera_models/equirectangular.h:59-72 + ensure_invariant :104-114
    const float pi_f = float(3.141592653589793238462);
    if (!(w > 0 && h > 0 && theta_min >= 0.f && theta_max <= pi_f && theta_min < theta_max))
        return fail(c, DC_ERR_INVALID,
                    "equirectangular needs w, h > 0 and 0 <= theta_min < theta_max <= pi");
    dc::camera cam;
    cam.model = dc::camera_model::equirectangular;
    cam.w     = w;
    cam.h     = h;
    cam.p[0] = theta_min, cam.p[1] = theta_max;
    return set_camera(c, cam);
}

int dc_set_frame_size(dc_ctx* c, int w, int h) {
    if (c == nullptr)
        return DC_ERR_INVALID;
    if (c->sticky)
        return DC_ERR_CUDA;
    if (!(w > 0 && h > 0))
        return fail(c, DC_ERR_INVALID, "frame size needs w, h > 0");
    dc::camera cam;
    cam.model = dc::camera_model::none;
    cam.w     = w;
    cam.h     = h;
    return set_camera(c, cam);
}

int dc_set_pinhole_f64(dc_ctx* c, int w, int h, double fx, double fy, double cx, double cy) {
    if (c == nullptr)
        return DC_ERR_INVALID;
    if (c->sticky)
        return DC_ERR_CUDA;
    if (!(w > 0 && h > 0 && fx > 0. && fy > 0. && cx > 0. && cy > 0.))
        return fail(c, DC_ERR_INVALID, "pinhole needs w, h > 0 and fx, fy, cx, cy > 0");
    dc::camera cam;
    cam.model = dc::camera_model::pinhole;
    cam.w = w, cam.h = h;
    cam.p[0] = float(fx), cam.p[1] = float(fy), cam.p[2] = float(cx), cam.p[3] = float(cy);
    const double precise[4] = {fx, fy, cx, cy};
    return set_camera(c, cam, precise);
}

int dc_set_equirectangular_f64(dc_ctx* c, int w, int h, double theta_min, double theta_max) {
    if (c == nullptr)
        return DC_ERR_INVALID;
    if (c->sticky)
        return DC_ERR_CUDA;
    if (!(w > 0 && h > 0 && theta_min >= 0. && theta_max <= 3.141592653589793238462 &&
          theta_min < theta_max))
        return fail(c, DC_ERR_INVALID,
                    "equirectangular needs w, h > 0 and 0 <= theta_min < theta_max <= pi");
    dc::camera cam;
    cam.model = dc::camera_model::equirectangular;
    cam.w = w, cam.h = h;
    cam.p[0] = float(theta_min), cam.p[1] = float(theta_max);
    const double precise[4] = {theta_min, theta_max, 0., 0.};
    return set_camera(c, cam, precise);
}

int dc_width(const dc_ctx* c) { return c != nullptr ? c->tabs.cam.w : 0; }
int dc_height(const dc_ctx* c) { return c != nullptr ? c->tabs.cam.h : 0; }

int dc_convert_device(dc_ctx* c, const dc_job* j, void* cuda_stream) {
    if (c == nullptr)
        return DC_ERR_INVALID;
    if (c->sticky)
        return DC_ERR_CUDA;
    const int rc = validate_job(c, j);
    if (rc != DC_OK)
        return rc;
    DC_CUDA(c, cudaSetDevice(c->device));
    cudaStream_t s = cuda_stream != nullptr ? static_cast<cudaStream_t>(cuda_stream) : c->stream;
    return launch_job(c, j, s);
}

int dc_convert_host(dc_ctx* c, const dc_job* j) {
    if (c == nullptr)
        return DC_ERR_INVALID;
    if (c->sticky)
        return DC_ERR_CUDA;
    int rc = validate_job(c, j);
    if (rc != DC_OK)
        return rc;
    if (j->n_frames == 0)
        return DC_OK;
    DC_CUDA(c, cudaSetDevice(c->device));
    const int w = c->tabs.cam.w, h = c->tabs.cam.h;
    if (w <= 0 || h <= 0)
        return fail(c, DC_ERR_NO_CAMERA, "frame size unknown: install a camera first");
    const size_t px     = size_t(w) * h;
    const size_t in_fb  = px * in_elem_bytes(j->in_kind);
    const size_t out_fb = px * out_elem_bytes(j->out_kind);
    int          n_out  = 0;
    for (int k = 0; k < DC_MAX_OUTPUTS; ++k)
        n_out += j->out[k] != nullptr;
    const size_t per_frame = in_fb + size_t(n_out) * out_fb;
    const int    chunk =
        int(std::max<size_t>(1, std::min<size_t>(size_t(j->n_frames), kChunkTargetBytes / per_frame)));

    const bool pin_in = is_pinned(j->in);
    bool       pin_out[DC_MAX_OUTPUTS];
    for (int k = 0; k < DC_MAX_OUTPUTS; ++k)
        pin_out[k] = j->out[k] != nullptr && is_pinned(j->out[k]);

    int ci = 0;
    for (int f0 = 0; f0 < j->n_frames; f0 += chunk, ++ci) {
        const int nf = std::min(chunk, j->n_frames - f0);
        slot&     s  = c->slots[ci % kSlots];
        rc           = finish_slot(c, s);
        if (rc != DC_OK)
            return rc;
        DC_CUDA(c, reserve_dev(s.d_in, size_t(chunk) * in_fb));
        const char* src = static_cast<const char*>(j->in) + size_t(f0) * in_fb;
        if (!pin_in) {
            DC_CUDA(c, reserve_pinned(s.h_in, size_t(chunk) * in_fb));
            std::memcpy(s.h_in.p, src, size_t(nf) * in_fb);
            src = static_cast<const char*>(s.h_in.p);
        }
        DC_CUDA(c, cudaMemcpyAsync(s.d_in.p, src, size_t(nf) * in_fb, cudaMemcpyHostToDevice,
                                   s.stream));
        dc_job dj   = *j;
        dj.in       = s.d_in.p;
        dj.n_frames = nf;
        for (int k = 0; k < DC_MAX_OUTPUTS; ++k) {
            dj.out[k] = nullptr;
            if (j->out[k] == nullptr)
                continue;
            DC_CUDA(c, reserve_dev(s.d_out[k], size_t(chunk) * out_fb));
            dj.out[k] = s.d_out[k].p;
        }
        rc = launch_job(c, &dj, s.stream);
        if (rc != DC_OK)
            return rc;
        s.out_bytes = size_t(nf) * out_fb;
        for (int k = 0; k < DC_MAX_OUTPUTS; ++k) {
            s.user_out[k] = nullptr;
            if (j->out[k] == nullptr)
                continue;
            char* dst = static_cast<char*>(j->out[k]) + size_t(f0) * out_fb;
            if (!pin_out[k]) {
                DC_CUDA(c, reserve_pinned(s.h_out[k], size_t(chunk) * out_fb));
                s.user_out[k] = dst;
                dst           = static_cast<char*>(s.h_out[k].p);
            }
            DC_CUDA(c, cudaMemcpyAsync(dst, s.d_out[k].p, s.out_bytes, cudaMemcpyDeviceToHost,
                                       s.stream));
        }
        DC_CUDA(c, cudaEventRecord(s.done, s.stream));
        s.busy = true;
    }
    for (auto& s : c->slots) {
        rc = finish_slot(c, s);
        if (rc != DC_OK)
            return rc;
    }
    return DC_OK;
}

int dc_quantise_host_ex(dc_ctx* c, int quant, int pixel, int real_kind, const void* feature,
                        const void* mask, size_t n_px, void* out, const double* clamp_lo,
                        const double* clamp_hi) {
    if (c == nullptr)
        return DC_ERR_INVALID;
    if (c->sticky)
        return DC_ERR_CUDA;
    if (quant < DC_Q_RANGE || quant > DC_Q_MAX_CURVE)
        return fail(c, DC_ERR_INVALID, "unknown quantiser");
    if (pixel < DC_PX_U16 || pixel > DC_PX_I16)
        return fail(c, DC_ERR_INVALID, "unknown pixel type");
    if (real_kind != DC_OUT_F32 && real_kind != DC_OUT_F64)
        return fail(c, DC_ERR_INVALID, "real_kind is DC_OUT_F32 or DC_OUT_F64");
    if (n_px == 0)
        return DC_OK;
    if (feature == nullptr || out == nullptr)
        return fail(c, DC_ERR_INVALID, "quantiser buffers are NULL");
    if (quant == DC_Q_CURVATURE && mask == nullptr)
        return fail(c, DC_ERR_INVALID, "curvature quantiser needs the range image as mask");
    if (quant == DC_Q_CURVATURE && pixel != DC_PX_U16)
        return fail(c, DC_ERR_UNSUPPORTED,
                    "curvature_to_image yields 16 bit images only (depth_to_curvature.h:263-275: "
                    "the ushort intermediate re-types the result)");
    if (quant == DC_Q_RANGE && pixel != DC_PX_U16)
        return fail(c, DC_ERR_UNSUPPORTED, "the range image is written as 16 bit (converter_laserscan.h.inl:9-10)");
    DC_CUDA(c, cudaSetDevice(c->device));
    const bool   f64 = real_kind == DC_OUT_F64;
    const size_t rb = f64 ? 8 : 4, pb = (pixel == DC_PX_U8 || pixel == DC_PX_I8) ? 1 : 2;
    slot&        s  = c->slots[0];
    int          rc = finish_slot(c, s);
    if (rc != DC_OK)
        return rc;
    DC_CUDA(c, reserve_dev(s.d_in, n_px * rb));
    DC_CUDA(c, reserve_dev(s.d_out[0], n_px * rb));
    DC_CUDA(c, reserve_dev(s.d_out[1], n_px * pb));
    DC_CUDA(c, cudaMemcpyAsync(s.d_in.p, feature, n_px * rb, cudaMemcpyHostToDevice, s.stream));
    double lo = clamp_lo != nullptr ? *clamp_lo : 0., hi = clamp_hi != nullptr ? *clamp_hi : 0.;
    if (quant == DC_Q_CURVATURE) {
        DC_CUDA(c, cudaMemcpyAsync(s.d_out[0].p, mask, n_px * rb, cudaMemcpyHostToDevice, s.stream));
        if (clamp_lo == nullptr || clamp_hi == nullptr) {
            // depth_to_curvature.h:218-228: a missing clamp is the image's own
            // minimum / maximum (std::minmax_element over every pixel)
            DC_CUDA(c, reserve_dev(s.d_out[2], 16));
            const unsigned long long init64[2] = {~0ull, 0ull};
            const unsigned           init32[2] = {~0u, 0u};
            DC_CUDA(c, cudaMemcpyAsync(s.d_out[2].p, f64 ? static_cast<const void*>(init64) : init32,
                                       f64 ? 16 : 8, cudaMemcpyHostToDevice, s.stream));
            const unsigned blocks = unsigned(std::min<size_t>((n_px + 255) / 256, 148 * 8));
            if (f64)
                dc::minmax_kernel<double><<<blocks, 256, 0, s.stream>>>(
                    static_cast<const double*>(s.d_in.p), n_px, static_cast<unsigned long long*>(s.d_out[2].p));
            else
                dc::minmax_kernel<float><<<blocks, 256, 0, s.stream>>>(
                    static_cast<const float*>(s.d_in.p), n_px, static_cast<unsigned*>(s.d_out[2].p));
            const cudaError_t e = cudaGetLastError();
            if (e != cudaSuccess)
                return cuda_fail(c, e, "min/max kernel launch");
            c->launches += 1;
            unsigned long long r64[2];
            unsigned           r32[2];
            DC_CUDA(c, cudaMemcpyAsync(f64 ? static_cast<void*>(r64) : r32, s.d_out[2].p, f64 ? 16 : 8,
                                       cudaMemcpyDeviceToHost, s.stream));
            DC_CUDA(c, cudaStreamSynchronize(s.stream));
            const double mn = f64 ? dec_ordered(r64[0]) : double(dec_ordered(r32[0]));
            const double mx = f64 ? dec_ordered(r64[1]) : double(dec_ordered(r32[1]));
            if (clamp_lo == nullptr)
                lo = mn;
            if (clamp_hi == nullptr)
                hi = mx;
        }
        if (!(lo < hi))  // Expects(source_range.min < source_range.max), math/scaling.h:36
            return fail(c, DC_ERR_INVALID, "curvature quantiser needs lower bound < upper bound");
    }
    const double pi_f = double(float(3.141592653589793238462));
    // pi<Real> (math/constants.h:8-10); max-curve: Real(2. * pi<Real>) (depth_to_max_curve.h:166)
    const double pi_r = f64 ? 3.141592653589793238462 : pi_f;
    const double max_value = quant == DC_Q_BEARING     ? pi_r
                             : quant == DC_Q_MAX_CURVE ? (f64 ? 2. * pi_r : double(float(2. * pi_r)))
                                                       : 1.0;
    const cudaError_t e =
        f64 ? run_quantise_px<double>(quant, pixel, s.d_in.p, s.d_out[0].p, n_px, s.d_out[1].p, max_value, lo, hi, s.stream)
            : run_quantise_px<float>(quant, pixel, s.d_in.p, s.d_out[0].p, n_px, s.d_out[1].p, max_value, lo, hi, s.stream);
    if (e != cudaSuccess)
        return cuda_fail(c, e, "quantise kernel launch");
    c->launches += 1;
    DC_CUDA(c, cudaMemcpyAsync(out, s.d_out[1].p, n_px * pb, cudaMemcpyDeviceToHost, s.stream));
    DC_CUDA(c, cudaStreamSynchronize(s.stream));
    return DC_OK;
}

int dc_quantise_host(dc_ctx* c, int quant, const float* feature, const float* mask, size_t n_px,
                     uint16_t* out, float clamp_lo, float clamp_hi) {
    const double lo = clamp_lo, hi = clamp_hi;
    return dc_quantise_host_ex(c, quant, DC_PX_U16, DC_OUT_F32, feature, mask, n_px, out, &lo, &hi);
}

int dc_quantise_host_f64(dc_ctx* c, int quant, const double* feature, const double* mask,
                         size_t n_px, uint16_t* out, double clamp_lo, double clamp_hi) {
    return dc_quantise_host_ex(c, quant, DC_PX_U16, DC_OUT_F64, feature, mask, n_px, out, &clamp_lo,
                               &clamp_hi);
}

static int median_check(dc_ctx* c, int elem_kind, const void* in, int n_frames, int ksize,
                        const void* out) {
    if (c == nullptr)
        return DC_ERR_INVALID;
    if (c->sticky)
        return DC_ERR_CUDA;
    if (elem_kind != DC_IN_U16_RANGE && elem_kind != DC_IN_F32_RANGE)
        return fail(c, DC_ERR_INVALID, "median: elem_kind is DC_IN_U16_RANGE or DC_IN_F32_RANGE");
    if (ksize != 3 && ksize != 5)
        return fail(c, DC_ERR_INVALID, "median: ksize must be 3 or 5 (filter_functor.h:57-60)");
    if (n_frames < 0 || (n_frames > 0 && (in == nullptr || out == nullptr)) || in == out)
        return fail(c, DC_ERR_INVALID, "median: bad buffers");
    if (c->tabs.cam.w <= 0 || c->tabs.cam.h <= 0)
        return fail(c, DC_ERR_NO_CAMERA, "frame size unknown: dc_set_frame_size");
    return DC_OK;
}

int dc_median_device(dc_ctx* c, int elem_kind, const void* in, int n_frames, int ksize, void* out,
                     void* cuda_stream) {
    const int rc = median_check(c, elem_kind, in, n_frames, ksize, out);
    if (rc != DC_OK || n_frames == 0)
        return rc;
    DC_CUDA(c, cudaSetDevice(c->device));
    cudaStream_t s = cuda_stream != nullptr ? static_cast<cudaStream_t>(cuda_stream) : c->stream;
    const int    w = c->tabs.cam.w, h = c->tabs.cam.h;
    const cudaError_t e =
        elem_kind == DC_IN_U16_RANGE
            ? dc::med::launch_median(static_cast<const unsigned short*>(in),
                                     static_cast<unsigned short*>(out), w, h, n_frames, ksize, s)
            : dc::med::launch_median(static_cast<const float*>(in), static_cast<float*>(out), w, h,
                                     n_frames, ksize, s);
    if (e != cudaSuccess)
        return cuda_fail(c, e, "median kernel launch");
    c->launches += uint64_t((n_frames + 65534) / 65535);
    return DC_OK;
}

int dc_median_host(dc_ctx* c, int elem_kind, const void* in, int n_frames, int ksize, void* out) {
    int rc = median_check(c, elem_kind, in, n_frames, ksize, out);
    if (rc != DC_OK || n_frames == 0)
        return rc;
    DC_CUDA(c, cudaSetDevice(c->device));
    const size_t fb = size_t(c->tabs.cam.w) * c->tabs.cam.h * (elem_kind == DC_IN_U16_RANGE ? 2 : 4);
    const int chunk = int(std::max<size_t>(1, std::min<size_t>(size_t(n_frames), kChunkTargetBytes / (2 * fb))));
    slot&     s     = c->slots[0];
    for (int f0 = 0; f0 < n_frames; f0 += chunk) {
        const int nf = std::min(chunk, n_frames - f0);
        rc           = finish_slot(c, s);
        if (rc != DC_OK)
            return rc;
        DC_CUDA(c, reserve_dev(s.d_in, size_t(chunk) * fb));
        DC_CUDA(c, reserve_dev(s.d_out[0], size_t(chunk) * fb));
        DC_CUDA(c, cudaMemcpyAsync(s.d_in.p, static_cast<const char*>(in) + size_t(f0) * fb,
                                   size_t(nf) * fb, cudaMemcpyHostToDevice, s.stream));
        rc = dc_median_device(c, elem_kind, s.d_in.p, nf, ksize, s.d_out[0].p, s.stream);
        if (rc != DC_OK)
            return rc;
        DC_CUDA(c, cudaMemcpyAsync(static_cast<char*>(out) + size_t(f0) * fb, s.d_out[0].p,
                                   size_t(nf) * fb, cudaMemcpyDeviceToHost, s.stream));
        DC_CUDA(c, cudaStreamSynchronize(s.stream));
    }
    return DC_OK;
}

// ---- preprocess::gaussian_blur / preprocess::bilateral_filter (dc_filters.cuh) ----
static int filter_check(dc_ctx* c, const char* what, int elem_kind, const void* in, int n_frames,
                        const void* out) {
    if (c == nullptr)
        return DC_ERR_INVALID;
    if (c->sticky)
        return DC_ERR_CUDA;
    if (elem_kind != DC_IN_U16_RANGE && elem_kind != DC_IN_F32_RANGE)
        return fail(c, DC_ERR_INVALID, std::string(what) + ": elem_kind is DC_IN_U16_RANGE or DC_IN_F32_RANGE");
    if (n_frames < 0 || (n_frames > 0 && (in == nullptr || out == nullptr)) || in == out)
        return fail(c, DC_ERR_INVALID, std::string(what) + ": bad buffers");
    if (c->tabs.cam.w <= 0 || c->tabs.cam.h <= 0)
        return fail(c, DC_ERR_NO_CAMERA, "frame size unknown: dc_set_frame_size");
    return DC_OK;
}

// cv::GaussianBlur's kernel sizes: a size <= 0 is derived from sigma (float / 16 bit
// images: round(sigma * 8 + 1) | 1), then both must be positive and odd.
static int gaussian_setup(dc_ctx* c, int kw, int kh, double sx, double sy, dc::flt::gauss_params& p) {
    if (sy <= 0)
        sy = sx;
    if (kw <= 0 && sx > 0)
        kw = int(std::nearbyint(sx * 4 * 2 + 1)) | 1;
    if (kh <= 0 && sy > 0)
        kh = int(std::nearbyint(sy * 4 * 2 + 1)) | 1;
    if (kw <= 0 || kh <= 0 || kw % 2 != 1 || kh % 2 != 1)
        return fail(c, DC_ERR_INVALID, "gaussian_blur: kernel sizes must be positive and odd (or 0 with sigma > 0)");
    if (kw / 2 > dc::flt::kMaxRadius || kh / 2 > dc::flt::kMaxRadius)
        return fail(c, DC_ERR_UNSUPPORTED, "gaussian_blur: kernel sizes up to 31");
    p.w  = c->tabs.cam.w;
    p.h  = c->tabs.cam.h;
    p.rx = kw / 2;
    p.ry = kh / 2;
    dc::flt::gaussian_kernel_host(kw, sx, p.kx);
    dc::flt::gaussian_kernel_host(kh, sy, p.ky);
    return DC_OK;
}

int dc_gaussian_blur_device(dc_ctx* c, int elem_kind, const void* in, int n_frames, int ksize_w,
                            int ksize_h, double sigma_x, double sigma_y, void* out,
                            void* cuda_stream) {
    int rc = filter_check(c, "gaussian_blur", elem_kind, in, n_frames, out);
    if (rc != DC_OK)
        return rc;
    dc::flt::gauss_params p{};
    rc = gaussian_setup(c, ksize_w, ksize_h, sigma_x, sigma_y, p);
    if (rc != DC_OK || n_frames == 0)
        return rc;
    DC_CUDA(c, cudaSetDevice(c->device));
    cudaStream_t s = cuda_stream != nullptr ? static_cast<cudaStream_t>(cuda_stream) : c->stream;
    const cudaError_t e =
        elem_kind == DC_IN_U16_RANGE
            ? dc::flt::launch_gaussian(static_cast<const unsigned short*>(in),
                                       static_cast<unsigned short*>(out), n_frames, p, s)
            : dc::flt::launch_gaussian(static_cast<const float*>(in), static_cast<float*>(out),
                                       n_frames, p, s);
    if (e != cudaSuccess)
        return cuda_fail(c, e, "gaussian blur kernel launch");
    c->launches += uint64_t((n_frames + 65534) / 65535);
    c->last_kernel = "dc::flt::gaussian_blur_kernel";
    return DC_OK;
}

// radius = d <= 0 ? round(1.5 sigma_space) : d / 2, at least 1 (bilateralFilter_32f)
static int bilateral_radius(int d, double sigma_space) {
    const int r = d <= 0 ? int(std::nearbyint(sigma_space * 1.5)) : d / 2;
    return std::max(r, 1);
}

int dc_bilateral_device(dc_ctx* c, int elem_kind, const void* in, int n_frames, int d,
                        double sigma_color, double sigma_space, void* out, void* cuda_stream) {
    const int rc = filter_check(c, "bilateral_filter", elem_kind, in, n_frames, out);
    if (rc != DC_OK)
        return rc;
    if (!(sigma_color == sigma_color) || !(sigma_space == sigma_space))
        return fail(c, DC_ERR_INVALID, "bilateral_filter: sigma is NaN");
    if (sigma_color <= 0)
        sigma_color = 1;
    if (sigma_space <= 0)
        sigma_space = 1;
    const int radius = bilateral_radius(d, sigma_space);
    if (radius > dc::flt::kMaxRadius)
        return fail(c, DC_ERR_UNSUPPORTED, "bilateral_filter: radius up to 15 (d <= 31)");
    if (n_frames == 0)
        return DC_OK;
    DC_CUDA(c, cudaSetDevice(c->device));
    cudaStream_t s = cuda_stream != nullptr ? static_cast<cudaStream_t>(cuda_stream) : c->stream;
    const size_t bytes = size_t(n_frames) * dc::flt::kBilatScratchWords * sizeof(unsigned);
    dev_buf&     tmp   = c->d_bilateral[s];
    if (tmp.cap < bytes) {
        // (a growing buffer is freed only after everything that reads it ran)
        DC_CUDA(c, cudaStreamSynchronize(s));
        DC_CUDA(c, reserve_dev(tmp, bytes));
    }
    const int         w = c->tabs.cam.w, h = c->tabs.cam.h;
    unsigned* const   sc = static_cast<unsigned*>(tmp.p);
    const cudaError_t e =
        elem_kind == DC_IN_U16_RANGE
            ? dc::flt::launch_bilateral(static_cast<const unsigned short*>(in),
                                        static_cast<unsigned short*>(out), w, h, n_frames, radius,
                                        sigma_color, sigma_space, sc, s)
            : dc::flt::launch_bilateral(static_cast<const float*>(in), static_cast<float*>(out), w, h,
                                        n_frames, radius, sigma_color, sigma_space, sc, s);
    if (e != cudaSuccess)
        return cuda_fail(c, e, "bilateral filter kernel launch");
    c->launches += 4 * uint64_t((n_frames + 65534) / 65535);
    c->last_kernel = "dc::flt::bilateral_kernel";
    return DC_OK;
}

// HOST buffers: chunks of frames through staging slot 0, synchronous.
static int filter_host(dc_ctx* c, int elem_kind, const void* in, int n_frames, void* out,
                       const std::function<int(const void*, int, void*, cudaStream_t)>& launch) {
    DC_CUDA(c, cudaSetDevice(c->device));
    const size_t fb = size_t(c->tabs.cam.w) * c->tabs.cam.h * (elem_kind == DC_IN_U16_RANGE ? 2 : 4);
    const int chunk = int(std::max<size_t>(1, std::min<size_t>(size_t(n_frames), kChunkTargetBytes / (2 * fb))));
    slot&     s     = c->slots[0];
    for (int f0 = 0; f0 < n_frames; f0 += chunk) {
        const int nf = std::min(chunk, n_frames - f0);
        int       rc = finish_slot(c, s);
        if (rc != DC_OK)
            return rc;
        DC_CUDA(c, reserve_dev(s.d_in, size_t(chunk) * fb));
        DC_CUDA(c, reserve_dev(s.d_out[0], size_t(chunk) * fb));
        DC_CUDA(c, cudaMemcpyAsync(s.d_in.p, static_cast<const char*>(in) + size_t(f0) * fb,
                                   size_t(nf) * fb, cudaMemcpyHostToDevice, s.stream));
        rc = launch(s.d_in.p, nf, s.d_out[0].p, s.stream);
        if (rc != DC_OK)
            return rc;
        DC_CUDA(c, cudaMemcpyAsync(static_cast<char*>(out) + size_t(f0) * fb, s.d_out[0].p,
                                   size_t(nf) * fb, cudaMemcpyDeviceToHost, s.stream));
        DC_CUDA(c, cudaStreamSynchronize(s.stream));
    }
    return DC_OK;
}

int dc_gaussian_blur_host(dc_ctx* c, int elem_kind, const void* in, int n_frames, int ksize_w,
                          int ksize_h, double sigma_x, double sigma_y, void* out) {
    int rc = filter_check(c, "gaussian_blur", elem_kind, in, n_frames, out);
    if (rc != DC_OK)
        return rc;
    dc::flt::gauss_params p{};
    rc = gaussian_setup(c, ksize_w, ksize_h, sigma_x, sigma_y, p);
    if (rc != DC_OK || n_frames == 0)
        return rc;
    return filter_host(c, elem_kind, in, n_frames, out, [&](const void* di, int nf, void* dout, cudaStream_t st) {
        return dc_gaussian_blur_device(c, elem_kind, di, nf, ksize_w, ksize_h, sigma_x, sigma_y, dout, st);
    });
}

int dc_bilateral_host(dc_ctx* c, int elem_kind, const void* in, int n_frames, int d,
                      double sigma_color, double sigma_space, void* out) {
    const int rc = filter_check(c, "bilateral_filter", elem_kind, in, n_frames, out);
    if (rc != DC_OK || n_frames == 0)
        return rc;
    return filter_host(c, elem_kind, in, n_frames, out, [&](const void* di, int nf, void* dout, cudaStream_t st) {
        return dc_bilateral_device(c, elem_kind, di, nf, d, sigma_color, sigma_space, dout, st);
    });
}

int dc_synchronize(dc_ctx* c) {
    if (c == nullptr)
        return DC_ERR_INVALID;
    DC_CUDA(c, cudaSetDevice(c->device));
    DC_CUDA(c, cudaStreamSynchronize(c->stream));
    for (auto& s : c->slots)
        DC_CUDA(c, cudaStreamSynchronize(s.stream));
    return DC_OK;
}

void* dc_host_alloc(size_t bytes) {
    void* p = nullptr;
    if (cudaHostAlloc(&p, bytes, cudaHostAllocDefault) != cudaSuccess) {
        cudaGetLastError();
        return nullptr;
    }
    return p;
}
void dc_host_free(void* p) {
    if (p != nullptr)
        cudaFreeHost(p);
}

int dc_synth_depth_device(dc_ctx* c, uint64_t seed, int frame0, int n_frames,
                          uint16_t* out_device, void* cuda_stream) {
    if (c == nullptr)
        return DC_ERR_INVALID;
    if (c->sticky)
        return DC_ERR_CUDA;
    if (out_device == nullptr || n_frames < 0)
        return fail(c, DC_ERR_INVALID, "bad synth arguments");
    const int w = c->tabs.cam.w, h = c->tabs.cam.h;
    if (w <= 0 || h <= 0)
        return fail(c, DC_ERR_NO_CAMERA, "frame size unknown: install a camera first");
    DC_CUDA(c, cudaSetDevice(c->device));
    cudaStream_t s = cuda_stream != nullptr ? static_cast<cudaStream_t>(cuda_stream) : c->stream;
    const dim3   block(32, 8, 1);
    for (int f0 = 0; f0 < n_frames; f0 += 65535) {
        const int  nf = std::min(65535, n_frames - f0);
        const dim3 grid((w + 31) / 32, (h + 7) / 8, unsigned(nf));
        dc::synth_depth_kernel<<<grid, block, 0, s>>>(
            seed, frame0 + f0, w, h, out_device + size_t(f0) * w * h);
        const cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess)
            return cuda_fail(c, e, "synth kernel launch");
        c->launches += 1;
    }
    return DC_OK;
}

uint64_t dc_kernel_launches(const dc_ctx* c) { return c != nullptr ? c->launches : 0; }

const char* dc_last_kernel(const dc_ctx* c) { return c != nullptr ? c->last_kernel : ""; }

uint64_t dc_f32_fallbacks(dc_ctx* c) {
    if (c == nullptr || c->d_fallbacks == nullptr)
        return 0;
    unsigned long long n = 0;
    if (cudaSetDevice(c->device) != cudaSuccess || cudaDeviceSynchronize() != cudaSuccess ||
        cudaMemcpy(&n, c->d_fallbacks, sizeof n, cudaMemcpyDeviceToHost) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}

}  // extern "C"
