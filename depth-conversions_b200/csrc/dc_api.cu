// C ABI of libdepthconv_b200.so: context, per-camera device tables, kernel
// dispatch and the host-buffer pipeline. See include/depthconv_b200.h.
#include "../../include/depthconv_b200.h"
#include "dc_kernels.cuh"
#include "dc_median.cuh"
#include "dc_pair_kernels.cuh"
#include "dc_f64.h"
#include "dc_tables.h"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
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
    const char*  last_kernel = "";  // kernel family of the most recent float-path launch
    // Real = double (dc_f64.h): camera parameters in double, tables on demand
    dc64::camera cam64;
    dc64::state* f64 = nullptr;
    // filtered copy of the input frames (dc_job::median_ksize), one per stream:
    // the chunks of dc_convert_host run concurrently on the slots' streams
    std::map<cudaStream_t, dev_buf> d_median;
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

// ---------------------------------------------------------------------------
bool aligned(const void* p, size_t a) { return (reinterpret_cast<uintptr_t>(p) % a) == 0; }

// Which kernel family the launch helpers below picked (dc_last_kernel: tests and
// bench.py name the kernel they measured instead of guessing it).
thread_local const char* g_kernel = "";

// Rows a thread of the frame-pair kernels marches over (tuning knob:
// DC_ROWS_PER_STRIP in the environment, read once).
int rows_per_strip() {
    static const int v = [] {
        const char* e = std::getenv("DC_ROWS_PER_STRIP");
        const int   n = e != nullptr ? std::atoi(e) : 0;
        return n >= 3 ? n : 18;
    }();
    return v;
}

template <int FEAT>
constexpr bool has_pair_kernel() {
    return FEAT == dc::FEAT_BEARING || FEAT == dc::FEAT_MAX_CURVE || FEAT == dc::FEAT_CURVATURE;
}

// The frame-pair row-marching kernels (dc_pair_kernels.cuh): u16 input,
// width % 4 == 0, aligned frames, per-camera range proof.
// What is known about a job before the launch (dc_pair.cuh SPEC_*): with all
// outputs wanted and the camera tables of the feature either all per-pixel
// arrays or all scalars, the row loop is compiled without pointer tests.
template <int FEAT, int IN>
int pair_spec(const dc::kparams& p) {
    if (IN == dc::IN_U16_DEPTH && p.range_factor.p == nullptr)
        return dc::pr::SPEC_GENERIC;
    int tables = 0, scalars = 0;
    auto see = [&](const dc::tab_ref& t) { (t.p != nullptr ? tables : scalars) += 1; };
    if (FEAT == dc::FEAT_BEARING || FEAT == dc::FEAT_MAX_CURVE) {
        for (int d = 0; d < 4; ++d) {
            if (FEAT == dc::FEAT_BEARING && p.out[d] == nullptr)
                return dc::pr::SPEC_GENERIC;
            see(p.cosp[d]);
        }
    } else if (FEAT == dc::FEAT_CURVATURE) {
        if (p.out[0] == nullptr || p.out[1] == nullptr)
            return dc::pr::SPEC_GENERIC;
        see(p.d_phi), see(p.d_theta), see(p.d_phi_theta);
    }
    return scalars == 0 ? dc::pr::SPEC_TABLES : tables == 0 ? dc::pr::SPEC_UNIFORM : dc::pr::SPEC_GENERIC;
}

template <int FEAT, int IN, int OUT>
cudaError_t launch_pair(const dc::kparams& p, int n_frames, cudaStream_t s) {
    const int       rps    = rows_per_strip();
    const long long strips = (p.h + rps - 1) / rps;
    const long long pairs  = (n_frames + 1) / 2;
    const long long threads = pairs * strips * (p.w / 4);
    const long long blocks  = (threads + dc::pr::kPairThreads - 1) / dc::pr::kPairThreads;
    if (blocks > 0x7fffffffLL)
        return cudaErrorInvalidConfiguration;
    const dim3   grid{unsigned(blocks)}, block{unsigned(dc::pr::kPairThreads)};
    static const int forced = [] {
        const char* e = std::getenv("DC_PAIR_SPEC");  // debugging: 0 forces the generic kernel
        return e != nullptr ? std::atoi(e) : -1;
    }();
    const int spec = forced == 0 ? 0 : pair_spec<FEAT, IN>(p);
    if constexpr (FEAT == dc::FEAT_MAX_CURVE) {
        // The kernel that computes every pixel-pair edge once for both of its ends
        // (dc_pair_kernels.cuh maxcurve_edge_kernel): one scalar for all four
        // neighbour-angle tables (1080p and larger with Kinect-like intrinsics), or
        // four per-pixel tables (the anti-diagonal pair is not shared there).
        static const int edge_rows = [] {
            const char* e = std::getenv("DC_MCE_ROWS");  // 0 disables the kernel
            return e != nullptr ? std::atoi(e) : 36;
        }();
        auto bits = [](float v) {
            unsigned u;
            std::memcpy(&u, &v, sizeof u);
            return u;
        };
        const unsigned c0 = bits(p.cosp[0].value);
        const bool one_cos = spec == dc::pr::SPEC_UNIFORM && c0 == bits(p.cosp[1].value) &&
                             c0 == bits(p.cosp[2].value) && c0 == bits(p.cosp[3].value);
        if (edge_rows >= 3 && (one_cos || spec == dc::pr::SPEC_TABLES)) {
            const long long estrips  = (p.h + edge_rows - 1) / edge_rows;
            const long long ethreads = pairs * estrips * (p.w / 4);
            const long long eblocks  = (ethreads + dc::pr::kPairThreads - 1) / dc::pr::kPairThreads;
            if (eblocks > 0x7fffffffLL)
                return cudaErrorInvalidConfiguration;
            if (one_cos) {
                dc::pr::maxcurve_edge_kernel<IN, OUT, dc::pr::SPEC_UNIFORM>
                    <<<unsigned(eblocks), block, dc::pr::mce_smem_bytes<dc::pr::SPEC_UNIFORM>(), s>>>(
                        p, n_frames, edge_rows);
                g_kernel = "dc::pr::maxcurve_edge_kernel<SPEC_UNIFORM>";
            } else {
                constexpr size_t smem = dc::pr::mce_smem_bytes<dc::pr::SPEC_TABLES>();  // 54 KB
                // per device, hence not cached in a static (depth2x --devices N runs
                // one context per GPU in one process); a few microseconds per launch
                const cudaError_t attr = cudaFuncSetAttribute(
                    dc::pr::maxcurve_edge_kernel<IN, OUT, dc::pr::SPEC_TABLES>,
                    cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
                if (attr != cudaSuccess)
                    return attr;
                dc::pr::maxcurve_edge_kernel<IN, OUT, dc::pr::SPEC_TABLES>
                    <<<unsigned(eblocks), block, smem, s>>>(p, n_frames, edge_rows);
                g_kernel = "dc::pr::maxcurve_edge_kernel<SPEC_TABLES>";
            }
            return cudaGetLastError();
        }
    }
    g_kernel = spec == dc::pr::SPEC_TABLES    ? "dc::pr::pair_kernel<SPEC_TABLES>"
               : spec == dc::pr::SPEC_UNIFORM ? "dc::pr::pair_kernel<SPEC_UNIFORM>"
                                              : "dc::pr::pair_kernel<SPEC_GENERIC>";
    switch (spec) {
    case dc::pr::SPEC_TABLES:
        dc::pr::pair_kernel<FEAT, IN, OUT, dc::pr::SPEC_TABLES>
            <<<grid, block, dc::pr::pair_smem_bytes<FEAT, dc::pr::SPEC_TABLES>(), s>>>(p, n_frames, rps);
        break;
    case dc::pr::SPEC_UNIFORM:
        dc::pr::pair_kernel<FEAT, IN, OUT, dc::pr::SPEC_UNIFORM>
            <<<grid, block, dc::pr::pair_smem_bytes<FEAT, dc::pr::SPEC_UNIFORM>(), s>>>(p, n_frames, rps);
        break;
    default:
        dc::pr::pair_kernel<FEAT, IN, OUT, dc::pr::SPEC_GENERIC>
            <<<grid, block, dc::pr::pair_smem_bytes<FEAT, dc::pr::SPEC_GENERIC>(), s>>>(p, n_frames, rps);
    }
    return cudaGetLastError();
}

template <int IN, int OUT>
cudaError_t launch_flexion_pair(const dc::kparams& p, int n_frames, cudaStream_t s) {
    // a longer march amortises the point window (three rows of camera-frame points
    // per strip start): 72 rows measured 1.5 % faster than 36 (DC_FLEX_ROWS, read once)
    static const int flex_rows = [] {
        const char* e = std::getenv("DC_FLEX_ROWS");
        const int   n = e != nullptr ? std::atoi(e) : 0;
        return n >= 3 ? n : 4 * rows_per_strip();
    }();
    const int       rps     = flex_rows;
    const long long strips  = (p.h + rps - 1) / rps;
    const long long pairs   = (n_frames + 1) / 2;
    const long long threads = pairs * strips * (p.w / 2);
    const long long blocks  = (threads + dc::pr::kPairThreads - 1) / dc::pr::kPairThreads;
    if (blocks > 0x7fffffffLL)
        return cudaErrorInvalidConfiguration;
    // per device (one context per GPU in a depth2x --devices N process): not cached
    const cudaError_t carve = cudaFuncSetAttribute(
        dc::pr::flexion_pair_kernel<IN, OUT>, cudaFuncAttributePreferredSharedMemoryCarveout,
        int(cudaSharedmemCarveoutMaxShared));
    if (carve != cudaSuccess)
        return carve;
    const size_t smem = size_t(3) * dc::pr::kFlexSlots * dc::pr::kPairThreads * sizeof(uint4);
    g_kernel = "dc::pr::flexion_pair_kernel";
    dc::pr::flexion_pair_kernel<IN, OUT>
        <<<unsigned(blocks), dc::pr::kPairThreads, smem, s>>>(p, n_frames, rps);
    return cudaGetLastError();
}

// DC_RANGE / DC_SCALE with u16 input: the streaming kernel (8 pixels per thread,
// frames walked inside the thread). Needs frames that are multiples of 8 pixels
// and 16-byte aligned buffers.
template <int FEAT, int IN, int OUT>
cudaError_t launch_stream(const dc::kparams& p, int n_frames, cudaStream_t s) {
    const int groups = int(p.frame_px / 8);
    const int bx     = (groups + 255) / 256;
    // about 8 resident blocks per SM; at least 4 frames per thread when possible
    int by = std::max(1, std::min((148 * 8 + bx - 1) / bx, (n_frames + 3) / 4));
    by     = std::min(by, 65535);
    g_kernel = "dc::pr::stream_kernel";
    dc::pr::stream_kernel<FEAT, IN, OUT><<<dim3(bx, by, 1), 256, 0, s>>>(p, n_frames, groups);
    return cudaGetLastError();
}

template <int FEAT, int IN, int OUT>
cudaError_t launch_px(const dc::kparams& p, int n_frames, cudaStream_t s, bool vec4, bool fast) {
    if constexpr (IN != dc::IN_F32_RANGE && (FEAT == dc::FEAT_RANGE || FEAT == dc::FEAT_SCALE)) {
        bool ok = vec4 && (p.frame_px % 8) == 0 && aligned(p.in, 16) && aligned(p.out[0], 16);
        if (ok)
            return launch_stream<FEAT, IN, OUT>(p, n_frames, s);
    }
    if constexpr (IN != dc::IN_F32_RANGE && has_pair_kernel<FEAT>()) {
        if (vec4 && fast)
            return launch_pair<FEAT, IN, OUT>(p, n_frames, s);
    }
    if constexpr (IN != dc::IN_F32_RANGE && FEAT == dc::FEAT_FLEXION) {
        if (vec4 && fast)
            return launch_flexion_pair<IN, OUT>(p, n_frames, s);
    }
    g_kernel = vec4 && fast ? "dc::feature_kernel<4, fast>" : vec4 ? "dc::feature_kernel<4>" : "dc::feature_kernel<1>";
    const int px = vec4 ? 4 : 1;
    const int qw = (p.w + px - 1) / px;
    // pick the block shape that wastes the fewest threads
    static const int shapes[4][2] = {{32, 8}, {16, 16}, {64, 4}, {8, 32}};
    int       best = 0;
    long long best_cost = -1;
    for (int i = 0; i < 4; ++i) {
        const long long gx = (qw + shapes[i][0] - 1) / shapes[i][0];
        const long long gy = (p.h + shapes[i][1] - 1) / shapes[i][1];
        const long long cost = gx * gy;
        if (best_cost < 0 || cost < best_cost) {
            best_cost = cost;
            best      = i;
        }
    }
    const dim3 block(shapes[best][0], shapes[best][1], 1);
    const unsigned gx = (qw + block.x - 1) / block.x, gy = (p.h + block.y - 1) / block.y;
    const size_t   in_b = IN == dc::IN_F32_RANGE ? 4 : 2, out_b = OUT == dc::OUT_F32 ? 4 : 2;
    for (int f0 = 0; f0 < n_frames; f0 += 65535) {
        const int   nf = std::min(65535, n_frames - f0);
        dc::kparams q  = p;
        q.in           = static_cast<const char*>(p.in) + size_t(f0) * p.frame_px * in_b;
        for (int k = 0; k < 4; ++k)
            if (p.out[k] != nullptr)
                q.out[k] = static_cast<char*>(p.out[k]) + size_t(f0) * p.frame_px * out_b;
        const dim3 grid(gx, gy, unsigned(nf));
        if (vec4 && fast)
            dc::feature_kernel<FEAT, IN, OUT, 4, true><<<grid, block, 0, s>>>(q);
        else if (vec4)
            dc::feature_kernel<FEAT, IN, OUT, 4, false><<<grid, block, 0, s>>>(q);
        else
            dc::feature_kernel<FEAT, IN, OUT, 1, false><<<grid, block, 0, s>>>(q);
        const cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess)
            return e;
    }
    return cudaSuccess;
}

template <int FEAT, int IN>
cudaError_t launch_out(const dc::kparams& p, int out_kind, int n, cudaStream_t s, bool v4,
                       bool fast) {
    return out_kind == DC_OUT_F32 ? launch_px<FEAT, IN, dc::OUT_F32>(p, n, s, v4, fast)
                                  : launch_px<FEAT, IN, dc::OUT_U16>(p, n, s, v4, fast);
}
template <int FEAT>
cudaError_t launch_in(const dc::kparams& p, int in_kind, int out_kind, int n, cudaStream_t s,
                      bool v4, bool fast) {
    switch (in_kind) {
    case DC_IN_U16_DEPTH: return launch_out<FEAT, dc::IN_U16_DEPTH>(p, out_kind, n, s, v4, fast);
    case DC_IN_U16_RANGE: return launch_out<FEAT, dc::IN_U16_RANGE>(p, out_kind, n, s, v4, fast);
    default: return launch_out<FEAT, dc::IN_F32_RANGE>(p, out_kind, n, s, v4, false);
    }
}


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
    p.lo          = j->clamp_lo;
    p.hi          = j->clamp_hi;
    // conversion/util.h:71-79 scaling_factor<float, ushort>(max_angle) =
    // (65535 - 0) / max_angle; pi<float> and float(2. * pi<float>)
    const float pi_f  = float(3.141592653589793238462);
    p.q_bearing       = 65535.0f / pi_f;
    p.q_max_curve     = 65535.0f / float(2. * double(pi_f));
    p.s_scale         = float(j->scale);
    p.s_offset        = float(j->offset);

    // vector path: rows and frames must keep 8-byte (u16) / 16-byte (f32) alignment
    bool v4 = (w % 4) == 0;
    v4      = v4 && aligned(j->in, in_kind == DC_IN_F32_RANGE ? 16 : 8);
    for (int k = 0; k < 4; ++k)
        if (j->out[k] != nullptr)
            v4 = v4 && aligned(j->out[k], j->out_kind == DC_OUT_F32 ? 16 : 8);

    // guard-free exact arithmetic needs the per-camera range proof (dc_tables)
    const bool fast = c->tabs.fast_ok;
    cudaError_t e = cudaSuccess;
    switch (j->feature) {
    case DC_RANGE: e = launch_in<dc::FEAT_RANGE>(p, in_kind, j->out_kind, j->n_frames, s, v4, fast); break;
    case DC_BEARING: e = launch_in<dc::FEAT_BEARING>(p, in_kind, j->out_kind, j->n_frames, s, v4, fast); break;
    case DC_FLEXION: e = launch_in<dc::FEAT_FLEXION>(p, in_kind, j->out_kind, j->n_frames, s, v4, fast); break;
    case DC_CURVATURE: e = launch_in<dc::FEAT_CURVATURE>(p, in_kind, j->out_kind, j->n_frames, s, v4, fast); break;
    case DC_MAX_CURVE: e = launch_in<dc::FEAT_MAX_CURVE>(p, in_kind, j->out_kind, j->n_frames, s, v4, fast); break;
    case DC_SCALE:
        e = launch_px<dc::FEAT_SCALE, dc::IN_U16_RANGE, dc::OUT_U16>(p, j->n_frames, s, v4, false);
        break;
    default: return fail(c, DC_ERR_INVALID, "unknown feature");
    }
    if (e != cudaSuccess)
        return cuda_fail(c, e, "kernel launch");
    c->launches += uint64_t((j->n_frames + 65534) / 65535);
    c->last_kernel = g_kernel;
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
    for (auto& kv : c->d_median)
        free_dev(kv.second);
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

int dc_quantise_host(dc_ctx* c, int quant, const float* feature, const float* mask, size_t n_px,
                     uint16_t* out, float clamp_lo, float clamp_hi) {
    if (c == nullptr)
        return DC_ERR_INVALID;
    if (c->sticky)
        return DC_ERR_CUDA;
    if (quant < DC_Q_RANGE || quant > DC_Q_MAX_CURVE)
        return fail(c, DC_ERR_INVALID, "unknown quantiser");
    if (n_px == 0)
        return DC_OK;
    if (feature == nullptr || out == nullptr)
        return fail(c, DC_ERR_INVALID, "quantiser buffers are NULL");
    if (quant == DC_Q_CURVATURE && (mask == nullptr || !(clamp_lo < clamp_hi)))
        return fail(c, DC_ERR_INVALID,
                    "curvature quantiser needs the range image as mask and lower < upper bound");
    DC_CUDA(c, cudaSetDevice(c->device));
    slot& s = c->slots[0];
    int   rc = finish_slot(c, s);
    if (rc != DC_OK)
        return rc;
    DC_CUDA(c, reserve_dev(s.d_in, n_px * sizeof(float)));
    DC_CUDA(c, reserve_dev(s.d_out[0], n_px * sizeof(float)));
    DC_CUDA(c, reserve_dev(s.d_out[1], n_px * sizeof(uint16_t)));
    DC_CUDA(c, cudaMemcpyAsync(s.d_in.p, feature, n_px * sizeof(float), cudaMemcpyHostToDevice,
                               s.stream));
    if (quant == DC_Q_CURVATURE)
        DC_CUDA(c, cudaMemcpyAsync(s.d_out[0].p, mask, n_px * sizeof(float),
                                   cudaMemcpyHostToDevice, s.stream));
    // conversion/util.h:71-79 scaling_factor<float, ushort>(max_angle)
    const float pi_f  = float(3.141592653589793238462);
    const float alpha = quant == DC_Q_BEARING     ? 65535.0f / pi_f
                        : quant == DC_Q_MAX_CURVE ? 65535.0f / float(2. * double(pi_f))
                                                  : 65535.0f;
    const unsigned blocks = unsigned(std::min<size_t>((n_px + 255) / 256, 148 * 16));
    dc::quantise_kernel<<<blocks, 256, 0, s.stream>>>(
        quant, static_cast<const float*>(s.d_in.p), static_cast<const float*>(s.d_out[0].p), n_px,
        static_cast<unsigned short*>(s.d_out[1].p), alpha, clamp_lo, clamp_hi);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess)
        return cuda_fail(c, e, "quantise kernel launch");
    c->launches += 1;
    DC_CUDA(c, cudaMemcpyAsync(out, s.d_out[1].p, n_px * sizeof(uint16_t), cudaMemcpyDeviceToHost,
                               s.stream));
    DC_CUDA(c, cudaStreamSynchronize(s.stream));
    return DC_OK;
}

int dc_quantise_host_f64(dc_ctx* c, int quant, const double* feature, const double* mask,
                         size_t n_px, uint16_t* out, double clamp_lo, double clamp_hi) {
    if (c == nullptr)
        return DC_ERR_INVALID;
    if (c->sticky)
        return DC_ERR_CUDA;
    if (quant < DC_Q_RANGE || quant > DC_Q_MAX_CURVE)
        return fail(c, DC_ERR_INVALID, "unknown quantiser");
    if (n_px == 0)
        return DC_OK;
    if (feature == nullptr || out == nullptr)
        return fail(c, DC_ERR_INVALID, "quantiser buffers are NULL");
    if (quant == DC_Q_CURVATURE && (mask == nullptr || !(clamp_lo < clamp_hi)))
        return fail(c, DC_ERR_INVALID,
                    "curvature quantiser needs the range image as mask and lower < upper bound");
    DC_CUDA(c, cudaSetDevice(c->device));
    slot& s  = c->slots[0];
    int   rc = finish_slot(c, s);
    if (rc != DC_OK)
        return rc;
    DC_CUDA(c, reserve_dev(s.d_in, n_px * sizeof(double)));
    DC_CUDA(c, reserve_dev(s.d_out[0], n_px * sizeof(double)));
    DC_CUDA(c, reserve_dev(s.d_out[1], n_px * sizeof(uint16_t)));
    DC_CUDA(c, cudaMemcpyAsync(s.d_in.p, feature, n_px * sizeof(double), cudaMemcpyHostToDevice,
                               s.stream));
    if (quant == DC_Q_CURVATURE)
        DC_CUDA(c, cudaMemcpyAsync(s.d_out[0].p, mask, n_px * sizeof(double),
                                   cudaMemcpyHostToDevice, s.stream));
    const cudaError_t e = dc64::quantise(
        quant, static_cast<const double*>(s.d_in.p), static_cast<const double*>(s.d_out[0].p), n_px,
        static_cast<uint16_t*>(s.d_out[1].p), clamp_lo, clamp_hi, s.stream);
    if (e != cudaSuccess)
        return cuda_fail(c, e, "quantise kernel launch");
    c->launches += 1;
    DC_CUDA(c, cudaMemcpyAsync(out, s.d_out[1].p, n_px * sizeof(uint16_t), cudaMemcpyDeviceToHost,
                               s.stream));
    DC_CUDA(c, cudaStreamSynchronize(s.stream));
    return DC_OK;
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

}  // extern "C"
