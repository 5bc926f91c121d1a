// Kernel selection and launch, shared by the dc_k_<feature>.cu translation units
// (each instantiates launch_feature<FEAT> for its own feature only).
#ifndef DC_LAUNCH_CUH
#define DC_LAUNCH_CUH

#include "dc_launch.h"
#include "dc_pair_kernels.cuh"

#include <algorithm>
#include <cstdint>
#include <cstdlib>
#include <cstring>

namespace dc {
namespace ln {

inline bool aligned(const void* p, size_t a) { return (reinterpret_cast<uintptr_t>(p) % a) == 0; }

inline int env_int(const char* name, int fallback) {
    const char* e = std::getenv(name);
    return e != nullptr ? std::atoi(e) : fallback;
}

// Rows a thread of the frame-pair kernels marches over (tuning knob:
// DC_ROWS_PER_STRIP in the environment, read once per translation unit).
inline int rows_per_strip() {
    static const int v = [] {
        const int n = env_int("DC_ROWS_PER_STRIP", 0);
        return n >= 3 ? n : 18;
    }();
    return v;
}

template <int FEAT>
constexpr bool has_pair_kernel() {
    return FEAT == FEAT_BEARING || FEAT == FEAT_MAX_CURVE || FEAT == FEAT_CURVATURE;
}

// What is known about a job before the launch (dc_pair.cuh SPEC_*): with all
// outputs wanted and the camera tables of the feature either all per-pixel
// arrays or all scalars, the row loop is compiled without pointer tests.
template <int FEAT, int IN>
int pair_spec(const kparams& p) {
    if (IN == IN_U16_DEPTH && p.range_factor.p == nullptr)
        return pr::SPEC_GENERIC;
    int  tables = 0, scalars = 0;
    auto see = [&](const tab_ref& t) { (t.p != nullptr ? tables : scalars) += 1; };
    if (FEAT == FEAT_BEARING || FEAT == FEAT_MAX_CURVE) {
        for (int d = 0; d < 4; ++d) {
            if (FEAT == FEAT_BEARING && p.out[d] == nullptr)
                return pr::SPEC_GENERIC;
            see(p.cosp[d]);
        }
    } else if (FEAT == FEAT_CURVATURE) {
        if (p.out[0] == nullptr || p.out[1] == nullptr)
            return pr::SPEC_GENERIC;
        see(p.d_phi), see(p.d_theta), see(p.d_phi_theta);
    }
    return scalars == 0 ? pr::SPEC_TABLES : tables == 0 ? pr::SPEC_UNIFORM : pr::SPEC_GENERIC;
}

// The frame-pair row-marching kernels (dc_pair_kernels.cuh). `half`: the width is
// even but not a multiple of four (SPEC_HALF: 4-pixel accesses in two halves).
template <int FEAT, int IN, int OUT>
cudaError_t launch_pair(const kparams& p, int n_frames, cudaStream_t s, launch_info& li, bool half) {
    const int       rps     = rows_per_strip();
    const long long strips  = (p.h + rps - 1) / rps;
    const long long pairs   = (n_frames + 1) / 2;
    const long long threads = pairs * strips * ((p.w + 3) / 4);
    const long long blocks  = (threads + pr::kPairThreads - 1) / pr::kPairThreads;
    if (blocks > 0x7fffffffLL)
        return cudaErrorInvalidConfiguration;
    const dim3       grid{unsigned(blocks)}, block{unsigned(pr::kPairThreads)};
    static const int forced = env_int("DC_PAIR_SPEC", -1);  // debugging: 0 forces the generic kernel
    const int        spec   = half ? int(pr::SPEC_HALF) : forced == 0 ? 0 : pair_spec<FEAT, IN>(p);
    li.launches += 1;
    if constexpr (FEAT == FEAT_BEARING && OUT == OUT_U16) {
        // the 16-bit bearing epilogue has convert_bearing's factor compiled in
        if (p.q_bearing != pr::kQBearing)
            return cudaErrorInvalidValue;
    }
    if constexpr (FEAT == FEAT_MAX_CURVE) {
        // The kernel that computes every pixel-pair edge once for both of its ends
        // (dc_pair_kernels.cuh maxcurve_edge_kernel): one scalar for all four
        // neighbour-angle tables (1080p and larger with Kinect-like intrinsics), or
        // four per-pixel tables (the anti-diagonal pair is not shared there).
        static const int edge_rows = env_int("DC_MCE_ROWS", 36);  // 0 disables the kernel
        auto             bits      = [](float v) {
            unsigned u;
            std::memcpy(&u, &v, sizeof u);
            return u;
        };
        const unsigned c0      = bits(p.cosp[0].value);
        const bool     one_cos = spec == pr::SPEC_UNIFORM && c0 == bits(p.cosp[1].value) &&
                             c0 == bits(p.cosp[2].value) && c0 == bits(p.cosp[3].value);
        if (edge_rows >= 3 && (one_cos || spec == pr::SPEC_TABLES)) {
            const long long estrips  = (p.h + edge_rows - 1) / edge_rows;
            const long long ethreads = pairs * estrips * (p.w / 4);
            const long long eblocks  = (ethreads + pr::kPairThreads - 1) / pr::kPairThreads;
            if (eblocks > 0x7fffffffLL)
                return cudaErrorInvalidConfiguration;
            if (one_cos) {
                pr::maxcurve_edge_kernel<IN, OUT, pr::SPEC_UNIFORM>
                    <<<unsigned(eblocks), block, pr::mce_smem_bytes<pr::SPEC_UNIFORM>(), s>>>(
                        p, n_frames, edge_rows);
                li.kernel = "dc::pr::maxcurve_edge_kernel<SPEC_UNIFORM>";
            } else {
                constexpr size_t smem = pr::mce_smem_bytes<pr::SPEC_TABLES>();  // 54 KB
                // per device, hence not cached in a static (depth2x --devices N runs
                // one context per GPU in one process); a few microseconds per launch
                const cudaError_t attr = cudaFuncSetAttribute(
                    pr::maxcurve_edge_kernel<IN, OUT, pr::SPEC_TABLES>,
                    cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
                if (attr != cudaSuccess)
                    return attr;
                pr::maxcurve_edge_kernel<IN, OUT, pr::SPEC_TABLES>
                    <<<unsigned(eblocks), block, smem, s>>>(p, n_frames, edge_rows);
                li.kernel = "dc::pr::maxcurve_edge_kernel<SPEC_TABLES>";
            }
            return cudaGetLastError();
        }
    }
    li.kernel = spec == pr::SPEC_TABLES    ? "dc::pr::pair_kernel<SPEC_TABLES>"
                : spec == pr::SPEC_UNIFORM ? "dc::pr::pair_kernel<SPEC_UNIFORM>"
                : spec == pr::SPEC_HALF    ? "dc::pr::pair_kernel<SPEC_HALF>"
                                           : "dc::pr::pair_kernel<SPEC_GENERIC>";
    switch (spec) {
    case pr::SPEC_TABLES:
        pr::pair_kernel<FEAT, IN, OUT, pr::SPEC_TABLES>
            <<<grid, block, pr::pair_smem_bytes<FEAT, pr::SPEC_TABLES>(), s>>>(p, n_frames, rps);
        break;
    case pr::SPEC_UNIFORM:
        pr::pair_kernel<FEAT, IN, OUT, pr::SPEC_UNIFORM>
            <<<grid, block, pr::pair_smem_bytes<FEAT, pr::SPEC_UNIFORM>(), s>>>(p, n_frames, rps);
        break;
    case pr::SPEC_HALF:
        pr::pair_kernel<FEAT, IN, OUT, pr::SPEC_HALF>
            <<<grid, block, pr::pair_smem_bytes<FEAT, pr::SPEC_HALF>(), s>>>(p, n_frames, rps);
        break;
    default:
        pr::pair_kernel<FEAT, IN, OUT, pr::SPEC_GENERIC>
            <<<grid, block, pr::pair_smem_bytes<FEAT, pr::SPEC_GENERIC>(), s>>>(p, n_frames, rps);
    }
    return cudaGetLastError();
}

// Flexion: two columns per thread, so every even width runs it unchanged.
template <int IN, int OUT, bool LSB>
cudaError_t launch_flexion_pair_as(const kparams& p, int n_frames, cudaStream_t s, launch_info& li) {
    // a longer march amortises the point window (three rows of camera-frame points
    // per strip start): 72 rows measured 1.5 % faster than 36 (DC_FLEX_ROWS, read once)
    static const int flex_rows = [] {
        const int n = env_int("DC_FLEX_ROWS", 0);
        return n >= 3 ? n : 4 * rows_per_strip();
    }();
    const int       rps     = flex_rows;
    const long long strips  = (p.h + rps - 1) / rps;
    const long long pairs   = (n_frames + 1) / 2;
    const long long threads = pairs * strips * (p.w / 2);
    const long long blocks  = (threads + pr::kPairThreads - 1) / pr::kPairThreads;
    if (blocks > 0x7fffffffLL)
        return cudaErrorInvalidConfiguration;
    // per device (one context per GPU in a depth2x --devices N process): not cached
    const cudaError_t carve = cudaFuncSetAttribute(pr::flexion_pair_kernel<IN, OUT, LSB>,
                                                   cudaFuncAttributePreferredSharedMemoryCarveout,
                                                   int(cudaSharedmemCarveoutMaxShared));
    if (carve != cudaSuccess)
        return carve;
    const size_t smem = size_t(3) * pr::kFlexSlots * pr::kPairThreads * sizeof(uint4);
    li.kernel   = LSB ? "dc::pr::flexion_pair_kernel<LSB>" : "dc::pr::flexion_pair_kernel";
    li.launches += 1;
    pr::flexion_pair_kernel<IN, OUT, LSB><<<unsigned(blocks), pr::kPairThreads, smem, s>>>(p, n_frames, rps);
    return cudaGetLastError();
}

// 16-bit output: the kernel that is within +-1 LSB of the reference instead of bit exact
// (dc_pair_kernels.cuh "flexion for 16-bit output"), unless DC_EXACT_U16=1.
inline bool exact_u16() {
    static const bool v = env_int("DC_EXACT_U16", 0) != 0;
    return v;
}
template <int IN, int OUT>
cudaError_t launch_flexion_pair(const kparams& p, int n_frames, cudaStream_t s, launch_info& li) {
    if constexpr (OUT == OUT_U16) {
        if (!exact_u16())
            return launch_flexion_pair_as<IN, OUT, true>(p, n_frames, s, li);
    }
    return launch_flexion_pair_as<IN, OUT, false>(p, n_frames, s, li);
}

// DC_RANGE / DC_SCALE with u16 input: the streaming kernel (8 pixels per thread,
// frames walked inside the thread). Needs frames that are multiples of 8 pixels
// and 16-byte aligned buffers.
template <int FEAT, int IN, int OUT>
cudaError_t launch_stream(const kparams& p, int n_frames, cudaStream_t s, launch_info& li) {
    const int groups = int(p.frame_px / 8);
    const int bx     = (groups + 255) / 256;
    // about 8 resident blocks per SM; at least 4 frames per thread when possible
    int by = std::max(1, std::min((148 * 8 + bx - 1) / bx, (n_frames + 3) / 4));
    by     = std::min(by, 65535);
    li.kernel   = "dc::pr::stream_kernel";
    li.launches += 1;
    pr::stream_kernel<FEAT, IN, OUT><<<dim3(bx, by, 1), 256, 0, s>>>(p, n_frames, groups);
    return cudaGetLastError();
}

// block shape of the one-frame kernels that wastes the fewest threads
inline dim3 scalar_block(int qw, int h, unsigned& gx, unsigned& gy) {
    static const int shapes[4][2] = {{32, 8}, {16, 16}, {64, 4}, {8, 32}};
    int              best         = 0;
    long long        best_cost    = -1;
    for (int i = 0; i < 4; ++i) {
        const long long cx   = (qw + shapes[i][0] - 1) / shapes[i][0];
        const long long cy   = (h + shapes[i][1] - 1) / shapes[i][1];
        const long long cost = cx * cy;
        if (best_cost < 0 || cost < best_cost) {
            best_cost = cost;
            best      = i;
        }
    }
    const dim3 block(shapes[best][0], shapes[best][1], 1);
    gx = (qw + block.x - 1) / block.x, gy = (h + block.y - 1) / block.y;
    return block;
}

// The one-frame-per-z-slice kernels (dc_kernels.cuh): odd widths, unaligned
// buffers, cameras that fail the range proof.
template <int FEAT, int IN, int OUT>
cudaError_t launch_scalar(const kparams& p, int n_frames, cudaStream_t s, launch_info& li) {
    const bool vec4 = li.vec4, fast = li.fast && IN != IN_F32_RANGE;
    li.kernel = vec4 && fast ? "dc::feature_kernel<4, fast>" : vec4 ? "dc::feature_kernel<4>" : "dc::feature_kernel<1>";
    const int  px = vec4 ? 4 : 1;
    unsigned   gx, gy;
    const dim3 block = scalar_block((p.w + px - 1) / px, p.h, gx, gy);
    const size_t in_b = IN == IN_F32_RANGE ? 4 : 2, out_b = OUT == OUT_F32 ? 4 : 2;
    for (int f0 = 0; f0 < n_frames; f0 += 65535) {
        const int nf = std::min(65535, n_frames - f0);
        kparams   q  = p;
        q.in         = static_cast<const char*>(p.in) + size_t(f0) * p.frame_px * in_b;
        for (int k = 0; k < 4; ++k)
            if (p.out[k] != nullptr)
                q.out[k] = static_cast<char*>(p.out[k]) + size_t(f0) * p.frame_px * out_b;
        const dim3 grid(gx, gy, unsigned(nf));
        if (vec4 && fast)
            feature_kernel<FEAT, IN, OUT, 4, true><<<grid, block, 0, s>>>(q);
        else if (vec4)
            feature_kernel<FEAT, IN, OUT, 4, false><<<grid, block, 0, s>>>(q);
        else
            feature_kernel<FEAT, IN, OUT, 1, false><<<grid, block, 0, s>>>(q);
        li.launches += 1;
        const cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess)
            return e;
    }
    return cudaSuccess;
}

// Behind a guard-free f32 launch: the IEEE kernel that runs only if that launch
// found a value outside the proven window (feature_fallback_kernel). About one
// wave of CTAs, so that the usual case (nothing to do) costs a few microseconds.
template <int FEAT, int OUT>
cudaError_t launch_fallback(const kparams& p, int n_frames, cudaStream_t s, launch_info& li) {
    const int  px = li.vec4 ? 4 : 1;
    unsigned   gx, gy;
    const dim3 block = scalar_block((p.w + px - 1) / px, p.h, gx, gy);
    const long long per_frame = (long long)gx * gy;
    const unsigned  gz = unsigned(std::max<long long>(1, std::min<long long>(n_frames, (148 * 8 + per_frame - 1) / per_frame)));
    const dim3 grid(gx, gy, gz);
    if (li.vec4)
        feature_fallback_kernel<FEAT, IN_F32_RANGE, OUT, 4><<<grid, block, 0, s>>>(p, n_frames);
    else
        feature_fallback_kernel<FEAT, IN_F32_RANGE, OUT, 1><<<grid, block, 0, s>>>(p, n_frames);
    li.launches += 1;
    return cudaGetLastError();
}

template <int FEAT, int IN, int OUT>
cudaError_t launch_px(const kparams& p, int n_frames, cudaStream_t s, launch_info& li) {
    if constexpr (IN != IN_F32_RANGE && (FEAT == FEAT_RANGE || FEAT == FEAT_SCALE)) {
        if (li.vec4 && (p.frame_px % 8) == 0 && aligned(p.in, 16) && aligned(p.out[0], 16))
            return launch_stream<FEAT, IN, OUT>(p, n_frames, s, li);
    }
    if constexpr (has_pair_kernel<FEAT>() || FEAT == FEAT_FLEXION) {
        // the guard-free frame-pair kernels: per-camera range proof, an even width;
        // f32 input additionally under the run-time window check (kparams::guard)
        const bool f32ok = IN != IN_F32_RANGE || (p.guard != nullptr && p.fallbacks != nullptr);
        if (li.fast && f32ok && (li.vec4 || li.even)) {
            if constexpr (IN == IN_F32_RANGE) {
                const cudaError_t e = cudaMemsetAsync(p.guard, 0, sizeof(int), s);
                if (e != cudaSuccess)
                    return e;
            }
            cudaError_t e;
            if constexpr (FEAT == FEAT_FLEXION)
                e = launch_flexion_pair<IN, OUT>(p, n_frames, s, li);
            else
                e = launch_pair<FEAT, IN, OUT>(p, n_frames, s, li, !li.vec4);
            if constexpr (IN == IN_F32_RANGE) {
                if (e == cudaSuccess) {
                    const char* fast_name = li.kernel;
                    e                     = launch_fallback<FEAT, OUT>(p, n_frames, s, li);
                    li.kernel             = fast_name;
                }
            }
            return e;
        }
    }
    return launch_scalar<FEAT, IN, OUT>(p, n_frames, s, li);
}

template <int FEAT, int IN>
cudaError_t launch_out(const kparams& p, int out_kind, int n, cudaStream_t s, launch_info& li) {
    return out_kind == OUT_F32 ? launch_px<FEAT, IN, OUT_F32>(p, n, s, li)
                               : launch_px<FEAT, IN, OUT_U16>(p, n, s, li);
}
template <int FEAT>
cudaError_t launch_feature(const kparams& p, int in_kind, int out_kind, int n, cudaStream_t s,
                           launch_info& li) {
    switch (in_kind) {
    case IN_U16_DEPTH: return launch_out<FEAT, IN_U16_DEPTH>(p, out_kind, n, s, li);
    case IN_U16_RANGE: return launch_out<FEAT, IN_U16_RANGE>(p, out_kind, n, s, li);
    default: return launch_out<FEAT, IN_F32_RANGE>(p, out_kind, n, s, li);
    }
}

}  // namespace ln
}  // namespace dc
#endif
