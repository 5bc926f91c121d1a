// GPU self-check of depth-conversions_b200/csrc/dc_fastmath.cuh against the
// IEEE intrinsics (__fsqrt_rn, __fdiv_rn) and double-precision acos.
// Prints one line per check: "<name> checked=<n> mismatches=<m> [max_rel=<e>]".
// Exit code 0 iff every exactness check has 0 mismatches and acos max_rel < 1e-6.
#include "../../depth-conversions_b200/csrc/dc_fastmath.cuh"
#include "../../depth-conversions_b200/csrc/dc_pair.cuh"

#include <cmath>
#include <cstdio>
#include <cstdlib>

__device__ unsigned long long splitmix(unsigned long long x) {
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}
// positive float with exponent uniform in [elo, ehi] (unbiased) and random mantissa
__device__ float rnd_float(unsigned long long h, int elo, int ehi) {
    const unsigned mant = unsigned(h) & 0x7fffffu;
    const int      e    = elo + int((h >> 32) % unsigned(ehi - elo + 1));
    return __uint_as_float((unsigned(e + 127) << 23) | mant);
}

__global__ void check_sqrt(unsigned long long* bad, unsigned lo, unsigned hi) {
    // exhaustive over bit patterns [lo, hi]
    const unsigned long long n = (unsigned long long)hi - lo + 1;
    for (unsigned long long i = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x; i < n;
         i += (unsigned long long)gridDim.x * blockDim.x) {
        const float x = __uint_as_float(lo + unsigned(i));
        if (__float_as_uint(dc::sqrt_rn_fast(x)) != __float_as_uint(__fsqrt_rn(x)))
            atomicAdd(bad, 1ull);
    }
}

__global__ void check_div(unsigned long long* bad, unsigned long long n, int mode) {
    for (unsigned long long i = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x; i < n;
         i += (unsigned long long)gridDim.x * blockDim.x) {
        const unsigned long long h1 = splitmix(i * 2 + 0x1234567ull * mode);
        const unsigned long long h2 = splitmix(h1 ^ 0xABCDEF12345ull);
        float a, b;
        if (mode == 0) {  // wide exponents
            a = rnd_float(h1, -60, 60);
            b = rnd_float(h2, -40, 40);
        } else if (mode == 1) {  // bearing: |nom| vs den, ratio near +-1
            b = rnd_float(h2, -12, 21);
            a = b * (1.0f + float(int(h1 & 0xffff) - 32768) * 1e-7f);
        } else {  // b with extreme mantissas (all ones / all zeros / one bit)
            a = rnd_float(h1, -30, 30);
            const unsigned pats[4] = {0x7fffffu, 0x000000u, 0x000001u, 0x7ffffeu};
            b = __uint_as_float((__float_as_uint(rnd_float(h2, -30, 30)) & 0xff800000u) |
                                pats[(h2 >> 60) & 3]);
        }
        if (h1 >> 63)
            a = -a;
        if (__float_as_uint(dc::div_rn_fast(a, b)) != __float_as_uint(__fdiv_rn(a, b)))
            atomicAdd(bad, 1ull);
    }
}

// the packed square root of the frame-pair kernels (dc_pair.cuh sqrt2; with
// -DDC_SQRT_MUFU_SEED=1 its first estimate comes from MUFU.SQRT), exhaustive
__global__ void check_sqrt_pair(unsigned long long* bad, unsigned lo, unsigned hi) {
    const unsigned long long n = ((unsigned long long)hi - lo + 2) / 2;
    for (unsigned long long i = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x; i < n;
         i += (unsigned long long)gridDim.x * blockDim.x) {
        const unsigned b0 = lo + unsigned(2 * i), b1 = min(b0 + 1, hi);
        const float2   x  = make_float2(__uint_as_float(b0), __uint_as_float(b1));
        float2         y;
        const float2   g = dc::pr::sqrt2(x, y);
        if (__float_as_uint(g.x) != __float_as_uint(__fsqrt_rn(x.x)))
            atomicAdd(bad, 1ull);
        if (__float_as_uint(g.y) != __float_as_uint(__fsqrt_rn(x.y)))
            atomicAdd(bad, 1ull);
    }
}
// acos of the frame-pair kernels (dc_pair.cuh acos_signed2), exhaustive
__global__ void check_acos_pair(unsigned long long* worst_bits) {
    const unsigned hi = __float_as_uint(1.0f);  // acos_abs2 clamps to 1 - 10 eps itself
    double         w  = 0.0;
    for (unsigned long long i = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x;
         i <= hi; i += (unsigned long long)gridDim.x * blockDim.x) {
        const float  x  = __uint_as_float(unsigned(i));
        const float2 q  = make_float2(x, -x);
        const float2 r  = dc::pr::acos_signed2(q);  // what the kernels call
        const float  c  = 1.0f - 10.0f * 1.1920928955078125e-07f;
        const double xc = double(fminf(x, c));
        w = fmax(w, fabs(double(r.x) - acos(xc)) / acos(xc));
        w = fmax(w, fabs(double(r.y) - acos(-xc)) / acos(-xc));
    }
    atomicMax(worst_bits, (unsigned long long)__double_as_longlong(w));
}

// the degree-4 polynomial of the 16-bit kernels (dc_pair.cuh acos_poly2<true>): ABSOLUTE
// error of the angle over every float in [-1, 1], in radians
__global__ void check_acos_pair_u16(unsigned long long* worst_bits) {
    const unsigned hi = __float_as_uint(1.0f);
    double         w  = 0.0;
    for (unsigned long long i = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x;
         i <= hi; i += (unsigned long long)gridDim.x * blockDim.x) {
        const float  x  = __uint_as_float(unsigned(i));
        const float2 r  = dc::pr::acos_signed2<true>(make_float2(x, -x));
        const float  c  = 1.0f - 10.0f * 1.1920928955078125e-07f;
        const double xc = double(fminf(x, c));
        w = fmax(w, fabs(double(r.x) - acos(xc)));
        w = fmax(w, fabs(double(r.y) - acos(-xc)));
    }
    atomicMax(worst_bits, (unsigned long long)__double_as_longlong(w));
}

// normalisation as flexion does it: component / sqrt(z), reciprocal shared and
// seeded from the rsqrt of the square root step.
__global__ void check_norm(unsigned long long* bad, unsigned long long n) {
    for (unsigned long long i = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x; i < n;
         i += (unsigned long long)gridDim.x * blockDim.x) {
        const unsigned long long h1 = splitmix(i * 3 + 77);
        const unsigned long long h2 = splitmix(h1 ^ 0x5555aaaa5555ull);
        const float z = rnd_float(h1, -60, 60);
        // |component| <= sqrt(z); pick a fraction of it, sometimes tiny, sometimes 0
        const float s_ref = __fsqrt_rn(z);
        float       a     = s_ref * __uint_as_float(0x3f800000u - unsigned(h2 & 0x3ffffffu));
        if (((h2 >> 40) & 15) == 0)
            a *= 1e-9f;
        if (((h2 >> 44) & 63) == 0)
            a = 0.0f;
        if (h2 >> 63)
            a = -a;
        float       hh;
        const float s = dc::sqrt_rn_fast(z, hh);
        const float y = dc::rcp_refined(s, __fadd_rn(hh, hh));
        const float q = dc::div_rn_with_rcp(a, s, y);
        const float want = __fdiv_rn(a, s_ref);
        if (__float_as_uint(s) != __float_as_uint(s_ref) ||
            (__float_as_uint(q) != __float_as_uint(want) && !(q == 0.0f && want == 0.0f)))
            atomicAdd(bad, 1ull);
    }
}

__global__ void check_acos(unsigned long long* worst_bits) {
    // exhaustive over |x| <= 1 - 10 eps, both signs: bit patterns [0, 0x3f7fffec]
    const unsigned hi = __float_as_uint(1.0f - 10.0f * 1.1920928955078125e-07f);
    double         w  = 0.0;
    for (unsigned long long i = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x;
         i <= 2ull * hi + 1; i += (unsigned long long)gridDim.x * blockDim.x) {
        const unsigned b = unsigned(i >> 1);
        float          x = __uint_as_float(b);
        if (i & 1)
            x = -x;
        const double ref = acos(double(x));
        const double got = double(dc::acos_rel(x));
        const double rel = fabs(got - ref) / ref;
        w                = fmax(w, rel);
    }
    atomicMax(worst_bits, (unsigned long long)__double_as_longlong(w));
}

int main() {
    unsigned long long* d;
    cudaMalloc(&d, 8);
    auto run = [&](const char* name, unsigned long long n, auto launch) {
        cudaMemset(d, 0, 8);
        launch();
        unsigned long long h = 0;
        const cudaError_t  e = cudaDeviceSynchronize();
        cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
        if (e != cudaSuccess) {
            printf("%s CUDA error %s\n", name, cudaGetErrorString(e));
            exit(2);
        }
        printf("%s checked=%llu mismatches=%llu\n", name, n, h);
        return h;
    };
    unsigned long long bad = 0;
    const unsigned lo = 0x0d800000u /* 2^-100 */, hi = 0x7f7fffffu;
    bad += run("sqrt_rn_fast_exhaustive", (unsigned long long)hi - lo + 1,
               [&] { check_sqrt<<<148 * 16, 256>>>(d, lo, hi); });
    bad += run("pair_sqrt2_exhaustive", (unsigned long long)hi - lo + 1,
               [&] { check_sqrt_pair<<<148 * 16, 256>>>(d, lo, hi); });
    const unsigned long long N = 1ull << 33;
    for (int m = 0; m < 3; ++m) {
        char nm[64];
        snprintf(nm, sizeof nm, "div_rn_fast_mode%d", m);
        bad += run(nm, N, [&] { check_div<<<148 * 16, 256>>>(d, N, m); });
    }
    bad += run("normalise_shared_rcp", N, [&] { check_norm<<<148 * 16, 256>>>(d, N); });
    cudaMemset(d, 0, 8);
    check_acos<<<148 * 16, 256>>>(d);
    cudaDeviceSynchronize();
    unsigned long long wb = 0;
    cudaMemcpy(&wb, d, 8, cudaMemcpyDeviceToHost);
    double worst;
    memcpy(&worst, &wb, 8);
    printf("acos_rel exhaustive max_rel=%.3e\n", worst);
    cudaMemset(d, 0, 8);
    check_acos_pair<<<148 * 16, 256>>>(d);
    cudaDeviceSynchronize();
    cudaMemcpy(&wb, d, 8, cudaMemcpyDeviceToHost);
    double worst_pair;
    memcpy(&worst_pair, &wb, 8);
    printf("pair acos_signed2 exhaustive max_rel=%.3e\n", worst_pair);
    cudaMemset(d, 0, 8);
    check_acos_pair_u16<<<148 * 16, 256>>>(d);
    cudaDeviceSynchronize();
    cudaMemcpy(&wb, d, 8, cudaMemcpyDeviceToHost);
    double worst_u16;
    memcpy(&worst_u16, &wb, 8);
    // 0.12 LSB of convert_bearing<ushort> (pi / 65535 = 4.79e-5 rad)
    printf("pair acos_signed2<u16> exhaustive max_abs_rad=%.3e\n", worst_u16);
    const bool ok = bad == 0 && worst < 1e-6 && worst_pair < 2e-6 && worst_u16 < 6e-6;
    printf("%s\n", ok ? "MATHCHECK OK" : "MATHCHECK FAILED");
    return ok ? 0 : 1;
}
