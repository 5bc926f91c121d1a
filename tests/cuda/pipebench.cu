// Issue-rate microbenchmark for the FP32 paths the depth kernels are bound by
// (sm_100a): scalar FFMA / FMUL / FADD vs the packed FFMA2 / FMUL2 / FADD2
// (PTX fma.rn.f32x2 ...), alone and mixed with ALU-pipe (FMNMX) and MUFU work.
// Prints warp-instructions per cycle per SM sub-partition (SMSP) and the
// resulting FP32 lane-operations per cycle per SM.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pipebench pipebench.cu
#include <cstdio>
#include <cuda_runtime.h>

typedef unsigned long long u64;
__device__ __forceinline__ u64 pk(float a, float b) {
    u64 r;
    asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(a), "f"(b));
    return r;
}
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) {
    u64 d;
    asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}
__device__ __forceinline__ u64 mul2(u64 a, u64 b) {
    u64 d;
    asm volatile("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
__device__ __forceinline__ u64 add2(u64 a, u64 b) {
    u64 d;
    asm volatile("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
__device__ __forceinline__ float ffma(float a, float b, float c) {
    float d;
    asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
    return d;
}
__device__ __forceinline__ float fmul(float a, float b) {
    float d;
    asm volatile("mul.rn.f32 %0, %1, %2;" : "=f"(d) : "f"(a), "f"(b));
    return d;
}
__device__ __forceinline__ float fadd(float a, float b) {
    float d;
    asm volatile("add.rn.f32 %0, %1, %2;" : "=f"(d) : "f"(a), "f"(b));
    return d;
}
__device__ __forceinline__ float fmnmx(float a, float b) {
    float d;
    asm volatile("min.f32 %0, %1, %2;" : "=f"(d) : "f"(a), "f"(b));
    return d;
}
__device__ __forceinline__ float rsq(float a) {
    float d;
    asm volatile("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(d) : "f"(a));
    return d;
}

constexpr int CH = 8;  // independent chains per thread

template <int MODE>
__global__ void __launch_bounds__(512) bench(float* out, int iters, long long* cyc) {
    float x[CH], y[CH];
    u64   p[CH], q[CH];
    const float s = 1.0f + threadIdx.x * 1e-7f, t = 0.5f;
#pragma unroll
    for (int i = 0; i < CH; ++i) {
        x[i] = s + i;
        y[i] = t + i;
        p[i] = pk(x[i], y[i]);
        q[i] = pk(y[i], x[i]);
    }
    const u64 cs = pk(s, s), ct = pk(t, t);
    __syncthreads();
    const long long c0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < CH; ++i) {
            if (MODE == 0) x[i] = ffma(x[i], s, t);
            if (MODE == 1) p[i] = fma2(p[i], cs, ct);
            if (MODE == 2) x[i] = fmul(x[i], s);
            if (MODE == 3) p[i] = mul2(p[i], cs);
            if (MODE == 4) x[i] = fadd(x[i], s);
            if (MODE == 5) p[i] = add2(p[i], cs);
            if (MODE == 6) { x[i] = ffma(x[i], s, t); y[i] = fmnmx(y[i], x[i]); }         // FFMA + ALU
            if (MODE == 7) { p[i] = fma2(p[i], cs, ct); y[i] = fmnmx(y[i], s); x[i] = fmnmx(x[i], t); }  // FFMA2 + 2 ALU
            if (MODE == 8) { p[i] = fma2(p[i], cs, ct); q[i] = fma2(q[i], ct, cs); y[i] = fmnmx(y[i], s); }  // 2 FFMA2 + 1 ALU
            if (MODE == 9) x[i] = rsq(x[i]);
            if (MODE == 10) { x[i] = rsq(x[i]); p[i] = fma2(p[i], cs, ct); q[i] = fma2(q[i], ct, cs); y[i] = fmnmx(y[i], s); }
            if (MODE == 11) { x[i] = ffma(x[i], s, t); y[i] = ffma(y[i], t, s); }          // 2 scalar FFMA chains
            if (MODE == 12) { p[i] = fma2(p[i], cs, ct); x[i] = ffma(x[i], s, t); }        // FFMA2 + FFMA
        }
    }
    const long long c1 = clock64();
    float acc = 0.f;
#pragma unroll
    for (int i = 0; i < CH; ++i) {
        float a, b, c, d;
        asm("mov.b64 {%0,%1}, %2;" : "=f"(a), "=f"(b) : "l"(p[i]));
        asm("mov.b64 {%0,%1}, %2;" : "=f"(c), "=f"(d) : "l"(q[i]));
        acc += x[i] + y[i] + a + b + c + d;
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = c1 - c0;
}

template <int MODE>
void run(const char* name, int inst_per_chain, int lanes_per_chain) {
    const int  blocks = 148 * 2, threads = 512, iters = 4096;
    float*     out;
    long long* cyc;
    cudaMalloc(&out, sizeof(float) * blocks * threads);
    cudaMalloc(&cyc, 8);
    bench<MODE><<<blocks, threads>>>(out, 16, cyc);
    cudaEvent_t a, b;
    cudaEventCreate(&a);
    cudaEventCreate(&b);
    cudaEventRecord(a);
    bench<MODE><<<blocks, threads>>>(out, iters, cyc);
    cudaEventRecord(b);
    cudaDeviceSynchronize();
    float ms;
    cudaEventElapsedTime(&ms, a, b);
    long long c;
    cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
    // warps per SMSP: 2 blocks * 16 warps / 4 SMSP = 8
    const double winst_smsp = 8.0 * iters * CH * inst_per_chain;
    const double ipc        = winst_smsp / double(c);
    printf("%-28s %8.3f ms  cycles %9lld  warp-inst/clk/SMSP %.3f  fp32 lane-ops/clk/SM %.1f\n", name,
           ms, c, ipc, ipc * 4 * 32 * lanes_per_chain / inst_per_chain);
    cudaFree(out);
    cudaFree(cyc);
}

int main() {
    run<0>("FFMA", 1, 1);
    run<1>("FFMA2", 1, 2);
    run<2>("FMUL", 1, 1);
    run<3>("FMUL2", 1, 2);
    run<4>("FADD", 1, 1);
    run<5>("FADD2", 1, 2);
    run<11>("FFMA x2 chains", 2, 2);
    run<12>("FFMA2 + FFMA", 2, 3);
    run<6>("FFMA + FMNMX", 2, 1);
    run<7>("FFMA2 + 2 FMNMX", 3, 2);
    run<8>("2 FFMA2 + FMNMX", 3, 4);
    run<9>("MUFU.RSQ", 1, 1);
    run<10>("RSQ + 2 FFMA2 + FMNMX", 4, 4);
    return 0;
}
