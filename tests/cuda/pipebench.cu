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

__constant__ float k_nz = -0.0f;
__device__ __forceinline__ float2 un(u64 v) {
    float2 r;
    asm("mov.b64 {%0,%1}, %2;" : "=f"(r.x), "=f"(r.y) : "l"(v));
    return r;
}
constexpr int CH = 8;  // independent chains per thread

template <int MODE>
__global__ void __launch_bounds__(512, 1) bench(float* out, int iters, long long* cyc, float upar) {
    float x[CH], y[CH], z[CH];
    u64   p[CH], q[CH], r[CH];
    const float s = 1.0f + threadIdx.x * 1e-7f, t = 0.5f;
#pragma unroll
    for (int i = 0; i < CH; ++i) {
        x[i] = s + i;
        y[i] = s * 0.5f + i;
        z[i] = s * 0.3f + i;
        p[i] = pk(x[i], y[i]);
        q[i] = pk(y[i], x[i]);
        r[i] = pk(y[i] * 0.25f, x[i] * 0.125f);
    }
    const u64 cs = pk(s, s), ct = pk(t, t);
    const float nzr = out[blockIdx.x * blockDim.x + threadIdx.x];  // loop invariant, per thread (not uniform)
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
            // operand forms of FFMA2 as the product kernels use them (C intrinsics)
            if (MODE == 13) { float2 r = __ffma2_rn(un(p[i]), un(cs), make_float2(k_nz, k_nz)); p[i] = pk(r.x, r.y); }  // c = UR.F32
            if (MODE == 14) { float2 r = __ffma2_rn(un(p[i]), make_float2(s, s), un(ct)); p[i] = pk(r.x, r.y); }      // b = R.F32 broadcast
            if (MODE == 15) { float2 r = __ffma2_rn(un(p[i]), un(cs), make_float2(0.0878f, 0.0878f)); p[i] = pk(r.x, r.y); }  // c = immediate
            if (MODE == 16) { float2 r = __ffma2_rn(un(p[i]), make_float2(s, s), make_float2(k_nz, k_nz)); p[i] = pk(r.x, r.y); }  // b = R.F32, c = UR.F32
            if (MODE == 17) { float2 r = __ffma2_rn(un(p[i]), un(p[i]), make_float2(k_nz, k_nz)); p[i] = pk(r.x, r.y); }  // square, c = UR.F32
            // register operand patterns (4 warps per SMSP, as the product kernels run)
            if (MODE == 19) p[i] = fma2(p[i], q[i], r[i]);                       // three distinct register pairs
            if (MODE == 20) p[i] = fma2(p[i], p[i], r[i]);                       // a == b
            if (MODE == 21) { float2 v = __ffma2_rn(un(p[i]), un(q[i]), make_float2(k_nz, k_nz)); p[i] = pk(v.x, v.y); }  // a, b distinct, c = UR
            if (MODE == 22) { float2 v = __ffma2_rn(un(p[i]), un(q[i]), make_float2(0.0878f, 0.0878f)); p[i] = pk(v.x, v.y); }  // c = imm
            if (MODE == 23) { float2 v = __ffma2_rn(un(p[i]), make_float2(x[i], x[i]), un(r[i])); p[i] = pk(v.x, v.y); }  // b = R.F32 broadcast, distinct
            if (MODE == 24) p[i] = add2(p[i], q[i]);                              // FADD2 distinct
            if (MODE == 25) { p[i] = fma2(p[i], q[i], r[i]); q[i] = fma2(q[i], r[i], p[i]); }  // rotating operands, no reuse
            if (MODE == 26) x[i] = ffma(x[i], y[i], s);                           // scalar FFMA distinct
            if (MODE == 27) { float2 v = __ffma2_rn(un(p[i]), make_float2(x[i], x[i]), make_float2(k_nz, k_nz)); p[i] = pk(v.x, v.y); }  // b = R.F32 varying, c = UR
            if (MODE == 28) x[i] = ffma(x[i], y[i], z[i]);                        // scalar, three distinct
            if (MODE == 29) { float2 v = __ffma2_rn(un(p[i]), make_float2(0.0878f, 0.0878f), un(r[i])); p[i] = pk(v.x, v.y); }  // b = imm, a, c distinct
            if (MODE == 30) { float2 a = un(p[i]), b = un(q[i]), c = un(r[i]);    // the packed op as two scalar FFMAs on register pairs
                              p[i] = pk(ffma(a.x, b.x, c.x), ffma(a.y, b.y, c.y)); }
            if (MODE == 31) { float2 a = un(p[i]), b = un(q[i]);                  // two scalar FMULs on register pairs
                              p[i] = pk(fmul(a.x, b.x), fmul(a.y, b.y)); }
            if (MODE == 32) { float2 a = un(p[i]); p[i] = pk(ffma(a.x, x[i], t), ffma(a.y, x[i], t)); }  // scalar pair, shared b
            if (MODE == 33) { p[i] = fma2(p[i], q[i], r[i]); x[i] = ffma(x[i], y[i], z[i]); y[i] = fmnmx(y[i], s); }  // FFMA2 + FFMA + ALU, distinct
            if (MODE == 34) { float2 a = un(p[i]), b = un(q[i]), c = un(r[i]);    // 2 scalar + 1 ALU
                              p[i] = pk(ffma(a.x, b.x, c.x), ffma(a.y, b.y, c.y)); y[i] = fmnmx(y[i], s); }
            // Is the register file's read bandwidth shared between the pipes? FFMA2 with
            // two distinct register pairs (FMA pipe 2.07 clk) next to ALU-pipe work
            // (2 clk per instruction) that reads further registers: independent pipes
            // predict max(FMA, ALU), a shared operand bandwidth of two 32-bit registers
            // per lane and clock predicts (registers read) / 2.
            if (MODE == 40) { float2 v = __ffma2_rn(un(p[i]), un(q[i]), make_float2(k_nz, k_nz)); p[i] = pk(v.x, v.y); y[i] = fmnmx(y[i], z[i]); }  // 4 + 2 regs
            if (MODE == 41) { p[i] = fma2(p[i], q[i], r[i]); y[i] = fmnmx(y[i], z[i]); }  // 6 + 2 regs
            if (MODE == 43) { float2 v = __ffma2_rn(un(p[i]), un(q[i]), make_float2(k_nz, k_nz)); p[i] = pk(v.x, v.y); y[i] = fmnmx(y[i], 0.75f); }  // 4 + 1 regs
            if (MODE == 44) { float2 v = __ffma2_rn(un(p[i]), un(q[i]), make_float2(k_nz, k_nz)); p[i] = pk(v.x, v.y);
                              float2 u = __ffma2_rn(un(q[i]), un(r[i]), make_float2(k_nz, k_nz)); q[i] = pk(u.x, u.y); y[i] = fmnmx(y[i], z[i]); }  // 2 FFMA2 (8 regs) + ALU (2)
            if (MODE == 45) { float2 v = __ffma2_rn(un(p[i]), un(q[i]), make_float2(k_nz, k_nz)); p[i] = pk(v.x, v.y);
                              float2 u = __ffma2_rn(un(q[i]), un(r[i]), make_float2(k_nz, k_nz)); q[i] = pk(u.x, u.y);
                              y[i] = fmnmx(y[i], z[i]); x[i] = fmnmx(x[i], z[i]); }  // 2 FFMA2 + 2 ALU: the product kernels' mix
            if (MODE == 46) { float2 v = __ffma2_rn(un(p[i]), un(q[i]), make_float2(k_nz, k_nz)); p[i] = pk(v.x, v.y);
                              float2 u = __ffma2_rn(un(q[i]), un(r[i]), make_float2(k_nz, k_nz)); q[i] = pk(u.x, u.y);
                              y[i] = fmnmx(y[i], 0.75f); x[i] = fmnmx(x[i], 0.5f); }  // same, the ALU work reads one register each
            // the addend as a scalar-broadcast REGISTER (c = R.F32): what ptxas emits for the
            // -0.0 of dc::pr::mul2 inside loops with several exits (flexion, curvature,
            // shared-edge max-curve), where it does not keep the constant in a uniform register
            if (MODE == 47) { float2 v = __ffma2_rn(un(p[i]), un(q[i]), make_float2(nzr, nzr)); p[i] = pk(v.x, v.y); }
            if (MODE == 48) { float2 v = __ffma2_rn(un(p[i]), un(p[i]), make_float2(nzr, nzr)); p[i] = pk(v.x, v.y); }
            if (MODE == 49) { float2 v = __ffma2_rn(un(p[i]), make_float2(x[i], x[i]), make_float2(nzr, nzr)); p[i] = pk(v.x, v.y); }
            // division residuals with a camera-only divisor (curvature): the scalar as an
            // ordinary register (mode 23) against the same scalar as a uniform operand
            if (MODE == 50) { float2 v = __ffma2_rn(un(p[i]), make_float2(upar, upar), un(r[i])); p[i] = pk(v.x, v.y); }
            if (MODE == 18) { x[i] = rsq(x[i]); y[i] = rsq(y[i]); p[i] = fma2(p[i], cs, ct); q[i] = fma2(q[i], ct, cs); }  // 2 MUFU back to back + 2 FFMA2
        }
    }
    const long long c1 = clock64();
    float acc = 0.f;
#pragma unroll
    for (int i = 0; i < CH; ++i) {
        float a, b, c, d;
        asm("mov.b64 {%0,%1}, %2;" : "=f"(a), "=f"(b) : "l"(p[i]));
        asm("mov.b64 {%0,%1}, %2;" : "=f"(c), "=f"(d) : "l"(q[i]));
        acc += x[i] + y[i] + z[i] + a + b + c + d + un(r[i]).x;
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = c1 - c0;
}

template <int MODE>
void run(const char* name, int inst_per_chain, int lanes_per_chain, int threads = 512) {
    // one 512-thread CTA per SM: 4 warps per SMSP (384: 3, 256: 2)
    const int  blocks = 148, iters = 4096;
    float*     out;
    long long* cyc;
    cudaMalloc(&out, sizeof(float) * blocks * threads);
    cudaMemset(out, 0, sizeof(float) * blocks * threads);
    cudaMalloc(&cyc, 8);
    bench<MODE><<<blocks, threads>>>(out, 16, cyc, 0.999f);
    cudaEvent_t a, b;
    cudaEventCreate(&a);
    cudaEventCreate(&b);
    cudaEventRecord(a);
    bench<MODE><<<blocks, threads>>>(out, iters, cyc, 0.999f);
    cudaEventRecord(b);
    cudaDeviceSynchronize();
    float ms;
    cudaEventElapsedTime(&ms, a, b);
    long long c;
    cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
    const double winst_smsp = (threads / 128.0) * iters * CH * inst_per_chain;
    const double ipc        = winst_smsp / double(c);
    printf("%-28s %8.3f ms  cycles %9lld  clk/warp-inst/SMSP %.3f  fp32 lane-ops/clk/SM %.1f\n", name,
           ms, c, 1.0 / ipc, ipc * 4 * 32 * lanes_per_chain / inst_per_chain);
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
    run<13>("FFMA2 c=UR.F32", 1, 2);
    run<14>("FFMA2 b=R.F32 bcast", 1, 2);
    run<15>("FFMA2 c=imm", 1, 2);
    run<16>("FFMA2 b=R.F32 c=UR.F32", 1, 2);
    run<17>("FFMA2 square c=UR.F32", 1, 2);
    run<18>("2 RSQ + 2 FFMA2", 4, 6);
    run<19>("FFMA2 a,b,c distinct", 1, 2);
    run<20>("FFMA2 a==b, c distinct", 1, 2);
    run<21>("FFMA2 a,b distinct c=UR", 1, 2);
    run<22>("FFMA2 a,b distinct c=imm", 1, 2);
    run<23>("FFMA2 b=R.F32 distinct", 1, 2);
    run<24>("FADD2 distinct", 1, 2);
    run<25>("FFMA2 rotating operands", 2, 4);
    run<26>("FFMA a,b distinct", 1, 1);
    run<27>("FFMA2 b=R.F32 varying c=UR", 1, 2);
    run<28>("FFMA a,b,c distinct", 1, 1);
    run<29>("FFMA2 b=imm a,c distinct", 1, 2);
    run<30>("2x FFMA on pairs distinct", 2, 2);
    run<31>("2x FMUL on pairs distinct", 2, 2);
    run<32>("2x FFMA pair, shared b", 2, 2);
    run<33>("FFMA2+FFMA+FMNMX distinct", 3, 3);
    run<34>("2x FFMA + FMNMX", 3, 2);
    for (int thr : {512, 384, 256}) {
        printf("-- %d warps per SMSP\n", thr / 128);
        run<21>("FFMA2 a,b distinct c=UR", 1, 2, thr);
        run<40>("FFMA2(4 regs) + ALU(2 regs)", 2, 2, thr);
        run<43>("FFMA2(4 regs) + ALU(1 reg)", 2, 2, thr);
        run<41>("FFMA2(6 regs) + ALU(2 regs)", 2, 2, thr);
        run<44>("2 FFMA2(8) + ALU(2)", 3, 4, thr);
        run<45>("2 FFMA2(8) + 2 ALU(4)", 4, 4, thr);
        run<46>("2 FFMA2(8) + 2 ALU(2)", 4, 4, thr);
        run<47>("FFMA2 a,b distinct c=R.F32", 1, 2, thr);
        run<48>("FFMA2 square c=R.F32", 1, 2, thr);
        run<49>("FFMA2 b=R.F32 c=R.F32", 1, 2, thr);
        run<23>("FFMA2 b=R.F32, a,c pairs", 1, 2, thr);
        run<50>("FFMA2 b=uniform, a,c pairs", 1, 2, thr);
        run<19>("FFMA2 a,b,c pairs", 1, 2, thr);
    }
    return 0;
}
