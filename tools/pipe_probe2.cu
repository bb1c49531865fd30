// pipe_probe2.cu -- developer microbenchmark (not part of the product): packed FFMA2 with three DISTINCT register-pair
// operands (what the EKF update issues), its dependent-issue latency, and mixes with MUFU at the rollout kernel's
// occupancy (5 warps per scheduler).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/pipe_probe2 tools/pipe_probe2.cu && tools/pipe_probe2
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 ffma2(u64 a, u64 b, u64 c) { u64 r; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
__device__ __forceinline__ u64 fmul2(u64 a, u64 b) { u64 r; asm volatile("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ float ffma(float a, float b, float c) { float r; asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c)); return r; }
__device__ __forceinline__ u64 fadd2(u64 a, u64 b) { u64 r; asm volatile("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ u64 bc2(float s) { u64 r; asm("mov.b64 %0, {%1, %1};" : "=l"(r) : "f"(s)); return r; }
__device__ __forceinline__ float rcp(float a) { float r; asm volatile("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(a)); return r; }

// MODE 0: 8 independent FFMA2 chains, operands shared (p, q)        MODE 1: operands all distinct, rotating
// MODE 2: ONE dependent FFMA2 chain (latency)                       MODE 3: one dependent scalar FFMA chain
// MODE 4: 8 FFMA2 distinct + 2 MUFU per iteration                   MODE 5: 8 scalar FFMA distinct operands
// MODE 6: 8 FMUL2 distinct                                          MODE 7: 16 scalar FFMA distinct (same flops as MODE 1)
// MODE 8: FFMA2 with a broadcast 32-bit register operand   MODE 9: FFMA2 a * b + a (two distinct pairs)
// MODE 10: FFMA2 with a kernel-parameter (uniform) operand  MODE 11: 8 FADD2 distinct   MODE 12: 4 FFMA2 distinct + 8 FFMA distinct
template <int MODE>
__global__ void __launch_bounds__(128) k(u64* out, int iters, u64 seed, float cst = 1.5f) {
    u64 x[8], y[8], z[8];
    float f[16], g[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) { x[i] = seed + i + threadIdx.x; y[i] = seed * 3 + i; z[i] = seed * 7 + i; g[i] = 1.0f + i; }
#pragma unroll
    for (int i = 0; i < 16; ++i) f[i] = float(i) + threadIdx.x;
    for (int it = 0; it < iters; ++it) {
        if (MODE == 0) {
#pragma unroll
            for (int i = 0; i < 8; ++i) x[i] = ffma2(x[i], y[0], z[0]);
        } else if (MODE == 1 || MODE == 4) {
#pragma unroll
            for (int i = 0; i < 8; ++i) x[i] = ffma2(x[i], y[(i + 1) & 7], z[(i + 3) & 7]);
            if (MODE == 4) { g[0] = rcp(g[0]); g[1] = rcp(g[1]); }
        } else if (MODE == 2) {
#pragma unroll
            for (int i = 0; i < 8; ++i) x[0] = ffma2(x[0], y[i], z[i]);
        } else if (MODE == 3) {
#pragma unroll
            for (int i = 0; i < 8; ++i) f[0] = ffma(f[0], g[i], f[i + 1]);
        } else if (MODE == 5) {
#pragma unroll
            for (int i = 0; i < 8; ++i) f[i] = ffma(f[i], g[(i + 1) & 7], f[8 + ((i + 3) & 7)]);
        } else if (MODE == 6) {
#pragma unroll
            for (int i = 0; i < 8; ++i) x[i] = fmul2(x[i], y[(i + 1) & 7]);
        } else if (MODE == 8) {
#pragma unroll
            for (int i = 0; i < 8; ++i) x[i] = ffma2(x[i], bc2(g[(i + 1) & 7]), z[(i + 3) & 7]);
        } else if (MODE == 9) {
#pragma unroll
            for (int i = 0; i < 8; ++i) x[i] = ffma2(x[i], y[(i + 1) & 7], x[i]);
        } else if (MODE == 10) {
#pragma unroll
            for (int i = 0; i < 8; ++i) x[i] = ffma2(x[i], y[(i + 1) & 7], bc2(cst));
        } else if (MODE == 11) {
#pragma unroll
            for (int i = 0; i < 8; ++i) x[i] = fadd2(x[i], y[(i + 1) & 7]);
        } else if (MODE == 12) {
#pragma unroll
            for (int i = 0; i < 4; ++i) x[i] = ffma2(x[i], y[(i + 1) & 7], z[(i + 3) & 7]);
#pragma unroll
            for (int i = 0; i < 8; ++i) f[i] = ffma(f[i], g[(i + 1) & 7], f[8 + ((i + 3) & 7)]);
        } else if (MODE == 7) {
#pragma unroll
            for (int i = 0; i < 16; ++i) f[i] = ffma(f[i], g[(i + 1) & 7], f[(i + 5) & 15]);
        }
    }
    u64 s = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) s ^= x[i] ^ y[i] ^ z[i] ^ (u64)__float_as_uint(g[i]);
#pragma unroll
    for (int i = 0; i < 16; ++i) s ^= (u64)__float_as_uint(f[i]);
    if (s == 0x1234567ull) out[0] = s;
}

template <int MODE>
static void run(const char* name, int sms, double ghz, int ctas_per_sm, int n_instr) {
    u64* d; cudaMalloc(&d, 64);
    const int grid = sms * ctas_per_sm, iters = 1 << 15;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<MODE><<<grid, 128>>>(d, iters, 5);
    cudaEventRecord(e0);
    k<MODE><<<grid, 128>>>(d, iters, 5);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double warps_per_sched = ctas_per_sm;   // 4 warps per CTA over 4 schedulers
    const double cyc = ms * 1e-3 * ghz * 1e9 / double(iters);   // cycles per iteration (all warps of a scheduler in flight)
    printf("%-44s %d warps/sched: %7.2f cycles/iter, %5.2f cycles per instr per warp (pipe view: %5.2f)\n", name, ctas_per_sm, cyc, cyc / n_instr, cyc / n_instr / warps_per_sched);
    cudaFree(d);
}

int main() {
    cudaDeviceProp pr; cudaGetDeviceProperties(&pr, 0);
    int khz = 0; cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    const double ghz = khz * 1e-6; const int sms = pr.multiProcessorCount;
    printf("%s, %d SMs, %.3f GHz\n", pr.name, sms, ghz);
    for (int w : {5}) {
        run<0>("8 FFMA2, shared operands", sms, ghz, w, 8);
        run<1>("8 FFMA2, distinct operands", sms, ghz, w, 8);
        run<6>("8 FMUL2, distinct operands", sms, ghz, w, 8);
        run<5>("8 FFMA, distinct operands", sms, ghz, w, 8);
        run<7>("16 FFMA, distinct operands", sms, ghz, w, 16);
        run<2>("8 FFMA2, ONE dependent chain", sms, ghz, w, 8);
        run<3>("8 FFMA, ONE dependent chain", sms, ghz, w, 8);
        run<4>("8 FFMA2 distinct + 2 MUFU", sms, ghz, w, 10);
        run<8>("8 FFMA2, broadcast 32-bit register operand", sms, ghz, w, 8);
        run<9>("8 FFMA2, a * b + a", sms, ghz, w, 8);
        run<10>("8 FFMA2, uniform (parameter) addend", sms, ghz, w, 8);
        run<11>("8 FADD2, distinct operands", sms, ghz, w, 8);
        run<12>("4 FFMA2 distinct + 8 FFMA distinct", sms, ghz, w, 12);
    }
    return 0;
}
