// pipe_probe.cu -- developer microbenchmark (not part of the product): how packed FFMA2, scalar FFMA with three
// register operands, and ALU-pipe instructions (LOP3 / FSETP+SEL) share an SM sub-partition on sm_100a.
// Prints cycles per loop iteration per warp and per scheduler for several instruction mixes.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/pipe_probe tools/pipe_probe.cu && tools/pipe_probe
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ unsigned long long ffma2(unsigned long long a, unsigned long long b, unsigned long long c) {
    unsigned long long r;
    asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
    return r;
}
__device__ __forceinline__ unsigned lop(unsigned a, unsigned b, unsigned c) {
    unsigned r;
    asm volatile("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
    return r;
}
__device__ __forceinline__ float ffma(float a, float b, float c) {
    float r;
    asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c));
    return r;
}

// NF2 packed FMAs, NF scalar 3-register FMAs, NA LOP3 per iteration; 8 independent chains each
template <int NF2, int NF, int NA>
__global__ void __launch_bounds__(256) mix(unsigned long long* out, int iters, unsigned long long seed) {
    unsigned long long x[8], p = seed + threadIdx.x, q = seed * 3 + threadIdx.x;
    float f[8], fa = float(threadIdx.x) * 1e-3f, fb = fa + 0.5f;
    unsigned u[8], ua = threadIdx.x * 7u, ub = threadIdx.x * 13u;
#pragma unroll
    for (int k = 0; k < 8; ++k) { x[k] = seed + k; f[k] = float(k); u[k] = k * 2654435761u; }
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            if (k < NF2) x[k] = ffma2(x[k], p, q);
            if (k < NF) f[k] = ffma(f[k], fa, fb);
            if (k < NA) u[k] = lop(u[k], ua, ub);
        }
    }
    unsigned long long s = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) s ^= x[k] ^ (unsigned long long)__float_as_uint(f[k]) ^ u[k];
    if (s == 0x1234567ull) out[0] = s;
}

template <int NF2, int NF, int NA>
static void run(const char* name, int sms, double ghz) {
    unsigned long long* d;
    cudaMalloc(&d, 64);
    const int grid = sms * 8, block = 256, iters = 1 << 16;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    mix<NF2, NF, NA><<<grid, block>>>(d, iters, 5);
    cudaEventRecord(e0);
    mix<NF2, NF, NA><<<grid, block>>>(d, iters, 5);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    // 16 warps per scheduler (8 CTAs x 8 warps / 4): cycles one scheduler spends per iteration of ONE warp
    const double cyc = ms * 1e-3 * ghz * 1e9 / (double(iters) * 16.0);
    printf("%-34s %7.3f ms  %6.2f scheduler-cycles per warp-iteration (%d FFMA2 + %d FFMA + %d LOP3)\n", name, ms, cyc, NF2, NF, NA);
    cudaFree(d);
}

int main() {
    cudaDeviceProp pr;
    cudaGetDeviceProperties(&pr, 0);
    int khz = 0;
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    const double ghz = khz * 1e-6;
    printf("%s, %d SMs, %.3f GHz\n", pr.name, pr.multiProcessorCount, ghz);
    run<8, 0, 0>("FFMA2 only", pr.multiProcessorCount, ghz);
    run<0, 8, 0>("FFMA (3 registers) only", pr.multiProcessorCount, ghz);
    run<0, 0, 8>("LOP3 only", pr.multiProcessorCount, ghz);
    run<8, 0, 8>("FFMA2 + LOP3 1:1", pr.multiProcessorCount, ghz);
    run<8, 0, 4>("FFMA2 + LOP3 2:1", pr.multiProcessorCount, ghz);
    run<0, 8, 8>("FFMA + LOP3 1:1", pr.multiProcessorCount, ghz);
    run<4, 4, 0>("FFMA2 + FFMA 1:1", pr.multiProcessorCount, ghz);
    run<4, 4, 4>("FFMA2 + FFMA + LOP3", pr.multiProcessorCount, ghz);
    return 0;
}
