// peaks_sm100.cuh -- FP32 FMA-pipe and MUFU (SFU) peak microbenchmarks.
//
// MEASURED_PEAKS.json holds HBM and bf16 tensor peaks only; the MC path is bound by the FP32 pipe
// (SAT) and, second, the SFU (sincos / log / sqrt / rcp), so bench.py measures both denominators on
// the box, at the clock the kernel actually runs at (SURVEY.md 8d "Peaks").
#ifndef MCSIM_PEAKS_CUH
#define MCSIM_PEAKS_CUH

// 8 independent FFMA chains per thread; iters x 8 x 2 flops per thread
__global__ void __launch_bounds__(256) fma_peak_kernel(float* out, int iters, float a, float b) {
    float x0 = threadIdx.x, x1 = x0 + 1.f, x2 = x0 + 2.f, x3 = x0 + 3.f;
    float x4 = x0 + 4.f, x5 = x0 + 5.f, x6 = x0 + 6.f, x7 = x0 + 7.f;
#pragma unroll 8
    for (int i = 0; i < iters; ++i) {
        x0 = fmaf(x0, a, b); x1 = fmaf(x1, a, b); x2 = fmaf(x2, a, b); x3 = fmaf(x3, a, b);
        x4 = fmaf(x4, a, b); x5 = fmaf(x5, a, b); x6 = fmaf(x6, a, b); x7 = fmaf(x7, a, b);
    }
    const float s = ((x0 + x1) + (x2 + x3)) + ((x4 + x5) + (x6 + x7));
    if (s == 12345.678f) out[0] = s;  // never true in practice; keeps the chains alive
}

__device__ __forceinline__ float mufu_ex2_chain(float x) {   // volatile: the peak kernel's chain must not be folded
    float y;
    asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// 8 independent MUFU chains per thread (MUFU.EX2); iters x 8 SFU ops per thread
__global__ void __launch_bounds__(256) mufu_peak_kernel(float* out, int iters) {
    float x0 = threadIdx.x * 1e-3f, x1 = x0 + .1f, x2 = x0 + .2f, x3 = x0 + .3f;
    float x4 = x0 + .4f, x5 = x0 + .5f, x6 = x0 + .6f, x7 = x0 + .7f;
#pragma unroll 8
    for (int i = 0; i < iters; ++i) {
        x0 = mufu_ex2_chain(x0); x1 = mufu_ex2_chain(x1); x2 = mufu_ex2_chain(x2); x3 = mufu_ex2_chain(x3);
        x4 = mufu_ex2_chain(x4); x5 = mufu_ex2_chain(x5); x6 = mufu_ex2_chain(x6); x7 = mufu_ex2_chain(x7);
    }
    const float s = ((x0 + x1) + (x2 + x3)) + ((x4 + x5) + (x6 + x7));
    if (s == 12345.678f) out[0] = s;
}

// packed FP32 pairs (fma.rn.f32x2 -> SASS FFMA2): 8 independent chains of 2 lanes; iters x 8 x 4 flops per thread
__device__ __forceinline__ unsigned long long ffma2(unsigned long long x, unsigned long long a, unsigned long long b) {
    asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(x) : "l"(a), "l"(b));
    return x;
}
__global__ void __launch_bounds__(256) fma2_peak_kernel(float* out, int iters, float a, float b) {
    float2 av = make_float2(a, a), bv = make_float2(b, b), v = make_float2(threadIdx.x, threadIdx.x + 0.5f);
    const unsigned long long aa = *reinterpret_cast<unsigned long long*>(&av), bb = *reinterpret_cast<unsigned long long*>(&bv);
    unsigned long long x0 = *reinterpret_cast<unsigned long long*>(&v), x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3;
    unsigned long long x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
#pragma unroll 8
    for (int i = 0; i < iters; ++i) {
        x0 = ffma2(x0, aa, bb); x1 = ffma2(x1, aa, bb); x2 = ffma2(x2, aa, bb); x3 = ffma2(x3, aa, bb);
        x4 = ffma2(x4, aa, bb); x5 = ffma2(x5, aa, bb); x6 = ffma2(x6, aa, bb); x7 = ffma2(x7, aa, bb);
    }
    const unsigned long long s = x0 ^ x1 ^ x2 ^ x3 ^ x4 ^ x5 ^ x6 ^ x7;
    if (s == 12345ull) out[0] = float(s);
}

#endif
