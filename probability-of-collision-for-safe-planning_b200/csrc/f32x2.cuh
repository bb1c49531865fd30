// f32x2.cuh -- packed FP32 pairs for sm_100a (PTX fma/mul/add.rn.f32x2 -> SASS FFMA2 / FMUL2 / FADD2).
//
// One FFMA2 does two IEEE FP32 FMAs for one issue slot; the FMA pipe delivers the same flop rate
// as scalar FFMA (measured: csrc/peaks_sm100.cuh), so packing two rollouts into the two lanes of a
// thread halves the issue-slot cost of all FP32 arithmetic.  ptxas folds |x| and -x of a packed
// register and the broadcast of a scalar (register, uniform register or constant) into the
// instruction's operand modifiers (checked with cuobjdump: "FFMA2 R, |R|.F32x2.HI_LO, UR.F32, -R"),
// so abs2/neg2/bc below cost no instruction.
//
// MUFU, compares, selects and integer work stay per lane.
#ifndef MCSIM_F32X2_CUH
#define MCSIM_F32X2_CUH

struct f2 {
    unsigned long long v;
};

__device__ __forceinline__ f2 mk2(float lo, float hi) {
    f2 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r.v) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ f2 bc(float s) { return mk2(s, s); }
__device__ __forceinline__ void un2(f2 x, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(x.v)); }
__device__ __forceinline__ float lo(f2 x) { float a, b; un2(x, a, b); return a; }
__device__ __forceinline__ float hi(f2 x) { float a, b; un2(x, a, b); return b; }

__device__ __forceinline__ f2 fma2(f2 a, f2 b, f2 c) {
    f2 r;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r.v) : "l"(a.v), "l"(b.v), "l"(c.v));
    return r;
}
__device__ __forceinline__ f2 mul2(f2 a, f2 b) {
    f2 r;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r.v) : "l"(a.v), "l"(b.v));
    return r;
}
__device__ __forceinline__ f2 add2(f2 a, f2 b) {
    f2 r;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r.v) : "l"(a.v), "l"(b.v));
    return r;
}
__device__ __forceinline__ f2 neg2(f2 a) { float x, y; un2(a, x, y); return mk2(-x, -y); }
__device__ __forceinline__ f2 abs2(f2 a) { float x, y; un2(a, x, y); return mk2(fabsf(x), fabsf(y)); }
__device__ __forceinline__ f2 sub2(f2 a, f2 b) { return add2(a, neg2(b)); }

// scalar-broadcast forms
__device__ __forceinline__ f2 fma2(f2 a, float b, f2 c) { return fma2(a, bc(b), c); }
__device__ __forceinline__ f2 fma2(f2 a, f2 b, float c) { return fma2(a, b, bc(c)); }
__device__ __forceinline__ f2 fma2(f2 a, float b, float c) { return fma2(a, bc(b), bc(c)); }
__device__ __forceinline__ f2 mul2(f2 a, float b) { return mul2(a, bc(b)); }
__device__ __forceinline__ f2 add2(f2 a, float b) { return add2(a, bc(b)); }

// ---- per-lane special functions (MUFU) ----------------------------------------------------------
__device__ __forceinline__ float mufu_rcp(float x) { float y; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float mufu_rsqrt(float x) { float y; asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float mufu_sqrt(float x) { float y; asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float mufu_ex2(float x) { float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float mufu_lg2(float x) { float y; asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float mufu_sin(float x) { float y; asm("sin.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float mufu_cos(float x) { float y; asm("cos.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }

__device__ __forceinline__ f2 rcp2(f2 a) { float x, y; un2(a, x, y); return mk2(mufu_rcp(x), mufu_rcp(y)); }
__device__ __forceinline__ f2 rsqrt2(f2 a) { float x, y; un2(a, x, y); return mk2(mufu_rsqrt(x), mufu_rsqrt(y)); }
__device__ __forceinline__ f2 sqrt2(f2 a) { float x, y; un2(a, x, y); return mk2(mufu_sqrt(x), mufu_sqrt(y)); }
// sin/cos of a bounded angle (|a| <= ~4 pi on this path): sin.approx / cos.approx = one FMUL by
// 1/(2 pi) + MUFU.SIN / MUFU.COS; absolute error ~1e-6, far below the Monte-Carlo resolution.
__device__ __forceinline__ void sincos2(f2 a, f2& s, f2& c) {
    float x, y;
    un2(a, x, y);
    s = mk2(mufu_sin(x), mufu_sin(y));
    c = mk2(mufu_cos(x), mufu_cos(y));
}

// a mod 2 pi into [0, 2 pi] (angleWrap, MCSimulator.h:56-65; both ends kept) without branches:
// k = rint(a / 2 pi - 1/2) by the 1.5 * 2^23 trick, a - 2 pi k in two FMAs (hi + lo of 2 pi).
__device__ __forceinline__ f2 wrap2(f2 a) {  // @phase wrap
    const float M = 12582912.0f;
    const f2 t = fma2(a, 0.15915494309189535f, -0.5f);
    const f2 k = add2(add2(t, M), -M);
    return fma2(k, 1.7484555e-7f, fma2(k, -6.2831854820251465f, a));
}

// atan2(y, x) per lane pair: octant reduction + odd minimax polynomial of degree 17
// (Abramowitz & Stegun 4.4.49, |error| <= 1e-7 in FP32), result in [-pi, pi].
__device__ __forceinline__ f2 atan2_2(f2 y, f2 x) {  // @phase atan2
    float x0, x1, y0, y1;
    un2(x, x0, x1);
    un2(y, y0, y1);
    const float ax0 = fabsf(x0), ay0 = fabsf(y0), ax1 = fabsf(x1), ay1 = fabsf(y1);
    const float mx0 = fmaxf(ax0, ay0), mn0 = fminf(ax0, ay0), mx1 = fmaxf(ax1, ay1), mn1 = fminf(ax1, ay1);
    const f2 r = mul2(mk2(mn0, mn1), mk2(mufu_rcp(fmaxf(mx0, 1e-30f)), mufu_rcp(fmaxf(mx1, 1e-30f))));
    const f2 s = mul2(r, r);
    f2 p = fma2(s, 0.0028662257f, -0.0161657367f);
    p = fma2(p, s, 0.0429096138f);
    p = fma2(p, s, -0.0752896400f);
    p = fma2(p, s, 0.1065626393f);
    p = fma2(p, s, -0.1420889944f);
    p = fma2(p, s, 0.1999355085f);
    p = fma2(p, s, -0.3333314528f);
    p = fma2(mul2(p, s), r, r);
    float a0, a1;
    un2(p, a0, a1);
    a0 = ay0 > ax0 ? 1.5707963267948966f - a0 : a0;
    a1 = ay1 > ax1 ? 1.5707963267948966f - a1 : a1;
    a0 = x0 < 0.f ? 3.141592653589793f - a0 : a0;
    a1 = x1 < 0.f ? 3.141592653589793f - a1 : a1;
    return mk2(copysignf(a0, y0), copysignf(a1, y1));
}

#endif
