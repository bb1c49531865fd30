// em_sm100.cuh -- EM fit of K full-covariance 3-D Gaussian components (BASELINE.json config 5).
//
// Replaces the vendored Armadillo 8.400.0 EM the reference ships (SURVEY.md A13;
// armadillo_bits/gmm_full_meat.hpp): em_generate_acc :2646-2734 becomes em_estep_kernel (responsibilities
// are never materialised: log-likelihoods, log-sum-exp and the weighted sums stay in registers),
// em_update_params :2542-2639 + em_fix_params :2741-2800 + the convergence test of em_iterate
// :2447-2465 + init_constants :980-1046 become em_mstep_kernel (one CTA per fit, FP64).
//
// Samples: FP32, component-major per fit (x[n], y[n], theta[n]) -- 12 B per sample and iteration
// from HBM.  Sufficient statistics are accumulated about a reference point near the data (d = x - ref):
// sum r, sum r d, sum r d d^T; new mean = ref + sum r d / sum r, new cov = sum r d d^T / sum r - delta delta^T.
// Algebraically this is the reference's sum r x x^T / sum r - mean mean^T (:2591-2600) without its cancellation,
// which FP32 partial sums could not afford (|x|^2 ~ 10, cov ~ 1e-3).  ref is ONE point per fit while the fit's
// components lie close together (the usual case: 11 instead of 16 instructions per sample and component, no
// per-component d to keep in registers), else each component's own mean -- see "reference point" below.
#ifndef MCSIM_EM_CUH
#define MCSIM_EM_CUH

#include <float.h>

#include "f32x2.cuh"
#include "mcsim_cuda.h"

#define EM_COMP_STRIDE 12  // floats per component: mu[3], b00 b01 b02 b11 b12 b22, cst, pad[2]
#define EM_NSTAT 10        // sum r, sum r d[3], sum r dd^T [xx xy xz yy yz zz]

struct EmLaunch {
    const float* x;        // [n_fits][3][n]
    unsigned long long n;  // samples per fit in this shard
    double n_total;        // samples per fit over all shards
    int K, n_fits, iters;
    int iter;              // index of the iteration being run
    double var_floor;
    float* comps;          // [n_fits][K][EM_COMP_STRIDE] E-step constants (FP32)
    double* hefts;         // [n_fits][K]
    double* means;         // [n_fits][K][3]
    double* fcovs;         // [n_fits][K][9]
    double* stats;         // [n_fits][K * 10 + 1]
    double* old_avg;       // [n_fits]
    int* done;             // [n_fits] 0 running, 1 converged, 2 non-finite likelihood (reference: learn() fails)
    int* iters_run;        // [n_fits]
    double* trace;         // [n_fits][iters] avg_log_p before each update
    // ---- sharded fit with the exchange fused into the M-step kernel (mcsim_group.cuh); n_peers <= 1: stats are the totals
    int n_peers, my_rank;                      // devices of the group, this device's position
    double* peer_slots[8];                     // device q's gather buffer of this iteration (peer-mapped memory, q == my_rank
                                               // included): [n_peers slots][n_fits][K * 10 + 1]; this device writes slot [my_rank]
    unsigned long long* peer_flags[8];         // device q's arrival counters [n_fits][n_peers]; this device writes column [my_rank]
    const double* my_slots;                    // = peer_slots[my_rank], my_flags = peer_flags[my_rank] (kept apart: a run-time
    const unsigned long long* my_flags;        //   index into a kernel-parameter array would put the array on the stack)
    unsigned long long wait_value;             // value a counter holds once the block of this iteration has arrived
    int moments;                               // EM_MOMENTS_AUTO / _CENTRED / _SHIFTED: reference point of the statistics (see em_header)
};
#define EM_MAX_PEERS 8   // = the extents of peer_slots / peer_flags above
#define EM_MSTEP_FUSED_THREADS 192   // block size of the M-step kernel when it carries the exchange (>= 10 * 16 + 1); 32 otherwise
#define EM_DONE_EXCHANGE_TIMEOUT 3     // done[] value: a peer's statistics did not arrive within EM_PEER_TIMEOUT_CYCLES
#define EM_PEER_TIMEOUT_CYCLES (40ll * 1000 * 1000 * 1000)   // ~20 s: a stuck peer ends the fit with an error instead of hanging the GPU

// ---- reference point of the sufficient statistics ---------------------------------------------------------------------
// The statistics of component g are sums of r, r d and r d d^T with d = x - (a reference point); the M-step turns them
// into the new mean (reference + sum r d / sum r) and covariance (sum r d d^T / sum r - the outer product of that step)
// in FP64 (em_update_params, :2583-2638, computes the same quantities about zero in FP64).  Two choices:
//  * CENTRED: the component's own current mean.  d is small (a few sigma), FP32 products lose nothing -- but d and its
//    products are per component: 3 + 3 + 10 instructions per sample and component for the accumulation alone.
//  * SHIFTED: ONE point c per fit (the heft-weighted mean of the current means, rounded to FP32).  e = x - c and its six
//    products are computed once per sample and every component accumulates r * (1, e, e e^T): 1 + 10 instructions.  The
//    price is cancellation in the M-step's covariance, (|mean - c|^2 + sigma^2) / sigma^2 times the FP32 rounding of the
//    per-thread partial sums, so the form is taken only while T = 1/2 ||Sigma^-1||_inf (|mean_g - c| + 6 sigma_g)^2 stays
//    below EM_SHIFT_LIMIT for every component of the fit (propagated beliefs: T = 100 .. 1,500 -- components of one
//    mixture lie within a few sigma of each other); otherwise the fit uses the centred form for that iteration.
//    What it buys depends on the kernel: the scalar kernels for K >= 6 no longer keep (or recompute) 3 K differences
//    and stop spilling (+7 .. +20 %); the packed-pair kernels for K <= 5 have the registers anyway (+-2 %) and stay
//    centred under AUTO (EM_SHIFT_MIN_K).
// The log-likelihoods are computed about the component means in both forms.  Header of a fit, in the spare floats of
// its first two records: rec0[10], rec0[11], rec1[10] = c; rec1[11] = 1.f if SHIFTED (K = 1 is always centred: its
// responsibilities are 1 and it has no second record).
#ifndef EM_SHIFT_LIMIT
#define EM_SHIFT_LIMIT 2048.0
#endif
#ifndef EM_SHIFT_MIN_K
#define EM_SHIFT_MIN_K 6    // AUTO takes the shifted form from this many components on (the scalar kernels)
#endif
#define EM_MOMENTS_AUTO 0
#define EM_MOMENTS_CENTRED 1
#define EM_MOMENTS_SHIFTED 2
struct EmHeader { float cx, cy, ct; bool shifted; };
template <bool V> struct EmFormTag { static constexpr bool value = V; };
template <int K>
__device__ __forceinline__ EmHeader em_header(const float* s_c) {
    EmHeader h;
    if (K >= 2) { h.cx = s_c[10]; h.cy = s_c[11]; h.ct = s_c[EM_COMP_STRIDE + 10]; h.shifted = s_c[EM_COMP_STRIDE + 11] != 0.f; }
    else { h.cx = h.cy = h.ct = 0.f; h.shifted = false; }
    return h;
}

// ---- E-step ------------------------------------------------------------------------------------------
// K: components of the model; G0: first component whose statistics THIS launch accumulates (at most 8
// per launch: 80 FP32 accumulators per thread).  K > 8 takes a second launch with G0 = 8 that
// recomputes the log-likelihoods; the log-likelihood sum is accumulated by the G0 = 0 launch only.
// The one-warp kernel for EM_OW_LO <= K <= 8 carries 10 K accumulators per thread.  Left alone it takes 183 - 227
// registers (one 256-thread CTA per SM, 2 warps per scheduler, 56 % issue active); capped at 128 registers with one
// sample in flight and 128-thread CTAs it runs 4 warps per scheduler: K = 6 / 7 / 8 0.340 / 0.361 / 0.375 -> 0.445 /
// 0.452 / 0.454 of the FP32 peak with 0.1 - 0.5 KB of spills per thread (profiles/r02_em_k68_occupancy.log), and
// 0.507 / 0.555 / 0.536 without spills once the statistics took the shifted form and the exponentials lost their
// denormal handling (346 -> 278 instructions per sample at K = 8; profiles/r02_em_shifted_moments.log).
#ifndef EM_SPLIT_U1_KG
#define EM_SPLIT_U1_KG 6   // the split kernel runs one sample per thread between barriers when a member owns more components than this
#endif
#ifndef EM_OW_LO
#define EM_OW_LO 6
#endif
#ifndef EM_OW_HI
#define EM_OW_HI 9       // most components the one-warp kernel takes in one launch (10 accumulators each): K = 9 spills 0.25 KB per thread at
                         // 128 registers and still beats the group of two (0.559 vs 0.481 of the FP32 peak on config 5); K = 10 ties, K >= 11 loses
#endif
#ifndef EM_OW_U68
#define EM_OW_U68 1      // samples in flight per thread for EM_OW_LO <= K <= 8
#endif
#ifndef EM_OW_THREADS68
#define EM_OW_THREADS68 128   // CTA size for EM_OW_LO <= K <= 8 (the host launches it: em_pick in mcsim_cuda.cu)
#endif
#ifndef EM_OW_MINB68
#define EM_OW_MINB68 4   // resident CTAs of that size per SM the register budget is set for: 128 registers, 16 warps per SM
#endif
// The sample loop of em_estep_kernel for one form of the statistics.  The two forms are separate instantiations (chosen per
// fit at the top of the kernel): inside one loop ptxas hoists the centred form's d = x - mu_g into the shared part and
// the shifted form pays for it (measured: 346 against 356 instructions per sample at K = 8 instead of 275).
template <int K, int G0, int KG, int U, bool SHIFTED>
__device__ __forceinline__ void em_ow_loop(const float* s_c, const EmHeader hd, const float* __restrict__ xs, const float* __restrict__ ys,
                                           const float* __restrict__ ts, unsigned n, float (&acc)[KG][EM_NSTAT], float& lsum) {
    const unsigned stride = gridDim.x * blockDim.x;
    unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
    float bx[U], by[U], bt[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
        const unsigned k = i + u * stride;
        const bool ok = k < n;
        bx[u] = ok ? __ldg(xs + k) : 0.f; by[u] = ok ? __ldg(ys + k) : 0.f; bt[u] = ok ? __ldg(ts + k) : 0.f;
    }
    for (; i < n; i += U * stride) {
        float cx[U], cy[U], ct[U];
#pragma unroll
        for (int u = 0; u < U; ++u) { cx[u] = bx[u]; cy[u] = by[u]; ct[u] = bt[u]; }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const unsigned k = i + (U + u) * stride;
            const bool ok = k < n;
            bx[u] = ok ? __ldg(xs + k) : 0.f; by[u] = ok ? __ldg(ys + k) : 0.f; bt[u] = ok ? __ldg(ts + k) : 0.f;
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            if (i + u * stride >= n) break;
            const float x = cx[u], y = cy[u], t = ct[u];
            float ll[K];
            float m = -FLT_MAX;
#pragma unroll
            for (int g = 0; g < K; ++g) {
                const float4* c4 = reinterpret_cast<const float4*>(s_c + g * EM_COMP_STRIDE);  // 3 LDS.128
                const float4 ca = c4[0], cb = c4[1], cc = c4[2];   // mu0 mu1 mu2 b00 | b01 b02 b11 b12 | b22 cst - -
                const float d0 = x - ca.x, d1 = y - ca.y, d2 = t - ca.z;
                // -1/2 d^T A d + cst with b = -1/2 A (off-diagonals doubled)
                const float q0 = fmaf(ca.w, d0, fmaf(cb.x, d1, cb.y * d2));
                const float q1 = fmaf(cb.z, d1, cb.w * d2);
                ll[g] = fmaf(d0, q0, fmaf(d1, q1, fmaf(cc.x * d2, d2, cc.y)));
                m = fmaxf(m, ll[g]);
            }
            // exp(ll - m) as one FFMA + MUFU.EX2 per component (exp2f adds a compare and two scalings for denormal results:
            // a responsibility below 2^-126 is zero here), 1 / s as MUFU.RCP (1 ulp; __frcp_rn is a 9-instruction sequence)
            float s = 0.f;
            const float ms = -m * 1.4426950408889634f;
#pragma unroll
            for (int g = 0; g < K; ++g) {
                ll[g] = mufu_ex2(fmaf(ll[g], 1.4426950408889634f, ms));
                s += ll[g];
            }
            if (G0 == 0) lsum += fmaf(mufu_lg2(s), 0.6931471805599453f, m);
            const float inv = mufu_rcp(s);
            if (SHIFTED) {
                // statistics about the fit's common point: e and its products once per sample, 1 + 10 instructions per component
                const float e0 = x - hd.cx, e1 = y - hd.cy, e2 = t - hd.ct;
                const float p00 = e0 * e0, p01 = e0 * e1, p02 = e0 * e2, p11 = e1 * e1, p12 = e1 * e2, p22 = e2 * e2;
#pragma unroll
                for (int gg = 0; gg < KG; ++gg) {
                    const float r = ll[G0 + gg] * inv;
                    acc[gg][0] += r;
                    acc[gg][1] = fmaf(r, e0, acc[gg][1]);
                    acc[gg][2] = fmaf(r, e1, acc[gg][2]);
                    acc[gg][3] = fmaf(r, e2, acc[gg][3]);
                    acc[gg][4] = fmaf(r, p00, acc[gg][4]);
                    acc[gg][5] = fmaf(r, p01, acc[gg][5]);
                    acc[gg][6] = fmaf(r, p02, acc[gg][6]);
                    acc[gg][7] = fmaf(r, p11, acc[gg][7]);
                    acc[gg][8] = fmaf(r, p12, acc[gg][8]);
                    acc[gg][9] = fmaf(r, p22, acc[gg][9]);
                }
            } else {
#pragma unroll
                for (int gg = 0; gg < KG; ++gg) {
                    const float4 ca = *reinterpret_cast<const float4*>(s_c + (G0 + gg) * EM_COMP_STRIDE);
                    const float r = ll[G0 + gg] * inv;
                    const float d0 = x - ca.x, d1 = y - ca.y, d2 = t - ca.z;
                    const float r0 = r * d0, r1 = r * d1, r2 = r * d2;
                    acc[gg][0] += r;
                    acc[gg][1] += r0;
                    acc[gg][2] += r1;
                    acc[gg][3] += r2;
                    acc[gg][4] = fmaf(r0, d0, acc[gg][4]);
                    acc[gg][5] = fmaf(r0, d1, acc[gg][5]);
                    acc[gg][6] = fmaf(r0, d2, acc[gg][6]);
                    acc[gg][7] = fmaf(r1, d1, acc[gg][7]);
                    acc[gg][8] = fmaf(r1, d2, acc[gg][8]);
                    acc[gg][9] = fmaf(r2, d2, acc[gg][9]);
                }
            }
        }
    }
}

template <int K, int G0>
__global__ void __launch_bounds__((K >= EM_OW_LO && K <= EM_OW_HI && G0 == 0) ? EM_OW_THREADS68 : 256, (K >= EM_OW_LO && K <= EM_OW_HI && G0 == 0) ? EM_OW_MINB68 : 1) em_estep_kernel(const EmLaunch a) {
    constexpr int KG = (K - G0) < EM_OW_HI ? (K - G0) : EM_OW_HI;
    constexpr int NACC = KG * EM_NSTAT + 1;
    __shared__ __align__(16) float s_c[K * EM_COMP_STRIDE];
    __shared__ double s_acc[NACC];
    const int fit = blockIdx.y;
    if (a.done[fit]) return;
    for (int i = threadIdx.x; i < K * EM_COMP_STRIDE; i += blockDim.x) s_c[i] = a.comps[(size_t)fit * K * EM_COMP_STRIDE + i];
    for (int i = threadIdx.x; i < NACC; i += blockDim.x) s_acc[i] = 0.0;
    __syncthreads();

    const float* xs = a.x + (size_t)fit * 3 * a.n;
    const float* ys = xs + a.n;
    const float* ts = ys + a.n;
    const EmHeader hd = em_header<K>(s_c);
    float acc[KG][EM_NSTAT];
#pragma unroll
    for (int g = 0; g < KG; ++g)
#pragma unroll
        for (int k = 0; k < EM_NSTAT; ++k) acc[g][k] = 0.f;
    float lsum = 0.f;

    // U samples per pass, the next pass's 3 U loads issued before this pass's arithmetic: enough bytes
    // in flight per SM to stream from HBM while the FP32 pipe works on the previous samples
    constexpr int U = K <= 4 ? 4 : (K <= EM_OW_HI ? (K >= EM_OW_LO ? EM_OW_U68 : 2) : 1);
    // 32-bit sample indices (mcsim_cuda_em_prepare rejects n >= 2^31 per fit and shard)
    const unsigned n = unsigned(a.n);
    if (K >= 2 && hd.shifted) em_ow_loop<K, G0, KG, U, true>(s_c, hd, xs, ys, ts, n, acc, lsum);
    else em_ow_loop<K, G0, KG, U, false>(s_c, hd, xs, ys, ts, n, acc, lsum);

    // warp shuffle reduction (FP32 within the warp's <= 32 partial sums is widened to FP64 first)
#pragma unroll
    for (int g = 0; g < KG; ++g)
#pragma unroll
        for (int k = 0; k < EM_NSTAT; ++k) {
            double v = acc[g][k];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
            if ((threadIdx.x & 31) == 0) atomicAdd(&s_acc[g * EM_NSTAT + k], v);
        }
    if (G0 == 0) {
        double v = lsum;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if ((threadIdx.x & 31) == 0) atomicAdd(&s_acc[NACC - 1], v);
    }
    __syncthreads();
    double* st = a.stats + (size_t)fit * (K * EM_NSTAT + 1);
    for (int i = threadIdx.x; i < KG * EM_NSTAT; i += blockDim.x) atomicAdd(&st[G0 * EM_NSTAT + i], s_acc[i]);
    if (G0 == 0 && threadIdx.x == 0) atomicAdd(&st[K * EM_NSTAT], s_acc[NACC - 1]);
}

// The round loop of em_estep_split_kernel for one form of the statistics (separate instantiations, as em_ow_loop).
template <int K, int S, int KG, int NG, int U, bool SHIFTED>
__device__ __forceinline__ void em_split_loop(const EmLaunch& a, const float* s_c, float2 (&s_ex)[2][8][U][32], const EmHeader hd, const float* __restrict__ xs,
                                              const float* __restrict__ ys, const float* __restrict__ ts, unsigned lane, unsigned warp, unsigned grp, unsigned w,
                                              float (&acc)[KG][EM_NSTAT], float& lsum) {
    constexpr float LOG2E = 1.4426950408889634f;
    const unsigned n = unsigned(a.n);
    const unsigned stride = gridDim.x * NG * 32u;                 // samples per round of the whole grid, per u
    unsigned i = (blockIdx.x * NG + grp) * 32u + lane;            // sample u of this round: i + u * stride
    float bx[U], by[U], bt[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
        const unsigned k = i + u * stride;
        const bool ok = k < n;
        bx[u] = ok ? __ldg(xs + k) : 0.f; by[u] = ok ? __ldg(ys + k) : 0.f; bt[u] = ok ? __ldg(ts + k) : 0.f;
    }
    unsigned par = 0;
    // group-uniform trip count: the first lane of the group decides (all members run the same rounds and barriers)
    for (unsigned base = (blockIdx.x * NG + grp) * 32u; base < n; base += U * stride, i += U * stride, par ^= 1u) {
        float cx[U], cy[U], ct[U];
#pragma unroll
        for (int u = 0; u < U; ++u) { cx[u] = bx[u]; cy[u] = by[u]; ct[u] = bt[u]; }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const unsigned k = i + (U + u) * stride;
            const bool ok = k < n;
            bx[u] = ok ? __ldg(xs + k) : 0.f; by[u] = ok ? __ldg(ys + k) : 0.f; bt[u] = ok ? __ldg(ts + k) : 0.f;
        }
        float e[U][KG], m[U], sm[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            m[u] = -FLT_MAX;
#pragma unroll
            for (int gg = 0; gg < KG; ++gg) {
                const int g = int(w) + gg * S;
                if (g < K) {
                    const float4* c4 = reinterpret_cast<const float4*>(s_c + g * EM_COMP_STRIDE);
                    const float4 ca = c4[0], cb = c4[1], cc = c4[2];
                    const float d0 = cx[u] - ca.x, d1 = cy[u] - ca.y, d2 = ct[u] - ca.z;
                    const float q0 = fmaf(ca.w, d0, fmaf(cb.x, d1, cb.y * d2));
                    const float q1 = fmaf(cb.z, d1, cb.w * d2);
                    e[u][gg] = fmaf(d0, q0, fmaf(d1, q1, fmaf(cc.x * d2, d2, cc.y)));
                    m[u] = fmaxf(m[u], e[u][gg]);
                } else {
                    e[u][gg] = -FLT_MAX;
                }
            }
            sm[u] = 0.f;
            const float msu = -m[u] * LOG2E;
#pragma unroll
            for (int gg = 0; gg < KG; ++gg) {
                e[u][gg] = mufu_ex2(fmaf(e[u][gg], LOG2E, msu));   // a member without component gg: exp2(-huge) = 0
                sm[u] += e[u][gg];
            }
            s_ex[par][warp][u][lane] = make_float2(m[u], sm[u]);
        }
        asm volatile("bar.sync %0, %1;" ::"r"(1 + grp), "r"(S * 32) : "memory");
#pragma unroll
        for (int u = 0; u < U; ++u) {
            float om[S], os[S];
            float M = m[u];
#pragma unroll
            for (int q = 0; q < S; ++q) {
                const float2 v = s_ex[par][grp * S + q][u][lane];
                om[q] = v.x; os[q] = v.y;
                M = fmaxf(M, v.x);
            }
            float tot = 0.f;
#pragma unroll
            for (int q = 0; q < S; ++q) tot = fmaf(os[q], mufu_ex2((om[q] - M) * LOG2E), tot);
            const bool ok = i + u * stride < n;
            const float scale = ok ? mufu_ex2((m[u] - M) * LOG2E) * mufu_rcp(tot) : 0.f;
            if (w == 0 && ok) lsum += fmaf(mufu_lg2(tot), 0.6931471805599453f, M);
            if (SHIFTED) {
                const float e0 = cx[u] - hd.cx, e1 = cy[u] - hd.cy, e2 = ct[u] - hd.ct;
                const float p00 = e0 * e0, p01 = e0 * e1, p02 = e0 * e2, p11 = e1 * e1, p12 = e1 * e2, p22 = e2 * e2;
#pragma unroll
                for (int gg = 0; gg < KG; ++gg) {
                    if (int(w) + gg * S < K) {
                        const float r = e[u][gg] * scale;
                        acc[gg][0] += r;
                        acc[gg][1] = fmaf(r, e0, acc[gg][1]);
                        acc[gg][2] = fmaf(r, e1, acc[gg][2]);
                        acc[gg][3] = fmaf(r, e2, acc[gg][3]);
                        acc[gg][4] = fmaf(r, p00, acc[gg][4]);
                        acc[gg][5] = fmaf(r, p01, acc[gg][5]);
                        acc[gg][6] = fmaf(r, p02, acc[gg][6]);
                        acc[gg][7] = fmaf(r, p11, acc[gg][7]);
                        acc[gg][8] = fmaf(r, p12, acc[gg][8]);
                        acc[gg][9] = fmaf(r, p22, acc[gg][9]);
                    }
                }
            } else {
#pragma unroll
            for (int gg = 0; gg < KG; ++gg) {
                const int g = int(w) + gg * S;
                if (g < K) {
                    const float4 ca = *reinterpret_cast<const float4*>(s_c + g * EM_COMP_STRIDE);
                    const float r = e[u][gg] * scale;
                    const float d0 = cx[u] - ca.x, d1 = cy[u] - ca.y, d2 = ct[u] - ca.z;
                    const float r0 = r * d0, r1 = r * d1, r2 = r * d2;
                    acc[gg][0] += r;
                    acc[gg][1] += r0;
                    acc[gg][2] += r1;
                    acc[gg][3] += r2;
                    acc[gg][4] = fmaf(r0, d0, acc[gg][4]);
                    acc[gg][5] = fmaf(r0, d1, acc[gg][5]);
                    acc[gg][6] = fmaf(r0, d2, acc[gg][6]);
                    acc[gg][7] = fmaf(r1, d1, acc[gg][7]);
                    acc[gg][8] = fmaf(r1, d2, acc[gg][8]);
                    acc[gg][9] = fmaf(r2, d2, acc[gg][9]);
                }
            }
            }
        }
    }
}

// ---- E-step, components split over the warps of a group (K >= 5) ----------------------------------------------------
// One thread of em_estep_kernel carries 10 accumulators per component: 80 for K = 8 (227 registers, two warps per
// scheduler), and K = 16 needed a second launch that recomputed all 16 log-likelihoods.  Here S consecutive warps form a
// group that walks the SAME samples (lane l of every member holds sample l of the group's tile); member w owns the
// components g = w, w + S, w + 2 S, ... -- their quadratic forms, exponentials and 10 accumulators each.  What the
// members owe each other is the normaliser of the responsibilities: each writes its partial (max, sum of exponentials)
// per sample to shared memory, a named barrier over the S warps (bar.sync id, 32 S) makes them visible, and everyone
// combines them to the log-sum-exp of all K components (gmm_full_meat.hpp:2707-2723).  Per sample and member that is one
// STS.64, S - 1 LDS.64, S + 1 MUFU and a barrier shared by U samples; the accumulators per thread drop to 10 ceil(K / S),
// one launch covers any K <= 16, and twice to four times the warps are resident.  Member 0 also sums the log-likelihood.
template <int K, int S>
__global__ void __launch_bounds__(256, (K + S - 1) / S <= 2 ? 3 : 2) em_estep_split_kernel(const EmLaunch a) {
    constexpr int KG = (K + S - 1) / S;   // components per member (strided: counts differ by at most one)
    constexpr int NG = 8 / S;             // groups per CTA
    constexpr int U = KG > EM_SPLIT_U1_KG ? 1 : 2;     // samples per thread between two barriers
    constexpr int NACC = K * EM_NSTAT + 1;
    __shared__ __align__(16) float s_c[K * EM_COMP_STRIDE];
    __shared__ double s_acc[NACC];
    __shared__ float2 s_ex[2][8][U][32];  // [round parity][warp][sample of the round][lane]: partial (max, sum)
    const int fit = blockIdx.y;
    if (a.done[fit]) return;
    for (int i = threadIdx.x; i < K * EM_COMP_STRIDE; i += blockDim.x) s_c[i] = a.comps[(size_t)fit * K * EM_COMP_STRIDE + i];
    for (int i = threadIdx.x; i < NACC; i += blockDim.x) s_acc[i] = 0.0;
    __syncthreads();
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5, grp = warp / S, w = warp % S;
    const EmHeader hd = em_header<K>(s_c);
    const float* xs = a.x + (size_t)fit * 3 * a.n;
    const float* ys = xs + a.n;
    const float* ts = ys + a.n;
    float acc[KG][EM_NSTAT];
#pragma unroll
    for (int g = 0; g < KG; ++g)
#pragma unroll
        for (int k = 0; k < EM_NSTAT; ++k) acc[g][k] = 0.f;
    float lsum = 0.f;
    if (hd.shifted) em_split_loop<K, S, KG, NG, U, true>(a, s_c, s_ex, hd, xs, ys, ts, lane, warp, grp, w, acc, lsum);
    else em_split_loop<K, S, KG, NG, U, false>(a, s_c, s_ex, hd, xs, ys, ts, lane, warp, grp, w, acc, lsum);
#pragma unroll
    for (int gg = 0; gg < KG; ++gg) {
        const int g = int(w) + gg * S;
        if (g < K) {
#pragma unroll
            for (int k = 0; k < EM_NSTAT; ++k) {
                double v = acc[gg][k];
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
                if (lane == 0) atomicAdd(&s_acc[g * EM_NSTAT + k], v);
            }
        }
    }
    if (w == 0) {
        double v = lsum;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (lane == 0) atomicAdd(&s_acc[NACC - 1], v);
    }
    __syncthreads();
    double* st = a.stats + (size_t)fit * NACC;
    for (int k = threadIdx.x; k < NACC; k += blockDim.x) atomicAdd(&st[k], s_acc[k]);
}

// ---- E-step, packed: two samples per thread in the lanes of FFMA2 / FMUL2 / FADD2 (K <= 4) ----------------
// The E-step is almost pure FP32 arithmetic (no compares / selects to speak of), which is where packed
// pairs pay on sm_100a (tools/pipe_probe.cu): half the FMA-pipe issue slots per sample.  Same arithmetic
// per lane as em_estep_kernel; MUFU (ex2, lg2, rcp) and the running maximum stay per lane.
// one pair of samples (the two lanes of x, y, t) against the K components in s_c; v1 = 0 masks lane 1
template <int K, bool SHIFTED>
__device__ __forceinline__ void em_pair2(const float* s_c, const EmHeader& hd, f2 x, f2 y, f2 t, float v1, f2 (&acc)[K][EM_NSTAT], f2& lsum) {
    if (K == 1) {
        // one component: the responsibility is 1 and the log-sum-exp is the log-likelihood itself; the general
        // path below computes the same up to the rounding of its ex2 / rcp (1e-7), with 6 MUFU per pair more
        const float4* c4 = reinterpret_cast<const float4*>(s_c);
        const float4 ca = c4[0], cb = c4[1], cc = c4[2];
        const f2 valid = mk2(1.f, v1);
        const f2 d0 = add2(x, -ca.x), d1 = add2(y, -ca.y), d2 = add2(t, -ca.z);
        const f2 q0 = fma2(d0, ca.w, fma2(d1, cb.x, mul2(d2, cb.y)));
        const f2 q1 = fma2(d1, cb.z, mul2(d2, cb.w));
        const f2 l1 = fma2(d0, q0, fma2(d1, q1, fma2(mul2(d2, cc.x), d2, cc.y)));
        lsum = fma2(l1, valid, lsum);
        const f2 r0 = mul2(valid, d0), r1 = mul2(valid, d1), r2 = mul2(valid, d2);
        acc[0][0] = add2(acc[0][0], valid);
        acc[0][1] = add2(acc[0][1], r0);
        acc[0][2] = add2(acc[0][2], r1);
        acc[0][3] = add2(acc[0][3], r2);
        acc[0][4] = fma2(r0, d0, acc[0][4]);
        acc[0][5] = fma2(r0, d1, acc[0][5]);
        acc[0][6] = fma2(r0, d2, acc[0][6]);
        acc[0][7] = fma2(r1, d1, acc[0][7]);
        acc[0][8] = fma2(r1, d2, acc[0][8]);
        acc[0][9] = fma2(r2, d2, acc[0][9]);
        return;
    }
    f2 ll[K];
    float m0 = -FLT_MAX, m1 = -FLT_MAX;
#pragma unroll
    for (int g = 0; g < K; ++g) {
        const float4* c4 = reinterpret_cast<const float4*>(s_c + g * EM_COMP_STRIDE);
        const float4 ca = c4[0], cb = c4[1], cc = c4[2];
        const f2 d0 = add2(x, -ca.x), d1 = add2(y, -ca.y), d2 = add2(t, -ca.z);
        const f2 q0 = fma2(d0, ca.w, fma2(d1, cb.x, mul2(d2, cb.y)));
        const f2 q1 = fma2(d1, cb.z, mul2(d2, cb.w));
        ll[g] = fma2(d0, q0, fma2(d1, q1, fma2(mul2(d2, cc.x), d2, cc.y)));
        m0 = fmaxf(m0, lo(ll[g]));
        m1 = fmaxf(m1, hi(ll[g]));
    }
    const f2 mm = mk2(m0, m1);
    const f2 ms = mul2(mm, -1.4426950408889634f);
    f2 s = bc(0.f);
#pragma unroll
    for (int g = 0; g < K; ++g) {
        const f2 e = fma2(ll[g], 1.4426950408889634f, ms);
        float e0, e1;
        un2(e, e0, e1);
        asm("ex2.approx.ftz.f32 %0, %0;" : "+f"(e0));
        asm("ex2.approx.ftz.f32 %0, %0;" : "+f"(e1));
        ll[g] = mk2(e0, e1);
        s = add2(s, ll[g]);
    }
    const f2 lg = mk2(mufu_lg2(lo(s)), mufu_lg2(hi(s)));
    const f2 valid = mk2(1.f, v1);
    lsum = fma2(fma2(lg, 0.6931471805599453f, mm), valid, lsum);   // (m + ln s) per valid lane
    const f2 inv = mul2(rcp2(s), valid);
    if (SHIFTED) {
        const f2 e0 = add2(x, -hd.cx), e1 = add2(y, -hd.cy), e2 = add2(t, -hd.ct);
        const f2 p00 = mul2(e0, e0), p01 = mul2(e0, e1), p02 = mul2(e0, e2), p11 = mul2(e1, e1), p12 = mul2(e1, e2), p22 = mul2(e2, e2);
#pragma unroll
        for (int g = 0; g < K; ++g) {
            const f2 r = mul2(ll[g], inv);
            acc[g][0] = add2(acc[g][0], r);
            acc[g][1] = fma2(r, e0, acc[g][1]);
            acc[g][2] = fma2(r, e1, acc[g][2]);
            acc[g][3] = fma2(r, e2, acc[g][3]);
            acc[g][4] = fma2(r, p00, acc[g][4]);
            acc[g][5] = fma2(r, p01, acc[g][5]);
            acc[g][6] = fma2(r, p02, acc[g][6]);
            acc[g][7] = fma2(r, p11, acc[g][7]);
            acc[g][8] = fma2(r, p12, acc[g][8]);
            acc[g][9] = fma2(r, p22, acc[g][9]);
        }
    } else {
#pragma unroll
    for (int g = 0; g < K; ++g) {
        const float4 ca = *reinterpret_cast<const float4*>(s_c + g * EM_COMP_STRIDE);
        const f2 r = mul2(ll[g], inv);
        const f2 d0 = add2(x, -ca.x), d1 = add2(y, -ca.y), d2 = add2(t, -ca.z);
        const f2 r0 = mul2(r, d0), r1 = mul2(r, d1), r2 = mul2(r, d2);
        acc[g][0] = add2(acc[g][0], r);
        acc[g][1] = add2(acc[g][1], r0);
        acc[g][2] = add2(acc[g][2], r1);
        acc[g][3] = add2(acc[g][3], r2);
        acc[g][4] = fma2(r0, d0, acc[g][4]);
        acc[g][5] = fma2(r0, d1, acc[g][5]);
        acc[g][6] = fma2(r0, d2, acc[g][6]);
        acc[g][7] = fma2(r1, d1, acc[g][7]);
        acc[g][8] = fma2(r1, d2, acc[g][8]);
        acc[g][9] = fma2(r2, d2, acc[g][9]);
    }
    }
}

// VEC: the sample count is a multiple of 4, so every row of x is 16-byte aligned: one LDG.128 per row brings 4
// consecutive samples (two pairs) per thread, two such triples in flight while the previous one is processed
// (96 B per thread) -- the K <= 2 fits are HBM-bound and need the bytes in flight.
template <int K, bool VEC>
__global__ void __launch_bounds__(256, (VEC && K == 1) ? 3 : ((VEC && K == 2) ? 2 : 1)) em_estep2_kernel(const EmLaunch a) {
    constexpr int NACC = K * EM_NSTAT + 1;
    __shared__ __align__(16) float s_c[K * EM_COMP_STRIDE];
    __shared__ double s_acc[NACC];
    const int fit = blockIdx.y;
    if (a.done[fit]) return;
    for (int i = threadIdx.x; i < K * EM_COMP_STRIDE; i += blockDim.x) s_c[i] = a.comps[(size_t)fit * K * EM_COMP_STRIDE + i];
    for (int i = threadIdx.x; i < NACC; i += blockDim.x) s_acc[i] = 0.0;
    __syncthreads();
    const EmHeader hd = em_header<K>(s_c);
    const float* xs = a.x + (size_t)fit * 3 * a.n;
    const float* ys = xs + a.n;
    const float* ts = ys + a.n;
    f2 acc[K][EM_NSTAT];
#pragma unroll
    for (int g = 0; g < K; ++g)
#pragma unroll
        for (int k = 0; k < EM_NSTAT; ++k) acc[g][k] = bc(0.f);
    f2 lsum = bc(0.f);
    const unsigned n = unsigned(a.n);
    const unsigned stride = gridDim.x * blockDim.x;

    // the sample loops, instantiated once per form of the statistics (as em_ow_loop)
    auto loops = [&](auto form) {
    constexpr bool SH = decltype(form)::value;
    if (VEC) {
        const float4* x4 = reinterpret_cast<const float4*>(xs);
        const float4* y4 = reinterpret_cast<const float4*>(ys);
        const float4* t4 = reinterpret_cast<const float4*>(ts);
        const unsigned n4 = n >> 2;
        const float4 zero = make_float4(0.f, 0.f, 0.f, 0.f);
        unsigned j = blockIdx.x * blockDim.x + threadIdx.x;
        float4 bx[2], by[2], bt[2];
#pragma unroll
        for (int u = 0; u < 2; ++u) {
            const unsigned k = j + u * stride;
            const bool ok = k < n4;
            bx[u] = ok ? __ldg(x4 + k) : zero; by[u] = ok ? __ldg(y4 + k) : zero; bt[u] = ok ? __ldg(t4 + k) : zero;
        }
        for (; j < n4; j += 2 * stride) {
            float4 cx[2], cy[2], ct[2];
#pragma unroll
            for (int u = 0; u < 2; ++u) { cx[u] = bx[u]; cy[u] = by[u]; ct[u] = bt[u]; }
#pragma unroll
            for (int u = 0; u < 2; ++u) {
                const unsigned k = j + (2 + u) * stride;
                const bool ok = k < n4;
                bx[u] = ok ? __ldg(x4 + k) : zero; by[u] = ok ? __ldg(y4 + k) : zero; bt[u] = ok ? __ldg(t4 + k) : zero;
            }
#pragma unroll
            for (int u = 0; u < 2; ++u) {
                if (j + u * stride >= n4) break;
                em_pair2<K, SH>(s_c, hd, mk2(cx[u].x, cx[u].y), mk2(cy[u].x, cy[u].y), mk2(ct[u].x, ct[u].y), 1.f, acc, lsum);
                em_pair2<K, SH>(s_c, hd, mk2(cx[u].z, cx[u].w), mk2(cy[u].z, cy[u].w), mk2(ct[u].z, ct[u].w), 1.f, acc, lsum);
            }
        }
    } else {
        constexpr int U = 2;  // pairs per pass: 4 samples, 12 loads in flight while the previous 4 are processed
        unsigned i = blockIdx.x * blockDim.x + threadIdx.x;   // lane 0 of pair u: i + 2 u stride, lane 1: i + (2 u + 1) stride
        float bx[2 * U], by[2 * U], bt[2 * U];
#pragma unroll
        for (int u = 0; u < 2 * U; ++u) {
            const unsigned k = i + u * stride;
            const bool ok = k < n;
            bx[u] = ok ? __ldg(xs + k) : 0.f; by[u] = ok ? __ldg(ys + k) : 0.f; bt[u] = ok ? __ldg(ts + k) : 0.f;
        }
        for (; i < n; i += 2 * U * stride) {
            float cx[2 * U], cy[2 * U], ct[2 * U];
#pragma unroll
            for (int u = 0; u < 2 * U; ++u) { cx[u] = bx[u]; cy[u] = by[u]; ct[u] = bt[u]; }
#pragma unroll
            for (int u = 0; u < 2 * U; ++u) {
                const unsigned k = i + (2 * U + u) * stride;
                const bool ok = k < n;
                bx[u] = ok ? __ldg(xs + k) : 0.f; by[u] = ok ? __ldg(ys + k) : 0.f; bt[u] = ok ? __ldg(ts + k) : 0.f;
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                if (i + 2 * u * stride >= n) break;
                const float v1 = (i + (2 * u + 1) * stride < n) ? 1.f : 0.f;   // lane 1 may be past the end
                em_pair2<K, SH>(s_c, hd, mk2(cx[2 * u], cx[2 * u + 1]), mk2(cy[2 * u], cy[2 * u + 1]), mk2(ct[2 * u], ct[2 * u + 1]), v1, acc, lsum);
            }
        }
    }
    };
    if (K >= 2 && hd.shifted) loops(EmFormTag<true>{});
    else loops(EmFormTag<false>{});
#pragma unroll
    for (int g = 0; g < K; ++g)
#pragma unroll
        for (int k = 0; k < EM_NSTAT; ++k) {
            double v = double(lo(acc[g][k])) + double(hi(acc[g][k]));
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
            if ((threadIdx.x & 31) == 0) atomicAdd(&s_acc[g * EM_NSTAT + k], v);
        }
    {
        double v = double(lo(lsum)) + double(hi(lsum));
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if ((threadIdx.x & 31) == 0) atomicAdd(&s_acc[NACC - 1], v);
    }
    __syncthreads();
    double* st = a.stats + (size_t)fit * NACC;
    for (int k = threadIdx.x; k < NACC; k += blockDim.x) atomicAdd(&st[k], s_acc[k]);
}

// ---- the same E-step fed by the bulk-copy engine (K <= 2: the HBM-bound fits) ---------------------------------------
// Rows of x are contiguous 1-D arrays, so no tensor map is needed: one elected thread issues three
// cp.async.bulk (global -> shared, completion counted in bytes on an mbarrier) per tile of EM_TMA_TILE samples,
// EM_TMA_STAGES tiles ahead; the 256 threads wait on the stage's barrier, read their samples with LDS.128 and
// hand the stage back with one __syncthreads.  Nothing of the stream sits in registers, so three CTAs per SM keep
// up to 216 KB of loads in flight per SM (the register-prefetch kernel above: 74 KB).  Needs n % 4 == 0 (16-byte
// aligned rows and tile sizes).
#ifndef EM_TMA_TWO_CTAS_K
#define EM_TMA_TWO_CTAS_K 3   // the bulk-copy kernel is held to 128 registers (two CTAs per SM) up to this many components
#endif
#define EM_TMA_TILE 2048
#define EM_TMA_STAGES 3

__device__ __forceinline__ unsigned smem_u32(const void* p) { return unsigned(__cvta_generic_to_shared(p)); }
__device__ __forceinline__ void mbar_init(unsigned long long* bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, unsigned parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "EM_WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra EM_DONE_%=;\n"
        "bra EM_WAIT_%=;\n"
        "EM_DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, unsigned bytes, unsigned long long* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

template <int K>
__global__ void __launch_bounds__(256, K == 1 ? 3 : (K <= EM_TMA_TWO_CTAS_K ? 2 : 1)) em_estep_tma_kernel(const EmLaunch a) {
    constexpr int NACC = K * EM_NSTAT + 1;
    extern __shared__ __align__(128) float s_tile[];   // [EM_TMA_STAGES][3][EM_TMA_TILE]
    __shared__ __align__(8) unsigned long long s_bar[EM_TMA_STAGES];
    __shared__ __align__(16) float s_c[K * EM_COMP_STRIDE];
    __shared__ double s_acc[NACC];
    const int fit = blockIdx.y;
    if (a.done[fit]) return;
    const unsigned n = unsigned(a.n);
    const unsigned n_tiles = (n + EM_TMA_TILE - 1) / EM_TMA_TILE;
    const float* xs = a.x + (size_t)fit * 3 * a.n;
    // this CTA's tiles: blockIdx.x, blockIdx.x + gridDim.x, ...
    const unsigned my_tiles = blockIdx.x < n_tiles ? (n_tiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
    auto issue = [&](unsigned i) {   // tile number i of this CTA into stage i % STAGES (one thread)
        const unsigned tile = blockIdx.x + i * gridDim.x, first = tile * EM_TMA_TILE;
        const unsigned count = n - first < EM_TMA_TILE ? n - first : EM_TMA_TILE, bytes = count * 4u;
        unsigned long long* bar = &s_bar[i % EM_TMA_STAGES];
        float* dst = s_tile + (i % EM_TMA_STAGES) * 3 * EM_TMA_TILE;
        mbar_expect_tx(bar, 3 * bytes);
        bulk_g2s(dst, xs + first, bytes, bar);
        bulk_g2s(dst + EM_TMA_TILE, xs + a.n + first, bytes, bar);
        bulk_g2s(dst + 2 * EM_TMA_TILE, xs + 2 * a.n + first, bytes, bar);
    };
    if (threadIdx.x == 0) {
        for (int s = 0; s < EM_TMA_STAGES; ++s) mbar_init(&s_bar[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        for (unsigned i = 0; i < EM_TMA_STAGES && i < my_tiles; ++i) issue(i);
    }
    for (int i = threadIdx.x; i < K * EM_COMP_STRIDE; i += blockDim.x) s_c[i] = a.comps[(size_t)fit * K * EM_COMP_STRIDE + i];
    for (int i = threadIdx.x; i < NACC; i += blockDim.x) s_acc[i] = 0.0;
    __syncthreads();
    const EmHeader hd = em_header<K>(s_c);
    f2 acc[K][EM_NSTAT];
#pragma unroll
    for (int g = 0; g < K; ++g)
#pragma unroll
        for (int k = 0; k < EM_NSTAT; ++k) acc[g][k] = bc(0.f);
    f2 lsum = bc(0.f);

    auto tiles = [&](auto form) {   // the tile loop, instantiated once per form of the statistics (as em_ow_loop)
    constexpr bool SH = decltype(form)::value;
    for (unsigned i = 0; i < my_tiles; ++i) {
        const unsigned stage = i % EM_TMA_STAGES;
        const unsigned first = (blockIdx.x + i * gridDim.x) * EM_TMA_TILE;
        const unsigned count = n - first < EM_TMA_TILE ? n - first : EM_TMA_TILE;
        mbar_wait(&s_bar[stage], (i / EM_TMA_STAGES) & 1u);
        const float4* tx = reinterpret_cast<const float4*>(s_tile + stage * 3 * EM_TMA_TILE);
        const float4* ty = tx + EM_TMA_TILE / 4;
        const float4* tt = ty + EM_TMA_TILE / 4;
#pragma unroll 1
        for (unsigned q = threadIdx.x; 4 * q < count; q += 256) {
            const float4 vx = tx[q], vy = ty[q], vt = tt[q];
            em_pair2<K, SH>(s_c, hd, mk2(vx.x, vx.y), mk2(vy.x, vy.y), mk2(vt.x, vt.y), 1.f, acc, lsum);
            em_pair2<K, SH>(s_c, hd, mk2(vx.z, vx.w), mk2(vy.z, vy.w), mk2(vt.z, vt.w), 1.f, acc, lsum);
        }
        // every thread is done reading the stage before the bulk-copy engine may overwrite it
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncthreads();
        if (threadIdx.x == 0 && i + EM_TMA_STAGES < my_tiles) issue(i + EM_TMA_STAGES);
    }
    };
    if (K >= 2 && hd.shifted) tiles(EmFormTag<true>{});
    else tiles(EmFormTag<false>{});
#pragma unroll
    for (int g = 0; g < K; ++g)
#pragma unroll
        for (int k = 0; k < EM_NSTAT; ++k) {
            double v = double(lo(acc[g][k])) + double(hi(acc[g][k]));
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
            if ((threadIdx.x & 31) == 0) atomicAdd(&s_acc[g * EM_NSTAT + k], v);
        }
    {
        double v = double(lo(lsum)) + double(hi(lsum));
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if ((threadIdx.x & 31) == 0) atomicAdd(&s_acc[NACC - 1], v);
    }
    __syncthreads();
    double* st = a.stats + (size_t)fit * NACC;
    for (int k = threadIdx.x; k < NACC; k += blockDim.x) atomicAdd(&st[k], s_acc[k]);
}

// ---- M-step, parameter sanitising, convergence test and the next iteration's constants (FP64) --------
// lower Cholesky factor of a symmetric 3x3 (row-major 9); false when not positive definite
__device__ __forceinline__ bool em_chol3(const double* A, double* L) {
    if (!(A[0] > 0.0)) return false;
    L[0] = sqrt(A[0]);
    L[1] = A[3] / L[0];
    L[3] = A[6] / L[0];
    const double d1 = A[4] - L[1] * L[1];
    if (!(d1 > 0.0)) return false;
    L[2] = sqrt(d1);
    L[4] = (A[7] - L[3] * L[1]) / L[2];
    const double d2 = A[8] - L[3] * L[3] - L[4] * L[4];
    if (!(d2 > 0.0)) return false;
    L[5] = sqrt(d2);
    return isfinite(L[0]) && isfinite(L[2]) && isfinite(L[5]);
}

__device__ __forceinline__ double em_det3(const double* A) {
    return A[0] * (A[4] * A[8] - A[5] * A[7]) - A[1] * (A[3] * A[8] - A[5] * A[6]) + A[2] * (A[3] * A[7] - A[4] * A[6]);
}

// init_constants (:980-1046) for one component: inverse covariance (from the Cholesky factor, as
// inv_sympd does), log|Sigma|; "last resort" diagonal treatment when either fails.
// inv = Sigma^-1 (00 01 02 11 12 22), cst = log_det_etc + log heft.
__device__ __forceinline__ void em_precision(const double* C, double heft, double* inv, double& cst) {
    double Lo[6], log_det;
    const double det = em_det3(C);
    const bool ok = em_chol3(C, Lo) && isfinite(log(det)) && det > 0.0;
    if (ok) {
        // inverse of L (lower), then Sigma^-1 = L^-T L^-1
        const double i00 = 1.0 / Lo[0], i11 = 1.0 / Lo[2], i22 = 1.0 / Lo[5];
        const double i10 = -Lo[1] * i00 * i11;
        const double i21 = -Lo[4] * i11 * i22;
        const double i20 = -(Lo[3] * i00 + Lo[4] * i10) * i22;
        inv[0] = i00 * i00 + i10 * i10 + i20 * i20;  // 00
        inv[1] = i10 * i11 + i20 * i21;              // 01
        inv[2] = i20 * i22;                          // 02
        inv[3] = i11 * i11 + i21 * i21;              // 11
        inv[4] = i21 * i22;                          // 12
        inv[5] = i22 * i22;                          // 22
        log_det = 2.0 * (log(Lo[0]) + log(Lo[2]) + log(Lo[5]));
    } else {
        log_det = 0.0;
        inv[1] = inv[2] = inv[4] = 0.0;
        const int dg[3] = {0, 4, 8}, iv[3] = {0, 3, 5};
        for (int d = 0; d < 3; ++d) {
            const double v = fmax(C[dg[d]], DBL_MIN);
            inv[iv[d]] = 1.0 / v;
            log_det += log(v);
        }
    }
    cst = -(1.5 * log(2.0 * 3.14159265358979323846) + 0.5 * log_det) + log(fmax(heft, DBL_MIN));
}
// The FP32 E-step record of a component: mu, b = -1/2 Sigma^-1 (off-diagonals doubled), cst; the two spare floats of
// the records of components 0 and 1 carry the fit's header (below).
__device__ __forceinline__ void em_record(const double* mean, const double* inv, double cst, float pad0, float pad1, float* rec) {
    rec[0] = float(mean[0]); rec[1] = float(mean[1]); rec[2] = float(mean[2]);
    rec[3] = float(-0.5 * inv[0]); rec[4] = float(-inv[1]); rec[5] = float(-inv[2]);
    rec[6] = float(-0.5 * inv[3]); rec[7] = float(-inv[4]); rec[8] = float(-0.5 * inv[5]);
    rec[9] = float(cst);
    rec[10] = pad0; rec[11] = pad1;
}

// One CTA per fit, thread g = component g.  prep_only: compute the E-step constants of the initial
// parameters (the init_constants call that opens iteration 1) and nothing else.
__global__ void __launch_bounds__(EM_MSTEP_FUSED_THREADS) em_mstep_kernel(const EmLaunch a, int prep_only) {
    const int fit = blockIdx.x, g = threadIdx.x, K = a.K;
    __shared__ double s_heft[MCSIM_MAX_GAUSSIANS];
    __shared__ double s_mean[MCSIM_MAX_GAUSSIANS][3];
    double* hefts = a.hefts + (size_t)fit * K;
    double* means = a.means + (size_t)fit * K * 3;
    double* fcovs = a.fcovs + (size_t)fit * K * 9;
    double* st = a.stats + (size_t)fit * (K * EM_NSTAT + 1);
    if (!prep_only && a.done[fit]) return;
    // reference point the E-step that produced these statistics used (header written by the previous M-step / prep launch;
    // read by every thread here, overwritten at the end of this kernel behind two barriers)
    float* recs = a.comps + (size_t)fit * K * EM_COMP_STRIDE;
    const bool was_shifted = !prep_only && K >= 2 && recs[EM_COMP_STRIDE + 11] != 0.f;
    double ref_c[3] = {0.0, 0.0, 0.0};
    if (was_shifted) { ref_c[0] = recs[10]; ref_c[1] = recs[11]; ref_c[2] = recs[EM_COMP_STRIDE + 10]; }
    __shared__ double s_tot[MCSIM_MAX_GAUSSIANS * EM_NSTAT + 1];
    if (!prep_only && a.n_peers > 1) {
        // Fused exchange (no collective launch between E- and M-step).  (1) push: this CTA stores its fit's local block
        // into slot [my_rank] of EVERY device's gather buffer -- remote stores over NVLink peer memory, nobody waits for
        // them -- clears the local block for the next E-step, and after a system-scope fence raises its arrival counter
        // on every device.  (2) wait for the G counters of this fit on THIS device.  (3) add the G slots in rank order:
        // the same order on every device, so totals, parameters and convergence decisions are bit-identical everywhere.
        // Slots alternate between two buffers by iteration parity: a device can be one iteration ahead of a peer that
        // is still adding up, never two (it would be waiting for that peer's next counter).
        // Launched with EM_MSTEP_FUSED_THREADS >= 10 K + 1 threads in this mode: one statistic per thread, so the push and
        // the gather are one round of memory latency each (with one warp the 161 doubles of K = 16 took 6 rounds, 3 us apiece).
        const int nst = K * EM_NSTAT + 1, G = a.n_peers;
        const size_t per_dev = (size_t)a.n_fits * nst;
        for (int i = g; i < nst; i += blockDim.x) {
            const double v = __ldcg(st + i);
            st[i] = 0.0;
#pragma unroll
            for (int q = 0; q < EM_MAX_PEERS; ++q)
                if (q < G) a.peer_slots[q][(size_t)a.my_rank * per_dev + (size_t)fit * nst + i] = v;
        }
        __threadfence_system();
        __syncthreads();
        unsigned long long* pf = nullptr;
#pragma unroll
        for (int q = 0; q < EM_MAX_PEERS; ++q)
            if (q == g) pf = a.peer_flags[q];
        if (g < G) *(volatile unsigned long long*)(pf + (size_t)fit * G + a.my_rank) = a.wait_value;
        __shared__ int s_late;
        if (g == 0) s_late = 0;
        __syncthreads();
        if (g < G) {
            const volatile unsigned long long* f = a.my_flags + (size_t)fit * G + g;
            const long long t0 = clock64();
            while (*f < a.wait_value)
                if (clock64() - t0 > EM_PEER_TIMEOUT_CYCLES) { s_late = 1; break; }
            __threadfence_system();
        }
        __syncthreads();
        if (s_late) { if (g == 0) a.done[fit] = EM_DONE_EXCHANGE_TIMEOUT; return; }
        const double* slots = a.my_slots;
        for (int i = g; i < nst; i += blockDim.x) {
            double t = 0.0;
#pragma unroll
            for (int p = 0; p < EM_MAX_PEERS; ++p)
                if (p < G) t += __ldcv(slots + p * per_dev + (size_t)fit * nst + i);
            s_tot[i] = t;
        }
        __syncthreads();
        st = s_tot;   // (the loop that clears the consumed statistics below then clears this copy)
    }
    double C[9], mean[3], heft = 0.0;
    if (g < K) {
        heft = hefts[g];
        for (int d = 0; d < 3; ++d) mean[d] = means[g * 3 + d];
        for (int k = 0; k < 9; ++k) C[k] = fcovs[g * 9 + k];
    }
    if (!prep_only) {
        // ---- em_update_params (:2583-2638): conditional update of component g
        if (g < K) {
            const double* s = st + g * EM_NSTAT;
            const double sr = fmax(s[0], DBL_MIN);
            if (isfinite(sr)) {
                const double d0 = s[1] / sr, d1 = s[2] / sr, d2 = s[3] / sr;
                double N[9];
                N[0] = s[4] / sr - d0 * d0; N[1] = s[5] / sr - d0 * d1; N[2] = s[6] / sr - d0 * d2;
                N[4] = s[7] / sr - d1 * d1; N[5] = s[8] / sr - d1 * d2; N[8] = s[9] / sr - d2 * d2;
                N[3] = N[1]; N[6] = N[2]; N[7] = N[5];
                if (N[0] < a.var_floor) N[0] = a.var_floor;
                if (N[4] < a.var_floor) N[4] = a.var_floor;
                if (N[8] < a.var_floor) N[8] = a.var_floor;
                bool fin = true;
                for (int k = 0; k < 9; ++k) fin = fin && isfinite(N[k]);
                double Lo[6];
                const double det = em_det3(N);
                if (fin && det > 0.0 && isfinite(log(det)) && em_chol3(N, Lo)) {
                    heft = sr / a.n_total;
                    if (was_shifted) { mean[0] = ref_c[0] + d0; mean[1] = ref_c[1] + d1; mean[2] = ref_c[2] + d2; }
                    else { mean[0] += d0; mean[1] += d1; mean[2] += d2; }
                    for (int k = 0; k < 9; ++k) C[k] = N[k];
                }
            }
            // ---- em_fix_params (:2755-2765): variance floor / ceiling / NaN
            const int dg[3] = {0, 4, 8};
            for (int d = 0; d < 3; ++d) {
                double& v = C[dg[d]];
                if (v < a.var_floor) v = a.var_floor;
                else if (v > DBL_MAX) v = DBL_MAX;
                else if (isnan(v)) v = 1.0;
            }
            s_heft[g] = heft;
            for (int d = 0; d < 3; ++d) s_mean[g][d] = mean[d];
        }
        __syncthreads();
        if (g == 0) {
            // duplicate components (:2770-2787): equal heft (within epsilon) and identical mean -> heft 0
            for (int g1 = 0; g1 < K; ++g1) {
                if (!(s_heft[g1] > 0.0)) continue;
                for (int g2 = g1 + 1; g2 < K; ++g2)
                    if (s_heft[g2] > 0.0 && fabs(s_heft[g1] - s_heft[g2]) <= DBL_EPSILON &&
                        s_mean[g1][0] == s_mean[g2][0] && s_mean[g1][1] == s_mean[g2][1] && s_mean[g1][2] == s_mean[g2][2])
                        s_heft[g2] = 0.0;
            }
            // heft floor / clip / NaN, renormalise (:2789-2799)
            double sum = 0.0;
            for (int k = 0; k < K; ++k) {
                double& h = s_heft[k];
                if (h < DBL_MIN) h = DBL_MIN;
                else if (h > 1.0) h = 1.0;
                else if (isnan(h)) h = 1.0 / K;
                sum += h;
            }
            if (sum < 1.0 - DBL_EPSILON || sum > 1.0 + DBL_EPSILON)
                for (int k = 0; k < K; ++k) s_heft[k] /= sum;
            // convergence test of em_iterate (:2447-2465)
            const double avg = st[K * EM_NSTAT] / a.n_total;
            if (a.trace) a.trace[(size_t)fit * a.iters + a.iter] = avg;
            a.iters_run[fit] = a.iter + 1;
            if (!isfinite(avg)) { a.done[fit] = 2; }
            else if (fabs(a.old_avg[fit] - avg) <= DBL_EPSILON) { a.done[fit] = 1; }
            a.old_avg[fit] = avg;
        }
        __syncthreads();
        if (g < K) {
            heft = s_heft[g];
            hefts[g] = heft;
            for (int d = 0; d < 3; ++d) means[g * 3 + d] = mean[d];
            for (int k = 0; k < 9; ++k) fcovs[g * 9 + k] = C[k];
        }
        // statistics consumed: zero them for the next E-step
        for (int i = g; i < K * EM_NSTAT + 1; i += blockDim.x) st[i] = 0.0;
    }
    // ---- init_constants for the next E-step (also floors the hefts at DBL_MIN, :1040-1045), and the reference point of
    // its statistics: c = sum_g heft_g mean_g (every thread adds the same K terms in the same order)
    __shared__ double s_T[MCSIM_MAX_GAUSSIANS];
    double inv[6], cst = 0.0;
    if (g < K) {
        if (heft < DBL_MIN) { heft = DBL_MIN; hefts[g] = heft; }
        s_heft[g] = heft;
        for (int d = 0; d < 3; ++d) s_mean[g][d] = mean[d];
        em_precision(C, heft, inv, cst);
    }
    __syncthreads();
    double c[3] = {0.0, 0.0, 0.0}, hs = 0.0;
    for (int k = 0; k < K; ++k) {
        hs += s_heft[k];
        for (int d = 0; d < 3; ++d) c[d] += s_heft[k] * s_mean[k][d];
    }
    float c32[3];
    for (int d = 0; d < 3; ++d) c32[d] = float(c[d] / hs);
    if (g < K) {
        const double row0 = fabs(inv[0]) + fabs(inv[1]) + fabs(inv[2]), row1 = fabs(inv[1]) + fabs(inv[3]) + fabs(inv[4]),
                     row2 = fabs(inv[2]) + fabs(inv[4]) + fabs(inv[5]);
        const double e0 = mean[0] - c32[0], e1 = mean[1] - c32[1], e2 = mean[2] - c32[2];
        const double R = sqrt(e0 * e0 + e1 * e1 + e2 * e2) + 6.0 * sqrt(fmax(C[0], 0.0) + fmax(C[4], 0.0) + fmax(C[8], 0.0));
        const double T = 0.5 * fmax(row0, fmax(row1, row2)) * R * R;
        s_T[g] = isfinite(T) ? T : DBL_MAX;
    }
    __syncthreads();
    bool shifted = K >= 2 && a.moments != EM_MOMENTS_CENTRED && isfinite(c32[0]) && isfinite(c32[1]) && isfinite(c32[2]);
    if (shifted && a.moments == EM_MOMENTS_AUTO) {
        for (int k = 0; k < K; ++k) shifted = shifted && s_T[k] < EM_SHIFT_LIMIT;
        // K <= 5 runs on packed pairs with every d in registers: there the shifted form only trades FADD2 / FMUL2 for
        // three-operand FFMA2 (config 5: K = 4 0.583 against 0.606 of the FP32 peak, K = 2, 3, 5 +1 .. +2 %), while its
        // cancellation amplifies the run-to-run noise of a fit whose samples arrive in varying order (the EM-refit
        // truncation compacts its free samples with atomics) -- not worth it.  K >= 6: +7 .. +20 %.
        if (K < EM_SHIFT_MIN_K) shifted = false;
    }
    if (g < K) {
        const float pad0 = g == 0 ? c32[0] : (g == 1 ? c32[2] : 0.f);
        const float pad1 = g == 0 ? c32[1] : (g == 1 ? (shifted ? 1.f : 0.f) : 0.f);
        em_record(mean, inv, cst, pad0, pad1, recs + (size_t)g * EM_COMP_STRIDE);
    }
}

#endif
