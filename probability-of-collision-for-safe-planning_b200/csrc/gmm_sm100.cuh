// gmm_sm100.cuh -- GMM mode: sample -> collide -> per-component moment refit, device-resident over all
// waypoints of a plan.
//
// Replaces (mcsimplugin/, reference repository):
//   GM_Model::sampleNPoints  GM_Model.h:83-116   multinomial counts, then mvnrnd per component
//                                                 (armadillo glue_mvnrnd_meat.hpp:89-147: x = mu + chol_lower(S) z)
//   MCSimulator::truncateGMM MCSimulator.h:570-642  collision test of every sample (:587 -> :241-253), mean and
//                                                 unbiased covariance of the collision-free samples of each
//                                                 component (:592-605), weights = free share (:618-629), returns
//                                                 colliding fraction (:633-641)
//   the GMM branch of EKF_GaussProp             :672-678 (waypoint 0), :766-772 (EKFpredict per component),
//                                                 :804-817 (EKFupdate per component, truncate), :843-856 (P)
//
// Three kernels per waypoint, no host round trip between waypoints:
//   gmm_count_kernel    (Philox mode) one uniform per sample -> component histogram (the multinomial counts)
//   gmm_sample_kernel   one thread per sample (grid-stride): x = mu_g + L_g z, SAT test, per-component
//                       (n_free, n_coll, sum d, sum d d^T) with d = x - mu_g = L_g z  (shifted sums: the one-pass
//                       covariance of op_cov_meat.hpp:38-47 without its cancellation)
//   gmm_refit_kernel    one thread per component: truncated moments, weights, p_t; then EKFpredict + EKFupdate of
//                       the component for the next waypoint and the Cholesky factor the next sampling needs
// Arithmetic: FP64 for sampling and moments (N x 11 FP64 adds per waypoint are negligible next to the SAT);
// SAT in FP64 in replay mode (bit-exact flags against the CPU oracle), FP32 from the __constant__ table in
// Philox mode.
#ifndef MCSIM_GMM_CUH
#define MCSIM_GMM_CUH

#include <cooperative_groups.h>

#include "estimator_math.h"
#include "philox.h"
#include "sat_device.cuh"
#include "linear_truncation.h"
#include "mc_philox2_sm100.cuh"   // packed FP32 pairs: box_muller2, hull_frame2, class_frame2, class_pair_test2, robot_hits2

#define GMM_STAT 11        // n_free, n_coll, sum d[3], sum d d^T [6: xx xy xz yy yz zz]
#define GMM_COMP 19        // w, mu[3], S[9] row-major, L[6] lower Cholesky (l00 l10 l11 l20 l21 l22)
#define GMM_REC 13         // w, mu[3], S[9]: layout of the recorded mixtures (gmm_pre / gmm_post)

// FP64 scene for the replay-mode SAT (global memory): n_hulls x HullD, then pairs grouped by hull
struct HullD { double ox, oy, cp, sp, hx, hy; int first, count; };
struct PairD { double cx, cy, cb, sb, hx, hy; };

struct GmmStepTable {      // per step t (0..T-2), FP64: applied command, motion noise diag, observations
    double u[3];
    double M[3];
    double z[MCSIM_MAX_LANDMARKS];
    double xs[3], xn[3], us[3];   // nominal plan: x*_t, x*_{t+1}, u*_t (the deviation-state filter linearises about them)
};

struct GmmLaunch {
    int K;
    int t;                        // waypoint index of this truncation
    int T;
    int L;                        // landmarks
    unsigned long long n;         // samples per waypoint
    double* comps;                // [K][GMM_COMP] current mixture
    double* stats;                // [K][GMM_STAT] accumulators (zeroed by the refit kernel)
    unsigned* hist;               // [K] multinomial counts of this waypoint
    const int* counts_replay;     // replay: [T][K] injected counts, else NULL
    const double* normals;        // replay: 3 x n normals of this waypoint, sample-major
    unsigned k0, k1;              // Philox key
    unsigned stream_id;           // Philox subsequence (trajectory id)
    const HullD* hulls_d;         // FP64 scene (replay)
    const PairD* pairs_d;
    int n_hulls;
    const GmmStepTable* steps;    // [T-1]
    double Q;
    double lx[MCSIM_MAX_LANDMARKS], ly[MCSIM_MAX_LANDMARKS];
    double* probabilities;        // [T]
    double* rec_pre;              // [T][K][GMM_REC] mixture right before each truncation (optional)
    double* rec_post;             // [K][GMM_REC] mixture after the last truncation (optional)
    double* trunc_out;            // [K][GMM_REC] mixture right after THIS truncation (optional; gmm_truncate)
    unsigned long long* n_out;    // [2K] n_coll, n_free of this truncation (optional)
    int* degenerate;              // count of (waypoint, component) refits with < 2 free samples
    int do_ekf;                   // refit kernel: propagate to waypoint t+1 (0 for the single-shot API); 1: the reference's EKF
                                  // on the state, 2: deviation-state filter about the nominal plan (mcsim_mc_params.gmm_filter)
    int fast_filter;              // Philox mode: EKF updates with one division on the dependent chain (same update up to rounding)
    int B;                        // batch: independent trajectories in the same launches (Philox mode; 0/1 = single)
    const float4* sat2;           // orientation-grouped FP32 pair table (Philox mode SAT)
    const unsigned* cull;         // Philox mode: the plan's candidate lists (sat_device.cuh), NULL = every pair
    int cull_stride;              // batch: 32-bit words between the members' lists (0: one plan for all)
    const float2* nominal;        // [T] (x*_t, y*_t), the centres of the lists; batch: [B][T] when cull_stride != 0
    unsigned long long* tests;    // optional: executed pair tests (all members)
    int spread;                   // Philox sample kernel: one sample per thread instead of two (small n, single run)
    // EM-refit variant (8f-4): the poses of this waypoint's samples, FP32, rows x | y | theta at stride free_cap, written at
    // the sample's own index, and free_flag[i] = 1 for the collision-free ones (gmm_sample_kernel only); NULL: not kept.
    // gmm_compact_free_kernel then packs the free ones IN SAMPLE ORDER (an atomic append handed the EM fit a different
    // order every run, and the truncate -> EM -> EKF chain amplified the 1e-10 that made to a flipped hit flag or thirty).
    float* free_out;
    unsigned char* free_flag;
    unsigned* free_n;
    unsigned free_cap;
};

// Batch member b (blockIdx.y of the count / sample kernels, blockIdx.x of the refit kernel): its own mixture, statistics,
// counts, step table, per-waypoint probabilities, lists and Philox subsequence.  Only these few pointers are per member;
// everything else is read from the launch struct where it lies (the kernel parameter, constant bank).  (A by-value copy
// of the whole GmmLaunch with adjusted pointers -- round 1 -- put 824 bytes on every thread's stack: its landmark
// arrays are indexed at run time.)
struct GmmMember {
    double* comps;
    double* stats;
    unsigned* hist;
    const GmmStepTable* steps;
    double* probabilities;
    int* degenerate;
    const unsigned* cull;
    const float2* nominal;
    unsigned stream_id;
};
__device__ __forceinline__ GmmMember gmm_member(const GmmLaunch& in, int b) {
    GmmMember m;
    m.comps = in.comps + size_t(b) * MCSIM_MAX_GAUSSIANS * GMM_COMP;
    m.stats = in.stats + size_t(b) * MCSIM_MAX_GAUSSIANS * GMM_STAT;
    m.hist = in.hist + size_t(b) * MCSIM_MAX_GAUSSIANS;
    m.steps = in.steps + size_t(b) * (in.T > 1 ? in.T - 1 : 1);
    m.probabilities = in.probabilities + size_t(b) * in.T;
    m.degenerate = in.degenerate + b;
    m.stream_id = in.stream_id + unsigned(b);
    const bool own = in.cull && in.cull_stride;
    m.cull = own ? in.cull + size_t(b) * in.cull_stride : in.cull;
    m.nominal = own ? in.nominal + size_t(b) * in.T : in.nominal;
    return m;
}

// ---- FP64 SAT, statement order of oracle/sat_cpu.h (no contraction: _rn intrinsics) --------------------------
__device__ __forceinline__ double gap_d(double hcx, double hcy, double ch, double sh, double hhx, double hhy, const PairD& b) {
    const double dx = __dsub_rn(b.cx, hcx), dy = __dsub_rn(b.cy, hcy);
    const double c = fabs(__dadd_rn(__dmul_rn(ch, b.cb), __dmul_rn(sh, b.sb)));
    const double s = fabs(__dsub_rn(__dmul_rn(sh, b.cb), __dmul_rn(ch, b.sb)));
    const double drx = fabs(__dadd_rn(__dmul_rn(ch, dx), __dmul_rn(sh, dy)));
    const double dry = fabs(__dsub_rn(__dmul_rn(ch, dy), __dmul_rn(sh, dx)));
    const double dbx = fabs(__dadd_rn(__dmul_rn(b.cb, dx), __dmul_rn(b.sb, dy)));
    const double dby = fabs(__dsub_rn(__dmul_rn(b.cb, dy), __dmul_rn(b.sb, dx)));
    double g = __dsub_rn(drx, __dadd_rn(__dadd_rn(hhx, __dmul_rn(b.hx, c)), __dmul_rn(b.hy, s)));
    double t = __dsub_rn(dry, __dadd_rn(__dadd_rn(hhy, __dmul_rn(b.hx, s)), __dmul_rn(b.hy, c)));
    if (t > g) g = t;
    t = __dsub_rn(dbx, __dadd_rn(__dadd_rn(b.hx, __dmul_rn(hhx, c)), __dmul_rn(hhy, s)));
    if (t > g) g = t;
    t = __dsub_rn(dby, __dadd_rn(__dadd_rn(b.hy, __dmul_rn(hhx, s)), __dmul_rn(hhy, c)));
    if (t > g) g = t;
    return g;
}

__device__ __forceinline__ double clearance_d(const HullD* hulls, const PairD* pairs, int n_hulls, double x, double y, double th) {
    const double ct = cos(th), st = sin(th);
    double best = 1.7976931348623157e308;
    for (int h = 0; h < n_hulls; ++h) {
        const HullD hu = hulls[h];
        const double hcx = __dadd_rn(x, __dsub_rn(__dmul_rn(ct, hu.ox), __dmul_rn(st, hu.oy)));
        const double hcy = __dadd_rn(y, __dadd_rn(__dmul_rn(st, hu.ox), __dmul_rn(ct, hu.oy)));
        const double ch = __dsub_rn(__dmul_rn(ct, hu.cp), __dmul_rn(st, hu.sp));
        const double sh = __dadd_rn(__dmul_rn(st, hu.cp), __dmul_rn(ct, hu.sp));
        for (int b = hu.first; b < hu.first + hu.count; ++b) {
            const double g = gap_d(hcx, hcy, ch, sh, hu.hx, hu.hy, pairs[b]);
            if (g < best) best = g;
        }
    }
    return best;
}

// ---- multinomial counts (Philox mode): inverse CDF on one uniform per sample -----------------------------------
__global__ void __launch_bounds__(256) gmm_count_kernel(const GmmLaunch a_in) {
    const GmmLaunch& a = a_in;
    const GmmMember m = gmm_member(a_in, blockIdx.y);
    __shared__ float s_cdf[MCSIM_MAX_GAUSSIANS];
    __shared__ unsigned s_hist[MCSIM_MAX_GAUSSIANS];
    if (threadIdx.x < MCSIM_MAX_GAUSSIANS) s_hist[threadIdx.x] = 0;
    if (threadIdx.x == 0) {
        double acc = 0, tot = 0;
        for (int g = 0; g < a.K; ++g) tot += m.comps[g * GMM_COMP];
        for (int g = 0; g < a.K; ++g) { acc += m.comps[g * GMM_COMP]; s_cdf[g] = float(acc / tot); }
        // components after the last one with positive weight can never be drawn
        for (int g = a.K - 1; g >= 0 && !(m.comps[g * GMM_COMP] > 0); --g) s_cdf[g] = 2.0f;
        int last = a.K - 1;
        while (last > 0 && !(m.comps[last * GMM_COMP] > 0)) --last;
        s_cdf[last] = 2.0f;  // u in (0,1]: the last live component absorbs rounding
    }
    __syncthreads();
    // one Philox block serves 4 consecutive samples (block index i / 4, word i % 4): one thread per block,
    // block-uniform trip count, per-component ballots instead of per-sample shared atomics
    const unsigned long long n4 = (a.n + 3) >> 2;
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    const int K = a.K;
    for (unsigned long long base = (unsigned long long)blockIdx.x * blockDim.x; base < n4; base += stride) {
        const unsigned long long j = base + threadIdx.x;
        mcsim_rng::u32x4 c = {unsigned(j), m.stream_id, unsigned(2 * a.t + 1) | unsigned((j >> 32) << 16), unsigned(mcsim_rng::PURPOSE_GMM_SAMPLE)};
        c = mcsim_rng::philox4x32_10(c, a.k0, a.k1);
        const unsigned w[4] = {c.x, c.y, c.z, c.w};
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const bool valid = j < n4 && 4 * j + q < a.n;
            const float u = mcsim_rng::u01(w[q]);
            int g = 0;
            while (g < K - 1 && u > s_cdf[g]) ++g;
            for (int gg = 0; gg < K; ++gg) {
                const unsigned m = __ballot_sync(0xffffffffu, valid && g == gg);
                if (m && (threadIdx.x & 31) == 0) atomicAdd(&s_hist[gg], unsigned(__popc(m)));
            }
        }
    }
    __syncthreads();
    if (threadIdx.x < a.K && s_hist[threadIdx.x]) atomicAdd(&m.hist[threadIdx.x], s_hist[threadIdx.x]);
}

// ---- sample -> collide -> moments -------------------------------------------------------------------------------
template <int MODE>
__global__ void __launch_bounds__(256) gmm_sample_kernel(const GmmLaunch a_in) {
    const GmmLaunch& a = a_in;
    const GmmMember m = gmm_member(a_in, blockIdx.y);
    __shared__ double s_comp[MCSIM_MAX_GAUSSIANS][GMM_COMP];
    __shared__ double s_stat[MCSIM_MAX_GAUSSIANS][GMM_STAT];
    __shared__ unsigned long long s_prefix[MCSIM_MAX_GAUSSIANS + 1];
    __shared__ unsigned char s_ccnt[(MCSIM_MAX_HULLS + MCSIM_MAX_POLY_HULLS) * MC_CULL_LEVELS];
    __shared__ ushort2 s_cidx[(MCSIM_MAX_HULLS + MCSIM_MAX_POLY_HULLS) * MC_CULL_CAP];
    const int K = a.K;
    if (MODE != MCSIM_MODE_REPLAY && m.cull) {
        const int nh = c_scene.n_hulls + c_scene.n_poly;   // the lists' hull index: box hulls, then polygon hulls
        const unsigned char* gc = reinterpret_cast<const unsigned char*>(m.cull) + size_t(a.t) * nh * MC_CULL_LEVELS;
        const ushort2* gi = reinterpret_cast<const ushort2*>(m.cull + (size_t(a.T) * nh * MC_CULL_LEVELS + 3) / 4) + size_t(a.t) * nh * MC_CULL_CAP;
        for (int i = threadIdx.x; i < nh * MC_CULL_LEVELS; i += blockDim.x) s_ccnt[i] = gc[i];
        for (int i = threadIdx.x; i < nh * MC_CULL_CAP; i += blockDim.x) s_cidx[i] = gi[i];
    }
    unsigned long long my_tests = 0;
    for (int i = threadIdx.x; i < K * GMM_COMP; i += blockDim.x) s_comp[i / GMM_COMP][i % GMM_COMP] = m.comps[i];
    for (int i = threadIdx.x; i < K * GMM_STAT; i += blockDim.x) s_stat[i / GMM_STAT][i % GMM_STAT] = 0.0;
    if (threadIdx.x == 0) {
        unsigned long long acc = 0;
        for (int g = 0; g < K; ++g) {
            s_prefix[g] = acc;
            acc += (MODE == MCSIM_MODE_REPLAY) ? (unsigned long long)a.counts_replay[a.t * K + g] : (unsigned long long)m.hist[g];
        }
        s_prefix[K] = acc;
    }
    __syncthreads();

    // Block-uniform trip count.  Every thread keeps the sums of its free samples in registers for the component
    // its samples currently belong to (samples are grouped by component and a thread's sample index only grows, so
    // that component changes at most K - 1 times); when any lane of the warp moves on to another component -- and
    // after the last pass -- the warp reduces its lanes' sums, segmented by component, into ONE shared-memory atomic
    // per statistic and warp.  (Per-thread FP64 shared atomics are CAS loops: with every thread of the grid
    // flushing 11 of them onto K x 11 addresses they serialised the whole kernel; a reduction after every pass
    // costs 90 shuffles per sample.)
    int g = 0, g_acc = -1;
    unsigned nf = 0, nc = 0;
    double acc[9] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
    auto flush = [&]() {
        const int gmin = __reduce_min_sync(0xffffffffu, g_acc >= 0 ? g_acc : 0x7fffffff);
        const int gmax = __reduce_max_sync(0xffffffffu, g_acc);
        for (int gg = gmin; gg <= gmax; ++gg) {  // warp-uniform
            const bool sel = g_acc == gg;
            const unsigned wf = __reduce_add_sync(0xffffffffu, sel ? nf : 0u);
            const unsigned wc = __reduce_add_sync(0xffffffffu, sel ? nc : 0u);
            if (wf) {
                double v[9];
#pragma unroll
                for (int k = 0; k < 9; ++k) {
                    v[k] = sel ? acc[k] : 0.0;
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) v[k] += __shfl_xor_sync(0xffffffffu, v[k], o);
                }
                if ((threadIdx.x & 31) == 0) {
                    atomicAdd(&s_stat[gg][0], double(wf));
#pragma unroll
                    for (int k = 0; k < 9; ++k) atomicAdd(&s_stat[gg][2 + k], v[k]);
                }
            }
            if (wc && (threadIdx.x & 31) == 0) atomicAdd(&s_stat[gg][1], double(wc));
        }
        g_acc = -1;
        nf = nc = 0;
#pragma unroll
        for (int k = 0; k < 9; ++k) acc[k] = 0.0;
    };
    const unsigned long long n = s_prefix[K] < a.n ? s_prefix[K] : a.n;
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (unsigned long long base = (unsigned long long)blockIdx.x * blockDim.x; base < n; base += stride) {
        const unsigned long long i = base + threadIdx.x;
        const bool valid = i < n;
        bool hit = false;
        double d0 = 0.0, d1 = 0.0, d2 = 0.0;
        float xf = 0.f, yf = 0.f, sn = 0.f, cs = 1.f;
        float px = 0.f, py = 0.f, pth = 0.f;   // the sample's pose (kept for the EM-refit variant)
        if (valid) {
            while (g < K - 1 && i >= s_prefix[g + 1]) ++g;  // i only grows: the component index is monotone
            double z0, z1, z2;
            if (MODE == MCSIM_MODE_REPLAY) {
                z0 = a.normals[3 * i]; z1 = a.normals[3 * i + 1]; z2 = a.normals[3 * i + 2];
            } else {
                // samples 2 j and 2 j + 1 share one block (philox.h, "GMM sample draws")
                const unsigned long long j = i >> 1;
                const bool odd = (i & 1ull) != 0;
                mcsim_rng::u32x4 c = {unsigned(j), m.stream_id, unsigned(2 * a.t) | unsigned((j >> 32) << 16), unsigned(mcsim_rng::PURPOSE_GMM_SAMPLE)};
                c = mcsim_rng::philox4x32_10(c, a.k0, a.k1);
                float f0, f1, f2, f3;
                mcsim_rng::box_muller_bits(odd ? c.y >> 8 : c.x >> 8, odd ? c.z >> 16 : c.z & 0xffffu, f0, f1);
                mcsim_rng::box_muller_bits(c.w >> 8, mcsim_rng::gmm_pair_c_angle(c), f2, f3);
                z0 = f0; z1 = f1; z2 = odd ? f3 : f2;
            }
            const double* cp = s_comp[g];
            const double* Lc = cp + 13;
            // d = L z in the summation order of a 3x3 times 3x1 product (left to right), no contraction
            d0 = __dmul_rn(Lc[0], z0);
            d1 = __dadd_rn(__dmul_rn(Lc[1], z0), __dmul_rn(Lc[2], z1));
            d2 = __dadd_rn(__dadd_rn(__dmul_rn(Lc[3], z0), __dmul_rn(Lc[4], z1)), __dmul_rn(Lc[5], z2));
            const double x = __dadd_rn(d0, cp[1]), y = __dadd_rn(d1, cp[2]), th = __dadd_rn(d2, cp[3]);
            px = float(x); py = float(y); pth = float(th);
            if (MODE == MCSIM_MODE_REPLAY) {
                hit = clearance_d(a.hulls_d, a.pairs_d, a.n_hulls, x, y, th) <= 0.0;
            } else {
                // theta of a sample is mu_theta + O(sigma): reduce to [-pi, pi] before the FP32 sincos
                double tr = th;
                if (tr > MCSIM_PI) tr -= 2 * MCSIM_PI;
                if (tr < -MCSIM_PI) tr += 2 * MCSIM_PI;
                sincosf(float(tr), &sn, &cs);
                xf = float(x);
                yf = float(y);
            }
        }  // valid
        if (MODE != MCSIM_MODE_REPLAY) {
            // warp-uniform choice of the candidate list by the warp's largest deviation from x*_t (exact: sat_device.cuh)
            int lev = MC_CULL_LEVELS;
            if (m.cull) {
                const float2 nom = __ldg(m.nominal + a.t);
                const float ex = xf - nom.x, ey = yf - nom.y;
                const float dm = __uint_as_float(__reduce_max_sync(0xffffffffu, valid ? __float_as_uint(fmaf(ex, ex, ey * ey)) : 0u));
                lev = cull_level(dm);
                if (lev < MC_CULL_LEVELS && s_ccnt[lev] == MC_CULL_OVERFLOW) lev = MC_CULL_LEVELS;
            }
            unsigned tested = 0;
            if (lev < MC_CULL_LEVELS) {
                hit = robot_hits_listed(a.sat2, s_ccnt, s_cidx, lev, xf, yf, cs, sn, tested);   // box and polygon hulls
            } else {
                hit = robot_hits_grouped(a.sat2, xf, yf, cs, sn);
                tested = c_scene.n_pairs;
                if (c_scene.n_poly && poly_clearance(xf, yf, cs, sn, tested) <= 0.f) hit = true;   // polygon hulls: every live pair
            }
            if (valid) my_tests += tested;
        }
        if (a.free_out && valid) {   // kept at the sample's own index; packed in sample order by gmm_compact_free_kernel
            a.free_flag[i] = hit ? 0 : 1;
            if (!hit) {
                a.free_out[i] = px;
                a.free_out[size_t(a.free_cap) + i] = py;
                a.free_out[2 * size_t(a.free_cap) + i] = pth;
            }
        }
        if (__any_sync(0xffffffffu, valid && g_acc >= 0 && g != g_acc)) flush();
        if (valid) {
            g_acc = g;
            if (hit) {
                ++nc;
            } else {
                ++nf;
                acc[0] += d0; acc[1] += d1; acc[2] += d2;
                acc[3] += d0 * d0; acc[4] += d0 * d1; acc[5] += d0 * d2;
                acc[6] += d1 * d1; acc[7] += d1 * d2; acc[8] += d2 * d2;
            }
        }
    }
    flush();
    __syncthreads();
    for (int i = threadIdx.x; i < K * GMM_STAT; i += blockDim.x) {
        const double v = s_stat[i / GMM_STAT][i % GMM_STAT];
        if (v != 0.0) atomicAdd(&m.stats[i], v);
    }
    if (MODE != MCSIM_MODE_REPLAY && a.tests) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) my_tests += __shfl_xor_sync(0xffffffffu, my_tests, o);
        if ((threadIdx.x & 31) == 0 && my_tests) atomicAdd(a.tests, my_tests);
    }
}

// ---- Philox mode: the same step in FP32 ----------------------------------------------------------------------------
// Same Philox counters, draws and sample -> component assignment as gmm_sample_kernel<PHILOX>; what changes is
// the arithmetic: x = mu_g + L_g z, the angle reduction and sin / cos in FP32 (MUFU), 32-bit sample indices, the
// sums of the free samples' d = L_g z still in FP64.  There is no bit-exact bar in this mode (the bar is the
// reference's run-to-run distribution, tests/test_gmm_gpu.py); replay mode keeps the FP64 kernel above.

// per-thread sums of the free samples of the component the thread's samples currently belong to (see
// gmm_sample_kernel); flush = warp reduction segmented by component, one shared atomic per statistic and warp
struct GmmAcc {
    int g_acc = -1;
    unsigned nf = 0, nc = 0;
    double acc[9] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
    __device__ __forceinline__ void add(int g, bool hit, float d0, float d1, float d2) {
        g_acc = g;
        if (hit) {
            ++nc;
        } else {
            ++nf;
            const double e0 = d0, e1 = d1, e2 = d2;
            acc[0] += e0; acc[1] += e1; acc[2] += e2;
            acc[3] = fma(e0, e0, acc[3]); acc[4] = fma(e0, e1, acc[4]); acc[5] = fma(e0, e2, acc[5]);
            acc[6] = fma(e1, e1, acc[6]); acc[7] = fma(e1, e2, acc[7]); acc[8] = fma(e2, e2, acc[8]);
        }
    }
    __device__ __forceinline__ void flush(double (*s_stat)[GMM_STAT]) {
        const int gmin = __reduce_min_sync(0xffffffffu, g_acc >= 0 ? g_acc : 0x7fffffff);
        const int gmax = __reduce_max_sync(0xffffffffu, g_acc);
        for (int gg = gmin; gg <= gmax; ++gg) {  // warp-uniform
            const bool sel = g_acc == gg;
            const unsigned wf = __reduce_add_sync(0xffffffffu, sel ? nf : 0u);
            const unsigned wc = __reduce_add_sync(0xffffffffu, sel ? nc : 0u);
            if (wf) {
                double v[9];
#pragma unroll
                for (int k = 0; k < 9; ++k) {
                    v[k] = sel ? acc[k] : 0.0;
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) v[k] += __shfl_xor_sync(0xffffffffu, v[k], o);
                }
                if ((threadIdx.x & 31) == 0) {
                    atomicAdd(&s_stat[gg][0], double(wf));
#pragma unroll
                    for (int k = 0; k < 9; ++k) atomicAdd(&s_stat[gg][2 + k], v[k]);
                }
            }
            if (wc && (threadIdx.x & 31) == 0) atomicAdd(&s_stat[gg][1], double(wc));
        }
        g_acc = -1;
        nf = nc = 0;
#pragma unroll
        for (int k = 0; k < 9; ++k) acc[k] = 0.0;
    }
};

// Two samples of one thread, i0 and i1, in the two lanes of packed FP32 pairs (f32x2.cuh): Philox block -> 3 normals ->
// x = mu_g + L_g z -> packed SAT (mc_philox2_sm100.cuh: same arithmetic for listed and exhaustive tests).  L / mu: the
// components' Cholesky factors and means, lane 0 for i0 and lane 1 for i1 (the same value twice unless the pair sits on a
// component boundary).  Every thread of the warp calls it.  tab / ccnt / cidx: global or shared memory.
struct GmmDraw2 {
    f2 d0, d1, d2;   // L z of both samples (the moment sums are taken about the component mean)
    bool hit0, hit1;
};
template <class V, bool AX = false>   // V = f2: per-lane components; V = float: one component for both lanes (broadcast at use); AX: axis-aligned scene
__device__ __forceinline__ GmmDraw2 gmm_draw_test2(unsigned stream_id, unsigned ctr2, unsigned k0, unsigned k1, unsigned i0, unsigned i1, bool valid0, bool valid1,
                                                   const V (&L)[6], const V (&mu)[3], const float4* tab, const unsigned char* ccnt, const ushort2* cidx,
                                                   bool cull, float2 nom, unsigned& my_tests) {
    // ONE block for the thread's two samples (philox.h, "GMM sample draws"): lane 0 = sample i0, lane 1 = i1 = i0 + 1 with i0
    // even -- except in a launch that spreads one sample per thread (i1 invalid), where an odd i0 takes the odd sample's share
    const unsigned odd = i0 & 1u;
    const mcsim_rng::u32x4 c = mcsim_rng::philox4x32_10(mcsim_rng::u32x4{i0 >> 1, stream_id, ctr2, unsigned(mcsim_rng::PURPOSE_GMM_SAMPLE)}, k0, k1);
    f2 z0, z1, z2;
    box_muller2_bits(odd ? c.y >> 8 : c.x >> 8, half16f(c.z, int(odd)), c.y >> 8, half16f(c.z, 1), z0, z1);
    {
        const float rc = mufu_sqrt(mufu_lg2(fmaf(float(c.w >> 8), MCSIM_U24_SCALE, MCSIM_U24_HALF)) * -1.3862943611198906f);
        const float tc = fmaf(float(mcsim_rng::gmm_pair_c_angle(c)), MCSIM_A16_SCALE, MCSIM_A16_HALF);
        const float zs = rc * mufu_sin(tc);
        z2 = mk2(odd ? zs : rc * mufu_cos(tc), zs);
    }
    GmmDraw2 o;
    o.d0 = mul2(z0, L[0]);
    o.d1 = fma2(z1, L[2], mul2(z0, L[1]));
    o.d2 = fma2(z2, L[5], fma2(z1, L[4], mul2(z0, L[3])));
    const f2 x = add2(o.d0, mu[0]), y = add2(o.d1, mu[1]), th = add2(o.d2, mu[2]);
    f2 sn, cs;
    sincos2(th, sn, cs);   // |theta| < 3 pi: mu_theta in [0, 2 pi] + O(sigma)
    // warp-uniform choice of the candidate list by the warp's largest deviation from x*_t (exact: sat_device.cuh)
    int lev = MC_CULL_LEVELS;
    if (cull) {
        const f2 ex = add2(x, -nom.x), ey = add2(y, -nom.y);
        const f2 e2 = fma2(ex, ex, mul2(ey, ey));
        const unsigned u0 = valid0 ? __float_as_uint(lo(e2)) : 0u, u1 = valid1 ? __float_as_uint(hi(e2)) : 0u;
        lev = cull_level(__uint_as_float(__reduce_max_sync(0xffffffffu, u0 > u1 ? u0 : u1)));
        if (lev < MC_CULL_LEVELS && ccnt[lev] == MC_CULL_OVERFLOW) lev = MC_CULL_LEVELS;
    }
    unsigned tested = 0;
    o.hit0 = o.hit1 = false;
    if (lev < MC_CULL_LEVELS) {
        for (int hI = 0; hI < c_scene.n_hulls; ++hI) {
            const int nc = ccnt[hI * MC_CULL_LEVELS + lev];
            if (nc == 0) continue;
            const HullFrame2 h = hull_frame2<AX>(c_scene.hulls[hI], x, y, cs, sn);
            if (AX) {   // one class, and its frame is the hull's (mc_philox2_sm100.cuh, collide2)
                const ClassFrame2 cf = class_frame2<true>(make_float4(1.f, 0.f, 0.f, 0.f), h.ch, h.sh, h.hcx, h.hcy, h.hhx, h.hhy);
                for (int k = 0; k < nc; ++k) {
                    const unsigned ox = cidx[hI * MC_CULL_CAP + k].x;
                    class_pair_test2(cf, h.ch, h.sh, h.nph, h.nqh, h.hhx, h.hhy, tab[ox], *reinterpret_cast<const float2*>(tab + ox + 1), o.hit0, o.hit1);
                }
            } else {
                ClassFrame2 cf;
                unsigned cls = 0xffffffffu;
                for (int k = 0; k < nc; ++k) {
                    const ushort2 oo = cidx[hI * MC_CULL_CAP + k];  // x: pair record, y: class header
                    if (oo.y != cls) {
                        cls = oo.y;
                        cf = class_frame2<false>(tab[oo.y], h.ch, h.sh, h.hcx, h.hcy, h.hhx, h.hhy);
                    }
                    class_pair_test2(cf, h.ch, h.sh, h.nph, h.nqh, h.hhx, h.hhy, tab[oo.x], *reinterpret_cast<const float2*>(tab + oo.x + 1), o.hit0, o.hit1);
                }
            }
            tested += nc;
        }
    } else {
        robot_hits2<AX>(tab, x, y, cs, sn, o.hit0, o.hit1);
        tested = c_scene.n_pairs;
    }
    my_tests += tested * (unsigned(valid0) + unsigned(valid1));
    return o;
}

// one pass of the generic sample loop (per-waypoint and cluster kernels): the thread's two samples may belong to any
// component (per-thread monotone search in the prefix sums); per-thread sums are flushed when the component changes
template <bool AX = false>
__device__ __forceinline__ void gmm_philox_pass2(unsigned stream_id, unsigned ctr2, unsigned k0, unsigned k1, unsigned i0, unsigned i1, unsigned n, int K, int& g,
                                                 const unsigned* s_prefix, const float (*s_mu)[3], const float (*s_L)[6], double (*s_stat)[GMM_STAT],
                                                 const float4* tab, const unsigned char* ccnt, const ushort2* cidx, bool cull, float2 nom,
                                                 GmmAcc& acc, unsigned& my_tests) {
    const bool valid0 = i0 < n, valid1 = i1 < n;   // i1 = i0 + 1, or >= n when the launch spreads one sample per thread
    while (valid0 && g < K - 1 && i0 >= s_prefix[g + 1]) ++g;  // i0 only grows: the component index is monotone
    int g1 = g;
    while (valid1 && g1 < K - 1 && i1 >= s_prefix[g1 + 1]) ++g1;
    // component records per lane (the two samples share the component except across a boundary)
    float La[6], Lb[6], ma[3], mb[3];
#pragma unroll
    for (int k = 0; k < 6; ++k) La[k] = Lb[k] = s_L[g][k];
#pragma unroll
    for (int k = 0; k < 3; ++k) ma[k] = mb[k] = s_mu[g][k];
    if (g1 != g) {
#pragma unroll
        for (int k = 0; k < 6; ++k) Lb[k] = s_L[g1][k];
#pragma unroll
        for (int k = 0; k < 3; ++k) mb[k] = s_mu[g1][k];
    }
    f2 L[6], mu[3];
#pragma unroll
    for (int k = 0; k < 6; ++k) L[k] = mk2(La[k], Lb[k]);
#pragma unroll
    for (int k = 0; k < 3; ++k) mu[k] = mk2(ma[k], mb[k]);
    const GmmDraw2 o = gmm_draw_test2<f2, AX>(stream_id, ctr2, k0, k1, i0, i1, valid0, valid1, L, mu, tab, ccnt, cidx, cull, nom, my_tests);
    if (__any_sync(0xffffffffu, valid0 && acc.g_acc >= 0 && g != acc.g_acc)) acc.flush(s_stat);
    if (valid0) acc.add(g, o.hit0, lo(o.d0), lo(o.d1), lo(o.d2));
    if (__any_sync(0xffffffffu, valid1 && acc.g_acc >= 0 && g1 != acc.g_acc)) acc.flush(s_stat);
    if (valid1) acc.add(g1, o.hit1, hi(o.d0), hi(o.d1), hi(o.d2));
}

// ---- the one-CTA-per-run kernel's sample loop: warp-uniform components -------------------------------------------------
// A warp's window of a pass is 64 consecutive sample indices, and samples are grouped by component (prefix sums), so all
// but at most K - 1 windows per warp and waypoint lie inside ONE component.  Those "pure" windows run as tight loops per
// component: the component's factor and mean sit in registers, there is no per-thread component search, no vote on
// "has anybody's component changed", and the FP64 sums are reduced once at the end of the loop (the generic loop paid
// ~20 % of its instructions for that bookkeeping: register moves at the flush merge points, two votes and the search
// per pass).  Draws, FP32 arithmetic and hit flags are those of gmm_philox_pass2 (same counters per sample index); only
// the order of the FP64 additions differs.
// warp sums of the selected lanes' statistics -> one shared atomic per statistic (component gg)
__device__ __forceinline__ void gmm_warp_flush(double (*s_stat)[GMM_STAT], int gg, bool sel, unsigned nf, unsigned nc, const double (&acc)[9]) {
    const unsigned wf = __reduce_add_sync(0xffffffffu, sel ? nf : 0u);
    const unsigned wc = __reduce_add_sync(0xffffffffu, sel ? nc : 0u);
    if (wf) {
        double v[9];
#pragma unroll
        for (int k = 0; k < 9; ++k) {
            v[k] = sel ? acc[k] : 0.0;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) v[k] += __shfl_xor_sync(0xffffffffu, v[k], o);
        }
        if ((threadIdx.x & 31) == 0) {
            atomicAdd(&s_stat[gg][0], double(wf));
#pragma unroll
            for (int k = 0; k < 9; ++k) atomicAdd(&s_stat[gg][2 + k], v[k]);
        }
    }
    if (wc && (threadIdx.x & 31) == 0) atomicAdd(&s_stat[gg][1], double(wc));
}
__device__ __forceinline__ void gmm_acc_add(bool valid, bool hit, float d0, float d1, float d2, unsigned& nf, unsigned& nc, double (&acc)[9]) {
    if (valid) {
        if (hit) {
            ++nc;
        } else {
            ++nf;
            const double e0 = d0, e1 = d1, e2 = d2;
            acc[0] += e0; acc[1] += e1; acc[2] += e2;
            acc[3] = fma(e0, e0, acc[3]); acc[4] = fma(e0, e1, acc[4]); acc[5] = fma(e0, e2, acc[5]);
            acc[6] = fma(e1, e1, acc[6]); acc[7] = fma(e1, e2, acc[7]); acc[8] = fma(e2, e2, acc[8]);
        }
    }
}
// all samples of a waypoint that fall to one warp: its 64-sample windows start at wbase, wbase + stride, ... (one CTA per
// run: wbase = 64 warp, stride = 2 blockDim; a cluster per run: the CTAs' passes interleave, gmm_cluster_run_kernel)
template <bool AX = false>
__device__ __forceinline__ void gmm_member_samples(unsigned stream_id, unsigned ctr2, unsigned k0, unsigned k1, unsigned n, int K,
                                                   const unsigned* s_prefix, const float (*s_mu)[3], const float (*s_L)[6], double (*s_stat)[GMM_STAT],
                                                   const float4* tab, const unsigned char* ccnt, const ushort2* cidx, bool cull, float2 nom, unsigned& my_tests,
                                                   unsigned wbase, unsigned stride) {
    const unsigned lane2 = 2u * (threadIdx.x & 31u);
    int gw = 0;   // component of the window's first sample (warp-uniform)
    for (unsigned w0 = wbase; w0 < n;) {
        while (gw < K - 1 && w0 >= s_prefix[gw + 1]) ++gw;
        const unsigned end_g = gw == K - 1 ? n : (s_prefix[gw + 1] < n ? s_prefix[gw + 1] : n);   // first index beyond the component
        const unsigned w_end = w0 + 64u < n ? w0 + 64u : n;                                      // first index beyond the window
        if (w_end <= end_g) {
            // ---- pure windows: this one and every following one of the warp that still ends inside the component
            const unsigned q_n = end_g == n ? (n - 1u - w0) / stride + 1u : (end_g - 64u - w0) / stride + 1u;
            float L[6], mu[3];
#pragma unroll
            for (int k = 0; k < 6; ++k) L[k] = s_L[gw][k];
#pragma unroll
            for (int k = 0; k < 3; ++k) mu[k] = s_mu[gw][k];
            unsigned nf = 0, nc = 0;
            double acc[9] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
            for (unsigned q = 0; q < q_n; ++q, w0 += stride) {
                const unsigned i0 = w0 + lane2, i1 = i0 + 1u;
                const bool valid0 = i0 < n, valid1 = i1 < n;
                const GmmDraw2 o = gmm_draw_test2<float, AX>(stream_id, ctr2, k0, k1, i0, i1, valid0, valid1, L, mu, tab, ccnt, cidx, cull, nom, my_tests);
                gmm_acc_add(valid0, o.hit0, lo(o.d0), lo(o.d1), lo(o.d2), nf, nc, acc);
                gmm_acc_add(valid1, o.hit1, hi(o.d0), hi(o.d1), hi(o.d2), nf, nc, acc);
            }
            gmm_warp_flush(s_stat, gw, true, nf, nc, acc);
        } else {
            // ---- a window across one or more component boundaries: per-lane components, statistics reduced per component
            const unsigned i0 = w0 + lane2, i1 = i0 + 1u;
            const bool valid0 = i0 < n, valid1 = i1 < n;
            int g0 = gw;
            while (valid0 && g0 < K - 1 && i0 >= s_prefix[g0 + 1]) ++g0;
            int g1 = g0;
            while (valid1 && g1 < K - 1 && i1 >= s_prefix[g1 + 1]) ++g1;
            f2 L[6], mu[3];
#pragma unroll
            for (int k = 0; k < 6; ++k) L[k] = mk2(s_L[g0][k], s_L[g1][k]);
#pragma unroll
            for (int k = 0; k < 3; ++k) mu[k] = mk2(s_mu[g0][k], s_mu[g1][k]);
            const GmmDraw2 o = gmm_draw_test2<f2, AX>(stream_id, ctr2, k0, k1, i0, i1, valid0, valid1, L, mu, tab, ccnt, cidx, cull, nom, my_tests);
            const int g_last = __reduce_max_sync(0xffffffffu, valid1 ? g1 : (valid0 ? g0 : gw));
            for (int gg = gw; gg <= g_last; ++gg) {   // warp-uniform
                unsigned nf = 0, nc = 0;
                double acc[9] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
                gmm_acc_add(valid0 && g0 == gg, o.hit0, lo(o.d0), lo(o.d1), lo(o.d2), nf, nc, acc);
                gmm_acc_add(valid1 && g1 == gg, o.hit1, hi(o.d0), hi(o.d1), hi(o.d2), nf, nc, acc);
                gmm_warp_flush(s_stat, gg, true, nf, nc, acc);
            }
            w0 += stride;
        }
    }
}

template <bool AX>
__global__ void __launch_bounds__(256) gmm_sample_philox_kernel(const GmmLaunch a_in) {
    const GmmLaunch& a = a_in;
    const GmmMember m = gmm_member(a_in, blockIdx.y);
    __shared__ float s_mu[MCSIM_MAX_GAUSSIANS][3];
    __shared__ float s_L[MCSIM_MAX_GAUSSIANS][6];
    __shared__ double s_stat[MCSIM_MAX_GAUSSIANS][GMM_STAT];
    __shared__ unsigned s_prefix[MCSIM_MAX_GAUSSIANS + 1];
    __shared__ unsigned char s_ccnt[MCSIM_MAX_HULLS * MC_CULL_LEVELS];
    __shared__ ushort2 s_cidx[MCSIM_MAX_HULLS * MC_CULL_CAP];
    const int K = a.K;
    if (m.cull) {
        const int nh = c_scene.n_hulls;
        const unsigned char* gc = reinterpret_cast<const unsigned char*>(m.cull) + size_t(a.t) * nh * MC_CULL_LEVELS;
        const ushort2* gi = reinterpret_cast<const ushort2*>(m.cull + (size_t(a.T) * nh * MC_CULL_LEVELS + 3) / 4) + size_t(a.t) * nh * MC_CULL_CAP;
        for (int i = threadIdx.x; i < nh * MC_CULL_LEVELS; i += blockDim.x) s_ccnt[i] = gc[i];
        for (int i = threadIdx.x; i < nh * MC_CULL_CAP; i += blockDim.x) s_cidx[i] = gi[i];
    }
    for (int i = threadIdx.x; i < K * 3; i += blockDim.x) s_mu[i / 3][i % 3] = float(m.comps[(i / 3) * GMM_COMP + 1 + i % 3]);
    for (int i = threadIdx.x; i < K * 6; i += blockDim.x) s_L[i / 6][i % 6] = float(m.comps[(i / 6) * GMM_COMP + 13 + i % 6]);
    for (int i = threadIdx.x; i < K * GMM_STAT; i += blockDim.x) s_stat[i / GMM_STAT][i % GMM_STAT] = 0.0;
    if (threadIdx.x == 0) {
        unsigned acc = 0;
        for (int g = 0; g < K; ++g) { s_prefix[g] = acc; acc += m.hist[g]; }
        s_prefix[K] = acc;
    }
    __syncthreads();

    int g = 0;
    unsigned my_tests = 0;
    GmmAcc acc;
    const unsigned n = unsigned(s_prefix[K] < a.n ? s_prefix[K] : a.n);  // Philox mode: n <= 2^31 (checked on the host)
    const unsigned stride = gridDim.x * blockDim.x;
    const float2 nom = m.cull ? __ldg(m.nominal + a.t) : make_float2(0.f, 0.f);
    // two samples per thread and pass; a.spread (small sample counts, one run): one per thread, the second lane idle,
    // so that twice as many warps share the waypoint's latency
    const unsigned per = a.spread ? 1u : 2u;
    for (unsigned base = per * blockIdx.x * blockDim.x; base < n; base += per * stride)
        gmm_philox_pass2<AX>(m.stream_id, unsigned(2 * a.t), a.k0, a.k1, base + per * threadIdx.x, a.spread ? 0xffffffffu : base + per * threadIdx.x + 1, n, K, g, s_prefix, s_mu, s_L, s_stat, a.sat2, s_ccnt, s_cidx, m.cull != nullptr, nom, acc, my_tests);
    acc.flush(s_stat);
    __syncthreads();
    for (int i = threadIdx.x; i < K * GMM_STAT; i += blockDim.x) {
        const double v = s_stat[i / GMM_STAT][i % GMM_STAT];
        if (v != 0.0) atomicAdd(&m.stats[i], v);
    }
    if (a.tests) {
        unsigned long long t64 = my_tests;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) t64 += __shfl_xor_sync(0xffffffffu, t64, o);
        if ((threadIdx.x & 31) == 0 && t64) atomicAdd(a.tests, t64);
    }
}

// lower Cholesky factor of the LOWER triangle of a row-major 3x3 (what LAPACK dpotrf('L') reads)
__host__ __device__ inline bool gmm_chol3(const double* A, double* Lo) {
    if (!(A[0] > 0)) return false;
    const double l00 = sqrt(A[0]);
    const double l10 = A[3] / l00, l20 = A[6] / l00;
    const double d1 = A[4] - l10 * l10;
    if (!(d1 > 0)) return false;
    const double l11 = sqrt(d1);
    const double l21 = (A[7] - l20 * l10) / l11;
    const double d2 = A[8] - l20 * l20 - l21 * l21;
    if (!(d2 > 0)) return false;
    Lo[0] = l00; Lo[1] = l10; Lo[2] = l11; Lo[3] = l20; Lo[4] = l21; Lo[5] = sqrt(d2);
    return true;
}

// ---- refit + propagate: one thread per component ------------------------------------------------------------------
// cp: the component's record (GMM_COMP doubles), st: its statistics; both may live in global or shared memory.
// Part 1: truncated moments of the free samples; returns (n_free, n_coll).
__device__ __forceinline__ void gmm_refit_moments(double* cp, const double* st, int* degenerate, double& nf, double& nc) {
    nf = st[0];
    nc = st[1];
    if (nf >= 2.0) {
        // mean(free) = mu + sum d / n;  cov = (sum d d^T - (sum d)(sum d)^T / n) / (n - 1)   (unbiased)
        const double m0 = st[2] / nf, m1 = st[3] / nf, m2 = st[4] / nf;
        const double inv = 1.0 / (nf - 1.0);
        const double cxx = (st[5] - st[2] * m0) * inv, cxy = (st[6] - st[2] * m1) * inv, cxz = (st[7] - st[2] * m2) * inv;
        const double cyy = (st[8] - st[3] * m1) * inv, cyz = (st[9] - st[3] * m2) * inv, czz = (st[10] - st[4] * m2) * inv;
        cp[1] += m0; cp[2] += m1; cp[3] += m2;
        cp[4] = cxx; cp[5] = cxy; cp[6] = cxz;
        cp[7] = cxy; cp[8] = cyy; cp[9] = cyz;
        cp[10] = cxz; cp[11] = cyz; cp[12] = czz;
    } else {
        // reference: mean of a 3x0 matrix -> exception / undefined behaviour.  Here: moments kept, counted.
        atomicAdd(degenerate, 1);
    }
}
// one component's belief from waypoint t to t+1: EKFpredict (:766-772) + EKFupdate (:804-812) of the reference, or the
// deviation-state filter (estimator_math.h) when the launch asks for it
__device__ __forceinline__ void gmm_filter_step(const GmmLaunch& a, const GmmStepTable& s, const double* mu, const double* S, double* pmu, double* pS) {
    using namespace mcsim_math;
    if (a.do_ekf == 2) {
        lkf_predict(mu, S, s.xs, s.xn, s.us, s.u, s.M, pmu, pS);
        for (int l = 0; l < a.L; ++l) lkf_update_one(a.lx[l], a.ly[l], a.Q, s.xn, pmu, pS, s.z[l]);
    } else if (a.fast_filter) {   // Philox mode: the same update with one division on the serial chain (estimator_math.h)
        ekf_predict(mu, S, s.u, s.M, pmu, pS);
        for (int l = 0; l < a.L; ++l) ekf_update_one_chain1(a.lx[l], a.ly[l], a.Q, pmu, pS, s.z[l]);
    } else {
        ekf_predict(mu, S, s.u, s.M, pmu, pS);
        for (int l = 0; l < a.L; ++l) ekf_update_one(a.lx[l], a.ly[l], a.Q, pmu, pS, s.z[l]);   // landmarks straight from the launch struct
    }
}
// Part 2 (after every component's part 1): weight, p_t, records, EKFpredict + EKFupdate to waypoint t+1, Cholesky.
__device__ __forceinline__ void gmm_refit_finish(const GmmLaunch& a, const GmmMember& m, int t, int g, double* cp, const double* s_free, const double* s_coll,
                                                 const GmmStepTable* staged_step = nullptr) {
    using namespace mcsim_math;
    const int K = a.K;
    double tf = 0, tc = 0;
    for (int k = 0; k < K; ++k) { tf += s_free[k]; tc += s_coll[k]; }
    // normalise(collisionCounts, 1, 1).row(1): divides by the row's 1-norm, or by 1 if that is 0
    cp[0] = s_free[g] / (tf > 0 ? tf : 1.0);
    if (g == 0) m.probabilities[t] = tc / double(a.n);
    if (a.n_out) { a.n_out[g] = (unsigned long long)s_coll[g]; a.n_out[K + g] = (unsigned long long)s_free[g]; }
    if (a.trunc_out) for (int k = 0; k < GMM_REC; ++k) a.trunc_out[g * GMM_REC + k] = cp[k];
    if (t == a.T - 1 && a.rec_post) for (int k = 0; k < GMM_REC; ++k) a.rec_post[g * GMM_REC + k] = cp[k];
    // ---- move the component to waypoint t+1: EKFpredict (:766-772) then EKFupdate (:804-812)
    if (a.do_ekf && t < a.T - 1) {
        const GmmStepTable& s = staged_step ? *staged_step : m.steps[t];
        double pmu[3], pS[9];
        gmm_filter_step(a, s, cp + 1, cp + 4, pmu, pS);
        for (int k = 0; k < 3; ++k) cp[1 + k] = pmu[k];
        for (int k = 0; k < 9; ++k) cp[4 + k] = pS[k];
        if (a.rec_pre) for (int k = 0; k < GMM_REC; ++k) a.rec_pre[(size_t(t + 1) * K + g) * GMM_REC + k] = cp[k];
    }
    if (!gmm_chol3(cp + 4, cp + 13)) {
        // not SPD (the reference falls back to an eigen-decomposition, glue_mvnrnd_meat.hpp:98-132):
        // sample from the diagonal instead and count the event
        atomicAdd(m.degenerate, 1);
        cp[13] = sqrt(fmax(cp[4], 0.0)); cp[14] = 0; cp[15] = sqrt(fmax(cp[8], 0.0));
        cp[16] = 0; cp[17] = 0; cp[18] = sqrt(fmax(cp[12], 0.0));
    }
}

// The mixture, the statistics and the step record are staged in shared memory first: the refit is one long dependent
// FP64 chain per component, and every read-modify-write of the record in global memory was a round trip to L2 on it.
__global__ void __launch_bounds__(32) gmm_refit_kernel(const GmmLaunch a_in) {
    const GmmLaunch& a = a_in;
    const GmmMember m = gmm_member(a_in, blockIdx.x);
    __shared__ double s_free[MCSIM_MAX_GAUSSIANS], s_coll[MCSIM_MAX_GAUSSIANS];
    __shared__ double s_cp[MCSIM_MAX_GAUSSIANS * GMM_COMP], s_st[MCSIM_MAX_GAUSSIANS * GMM_STAT];
    __shared__ GmmStepTable s_step;
    const int g = threadIdx.x;
    const int K = a.K;
    const bool ekf = a.do_ekf && a.t < a.T - 1;
    for (int i = g; i < K * GMM_COMP; i += 32) s_cp[i] = m.comps[i];
    for (int i = g; i < K * GMM_STAT; i += 32) s_st[i] = m.stats[i];
    if (ekf) {
        const double* src = reinterpret_cast<const double*>(m.steps + a.t);
        double* dst = reinterpret_cast<double*>(&s_step);
        for (int i = g; i < int(sizeof(GmmStepTable) / sizeof(double)); i += 32) dst[i] = src[i];
    }
    __syncthreads();
    double* cp = s_cp + g * GMM_COMP;
    if (g < K) gmm_refit_moments(cp, s_st + g * GMM_STAT, m.degenerate, s_free[g], s_coll[g]);
    __syncthreads();
    if (g < K) gmm_refit_finish(a, m, a.t, g, cp, s_free, s_coll, ekf ? &s_step : nullptr);
    __syncthreads();
    for (int i = g; i < K * GMM_COMP; i += 32) m.comps[i] = s_cp[i];
    for (int i = g; i < K * GMM_STAT; i += 32) m.stats[i] = 0.0;
    if (g < K) m.hist[g] = 0;
}

// EM-refit variant: the mixture was refitted elsewhere (EM on the free samples); move it to waypoint t + 1 (EKFpredict +
// EKFupdate), record it, and factor the covariances for the next sampling -- the second half of gmm_refit_finish.
__global__ void __launch_bounds__(32) gmm_propagate_kernel(const GmmLaunch a_in) {
    using namespace mcsim_math;
    const GmmLaunch& a = a_in;
    const GmmMember m = gmm_member(a_in, blockIdx.x);
    const int g = threadIdx.x, K = a.K, t = a.t;
    if (g >= K) return;
    double* cp = m.comps + size_t(g) * GMM_COMP;
    if (t == a.T - 1 && a.rec_post) for (int k = 0; k < GMM_REC; ++k) a.rec_post[g * GMM_REC + k] = cp[k];
    if (a.do_ekf && t < a.T - 1) {
        const GmmStepTable& s = m.steps[t];
        double pmu[3], pS[9];
        gmm_filter_step(a, s, cp + 1, cp + 4, pmu, pS);
        for (int k = 0; k < 3; ++k) cp[1 + k] = pmu[k];
        for (int k = 0; k < 9; ++k) cp[4 + k] = pS[k];
        if (a.rec_pre) for (int k = 0; k < GMM_REC; ++k) a.rec_pre[(size_t(t + 1) * K + g) * GMM_REC + k] = cp[k];
    }
    if (!gmm_chol3(cp + 4, cp + 13)) {
        atomicAdd(m.degenerate, 1);
        cp[13] = sqrt(fmax(cp[4], 0.0)); cp[14] = 0; cp[15] = sqrt(fmax(cp[8], 0.0));
        cp[16] = 0; cp[17] = 0; cp[18] = sqrt(fmax(cp[12], 0.0));
    }
}

// ---- method variant: analytic linear-constraint (Patil) truncation instead of sampling --------
// The same belief propagation as the GMM branch of EKF_GaussProp (:672-678, :766-772, :804-817, :843-856) with
// truncateGMM replaced by the closed form of linear_truncation.h: at every waypoint each component is truncated against
// the half-planes A x <= b (the constraint list of gaussprop.py:58-73, windows expanded into half-planes on the host),
// its weight is multiplied by its collision-free mass, p_t = the mixture's lost mass, then EKFpredict + EKFupdate move
// it to the next waypoint.  No samples, no collision geometry: one CTA walks the whole plan (one thread per component,
// FP64), one launch per run; blockIdx.x = batch member.  Parity unpinned (no reference implementation).
__global__ void __launch_bounds__(32) gmm_linear_run_kernel(const GmmLaunch a_in, const double* __restrict__ A, const double* __restrict__ b, int n_con) {
    using namespace mcsim_math;
    const GmmLaunch& a = a_in;
    const GmmMember m = gmm_member(a_in, blockIdx.x);
    __shared__ double s_free[MCSIM_MAX_GAUSSIANS], s_coll[MCSIM_MAX_GAUSSIANS];
    __shared__ double s_cp[MCSIM_MAX_GAUSSIANS][GMM_COMP];
    const int g = threadIdx.x, K = a.K, T = a.T;
    for (int i = g; i < K * GMM_COMP; i += 32) s_cp[i / GMM_COMP][i % GMM_COMP] = m.comps[i];
    __syncthreads();
    for (int t = 0; t < T; ++t) {
        if (g < K) {
            double* cp = s_cp[g];
            if (a.rec_pre) for (int k = 0; k < GMM_REC; ++k) a.rec_pre[(size_t(t) * K + g) * GMM_REC + k] = cp[k];
            double mo[3], co[9], fm = 1.0;
            const int rc = n_con ? mcsim_lt::truncate(cp + 1, cp + 4, n_con, A, b, mo, co, nullptr, &fm) : 0;
            if (rc == 0 && n_con) {
                for (int k = 0; k < 3; ++k) cp[1 + k] = mo[k];
                for (int k = 0; k < 9; ++k) cp[4 + k] = co[k];
            } else if (rc != 0) {       // no feasible mass (the whole component collides) or a degenerate covariance
                atomicAdd(m.degenerate, 1);
                fm = rc == 5 ? 0.0 : 1.0;
            }
            s_free[g] = cp[0] * fm;
            s_coll[g] = cp[0] * (1.0 - fm);
        }
        __syncthreads();
        if (g < K) {
            double* cp = s_cp[g];
            double tf = 0, tc = 0;
            for (int k = 0; k < K; ++k) { tf += s_free[k]; tc += s_coll[k]; }
            cp[0] = s_free[g] / (tf > 0 ? tf : 1.0);
            if (g == 0) m.probabilities[t] = (tf + tc) > 0 ? tc / (tf + tc) : 0.0;
            if (t == T - 1 && a.rec_post) for (int k = 0; k < GMM_REC; ++k) a.rec_post[g * GMM_REC + k] = cp[k];
            if (t < T - 1) {
                const GmmStepTable& s = m.steps[t];
                double pmu[3], pS[9];
                gmm_filter_step(a, s, cp + 1, cp + 4, pmu, pS);
                for (int k = 0; k < 3; ++k) cp[1 + k] = pmu[k];
                for (int k = 0; k < 9; ++k) cp[4 + k] = pS[k];
            }
        }
        __syncthreads();
    }
}

// ---- batch: ONE CTA runs ONE member's whole runGMMEstimation (all T waypoints) ---------------------------------------
// The three kernels per waypoint above keep a single run's 10^4 samples spread over the GPU; a batch fills the GPU
// with members instead, and then a member's mixture, statistics, counts, the pair table and the plan's candidate
// lists can live in shared memory for the whole run: no global atomics, no per-waypoint prologue, one launch instead
// of 3 T.  Same draws, same component assignment, same refit code as the per-waypoint kernels (member b ==
// mcsim_cuda_gmm_run on Philox subsequence stream0 + b up to the order of the FP64 sums).
// Dynamic shared memory: pair table (a.sat2, sat2_n float4) | candidate lists (cull_words 32-bit words).
// cumulative distribution of the component draw from the weights (thread-serial; gmm_count_kernel's rule: entries behind
// the last live component are 2 = never exceeded)
__device__ __forceinline__ void gmm_member_cdf(const double* w, int K, float* s_cdf) {
    double acc = 0, tot = 0;
    for (int g = 0; g < K; ++g) tot += w[g];
    for (int g = 0; g < K; ++g) { acc += w[g]; s_cdf[g] = float(acc / tot); }
    for (int g = K - 1; g >= 0 && !(w[g] > 0); --g) s_cdf[g] = 2.0f;
    int last = K - 1;
    while (last > 0 && !(w[last] > 0)) --last;
    s_cdf[last] = 2.0f;
}
// multinomial counts of waypoint t (gmm_count_kernel's draws: Philox block j -> 4 uniforms -> components of samples
// 4 j .. 4 j + 3), by `n_threads` threads (whole warps) numbered `idx`; sums land in s_hist (zeroed by the caller)
__device__ __forceinline__ void gmm_member_count(int t, unsigned idx, unsigned n_threads, int K, unsigned n_all, unsigned stream_id, unsigned k0, unsigned k1,
                                                 const float* s_cdf, unsigned* s_hist) {
    // per-thread counts in four registers of four 8-bit fields (component g: field g % 4 of register g / 4),
    // flushed by warp sums before a field can overflow and after the last pass
    const unsigned n4 = (n_all + 3) >> 2;
    unsigned cnt[4] = {0u, 0u, 0u, 0u};
    unsigned held = 0;
    auto flush_counts = [&]() {
        for (int gg = 0; gg < K; ++gg) {
            const unsigned r = gg < 4 ? cnt[0] : (gg < 8 ? cnt[1] : (gg < 12 ? cnt[2] : cnt[3]));
            const unsigned v = __reduce_add_sync(0xffffffffu, (r >> (8 * (gg & 3))) & 0xffu);
            if (v && (threadIdx.x & 31) == 0) atomicAdd(&s_hist[gg], v);
        }
        cnt[0] = cnt[1] = cnt[2] = cnt[3] = 0u;
        held = 0;
    };
    const float cdf0 = K > 1 ? s_cdf[0] : 2.0f, cdf1 = K > 2 ? s_cdf[1] : 2.0f, cdf2 = K > 3 ? s_cdf[2] : 2.0f;
    for (unsigned base = 0; base < n4; base += n_threads) {
        const unsigned j = base + idx;
        mcsim_rng::u32x4 c = {j, stream_id, unsigned(2 * t + 1), unsigned(mcsim_rng::PURPOSE_GMM_SAMPLE)};
        c = mcsim_rng::philox4x32_10(c, k0, k1);
        const unsigned w[4] = {c.x, c.y, c.z, c.w};
        if (K <= 4) {
            // the usual mixtures (K = 1 .. 4, the reference runs 1 .. 3): the three thresholds in registers (2 = never
            // exceeded, like the entries behind the last live component), one counter register, no loop over k
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const bool valid = j < n4 && 4 * j + q < n_all;   // n <= 2^31: no overflow
                const float u = mcsim_rng::u01(w[q]);
                const int g = int(u > cdf0) + int(u > cdf1) + int(u > cdf2);
                cnt[0] += valid ? 1u << (8 * g) : 0u;
            }
        } else {
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const bool valid = j < n4 && 4 * j + q < n_all;   // n <= 2^31: no overflow
                const float u = mcsim_rng::u01(w[q]);
                int g = 0;
                for (int k = 0; k < K - 1; ++k) g += u > s_cdf[k];   // the cdf is non-decreasing: first k with u <= cdf[k]
                const unsigned inc = valid ? 1u << (8 * (g & 3)) : 0u;
                cnt[0] += g < 4 ? inc : 0u;
                cnt[1] += (g >= 4 && g < 8) ? inc : 0u;
                cnt[2] += (g >= 8 && g < 12) ? inc : 0u;
                cnt[3] += g >= 12 ? inc : 0u;
            }
        }
        held += 4;
        if (held > 248) flush_counts();   // warp-uniform
    }
    flush_counts();
}

// Register budget: 128 per thread (no spills: the draw + SAT + 9 FP64 sums need ~110), i.e. 512 resident threads per SM.
// How they are cut into CTAs (= runs in flight per SM) is the launcher's choice -- 128, 256 or 512 threads per run
// (gmm_member_threads in mcsim_cuda.cu): the kernel only assumes whole warps and at least two of them.
#define GMM_MEMBER_MAX_THREADS 512
template <bool AX>   // AX: axis-aligned scene (mc_philox2_sm100.cuh, class_frame2<AX>)
__global__ void __launch_bounds__(GMM_MEMBER_MAX_THREADS, 1) gmm_member_run_kernel(const GmmLaunch a_in, int sat2_n, int cull_words) {
    extern __shared__ __align__(16) float4 s_tab[];
    __shared__ double s_comp[MCSIM_MAX_GAUSSIANS][GMM_COMP];
    __shared__ double s_stat[MCSIM_MAX_GAUSSIANS][GMM_STAT];
    __shared__ double s_free[MCSIM_MAX_GAUSSIANS], s_coll[MCSIM_MAX_GAUSSIANS];
    __shared__ float s_mu[MCSIM_MAX_GAUSSIANS][3];
    __shared__ float s_L[MCSIM_MAX_GAUSSIANS][6];
    __shared__ float s_cdf[MCSIM_MAX_GAUSSIANS];
    __shared__ unsigned s_hist[MCSIM_MAX_GAUSSIANS];
    __shared__ unsigned s_prefix[MCSIM_MAX_GAUSSIANS + 1];
    __shared__ double s_wt[MCSIM_MAX_GAUSSIANS];
    const GmmLaunch& a = a_in;
    const GmmMember m = gmm_member(a_in, blockIdx.x);
    const int K = a.K, T = a.T, nh = c_scene.n_hulls;
    const unsigned stream_id = m.stream_id, k0 = a.k0, k1 = a.k1, n_all = unsigned(a.n);
    const bool cull = m.cull != nullptr;
    unsigned* s_cull = reinterpret_cast<unsigned*>(s_tab + sat2_n);
    for (int i = threadIdx.x; i < sat2_n; i += blockDim.x) s_tab[i] = a.sat2[i];
    if (cull) for (int i = threadIdx.x; i < cull_words; i += blockDim.x) s_cull[i] = m.cull[i];
    for (int i = threadIdx.x; i < K * GMM_COMP; i += blockDim.x) s_comp[i / GMM_COMP][i % GMM_COMP] = m.comps[i];
    for (int i = threadIdx.x; i < K * GMM_STAT; i += blockDim.x) s_stat[i / GMM_STAT][i % GMM_STAT] = 0.0;
    if (threadIdx.x < MCSIM_MAX_GAUSSIANS) s_hist[threadIdx.x] = 0;
    const unsigned char* s_ccnt = reinterpret_cast<const unsigned char*>(s_cull);
    const ushort2* s_cidx = reinterpret_cast<const ushort2*>(s_cull + (T * nh * MC_CULL_LEVELS + 3) / 4);
    unsigned my_tests = 0;
    __syncthreads();

    // The multinomial counts of waypoint t + 1 need the new WEIGHTS only, and those are known once every component's
    // moments are (free counts); the rest of the refit -- EKF predict + 8 updates + Cholesky, ~6 us of dependent FP64 on
    // K threads of warp 0 -- used to hold the other seven warps at a barrier (10 % of all stall samples, and with it a
    // third of the SM's resident warps idle).  Now warps 1..7 draw the next waypoint's counts meanwhile.
    if (threadIdx.x == 0) {
        for (int g = 0; g < K; ++g) s_wt[g] = s_comp[g][0];
        gmm_member_cdf(s_wt, K, s_cdf);
    }
    __syncthreads();
    gmm_member_count(0, threadIdx.x, blockDim.x, K, n_all, stream_id, k0, k1, s_cdf, s_hist);
    __syncthreads();
    for (int t = 0; t < T; ++t) {
        // ---- the counts of this waypoint are in s_hist; the mixture of this waypoint is final in s_comp
        if (threadIdx.x == 0) {
            unsigned acc = 0;
            for (int g = 0; g < K; ++g) { s_prefix[g] = acc; acc += s_hist[g]; s_hist[g] = 0; }
            s_prefix[K] = acc;
        }
        for (int i = threadIdx.x; i < K * 3; i += blockDim.x) s_mu[i / 3][i % 3] = float(s_comp[i / 3][1 + i % 3]);
        for (int i = threadIdx.x; i < K * 6; i += blockDim.x) s_L[i / 6][i % 6] = float(s_comp[i / 6][13 + i % 6]);
        __syncthreads();
        // ---- sample -> collide -> moments (the draws of gmm_sample_philox_kernel; loop structure: gmm_member_samples)
        {
            const unsigned n = s_prefix[K] < n_all ? s_prefix[K] : n_all;
            const float2 nom = cull ? __ldg(m.nominal + t) : make_float2(0.f, 0.f);
            const unsigned char* cc = s_ccnt + t * nh * MC_CULL_LEVELS;
            const ushort2* ci = s_cidx + t * nh * MC_CULL_CAP;
            gmm_member_samples<AX>(stream_id, unsigned(2 * t), k0, k1, n, K, s_prefix, s_mu, s_L, s_stat, s_tab, cc, ci, cull, nom, my_tests,
                               64u * (threadIdx.x >> 5), 2u * blockDim.x);
        }
        __syncthreads();
        // ---- refit + propagate (gmm_refit_kernel)
        const int g = threadIdx.x;
        if (g < K) gmm_refit_moments(s_comp[g], s_stat[g], m.degenerate, s_free[g], s_coll[g]);
        __syncthreads();
        if (threadIdx.x < 32) {
            if (g < K) {
                gmm_refit_finish(a, m, t, g, s_comp[g], s_free, s_coll);
                for (int k = 0; k < GMM_STAT; ++k) s_stat[g][k] = 0.0;
            }
        } else if (t < T - 1) {
            if (threadIdx.x == 32) {   // the weights gmm_refit_finish is writing into s_comp[.][0] right now, from the same operands
                double tf = 0;
                for (int k = 0; k < K; ++k) tf += s_free[k];
                for (int k = 0; k < K; ++k) s_wt[k] = s_free[k] / (tf > 0 ? tf : 1.0);
                gmm_member_cdf(s_wt, K, s_cdf);
            }
            asm volatile("bar.sync 1, %0;" ::"r"(int(blockDim.x) - 32) : "memory");   // warps 1 .. 7 only
            gmm_member_count(t + 1, threadIdx.x - 32, blockDim.x - 32, K, n_all, stream_id, k0, k1, s_cdf, s_hist);
        }
        __syncthreads();
    }
    if (a.tests) {
        unsigned long long t64 = my_tests;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) t64 += __shfl_xor_sync(0xffffffffu, t64, o);
        if ((threadIdx.x & 31) == 0 && t64) atomicAdd(a.tests, t64);
    }
}

// ---- few runs (one runGMMEstimation at a time, small batches): ONE THREAD-BLOCK CLUSTER runs one whole run -----------
// A single run's 10^4 samples per waypoint are one pass of a cluster of 16 x 512 threads (two samples per thread); what
// the per-waypoint kernels pay for instead is three dependent launches per waypoint (6 + 9 + 11 us of mostly launch and
// tail latency, 168 launches per run).  Here the CTAs of a cluster split the samples of a waypoint (same Philox
// counters: sample index -> draws, so the samples are the per-waypoint kernels' own), and exchange their partial
// multinomial counts and sufficient statistics through distributed shared memory: every CTA stores its partials into
// slot [its rank] of EVERY peer (remote stores, nobody waits for them), a cluster barrier makes them visible, and every
// CTA adds up the slots in rank order -- bit-identical totals everywhere, so each CTA runs the (serial, FP64) refit +
// EKF step redundantly on its own copy of the mixture and nothing has to be sent back.  A slot is rewritten only after
// the next barrier, which every CTA reaches after it has read the slot: two cluster barriers per waypoint, no double
// buffers.  The EKF chain of the refit occupies one warp for ~ 5 us; the other warps draw the NEXT waypoint's
// multinomial counts meanwhile (they need the new weights only, which the moments give before the EKF step).
// Outputs (p_t, recorded mixtures) are written by every CTA with identical values; the degenerate-refit counter only
// by rank 0.
#define GMM_CLUSTER 8          // CTAs per cluster for small batches; single runs take 16 (non-portable size)
#define GMM_CLUSTER_MAX 16
#define GMM_CLUSTER_THREADS 512
template <bool AX>   // AX: axis-aligned scene (mc_philox2_sm100.cuh, class_frame2<AX>)
__global__ void __launch_bounds__(GMM_CLUSTER_THREADS, 1) gmm_cluster_run_kernel(const GmmLaunch a_in, int sat2_n, int cull_words) {
    namespace cg = cooperative_groups;
    cg::cluster_group cl = cg::this_cluster();
    const unsigned rank = cl.block_rank(), CL = cl.num_blocks();
    extern __shared__ __align__(16) float4 s_tab[];
    __shared__ double s_comp[MCSIM_MAX_GAUSSIANS][GMM_COMP];
    __shared__ double s_stat[MCSIM_MAX_GAUSSIANS][GMM_STAT];                      // this CTA's partial statistics
    __shared__ double s_part[GMM_CLUSTER_MAX][MCSIM_MAX_GAUSSIANS * GMM_STAT];   // slot r: rank r's partials (written remotely)
    __shared__ double s_tot[MCSIM_MAX_GAUSSIANS][GMM_STAT];
    __shared__ double s_free[MCSIM_MAX_GAUSSIANS], s_coll[MCSIM_MAX_GAUSSIANS];
    __shared__ float s_mu[MCSIM_MAX_GAUSSIANS][3];
    __shared__ float s_L[MCSIM_MAX_GAUSSIANS][6];
    __shared__ float s_cdf[MCSIM_MAX_GAUSSIANS];
    __shared__ unsigned s_hist[MCSIM_MAX_GAUSSIANS];                              // this CTA's partial counts
    __shared__ unsigned s_hpart[GMM_CLUSTER_MAX][MCSIM_MAX_GAUSSIANS];           // slot r: rank r's counts (written remotely)
    __shared__ unsigned s_prefix[MCSIM_MAX_GAUSSIANS + 1];
    __shared__ int s_sink;
    const GmmLaunch& a = a_in;
    GmmMember m = gmm_member(a_in, blockIdx.x / CL);
    if (rank) m.degenerate = &s_sink;
    const int K = a.K, T = a.T, nh = c_scene.n_hulls;
    const unsigned stream_id = m.stream_id, k0 = a.k0, k1 = a.k1, n_all = unsigned(a.n);
    const bool cull = m.cull != nullptr;
    unsigned* s_cull = reinterpret_cast<unsigned*>(s_tab + sat2_n);
    for (int i = threadIdx.x; i < sat2_n; i += blockDim.x) s_tab[i] = a.sat2[i];
    if (cull) for (int i = threadIdx.x; i < cull_words; i += blockDim.x) s_cull[i] = m.cull[i];
    for (int i = threadIdx.x; i < K * GMM_COMP; i += blockDim.x) s_comp[i / GMM_COMP][i % GMM_COMP] = m.comps[i];
    for (int i = threadIdx.x; i < MCSIM_MAX_GAUSSIANS * GMM_STAT; i += blockDim.x) (&s_stat[0][0])[i] = 0.0;
    if (threadIdx.x < MCSIM_MAX_GAUSSIANS) s_hist[threadIdx.x] = 0;
    const unsigned char* s_ccnt = reinterpret_cast<const unsigned char*>(s_cull);
    const ushort2* s_cidx = reinterpret_cast<const ushort2*>(s_cull + (T * nh * MC_CULL_LEVELS + 3) / 4);
    unsigned my_tests = 0;
    __syncthreads();
    if (threadIdx.x < K) s_free[threadIdx.x] = s_comp[threadIdx.x][0];   // the initial weights play the free counts' part
    __syncthreads();
    cl.sync();   // every CTA is resident and initialised before anyone stores into it

    // multinomial counts of waypoint t from weights proportional to s_free[] (= the weights the refit of waypoint t - 1
    // stores: n_free,g / sum n_free); run by warps 1.. of the CTA only, warp 0 is busy with the EKF chain meanwhile
    const unsigned cw = blockDim.x - 32;   // counting threads per CTA
    auto count_waypoint = [&](int t) {
        const unsigned tid = threadIdx.x - 32;
        if (tid == 0) {
            // the weights exactly as gmm_refit_finish stores them (t = 0: the initial weights, already in s_free)
            double tf = 0;
            for (int g = 0; g < K; ++g) tf += s_free[g];
            const double den = t == 0 ? 1.0 : (tf > 0 ? tf : 1.0);
            double acc = 0, tot = 0;
            for (int g = 0; g < K; ++g) tot += s_free[g] / den;
            for (int g = 0; g < K; ++g) { acc += s_free[g] / den; s_cdf[g] = float(acc / tot); }
            for (int g = K - 1; g >= 0 && !(s_free[g] > 0); --g) s_cdf[g] = 2.0f;
            int last = K - 1;
            while (last > 0 && !(s_free[last] > 0)) --last;
            s_cdf[last] = 2.0f;
        }
        asm volatile("bar.sync 1, %0;" ::"r"(cw) : "memory");
        // per-thread counts in four registers of four 8-bit fields, flushed by warp sums before a field can overflow
        const unsigned n4 = (n_all + 3) >> 2;
        unsigned cnt[4] = {0u, 0u, 0u, 0u};
        unsigned held = 0;
        auto flush_counts = [&]() {
            for (int gg = 0; gg < K; ++gg) {
                const unsigned r = gg < 4 ? cnt[0] : (gg < 8 ? cnt[1] : (gg < 12 ? cnt[2] : cnt[3]));
                const unsigned v = __reduce_add_sync(0xffffffffu, (r >> (8 * (gg & 3))) & 0xffu);
                if (v && (threadIdx.x & 31) == 0) atomicAdd(&s_hist[gg], v);
            }
            cnt[0] = cnt[1] = cnt[2] = cnt[3] = 0u;
            held = 0;
        };
        for (unsigned base = rank * cw; base < n4; base += CL * cw) {
            const unsigned j = base + tid;
            mcsim_rng::u32x4 c = {j, stream_id, unsigned(2 * t + 1), unsigned(mcsim_rng::PURPOSE_GMM_SAMPLE)};
            c = mcsim_rng::philox4x32_10(c, k0, k1);
            const unsigned w[4] = {c.x, c.y, c.z, c.w};
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const bool valid = j < n4 && 4 * j + q < n_all;
                const float u = mcsim_rng::u01(w[q]);
                int g = 0;
                for (int k = 0; k < K - 1; ++k) g += u > s_cdf[k];
                const unsigned inc = valid ? 1u << (8 * (g & 3)) : 0u;
                cnt[0] += g < 4 ? inc : 0u;
                cnt[1] += (g >= 4 && g < 8) ? inc : 0u;
                cnt[2] += (g >= 8 && g < 12) ? inc : 0u;
                cnt[3] += g >= 12 ? inc : 0u;
            }
            held += 4;
            if (held > 248) flush_counts();
        }
        flush_counts();
    };
    if (threadIdx.x >= 32) count_waypoint(0);

    for (int t = 0; t < T; ++t) {
        __syncthreads();   // counts of waypoint t (warps 1..) and the refit of waypoint t - 1 (warp 0) are done
        // ---- partial counts -> slot [rank] of every CTA
        for (unsigned i = threadIdx.x; i < unsigned(K) * CL; i += blockDim.x) {
            const unsigned g = i / CL, r = i % CL;
            *cl.map_shared_rank(&s_hpart[rank][g], r) = s_hist[g];
        }
        for (int i = threadIdx.x; i < K * 3; i += blockDim.x) s_mu[i / 3][i % 3] = float(s_comp[i / 3][1 + i % 3]);
        for (int i = threadIdx.x; i < K * 6; i += blockDim.x) s_L[i / 6][i % 6] = float(s_comp[i / 6][13 + i % 6]);
        cl.sync();   // barrier 1: every CTA's counts of waypoint t have arrived
        if (threadIdx.x == 0) {
            unsigned acc = 0;
            for (int g = 0; g < K; ++g) {
                s_prefix[g] = acc;
                for (unsigned r = 0; r < CL; ++r) acc += s_hpart[r][g];
            }
            s_prefix[K] = acc;
        }
        if (threadIdx.x < MCSIM_MAX_GAUSSIANS) s_hist[threadIdx.x] = 0;
        __syncthreads();
        // ---- sample -> collide -> moments: this CTA's share of the sample pairs
        {
            const unsigned n = s_prefix[K] < n_all ? s_prefix[K] : n_all;
            const float2 nom = cull ? __ldg(m.nominal + t) : make_float2(0.f, 0.f);
            const unsigned char* cc = s_ccnt + t * nh * MC_CULL_LEVELS;
            const ushort2* ci = s_cidx + t * nh * MC_CULL_CAP;
            // the same sample -> thread mapping as before (pass p of CTA `rank` covers 2 blockDim samples from
            // 2 blockDim (rank + p CL)), walked as per-warp windows: component in registers, one reduction per component
            gmm_member_samples<AX>(stream_id, unsigned(2 * t), k0, k1, n, K, s_prefix, s_mu, s_L, s_stat, s_tab, cc, ci, cull, nom, my_tests,
                               2u * blockDim.x * rank + 64u * (threadIdx.x >> 5), 2u * blockDim.x * CL);
        }
        __syncthreads();
        // ---- partial statistics -> slot [rank] of every CTA
        for (unsigned i = threadIdx.x; i < unsigned(K * GMM_STAT) * CL; i += blockDim.x) {
            const unsigned v = i / CL, r = i % CL;
            *cl.map_shared_rank(&s_part[rank][v], r) = (&s_stat[0][0])[v];
        }
        cl.sync();   // barrier 2: every CTA's statistics of waypoint t have arrived
        for (int i = threadIdx.x; i < K * GMM_STAT; i += blockDim.x) {
            double v = 0.0;
            for (unsigned r = 0; r < CL; ++r) v += s_part[r][i];   // rank order: bit-identical in every CTA
            (&s_tot[0][0])[i] = v;
            (&s_stat[0][0])[i] = 0.0;
        }
        __syncthreads();
        // ---- refit, redundantly in every CTA: moments and weights first ...
        const int g = threadIdx.x;
        if (g < K) gmm_refit_moments(s_comp[g], s_tot[g], m.degenerate, s_free[g], s_coll[g]);
        __syncthreads();
        // ... then warp 0 walks the EKF chain to waypoint t + 1 while the other warps draw that waypoint's counts
        if (threadIdx.x < 32) {
            if (g < K) gmm_refit_finish(a, m, t, g, s_comp[g], s_free, s_coll);
        } else if (t + 1 < T) {
            count_waypoint(t + 1);
        }
    }
    if (a.tests) {
        unsigned long long t64 = my_tests;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) t64 += __shfl_xor_sync(0xffffffffu, t64, o);
        if ((threadIdx.x & 31) == 0 && t64) atomicAdd(a.tests, t64);
    }
    cl.sync();   // no CTA leaves while a peer may still store into its shared memory
}

// ---- EM-refit variant: the free samples of a waypoint, packed in sample order -------------------------------------------
// One CTA walks the n flags in chunks of its size (ballot + popc within a warp, a scan of the warp totals, a running
// offset): out rows x | y | theta at stride cap hold the flagged samples in increasing sample index, *n_out their number.
// n = 10^4 is ten chunks (3 us); 10^6 is ~0.5 ms -- next to the host round trip per waypoint this variant already makes.
__global__ void __launch_bounds__(1024) gmm_compact_free_kernel(const float* __restrict__ in, const unsigned char* __restrict__ flag, unsigned n, unsigned cap,
                                                                float* __restrict__ out, unsigned* __restrict__ n_out) {
    __shared__ unsigned s_warp[32];
    __shared__ unsigned s_base, s_total;
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    if (threadIdx.x == 0) s_base = 0;
    for (unsigned base = 0; base < n; base += blockDim.x) {
        const unsigned i = base + threadIdx.x;
        const bool keep = i < n && flag[i] != 0;
        const unsigned bal = __ballot_sync(0xffffffffu, keep);
        if (lane == 0) s_warp[warp] = __popc(bal);
        __syncthreads();
        if (warp == 0) {
            const unsigned v = lane < nw ? s_warp[lane] : 0u;
            unsigned incl = v;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const unsigned u = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= unsigned(o)) incl += u;
            }
            s_warp[lane] = incl - v;       // exclusive prefix of the warp totals
            if (lane == 31) s_total = incl;
        }
        __syncthreads();
        const unsigned at = s_base + s_warp[warp] + __popc(bal & ((1u << lane) - 1u));
        if (keep) {
            out[at] = in[i];
            out[size_t(cap) + at] = in[size_t(cap) + i];
            out[2 * size_t(cap) + at] = in[2 * size_t(cap) + i];
        }
        __syncthreads();                   // everyone has read s_base and s_warp of this chunk
        if (threadIdx.x == 0) s_base += s_total;
    }
    __syncthreads();
    if (threadIdx.x == 0) *n_out = s_base;
}

// ---- batch: the members' shared real-robot threads on the device -------------------------------------------------------
// host_thread.cpp (one closed-loop real robot per run, FP64: MCSimulator.h:692-830 without the particles) for every
// member of a batch at once: one warp per member.  The 32 lanes draw the member's (T-1) x (3 + L) standard normals
// (the counters, the FP64 Box-Muller and the order of mcsim_host_normals), lane 0 then walks the plan -- the chain
// is sequential in t -- and writes the member's step table (applied command, motion noise, observations) and its
// initial mixture straight into device memory: nothing per member crosses PCIe.
struct RobotThreadLaunch {
    int B, T, L, K;
    unsigned k0, k1, stream0;
    double alphas[4], Q, cov0[9], chol0[6];
    double lx[MCSIM_MAX_LANDMARKS], ly[MCSIM_MAX_LANDMARKS];
    const double* trajectories;   // [3][T] (per_member: [B][3][T])
    const double* odometry;       // [3][T-1]
    int per_member;
    double* normals;              // scratch [B][T-1][3 + L]
    GmmStepTable* steps;          // [B][T-1]
    double* comps;                // [B][MCSIM_MAX_GAUSSIANS][GMM_COMP]
};

__global__ void __launch_bounds__(32) gmm_robot_thread_kernel(const RobotThreadLaunch a) {
    using namespace mcsim_math;
    const int b = blockIdx.x, T = a.T, L = a.L, TS = T > 1 ? T - 1 : 1, per_step = 3 + L, blocks = (per_step + 3) / 4;
    double* zn = a.normals + size_t(b) * TS * per_step;
    for (int q = threadIdx.x; q < (T - 1) * blocks; q += 32) {
        const int t = q / blocks, j = q - t * blocks;
        mcsim_rng::u32x4 c = {unsigned(t), a.stream0 + unsigned(b), unsigned(j), unsigned(mcsim_rng::PURPOSE_GMM_HOST)};
        c = mcsim_rng::philox4x32_10(c, a.k0, a.k1);
        const unsigned w[4] = {c.x, c.y, c.z, c.w};
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const double u1 = (double(w[2 * h]) + 0.5) * (1.0 / 4294967296.0);
            const double u2 = (double(w[2 * h + 1]) + 0.5) * (1.0 / 4294967296.0);
            const double r = sqrt(-2.0 * log(u1));
            double sn, cs;
            sincos(2 * MCSIM_PI * u2, &sn, &cs);
            if (4 * j + 2 * h < per_step) zn[size_t(t) * per_step + 4 * j + 2 * h] = r * cs;
            if (4 * j + 2 * h + 1 < per_step) zn[size_t(t) * per_step + 4 * j + 2 * h + 1] = r * sn;
        }
    }
    __syncwarp();
    const double* tr = a.trajectories + (a.per_member ? size_t(b) * 3 * T : 0);
    const double* od = a.odometry + (a.per_member ? size_t(b) * 3 * (T - 1) : 0);
    // initial mixture: K copies of N(x*_0, Sigma_0), weights 1/K (GM_Model::initModel, GM_Model.h:57-77)
    for (int q = threadIdx.x; q < a.K * GMM_COMP; q += 32) {
        const int k = q / GMM_COMP, f = q - k * GMM_COMP;
        a.comps[(size_t(b) * MCSIM_MAX_GAUSSIANS + k) * GMM_COMP + f] =
            f == 0 ? 1.0 / a.K : (f < 4 ? tr[(f - 1) * T] : (f < 13 ? a.cov0[f - 4] : a.chol0[f - 13]));
    }
    if (threadIdx.x != 0) return;
    NoiseModel<double> nm;
    for (int i = 0; i < 4; ++i) nm.alpha[i] = a.alphas[i];
    nm.Q = a.Q;
    nm.L = L;
    for (int l = 0; l < L; ++l) { nm.lx[l] = a.lx[l]; nm.ly[l] = a.ly[l]; }
    double mu[3] = {tr[0], tr[T], tr[2 * T]}, S[9], real[3] = {tr[0], tr[T], tr[2 * T]};
    for (int i = 0; i < 9; ++i) S[i] = a.cov0[i];
    const double sq = sqrt(nm.Q);
    for (int t = 0; t < T - 1; ++t) {
        const double* z = zn + size_t(t) * per_step;
        GmmStepTable& st = a.steps[size_t(b) * TS + t];
        const double ustar[3] = {od[t], od[(T - 1) + t], od[2 * (T - 1) + t]};
        const double xs[3] = {tr[t], tr[T + t], tr[2 * T + t]};
        const double xg[3] = {tr[t + 1], tr[T + t + 1], tr[2 * T + t + 1]};
        double M[3], u[3], pmu[3], pS[9], un[3], nxt[3], zobs[MCSIM_MAX_LANDMARKS];
        motion_noise(nm, ustar, M);
        applied_control(xs, mu, xg, ustar, u);
        ekf_predict(mu, S, u, M, pmu, pS);
        sample_odometry(nm, u, z, un);
        predict_pose(real, un, nxt);
        for (int k = 0; k < 3; ++k) real[k] = nxt[k];
        for (int l = 0; l < L; ++l) zobs[l] = range_to(nm, real, l) + z[3 + l] * sq;
        // the 8 updates with one division each on the dependent chain (estimator_math.h; the reference's update up to rounding):
        // this lane's chain of FP64 square roots and divisions is what the kernel's 0.41 ms per batch were made of
        for (int l = 0; l < L; ++l) ekf_update_one_chain1(nm.lx[l], nm.ly[l], nm.Q, pmu, pS, zobs[l]);
        for (int k = 0; k < 3; ++k) { mu[k] = pmu[k]; st.u[k] = u[k]; st.M[k] = M[k]; st.xs[k] = xs[k]; st.xn[k] = xg[k]; st.us[k] = ustar[k]; }
        for (int k = 0; k < 9; ++k) S[k] = pS[k];
        for (int l = 0; l < L; ++l) st.z[l] = zobs[l];
    }
}
#endif
