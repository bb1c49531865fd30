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

#include "estimator_math.h"
#include "philox.h"
#include "sat_device.cuh"

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
    int do_ekf;                   // refit kernel: propagate to waypoint t+1 (0 for the single-shot API)
    int B;                        // batch: independent trajectories in the same launches (Philox mode; 0/1 = single)
    const float4* sat2;           // orientation-grouped FP32 pair table (Philox mode SAT)
};

// Batch member b (blockIdx.y of the count / sample kernels, blockIdx.x of the refit kernel): same plan-independent
// fields, its own mixture, statistics, counts, step table, per-waypoint probabilities and Philox subsequence.
__device__ __forceinline__ GmmLaunch gmm_member(const GmmLaunch& in, int b) {
    GmmLaunch a = in;
    if (b > 0) {
        a.comps += size_t(b) * MCSIM_MAX_GAUSSIANS * GMM_COMP;
        a.stats += size_t(b) * MCSIM_MAX_GAUSSIANS * GMM_STAT;
        a.hist += size_t(b) * MCSIM_MAX_GAUSSIANS;
        a.steps += size_t(b) * (in.T > 1 ? in.T - 1 : 1);
        a.probabilities += size_t(b) * in.T;
        a.degenerate += b;
        a.stream_id += unsigned(b);
    }
    return a;
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
    const GmmLaunch a = gmm_member(a_in, blockIdx.y);
    __shared__ float s_cdf[MCSIM_MAX_GAUSSIANS];
    __shared__ unsigned s_hist[MCSIM_MAX_GAUSSIANS];
    if (threadIdx.x < MCSIM_MAX_GAUSSIANS) s_hist[threadIdx.x] = 0;
    if (threadIdx.x == 0) {
        double acc = 0, tot = 0;
        for (int g = 0; g < a.K; ++g) tot += a.comps[g * GMM_COMP];
        for (int g = 0; g < a.K; ++g) { acc += a.comps[g * GMM_COMP]; s_cdf[g] = float(acc / tot); }
        // components after the last one with positive weight can never be drawn
        for (int g = a.K - 1; g >= 0 && !(a.comps[g * GMM_COMP] > 0); --g) s_cdf[g] = 2.0f;
        int last = a.K - 1;
        while (last > 0 && !(a.comps[last * GMM_COMP] > 0)) --last;
        s_cdf[last] = 2.0f;  // u in (0,1]: the last live component absorbs rounding
    }
    __syncthreads();
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < a.n; i += stride) {
        // one Philox block serves 4 consecutive samples: block index i/4, word i%4
        mcsim_rng::u32x4 c = {unsigned(i >> 2), a.stream_id, unsigned(2 * a.t + 1) | unsigned((i >> 34) << 16), unsigned(mcsim_rng::PURPOSE_GMM_SAMPLE)};
        c = mcsim_rng::philox4x32_10(c, a.k0, a.k1);
        const unsigned w = (i & 3) == 0 ? c.x : (i & 3) == 1 ? c.y : (i & 3) == 2 ? c.z : c.w;
        const float u = mcsim_rng::u01(w);
        int g = 0;
        while (g < a.K - 1 && u > s_cdf[g]) ++g;
        atomicAdd(&s_hist[g], 1u);
    }
    __syncthreads();
    if (threadIdx.x < a.K && s_hist[threadIdx.x]) atomicAdd(&a.hist[threadIdx.x], s_hist[threadIdx.x]);
}

// ---- sample -> collide -> moments -------------------------------------------------------------------------------
template <int MODE>
__global__ void __launch_bounds__(256) gmm_sample_kernel(const GmmLaunch a_in) {
    const GmmLaunch a = gmm_member(a_in, blockIdx.y);
    __shared__ double s_comp[MCSIM_MAX_GAUSSIANS][GMM_COMP];
    __shared__ double s_stat[MCSIM_MAX_GAUSSIANS][GMM_STAT];
    __shared__ unsigned long long s_prefix[MCSIM_MAX_GAUSSIANS + 1];
    const int K = a.K;
    for (int i = threadIdx.x; i < K * GMM_COMP; i += blockDim.x) s_comp[i / GMM_COMP][i % GMM_COMP] = a.comps[i];
    for (int i = threadIdx.x; i < K * GMM_STAT; i += blockDim.x) s_stat[i / GMM_STAT][i % GMM_STAT] = 0.0;
    if (threadIdx.x == 0) {
        unsigned long long acc = 0;
        for (int g = 0; g < K; ++g) {
            s_prefix[g] = acc;
            acc += (MODE == MCSIM_MODE_REPLAY) ? (unsigned long long)a.counts_replay[a.t * K + g] : (unsigned long long)a.hist[g];
        }
        s_prefix[K] = acc;
    }
    __syncthreads();

    // Block-uniform trip count; every pass ends with a warp-level reduction of the pass's contributions
    // (segmented by component: samples are grouped by component, so a warp almost always holds one),
    // then ONE shared-memory atomic per statistic and warp.  (Per-thread FP64 shared atomics are CAS
    // loops: with every thread of the grid flushing 11 of them onto K x 11 addresses they serialised the
    // whole kernel.)
    int g = 0;
    const unsigned long long n = s_prefix[K] < a.n ? s_prefix[K] : a.n;
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (unsigned long long base = (unsigned long long)blockIdx.x * blockDim.x; base < n; base += stride) {
        const unsigned long long i = base + threadIdx.x;
        const bool valid = i < n;
        bool hit = false;
        double d0 = 0.0, d1 = 0.0, d2 = 0.0;
        if (valid) {
            while (g < K - 1 && i >= s_prefix[g + 1]) ++g;  // i only grows: the component index is monotone
            double z0, z1, z2;
            if (MODE == MCSIM_MODE_REPLAY) {
                z0 = a.normals[3 * i]; z1 = a.normals[3 * i + 1]; z2 = a.normals[3 * i + 2];
            } else {
                mcsim_rng::u32x4 c = {unsigned(i), a.stream_id, unsigned(2 * a.t) | unsigned((i >> 32) << 16), unsigned(mcsim_rng::PURPOSE_GMM_SAMPLE)};
                c = mcsim_rng::philox4x32_10(c, a.k0, a.k1);
                float f0, f1, f2, f3;
                mcsim_rng::box_muller(c.x, c.y, f0, f1);
                mcsim_rng::box_muller(c.z, c.w, f2, f3);
                z0 = f0; z1 = f1; z2 = f2;
            }
            const double* cp = s_comp[g];
            const double* Lc = cp + 13;
            // d = L z in the summation order of a 3x3 times 3x1 product (left to right), no contraction
            d0 = __dmul_rn(Lc[0], z0);
            d1 = __dadd_rn(__dmul_rn(Lc[1], z0), __dmul_rn(Lc[2], z1));
            d2 = __dadd_rn(__dadd_rn(__dmul_rn(Lc[3], z0), __dmul_rn(Lc[4], z1)), __dmul_rn(Lc[5], z2));
            const double x = __dadd_rn(d0, cp[1]), y = __dadd_rn(d1, cp[2]), th = __dadd_rn(d2, cp[3]);
            if (MODE == MCSIM_MODE_REPLAY) {
                hit = clearance_d(a.hulls_d, a.pairs_d, a.n_hulls, x, y, th) <= 0.0;
            } else {
                float sn, cs;
                // theta of a sample is mu_theta + O(sigma): reduce to [-pi, pi] before the FP32 sincos
                double tr = th;
                if (tr > MCSIM_PI) tr -= 2 * MCSIM_PI;
                if (tr < -MCSIM_PI) tr += 2 * MCSIM_PI;
                sincosf(float(tr), &sn, &cs);
                hit = robot_hits_grouped(a.sat2, float(x), float(y), cs, sn);
            }
        }  // valid
        const int gmin = __reduce_min_sync(0xffffffffu, valid ? g : 0x7fffffff);
        const int gmax = __reduce_max_sync(0xffffffffu, valid ? g : -1);
        for (int gg = gmin; gg <= gmax; ++gg) {  // warp-uniform
            const bool sel = valid && g == gg;
            const unsigned nf = __reduce_add_sync(0xffffffffu, unsigned(sel && !hit));
            const unsigned nc = __reduce_add_sync(0xffffffffu, unsigned(sel && hit));
            double v[9];
            const bool use = sel && !hit;
            v[0] = use ? d0 : 0.0; v[1] = use ? d1 : 0.0; v[2] = use ? d2 : 0.0;
            v[3] = use ? d0 * d0 : 0.0; v[4] = use ? d0 * d1 : 0.0; v[5] = use ? d0 * d2 : 0.0;
            v[6] = use ? d1 * d1 : 0.0; v[7] = use ? d1 * d2 : 0.0; v[8] = use ? d2 * d2 : 0.0;
            if (nf) {
#pragma unroll
                for (int k = 0; k < 9; ++k)
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) v[k] += __shfl_xor_sync(0xffffffffu, v[k], o);
            }
            if ((threadIdx.x & 31) == 0) {
                if (nf) {
                    atomicAdd(&s_stat[gg][0], double(nf));
#pragma unroll
                    for (int k = 0; k < 9; ++k) atomicAdd(&s_stat[gg][2 + k], v[k]);
                }
                if (nc) atomicAdd(&s_stat[gg][1], double(nc));
            }
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < K * GMM_STAT; i += blockDim.x) {
        const double v = s_stat[i / GMM_STAT][i % GMM_STAT];
        if (v != 0.0) atomicAdd(&a.stats[i], v);
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
__global__ void __launch_bounds__(32) gmm_refit_kernel(const GmmLaunch a_in) {
    const GmmLaunch a = gmm_member(a_in, blockIdx.x);
    using namespace mcsim_math;
    __shared__ double s_free[MCSIM_MAX_GAUSSIANS], s_coll[MCSIM_MAX_GAUSSIANS];
    const int g = threadIdx.x;
    const int K = a.K;
    double* cp = a.comps + g * GMM_COMP;
    if (g < K) {
        const double* st = a.stats + g * GMM_STAT;
        const double nf = st[0], nc = st[1];
        s_free[g] = nf; s_coll[g] = nc;
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
            atomicAdd(a.degenerate, 1);
        }
    }
    __syncthreads();
    if (g < K) {
        double tf = 0, tc = 0;
        for (int k = 0; k < K; ++k) { tf += s_free[k]; tc += s_coll[k]; }
        // normalise(collisionCounts, 1, 1).row(1): divides by the row's 1-norm, or by 1 if that is 0
        cp[0] = s_free[g] / (tf > 0 ? tf : 1.0);
        if (g == 0) a.probabilities[a.t] = tc / double(a.n);
        if (a.n_out) { a.n_out[g] = (unsigned long long)s_coll[g]; a.n_out[K + g] = (unsigned long long)s_free[g]; }
        if (a.trunc_out) for (int k = 0; k < GMM_REC; ++k) a.trunc_out[g * GMM_REC + k] = cp[k];
        if (a.t == a.T - 1 && a.rec_post) for (int k = 0; k < GMM_REC; ++k) a.rec_post[g * GMM_REC + k] = cp[k];
        // ---- move the component to waypoint t+1: EKFpredict (:766-772) then EKFupdate (:804-812)
        if (a.do_ekf && a.t < a.T - 1) {
            const GmmStepTable& s = a.steps[a.t];
            NoiseModel<double> nm;
            nm.Q = a.Q; nm.L = a.L;
            for (int l = 0; l < a.L; ++l) { nm.lx[l] = a.lx[l]; nm.ly[l] = a.ly[l]; }
            double pmu[3], pS[9];
            ekf_predict(cp + 1, cp + 4, s.u, s.M, pmu, pS);
            ekf_update(nm, pmu, pS, s.z, a.L);
            for (int k = 0; k < 3; ++k) cp[1 + k] = pmu[k];
            for (int k = 0; k < 9; ++k) cp[4 + k] = pS[k];
            if (a.rec_pre) for (int k = 0; k < GMM_REC; ++k) a.rec_pre[(size_t(a.t + 1) * K + g) * GMM_REC + k] = cp[k];
        }
        if (!gmm_chol3(cp + 4, cp + 13)) {
            // not SPD (the reference falls back to an eigen-decomposition, glue_mvnrnd_meat.hpp:98-132):
            // sample from the diagonal instead and count the event
            atomicAdd(a.degenerate, 1);
            cp[13] = sqrt(fmax(cp[4], 0.0)); cp[14] = 0; cp[15] = sqrt(fmax(cp[8], 0.0));
            cp[16] = 0; cp[17] = 0; cp[18] = sqrt(fmax(cp[12], 0.0));
        }
        for (int k = 0; k < GMM_STAT; ++k) a.stats[g * GMM_STAT + k] = 0.0;
        a.hist[g] = 0;
    }
}

#endif
