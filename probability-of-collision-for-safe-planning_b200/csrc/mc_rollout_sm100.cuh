// mc_rollout_sm100.cuh -- Monte-Carlo rollout kernel: one thread per rollout, pose (and, in Philox
// mode, the rollout's own real-robot/EKF thread) held in registers across all waypoints.
//
// Replaces the particle loop of the reference (mcsimplugin/MCSimulator.h):
//   initParticles :287-297 (x0 = mu0 + chol(Sigma0) z), moveParticles :300-322,
//   checkParticleCollisions :333-347, getCollisionProportion :324-330,
//   and in Philox mode the per-step body of EKF_GaussProp :692-830 per rollout.
#ifndef MCSIM_MC_ROLLOUT_CUH
#define MCSIM_MC_ROLLOUT_CUH

#include "estimator_math.h"
#include "philox.h"
#include "sat_device.cuh"

// per-step plan record staged in shared memory (floats)
//   replay : u~ = (rot1, trans, rot2) as 3 FP64 values             6 floats
//   philox : x*_t (3), u*_t (3), M_t (3) = generateM_EKF(u*_t)      9 floats
#define MC_REPLAY_STRIDE 6
#define MC_PHILOX_STRIDE 9

struct McPlanConst {
    int T;
    int pad;
    double mu0d[3]; // FP64 copies for the replay-mode start pose (injected draws are FP64)
    double chold[6];
    float mu0[3];
    float chol[6];  // lower Cholesky factor of Sigma0: l00 l10 l11 l20 l21 l22
    float sqrtQ;
    mcsim_math::NoiseModel<float> nm;
};
__constant__ McPlanConst c_mc;

// counters block in global memory
enum { CNT_HITS = 0, CNT_NEAR = 1, CNT_STEPS = 2, CNT_TESTS = 3, CNT_HIST = 4 };

struct McLaunch {
    const float* plan;            // [T-1 or T][stride]
    const double* z0;             // replay: 3 x n standard normals (rollout-major) for THIS launch
    unsigned long long n;         // rollouts in this launch
    unsigned long long offset;    // global id of rollout 0
    unsigned k0, k1;              // Philox key
    unsigned long long* counters; // [4 + T]
    unsigned* counts;             // optional [n]
    float* poses;                 // optional audit [T][3][n]
    float* clearance;             // optional audit [T][n]
    float near_eps;
    const float4* sat2;           // orientation-grouped pair table (packed kernel), sat2_n float4 records
    int sat2_n;
    // exact broad phase of the packed kernel (mc_philox2_sm100.cuh); cull == nullptr: every live pair at every step
    const unsigned* cull;         // per-waypoint candidate lists, cull_words 32-bit words
    int cull_words;
    int smem_lists;               // scalar kernels: pair table and lists were given room in dynamic shared memory behind the plan
    const float2* nominal;        // [T] (x*_t, y*_t): centres of the lists for the replay kernel (the packed kernels read c_plan2)
};

// FP64 angle in [0, 4 pi] -> FP32 angle in [-pi, pi] (rounding error <= 1.2e-7 rad, not accumulated)
__device__ __forceinline__ float reduce_angle(double a) {
    if (a > MCSIM_PI) a -= 2 * MCSIM_PI;
    if (a > MCSIM_PI) a -= 2 * MCSIM_PI;
    return float(a);
}

template <int MODE, bool AUDIT, bool EARLY_EXIT, int EST, int NL>
__global__ void __launch_bounds__(256) mc_rollout_kernel(const McLaunch a) {
    using namespace mcsim_math;
    extern __shared__ __align__(16) float s_plan[];
    __shared__ unsigned s_cnt[4];
    __shared__ unsigned s_hist[MCSIM_MAX_WAYPOINTS];

    const int T = c_mc.T;
    constexpr int STRIDE = (MODE == MCSIM_MODE_REPLAY) ? MC_REPLAY_STRIDE : MC_PHILOX_STRIDE;
    const int n_plan = (MODE == MCSIM_MODE_REPLAY) ? (T - 1) * STRIDE : T * STRIDE;
    for (int i = threadIdx.x; i < n_plan; i += blockDim.x) s_plan[i] = a.plan[i];
    // broad phase: pair table and candidate lists behind the plan.  Every step walks count -> index -> table record, three
    // DEPENDENT loads; from global memory that chain was half of the stall samples of a 1,000-particle run
    // (long_scoreboard 5.6 per issue, profiles/r02_replay_ncu_summary.csv): the reference driver's default mode is latency-bound
    float4* s_tab = reinterpret_cast<float4*>(s_plan + ((n_plan + 3) & ~3));
    unsigned* s_cull = reinterpret_cast<unsigned*>(s_tab + a.sat2_n);
    if (a.smem_lists) {
        for (int i = threadIdx.x; i < a.sat2_n; i += blockDim.x) s_tab[i] = a.sat2[i];
        for (int i = threadIdx.x; i < a.cull_words; i += blockDim.x) s_cull[i] = a.cull[i];
    }
    for (int i = threadIdx.x; i < T; i += blockDim.x) s_hist[i] = 0;
    if (threadIdx.x < 4) s_cnt[threadIdx.x] = 0;
    __syncthreads();

    unsigned my_hits = 0, my_near = 0;
    unsigned long long my_steps = 0, my_tests = 0;
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    // exact broad phase of the replay mode (sat_device.cuh): the plan's candidate lists, chosen per warp and waypoint by
    // the largest deviation of its live rollouts from x*_t -- the reference driver's default mode (one shared real robot,
    // MCSimulation.py:238-247) spends its time in this kernel, 63 pair tests per pose without the lists
    // (Philox mode comes here with polygon robot hulls only: box-hull scenes run the packed kernel, which has its own lists)
    const bool listed = !AUDIT && a.cull != nullptr;
    const int nh = c_scene.n_hulls + c_scene.n_poly;   // the lists' hull index: box hulls, then polygon hulls
    const unsigned* lists = a.smem_lists ? s_cull : a.cull;
    const unsigned char* g_ccnt = reinterpret_cast<const unsigned char*>(lists);
    const ushort2* g_cidx = reinterpret_cast<const ushort2*>(lists + (size_t(T) * nh * MC_CULL_LEVELS + 3) / 4);
    // block-uniform trip count: lanes past the end are masked (the warp-wide reductions of the broad phase need every lane)
    for (unsigned long long base = (unsigned long long)blockIdx.x * blockDim.x; base < a.n; base += stride) {
        const unsigned long long i = base + threadIdx.x;
        const bool valid = i < a.n;
        const unsigned long long id = a.offset + i;
        // ---- initial particle: x0 = mu0 + L z (initParticles, glue_mvnrnd_meat.hpp:89-147)
        // injected draws (replay with a z0 table) or the rollout's own Philox block (Philox mode, and the
        // "shared" mode of the plugin: replayed host thread + device-generated particles, z0 == NULL)
        float z[3] = {0.f, 0.f, 0.f};
        const bool injected = (MODE == MCSIM_MODE_REPLAY) && a.z0 != nullptr;
        if (!injected) {
            mcsim_rng::u32x4 c = {unsigned(id), unsigned(id >> 32), 0u, unsigned(mcsim_rng::PURPOSE_INIT)};
            c = mcsim_rng::philox4x32_10(c, a.k0, a.k1);
            float z3;
            mcsim_rng::box_muller(c.x, c.y, z[0], z[1]);
            mcsim_rng::box_muller(c.z, c.w, z[2], z3);
        }
        float p[3];
        double th = 0.0;  // replay mode keeps the heading in FP64 (see below)
        if (MODE == MCSIM_MODE_REPLAY) {
            // FP64 for the start pose only (3 FMAs): the injected draws are FP64
            const double z0 = (injected && valid) ? a.z0[3 * i] : double(z[0]);
            const double z1 = (injected && valid) ? a.z0[3 * i + 1] : double(z[1]);
            const double z2 = (injected && valid) ? a.z0[3 * i + 2] : double(z[2]);
            p[0] = float(c_mc.mu0d[0] + c_mc.chold[0] * z0);
            p[1] = float(c_mc.mu0d[1] + c_mc.chold[1] * z0 + c_mc.chold[2] * z1);
            th = c_mc.mu0d[2] + c_mc.chold[3] * z0 + c_mc.chold[4] * z1 + c_mc.chold[5] * z2;
            p[2] = float(th);
        } else {
            p[0] = c_mc.mu0[0] + c_mc.chol[0] * z[0];
            p[1] = c_mc.mu0[1] + c_mc.chol[1] * z[0] + c_mc.chol[2] * z[1];
            p[2] = c_mc.mu0[2] + c_mc.chol[3] * z[0] + c_mc.chol[4] * z[1] + c_mc.chol[5] * z[2];
        }
        // Philox mode: this rollout's own real robot + EKF
        float real[3], mu[3], S[9];
        if (MODE == MCSIM_MODE_PHILOX) {
#pragma unroll
            for (int k = 0; k < 3; ++k) { mu[k] = c_mc.mu0[k]; real[k] = (EST == MCSIM_EST_CLOSEDLOOP) ? p[k] : c_mc.mu0[k]; }
            S[0] = c_mc.chol[0] * c_mc.chol[0];
            S[1] = S[3] = c_mc.chol[0] * c_mc.chol[1];
            S[2] = S[6] = c_mc.chol[0] * c_mc.chol[3];
            S[4] = c_mc.chol[1] * c_mc.chol[1] + c_mc.chol[2] * c_mc.chol[2];
            S[5] = S[7] = c_mc.chol[1] * c_mc.chol[3] + c_mc.chol[2] * c_mc.chol[4];
            S[8] = c_mc.chol[3] * c_mc.chol[3] + c_mc.chol[4] * c_mc.chol[4] + c_mc.chol[5] * c_mc.chol[5];
        }

        // Replay mode keeps the heading in FP64: theta is a running sum of angles near 2 pi, where an
        // FP32 state would lose 2.4e-7 rad per addition and, through trans * d(theta), ~2e-5 m over
        // 55 steps (measured); x and y stay FP32.  Only the reduced angle is rounded to FP32.
        unsigned count = 0;
        int first = -1;
        bool alive = valid;   // still being tested (early exit: until the first hit)
        for (int t = 0; t < T; ++t) {
            // ---- collision test at waypoint t (checkParticleCollisions)
            const float* q = (MODE == MCSIM_MODE_PHILOX && EST == MCSIM_EST_CLOSEDLOOP) ? real : p;
            float sn, cs;
            if (MODE == MCSIM_MODE_REPLAY) sincos_t(reduce_angle(th), sn, cs);
            else sincos_t(q[2], sn, cs);
            bool hit;
            unsigned tested = 0;
            if (AUDIT) {
                const float cl = robot_clearance(q[0], q[1], cs, sn);
                hit = cl <= 0.0f;
                tested = c_scene.n_pairs;
                if (valid) {
                    my_near += (fabsf(cl) < a.near_eps);
                    if (a.clearance) a.clearance[(size_t)t * a.n + i] = cl;
                    if (a.poses) {
                        a.poses[((size_t)t * 3 + 0) * a.n + i] = q[0];
                        a.poses[((size_t)t * 3 + 1) * a.n + i] = q[1];
                        a.poses[((size_t)t * 3 + 2) * a.n + i] = (MODE == MCSIM_MODE_REPLAY) ? float(th) : q[2];
                    }
                }
            } else if (listed) {
                const float2 nom = __ldg(a.nominal + t);
                const float ex = q[0] - nom.x, ey = q[1] - nom.y;
                const float dm = __uint_as_float(__reduce_max_sync(0xffffffffu, alive ? __float_as_uint(fmaf(ex, ex, ey * ey)) : 0u));
                int lev = cull_level(dm);
                const unsigned char* cc = g_ccnt + size_t(t) * nh * MC_CULL_LEVELS;
                if (lev < MC_CULL_LEVELS && cc[lev] == MC_CULL_OVERFLOW) lev = MC_CULL_LEVELS;
                if (lev < MC_CULL_LEVELS && a.smem_lists) hit = robot_hits_listed<true>(s_tab, cc, g_cidx + size_t(t) * nh * MC_CULL_CAP, lev, q[0], q[1], cs, sn, tested);
                else if (lev < MC_CULL_LEVELS) hit = robot_hits_listed(a.sat2, cc, g_cidx + size_t(t) * nh * MC_CULL_CAP, lev, q[0], q[1], cs, sn, tested);
                else hit = robot_hits(q[0], q[1], cs, sn, tested);
            } else {
                hit = robot_hits(q[0], q[1], cs, sn, tested);
            }
            if (alive) {
                ++my_steps;
                my_tests += tested;
                if (hit) {
                    if (first < 0) first = t;
                    ++count;
                    if (EARLY_EXIT) alive = false;
                }
            }
            if (EARLY_EXIT && !__any_sync(0xffffffffu, alive)) break;
            if (t == T - 1) break;
            // ---- move to waypoint t+1
            if (MODE == MCSIM_MODE_REPLAY) {
                const double* u = reinterpret_cast<const double*>(s_plan) + 3 * t;
                float s1, c1;
                sincos_t(reduce_angle(th + u[0]), s1, c1);
                const float tr = float(u[1]);
                p[0] = fmaf(tr, c1, p[0]);
                p[1] = fmaf(tr, s1, p[1]);
                th = angle_wrap(th + u[0] + u[2]);
            } else {
                const float* r = s_plan + t * STRIDE;
                const float xs[3] = {r[0], r[1], r[2]};
                const float us[3] = {r[3], r[4], r[5]};
                const float M[3] = {r[6], r[7], r[8]};
                const float xg[3] = {r[STRIDE], r[STRIDE + 1], r[STRIDE + 2]};
                const int L = NL > 0 ? NL : c_mc.nm.L;
                // noise draws of this step: 3 odometry + L observation normals
                float zn[NL > 0 ? 12 * ((3 + NL + 11) / 12) : 12 * ((3 + MCSIM_MAX_LANDMARKS + 11) / 12)];
                const int ngrp = (3 + L + 11) / 12;  // 12 normals per two Philox blocks (philox.h: step draws)
#pragma unroll
                for (int g = 0; g < ngrp; ++g) {
                    const mcsim_rng::u32x4 ca = {unsigned(id), unsigned(id >> 32), unsigned(t * 8 + 2 * g), unsigned(mcsim_rng::PURPOSE_STEP)};
                    const mcsim_rng::u32x4 cb = {unsigned(id), unsigned(id >> 32), unsigned(t * 8 + 2 * g + 1), unsigned(mcsim_rng::PURPOSE_STEP)};
                    unsigned rad[6], ang[6];
                    mcsim_rng::step_bits(mcsim_rng::philox4x32_10(ca, a.k0, a.k1), mcsim_rng::philox4x32_10(cb, a.k0, a.k1), rad, ang);
#pragma unroll
                    for (int j = 0; j < 6; ++j) mcsim_rng::box_muller_bits(rad[j], ang[j], zn[12 * g + 2 * j], zn[12 * g + 2 * j + 1]);
                }
                float u[3], pmu[3], pS[9], un[3], nx[3];
                applied_control(xs, mu, xg, us, u);
                ekf_predict(mu, S, u, M, pmu, pS);
                sample_odometry(c_mc.nm, u, zn, un);
                predict_pose(real, un, nx);
                real[0] = nx[0]; real[1] = nx[1]; real[2] = nx[2];
                if (EST == MCSIM_EST_REFERENCE) {
                    predict_pose(p, un, nx);
                    p[0] = nx[0]; p[1] = nx[1]; p[2] = nx[2];
                }
                float zobs[MCSIM_MAX_LANDMARKS];
#pragma unroll
                for (int l = 0; l < L; ++l) zobs[l] = range_to(c_mc.nm, real, l) + zn[3 + l] * c_mc.sqrtQ;
                ekf_update(c_mc.nm, pmu, pS, zobs, L);
#pragma unroll
                for (int k = 0; k < 3; ++k) mu[k] = pmu[k];
#pragma unroll
                for (int k = 0; k < 9; ++k) S[k] = pS[k];
            }
        }
        if (count) {
            ++my_hits;
            atomicAdd(&s_hist[first], 1u);
        }
        if (a.counts && valid) a.counts[i] = count;
    }

    // ---- aggregation: warp reduce -> shared -> one 64-bit atomic per counter per block
    my_hits = __reduce_add_sync(0xffffffffu, my_hits);
    my_near = __reduce_add_sync(0xffffffffu, my_near);
    unsigned long long w_steps = my_steps, w_tests = my_tests;   // 64-bit: a 10^10-rollout precise run must not wrap
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        w_steps += __shfl_xor_sync(0xffffffffu, w_steps, o);
        w_tests += __shfl_xor_sync(0xffffffffu, w_tests, o);
    }
    if ((threadIdx.x & 31) == 0) {
        atomicAdd(&s_cnt[CNT_HITS], my_hits);
        atomicAdd(&s_cnt[CNT_NEAR], my_near);
        atomicAdd(&a.counters[CNT_STEPS], w_steps);
        atomicAdd(&a.counters[CNT_TESTS], w_tests);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        if (s_cnt[CNT_HITS]) atomicAdd(&a.counters[CNT_HITS], (unsigned long long)s_cnt[CNT_HITS]);
        if (s_cnt[CNT_NEAR]) atomicAdd(&a.counters[CNT_NEAR], (unsigned long long)s_cnt[CNT_NEAR]);
    }
    for (int t = threadIdx.x; t < T; t += blockDim.x)
        if (s_hist[t]) atomicAdd(&a.counters[CNT_HIST + t], (unsigned long long)s_hist[t]);
}

#endif
