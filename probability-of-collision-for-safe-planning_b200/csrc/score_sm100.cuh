// score_sm100.cuh -- batched plan scoring (BASELINE.json config 4): P(collision) of many candidate
// plans against one obstacle field, `rollouts` Philox rollouts per plan.
//
// The per-rollout law is the one of mc_philox2_kernel (the reference's EKF_GaussProp body,
// mcsimplugin/MCSimulator.h:649-864, per rollout: SURVEY.md A0); what changes is the batching: one CTA
// works on ONE plan (its table of T records staged in shared memory), rollout ids are
// (plan id << 32 | rollout), and the collision query has an exact broad phase:
//
//   per CTA, once:   for every waypoint t the list of (hull, box) pairs whose box lies within
//                    reach + box circumradius + margin of the nominal position x*_t (at most
//                    SCORE_CAND_CAP, else the waypoint is marked exhaustive);
//   per rollout-step: lanes within `margin` of x*_t can only touch listed pairs (triangle inequality
//                    on the bounding circles), so the warp tests the list; if ANY lane of the warp
//                    has strayed further, the warp tests every live pair instead.
// The flags are therefore identical to the exhaustive test's (tests/test_score_gpu.py asserts it);
// executed pair tests are counted, nominal ones are rollouts x T x live pairs.
#ifndef MCSIM_SCORE_CUH
#define MCSIM_SCORE_CUH

#include "mc_philox2_sm100.cuh"

#define SCORE_CAND_CAP 12
#define SCORE_EXHAUSTIVE 255

struct ScoreLaunch {
    const float* plans;            // [n_plans][T][MC_PLAN2_STRIDE]
    const float4* pairs;           // [n_pairs][2]: (cx, cy, cos, sin), (hx, hy, -p, -q), grouped by hull
    int T, n_plans, chunks, n_pairs;
    unsigned long long rollouts;   // per plan
    unsigned long long plan_offset;  // global id of plan 0 (Philox stream = global plan id)
    unsigned k0, k1;
    float margin;                  // broad-phase margin in metres; < 0: no broad phase
    unsigned long long* hits;      // [n_plans]
    unsigned long long* counters;  // [2]: executed rollout-steps, executed pair tests
};

template <int EST, int NL, bool EARLY_EXIT>
__global__ void __launch_bounds__(128, 4) score_plans_kernel(const ScoreLaunch a) {
    extern __shared__ __align__(16) unsigned char s_raw[];
    float4* s_pairs = reinterpret_cast<float4*>(s_raw);
    float* s_plan = reinterpret_cast<float*>(s_pairs + 2 * a.n_pairs);
    unsigned short* s_cand = reinterpret_cast<unsigned short*>(s_plan + (size_t)a.T * MC_PLAN2_STRIDE);
    unsigned char* s_ncand = reinterpret_cast<unsigned char*>(s_cand + (size_t)a.T * SCORE_CAND_CAP);
    __shared__ unsigned s_hits;

    const int T = a.T;
    const int plan = blockIdx.x / a.chunks, chunk = blockIdx.x - plan * a.chunks;
    const float* gp = a.plans + (size_t)plan * T * MC_PLAN2_STRIDE;
    for (int i = threadIdx.x; i < T * MC_PLAN2_STRIDE; i += blockDim.x) s_plan[i] = gp[i];
    for (int i = threadIdx.x; i < 2 * a.n_pairs; i += blockDim.x) s_pairs[i] = a.pairs[i];
    if (threadIdx.x == 0) s_hits = 0;
    __syncthreads();
    const bool cull = a.margin >= 0.f;
    if (cull) {
        // broad phase: candidate pairs per waypoint (ascending pair index = grouped by hull)
        for (int t = threadIdx.x; t < T; t += blockDim.x) {
            const float px = s_plan[t * MC_PLAN2_STRIDE], py = s_plan[t * MC_PLAN2_STRIDE + 1];
            int n = 0;
            for (int b = 0; b < a.n_pairs; ++b) {
                const float4 lo4 = s_pairs[2 * b], hi4 = s_pairs[2 * b + 1];
                const float dx = lo4.x - px, dy = lo4.y - py;
                const float lim = c_scene.reach + sqrtf(hi4.x * hi4.x + hi4.y * hi4.y) + a.margin + 1e-3f;
                if (dx * dx + dy * dy <= lim * lim) {
                    if (n < SCORE_CAND_CAP) s_cand[t * SCORE_CAND_CAP + n] = (unsigned short)b;
                    ++n;
                }
            }
            s_ncand[t] = n <= SCORE_CAND_CAP ? (unsigned char)n : (unsigned char)SCORE_EXHAUSTIVE;
        }
        __syncthreads();
    }

    unsigned my_hits = 0;
    unsigned long long my_steps = 0, my_tests = 0;
    const float margin2 = a.margin * a.margin;
    const unsigned long long per_chunk = (a.rollouts + a.chunks - 1) / a.chunks;
    const unsigned long long r_begin = per_chunk * chunk;
    const unsigned long long r_end = (r_begin + per_chunk < a.rollouts) ? r_begin + per_chunk : a.rollouts;
    const unsigned long long plan_id = (a.plan_offset + plan) << 32;
    for (unsigned long long base = r_begin; base < r_end; base += 2ull * blockDim.x) {
        const unsigned long long i0 = base + 2ull * threadIdx.x, i1 = i0 + 1;
        const bool valid0 = i0 < r_end, valid1 = i1 < r_end;
        const unsigned long long id0 = plan_id | i0, id1 = plan_id | i1;
        Rollout2 st;
        init2<EST>(st, s_plan, id0, id1, a.k0, a.k1);
        bool hit0 = false, hit1 = false;
        for (int t = 0; t < T; ++t) {
            const f2* q = (EST == MCSIM_EST_CLOSEDLOOP) ? st.real : st.p;
            f2 sn, cs;
            sincos2(q[2], sn, cs);
            bool listed = false;
            int nc = 0;
            if (cull) {
                nc = s_ncand[t];
                if (nc != SCORE_EXHAUSTIVE) {
                    const f2 ex = add2(q[0], -s_plan[t * MC_PLAN2_STRIDE]), ey = add2(q[1], -s_plan[t * MC_PLAN2_STRIDE + 1]);
                    const f2 d2 = fma2(ex, ex, mul2(ey, ey));
                    const bool out = (valid0 && !(lo(d2) <= margin2)) || (valid1 && !(hi(d2) <= margin2));
                    listed = !__any_sync(0xffffffffu, out);
                }
            }
            const int live = int(valid0 && !(EARLY_EXIT && hit0)) + int(valid1 && !(EARLY_EXIT && hit1));
            my_steps += live;
            if (listed) {
                const unsigned short* cl = s_cand + t * SCORE_CAND_CAP;
                int k = 0;
                for (int hI = 0; hI < c_scene.n_hulls && k < nc; ++hI) {
                    const HullRec& hr = c_scene.hulls[hI];
                    const int end = hr.first + hr.count;
                    if (cl[k] >= end) continue;
                    const HullPose2 hp = place_hull2(hr, q[0], q[1], cs, sn);
                    for (; k < nc && cl[k] < end; ++k) pair_test2(hp, s_pairs[2 * cl[k]], s_pairs[2 * cl[k] + 1], hit0, hit1);
                }
                my_tests += (unsigned long long)live * nc;
            } else {
                for (int hI = 0; hI < c_scene.n_hulls; ++hI) {
                    const HullRec& hr = c_scene.hulls[hI];
                    const HullPose2 hp = place_hull2(hr, q[0], q[1], cs, sn);
                    const int end = hr.first + hr.count;
#pragma unroll 2
                    for (int b = hr.first; b < end; ++b) pair_test2(hp, s_pairs[2 * b], s_pairs[2 * b + 1], hit0, hit1);
                }
                my_tests += (unsigned long long)live * a.n_pairs;
            }
            if (EARLY_EXIT && __all_sync(0xffffffffu, (hit0 || !valid0) && (hit1 || !valid1))) break;
            if (t == T - 1) break;
            step2<EST, NL>(st, s_plan + t * MC_PLAN2_STRIDE, t, id0, id1, a.k0, a.k1);
        }
        my_hits += unsigned(hit0 && valid0) + unsigned(hit1 && valid1);
    }
    my_hits = __reduce_add_sync(0xffffffffu, my_hits);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        my_steps += __shfl_xor_sync(0xffffffffu, my_steps, o);
        my_tests += __shfl_xor_sync(0xffffffffu, my_tests, o);
    }
    if ((threadIdx.x & 31) == 0) {
        atomicAdd(&s_hits, my_hits);
        atomicAdd(&a.counters[0], my_steps);
        atomicAdd(&a.counters[1], my_tests);
    }
    __syncthreads();
    if (threadIdx.x == 0 && s_hits) atomicAdd(&a.hits[plan], (unsigned long long)s_hits);
}

#endif
