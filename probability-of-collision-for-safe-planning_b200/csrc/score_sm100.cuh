// score_sm100.cuh -- batched plan scoring (BASELINE.json config 4): P(collision) of many candidate
// plans against one obstacle field, `rollouts` Philox rollouts per plan.
//
// The per-rollout law is the one of mc_philox2_kernel (the reference's EKF_GaussProp body,
// mcsimplugin/MCSimulator.h:649-864, per rollout: SURVEY.md A0); what changes is the batching: one CTA
// works on ONE plan (its table of T records staged in shared memory), rollout ids are
// (plan id << 32 | rollout), and the collision query has an exact broad phase:
//
//   per CTA, once:   for every waypoint t the list of (hull, box) pairs whose box lies within
//                    reach + box circumradius + margin of the nominal position x*_t, grouped by hull
//                    and, within a hull, ordered by the smallest of SCORE_LEVELS margins
//                    m_k = margin * 2^((k - 7) / 2) that lists the pair; cumulative counts per
//                    (waypoint, hull, level).  At most SCORE_CAND_CAP entries: the levels from the
//                    first one that would need more are marked exhaustive;
//   per rollout-step: lanes within m_k of x*_t can only touch pairs listed at level k (triangle
//                    inequality on the bounding circles), so the warp takes the level of its
//                    furthest live lane and tests that prefix of every hull's entries; beyond the
//                    last level (or into an overflowed one) it tests every live pair instead.
// The flags are therefore identical to the exhaustive test's (tests/test_score_gpu.py asserts it);
// executed pair tests are counted, nominal ones are rollouts x T x live pairs.
#ifndef MCSIM_SCORE_CUH
#define MCSIM_SCORE_CUH

#include "mc_philox2_sm100.cuh"

#ifndef SCORE_CAND_CAP
#define SCORE_CAND_CAP 16
#endif
#define SCORE_LEVELS 8

// level of a scaled squared deviation x = d^2 / margin^2: smallest k with x <= 2^(k - 7), read off the
// exponent (x in (2^(e-1), 2^e] -> e + 7); SCORE_LEVELS = beyond the last level (also NaN)
__device__ __forceinline__ int score_level(float x) {
    const unsigned b = __float_as_uint(x);
    const unsigned u = ((b ? b : 1u) - 1u) >> 23;  // e + 126
    const unsigned k = u >= 119u ? u - 119u : 0u;
    return k < unsigned(SCORE_LEVELS) ? int(k) : SCORE_LEVELS;
}

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

// dynamic shared memory of one CTA: pair table | plan table | candidate entries | overflow level | counts
static inline size_t score_smem_bytes(int T, int n_pairs, int n_hulls) {
    return sizeof(float4) * 2 * size_t(n_pairs) + sizeof(float) * size_t(T) * MC_PLAN2_STRIDE + sizeof(unsigned short) * size_t(T) * SCORE_CAND_CAP +
           size_t((T + 3) & ~3) + size_t(T) * n_hulls * SCORE_LEVELS + 16;
}

template <int EST, int NL, bool EARLY_EXIT>
__global__ void __launch_bounds__(128, 4) score_plans_kernel(const ScoreLaunch a) {
    extern __shared__ __align__(16) unsigned char s_raw[];
    float4* s_pairs = reinterpret_cast<float4*>(s_raw);
    float* s_plan = reinterpret_cast<float*>(s_pairs + 2 * a.n_pairs);
    unsigned short* s_cand = reinterpret_cast<unsigned short*>(s_plan + (size_t)a.T * MC_PLAN2_STRIDE);
    unsigned char* s_ovf = reinterpret_cast<unsigned char*>(s_cand + (size_t)a.T * SCORE_CAND_CAP);  // [T]: first exhaustive level
    unsigned char* s_ccnt = s_ovf + ((a.T + 3) & ~3);  // [T][n_hulls][SCORE_LEVELS]: entries of the hull with level <= k
    __shared__ unsigned s_hits;

    const int T = a.T;
    const int plan = blockIdx.x / a.chunks, chunk = blockIdx.x - plan * a.chunks;
    const float* gp = a.plans + (size_t)plan * T * MC_PLAN2_STRIDE;
    for (int i = threadIdx.x; i < T * MC_PLAN2_STRIDE; i += blockDim.x) s_plan[i] = gp[i];
    for (int i = threadIdx.x; i < 2 * a.n_pairs; i += blockDim.x) s_pairs[i] = a.pairs[i];
    if (threadIdx.x == 0) s_hits = 0;
    __syncthreads();
    const bool cull = a.margin >= 0.f;
    const int nh = c_scene.n_hulls;
    if (cull) {
        // broad phase: candidate pairs per waypoint, by hull, within a hull by level (stable insertion)
        for (int t = threadIdx.x; t < T; t += blockDim.x) {
            const float px = s_plan[t * MC_PLAN2_STRIDE], py = s_plan[t * MC_PLAN2_STRIDE + 1];
            unsigned short* cl = s_cand + t * SCORE_CAND_CAP;
            int n = 0, ovf = SCORE_LEVELS;
            for (int hI = 0; hI < nh; ++hI) {
                const HullRec& hr = c_scene.hulls[hI];
                const int seg = n;
                for (int b = hr.first; b < hr.first + hr.count; ++b) {
                    const float4 lo4 = s_pairs[2 * b], hi4 = s_pairs[2 * b + 1];
                    const float dx = lo4.x - px, dy = lo4.y - py;
                    // smallest margin that lists the pair (1 mm of slack covers the rounding of both sides)
                    const float need = sqrtf(dx * dx + dy * dy) - (c_scene.reach + sqrtf(hi4.x * hi4.x + hi4.y * hi4.y) + 1e-3f);
                    if (!(need <= a.margin)) continue;
                    int k = 0;
                    if (need > 0.f) {
                        const float x = need / a.margin;
                        k = score_level(x * x * 0.999999f);  // rounded down: a pair is never listed later than its margin says
                        if (k >= SCORE_LEVELS) k = SCORE_LEVELS - 1;
                    }
                    if (n == SCORE_CAND_CAP) {
                        // full: keep the lower level (the hull's last entry has its highest); the levels from the
                        // dropped one up become exhaustive
                        if (n > seg && int(cl[n - 1] >> 12) > k) { ovf = min(ovf, int(cl[n - 1] >> 12)); --n; }
                        else { ovf = min(ovf, k); continue; }
                    }
                    // entries carry their level in the top 4 bits while the list is built (pair index < 4096)
                    int pos = n;  // after the last entry of the segment with level <= k
                    while (pos > seg && int(cl[pos - 1] >> 12) > k) { cl[pos] = cl[pos - 1]; --pos; }
                    cl[pos] = (unsigned short)(b | (k << 12));
                    ++n;
                }
                unsigned char* cc = s_ccnt + (t * nh + hI) * SCORE_LEVELS;
                for (int k = 0; k < SCORE_LEVELS; ++k) {
                    int m = 0;
                    for (int e = seg; e < n; ++e) m += int(cl[e] >> 12) <= k;
                    cc[k] = (unsigned char)m;
                }
            }
            for (int e = 0; e < n; ++e) cl[e] &= 0x0fffu;
            s_ovf[t] = (unsigned char)ovf;
        }
        __syncthreads();
    }

    unsigned my_hits = 0;
    unsigned long long my_steps = 0, my_tests = 0;
    const float inv_margin2 = cull ? 1.0f / (a.margin * a.margin) : 0.f;
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
            int lev = SCORE_LEVELS;
            if (cull) {
                // squared deviation from x*_t, max over the warp's live lanes (NaN orders above every finite value)
                const f2 ex = add2(q[0], -s_plan[t * MC_PLAN2_STRIDE]), ey = add2(q[1], -s_plan[t * MC_PLAN2_STRIDE + 1]);
                const f2 d2 = fma2(ex, ex, mul2(ey, ey));
                const unsigned u0 = valid0 ? __float_as_uint(lo(d2)) : 0u, u1 = valid1 ? __float_as_uint(hi(d2)) : 0u;
                const float dm = __uint_as_float(__reduce_max_sync(0xffffffffu, u0 > u1 ? u0 : u1));
                lev = score_level(dm * inv_margin2);
                if (lev >= s_ovf[t]) lev = SCORE_LEVELS;
            }
            const int live = int(valid0 && !(EARLY_EXIT && hit0)) + int(valid1 && !(EARLY_EXIT && hit1));
            my_steps += live;
            if (lev < SCORE_LEVELS) {
                const unsigned short* cl = s_cand + t * SCORE_CAND_CAP;
                const unsigned char* cc = s_ccnt + t * nh * SCORE_LEVELS;
                int tested = 0;
                for (int hI = 0; hI < nh; ++hI) {
                    const int n = cc[hI * SCORE_LEVELS + lev];
                    if (n) {
                        const HullPose2 hp = place_hull2(c_scene.hulls[hI], q[0], q[1], cs, sn);
                        for (int k = 0; k < n; ++k) pair_test2(hp, s_pairs[2 * cl[k]], s_pairs[2 * cl[k] + 1], hit0, hit1);
                        tested += n;
                    }
                    cl += cc[hI * SCORE_LEVELS + SCORE_LEVELS - 1];
                }
                my_tests += (unsigned long long)live * tested;
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
