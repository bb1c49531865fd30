// mc_philox2_sm100.cuh -- Philox-mode Monte-Carlo rollout kernel, two rollouts per thread in the two
// lanes of packed FP32 pairs (f32x2.cuh).
//
// Same law, same Philox counters and the same draw order as mc_rollout_kernel<PHILOX> (the precise
// scalar kernel, kept for audits and as this kernel's cross-check): every rollout owns its
// real-robot / EKF thread and its particle.  What it replaces in the reference
// (mcsimplugin/MCSimulator.h), per rollout and per waypoint:
//   initParticles :287-297, checkParticleCollisions :333-347 (SAT, sat_device.cuh),
//   generateL :532-553 + applied control :719-726, EKFpredict :868-881 (G :517-529, V :453-468,
//   M :495-513), sampleOdometry :391-410, prediction :413-431 / moveParticles :300-322,
//   sampleObservation :383-387, EKFupdate :883-929, getCollisionProportion :324-330.
//
// Differences from the scalar kernel, all below the Monte-Carlo resolution of this mode (the parity
// bar of Philox mode is the reference's 99 % confidence interval, tests/test_mc_gpu.py):
//   * MUFU approximations (rcp, rsqrt, sqrt, sin, cos, lg2) instead of IEEE div / sqrt / libm;
//   * Sigma is kept symmetric (6 values): (I - K H) Sigma = Sigma - K (Sigma H^T)^T exactly;
//   * the controller uses the identity u* + ((u_req - u*) / xhat) * xhat = u_req (xhat != 0), and
//     xhat = 0 exactly only at t = 0 where mu = x*_0 (MCSimulator.h:164), so u = u*_0 there.
#ifndef MCSIM_MC_PHILOX2_CUH
#define MCSIM_MC_PHILOX2_CUH

#include "f32x2.cuh"
#include "mc_rollout_sm100.cuh"

// per-step plan record in __constant__ memory (uniform loads, operands broadcast from uniform
// registers): x*_t (3), u*_t (3), M_t (3) = generateM_EKF(u*_t), M_0 + M_2
#define MC_PLAN2_STRIDE 12
#ifndef MC_P2_CTAS_PER_SM
#define MC_P2_CTAS_PER_SM 5  // resident 128-thread CTAs per SM the register budget is set for
#endif
#define MC_PLAN2_MAX_T 256
__constant__ float c_plan2[MC_PLAN2_MAX_T * MC_PLAN2_STRIDE];

// two standard normals per lane from two 32-bit words per lane (Box-Muller)
__device__ __forceinline__ void box_muller2(unsigned a0, unsigned b0, unsigned a1, unsigned b1, f2& n0, f2& n1) {
    // u in (0, 1]: x 2^-32 + 2^-33;  angle = 2 pi u
    const f2 ua = fma2(mk2(float(a0), float(a1)), 2.3283064365386963e-10f, 1.1641532182693481e-10f);
    const f2 th = fma2(mk2(float(b0), float(b1)), 1.4629180792671596e-9f, 7.314590396335798e-10f);
    // r = sqrt(-2 ln u) = sqrt(-2 ln2 * lg2 u)
    const f2 r = sqrt2(mul2(mk2(mufu_lg2(lo(ua)), mufu_lg2(hi(ua))), -1.3862943611198906f));
    f2 s, c;
    sincos2(th, s, c);
    n0 = mul2(r, c);
    n1 = mul2(r, s);
}

struct Philox2 {
    mcsim_rng::u32x4 a, b;
};
__device__ __forceinline__ Philox2 philox_pair(unsigned long long id0, unsigned long long id1, unsigned blk, unsigned purpose, unsigned k0, unsigned k1) {
    Philox2 o;
    const mcsim_rng::u32x4 c0 = {unsigned(id0), unsigned(id0 >> 32), blk, purpose};
    const mcsim_rng::u32x4 c1 = {unsigned(id1), unsigned(id1 >> 32), blk, purpose};
    o.a = mcsim_rng::philox4x32_10(c0, k0, k1);
    o.b = mcsim_rng::philox4x32_10(c1, k0, k1);
    return o;
}

// ---- packed SAT: both lanes against every live pair ------------------------------------------------
// The same separating-axis test as pair_hit (sat_device.cuh) on the orientation-grouped table
// (mcsim_cuda_set_scene): per class of boxes sharing an axis direction (mod 90 degrees) the relative
// rotation (c, s), the hull centre in the class frame (gx, gy) and the hull's radii on the class
// axes (rbx, rby) are computed once (12 packed instructions); per pair and per two rollouts that
// leaves 8 FFMA2 + 4 FADD2 and 8 FSETP.  Every live pair is still tested at every step.
// Table records (float4, shared memory): class header (cos, sin, count, -), then per pair
// (cx, cy, hx, hy), (-p, -q, -, -) with p = cb cx + sb cy, q = cb cy - sb cx.
__device__ __forceinline__ void robot_hits2(const float4* __restrict__ tab, f2 x, f2 y, f2 ct, f2 st, bool& hit0, bool& hit1) {
    const int nh = c_scene.n_hulls;
    for (int hI = 0; hI < nh; ++hI) {
        const HullRec& hr = c_scene.hulls[hI];
        const f2 hcx = fma2(ct, hr.ox, fma2(neg2(st), hr.oy, x));
        const f2 hcy = fma2(st, hr.ox, fma2(ct, hr.oy, y));
        const f2 ch = fma2(ct, hr.cp, mul2(neg2(st), hr.sp));
        const f2 sh = fma2(st, hr.cp, mul2(ct, hr.sp));
        const f2 nph = neg2(fma2(ch, hcx, mul2(sh, hcy)));
        const f2 nqh = neg2(fma2(ch, hcy, mul2(neg2(sh), hcx)));
        const f2 hhx = bc(hr.hx), hhy = bc(hr.hy);
        const float4* g = tab + hr.g_first;
        for (int gi = 0; gi < hr.n_groups; ++gi) {
            const float4 head = g[0];
            const int cnt = __float_as_int(head.z);
            const f2 c = abs2(fma2(ch, head.x, mul2(sh, head.y)));
            const f2 s = abs2(fma2(sh, head.x, mul2(neg2(ch), head.y)));
            const f2 gx = fma2(hcx, head.x, mul2(hcy, head.y));
            const f2 gy = fma2(hcy, head.x, mul2(neg2(hcx), head.y));
            const f2 rbx = fma2(c, hhx, mul2(s, hhy));
            const f2 rby = fma2(s, hhx, mul2(c, hhy));
            ++g;
#pragma unroll 4
            for (int b = 0; b < cnt; ++b) {
                const float4 A = g[2 * b];                                           // cx, cy, hx, hy
                const float2 B = *reinterpret_cast<const float2*>(g + 2 * b + 1);    // -p, -q
                const f2 drx = fma2(ch, A.x, fma2(sh, A.y, nph));
                const f2 dry = fma2(ch, A.y, fma2(neg2(sh), A.x, nqh));
                const f2 r0 = fma2(c, A.z, fma2(s, A.w, hhx));
                const f2 r1 = fma2(s, A.z, fma2(c, A.w, hhy));
                const f2 dbx = add2(gx, B.x);
                const f2 dby = add2(gy, B.y);
                const f2 r2 = add2(rbx, A.z);
                const f2 r3 = add2(rby, A.w);
                if ((fabsf(lo(drx)) <= lo(r0)) & (fabsf(lo(dry)) <= lo(r1)) & (fabsf(lo(dbx)) <= lo(r2)) & (fabsf(lo(dby)) <= lo(r3))) hit0 = true;
                if ((fabsf(hi(drx)) <= hi(r0)) & (fabsf(hi(dry)) <= hi(r1)) & (fabsf(hi(dbx)) <= hi(r2)) & (fabsf(hi(dby)) <= hi(r3))) hit1 = true;
            }
            g += 2 * cnt;
        }
    }
}

// prediction (:413-431) on a packed pose
__device__ __forceinline__ void move2(f2 p[3], const f2 u[3]) {
    f2 sn, cs;
    sincos2(add2(p[2], u[0]), sn, cs);
    p[0] = fma2(u[1], cs, p[0]);
    p[1] = fma2(u[1], sn, p[1]);
    p[2] = wrap2(add2(add2(p[2], u[0]), u[2]));
}

// state of the two rollouts of a thread: particle, real robot, EKF mean and covariance
// (symmetric: s00 s01 s02 s11 s12 s22)
struct Rollout2 {
    f2 p[3], real[3], mu[3], S[6];
};

// x0 = mu0 + L z (initParticles), real robot and EKF at the nominal start (MCSimulator.h:656-657)
template <int EST>
__device__ __forceinline__ void init2(Rollout2& st, const float* mu0, unsigned long long id0, unsigned long long id1, unsigned k0, unsigned k1) {
    const Philox2 r = philox_pair(id0, id1, 0u, unsigned(mcsim_rng::PURPOSE_INIT), k0, k1);
    f2 z0, z1, z2, z3;
    box_muller2(r.a.x, r.a.y, r.b.x, r.b.y, z0, z1);
    box_muller2(r.a.z, r.a.w, r.b.z, r.b.w, z2, z3);
    st.p[0] = fma2(z0, c_mc.chol[0], mu0[0]);
    st.p[1] = fma2(z0, c_mc.chol[1], fma2(z1, c_mc.chol[2], mu0[1]));
    st.p[2] = fma2(z0, c_mc.chol[3], fma2(z1, c_mc.chol[4], fma2(z2, c_mc.chol[5], mu0[2])));
#pragma unroll
    for (int k = 0; k < 3; ++k) { st.mu[k] = bc(mu0[k]); st.real[k] = (EST == MCSIM_EST_CLOSEDLOOP) ? st.p[k] : bc(mu0[k]); }
    st.S[0] = bc(c_mc.chol[0] * c_mc.chol[0]);
    st.S[1] = bc(c_mc.chol[0] * c_mc.chol[1]);
    st.S[2] = bc(c_mc.chol[0] * c_mc.chol[3]);
    st.S[3] = bc(c_mc.chol[1] * c_mc.chol[1] + c_mc.chol[2] * c_mc.chol[2]);
    st.S[4] = bc(c_mc.chol[1] * c_mc.chol[3] + c_mc.chol[2] * c_mc.chol[4]);
    st.S[5] = bc(c_mc.chol[3] * c_mc.chol[3] + c_mc.chol[4] * c_mc.chol[4] + c_mc.chol[5] * c_mc.chol[5]);
}

// one step t -> t+1 of both rollouts' real robot + EKF (+ particle); r = plan record of waypoint t
// (MC_PLAN2_STRIDE floats, the next record follows it)
template <int EST, int NL>
__device__ __forceinline__ void step2(Rollout2& st, const float* r, int t, unsigned long long id0, unsigned long long id1, unsigned k0, unsigned k1) {
    f2 (&p)[3] = st.p;
    f2 (&real)[3] = st.real;
    f2 (&mu)[3] = st.mu;
    f2 (&S)[6] = st.S;
    // ---- step t -> t+1 of this rollout's own real robot + EKF
    // noise: block 0 = (rot1, trans, rot2, obs 0), block 1 = obs 1..4, block 2 = obs 5..7
    f2 zn[4 * ((3 + NL + 3) / 4)];
#pragma unroll
    for (int jb = 0; jb < (3 + NL + 3) / 4; ++jb) {
        const Philox2 w = philox_pair(id0, id1, unsigned(t * 8 + jb), unsigned(mcsim_rng::PURPOSE_STEP), k0, k1);
        box_muller2(w.a.x, w.a.y, w.b.x, w.b.y, zn[4 * jb], zn[4 * jb + 1]);
        box_muller2(w.a.z, w.a.w, w.b.z, w.b.w, zn[4 * jb + 2], zn[4 * jb + 3]);
    }
    // applied control (generateL + :719-726): u = u_req = inverseOdometry(mu, x*_{t+1}); u*_0 at t = 0
    f2 u[3];
    if (t == 0) {
        u[0] = bc(r[3]); u[1] = bc(r[4]); u[2] = bc(r[5]);
    } else {
        const f2 dx = add2(neg2(mu[0]), r[MC_PLAN2_STRIDE + 0]);
        const f2 dy = add2(neg2(mu[1]), r[MC_PLAN2_STRIDE + 1]);
        u[0] = wrap2(sub2(atan2_2(dy, dx), mu[2]));
        u[1] = sqrt2(fma2(dx, dx, mul2(dy, dy)));
        u[2] = wrap2(sub2(add2(neg2(mu[2]), r[MC_PLAN2_STRIDE + 2]), u[0]));
    }
    // EKF predict: Sigma' = G Sigma G^T + V M V^T (M from the nominal command), mu' = g(mu, u)
    {
        f2 sm, cm;
        sincos2(add2(mu[2], u[0]), sm, cm);
        const f2 g02 = mul2(neg2(u[1]), sm), g12 = mul2(u[1], cm);
        const f2 t0 = fma2(g02, S[5], S[2]);  // new s02 before R
        const f2 t1 = fma2(g12, S[5], S[4]);  // new s12 before R
        f2 n00 = fma2(g02, add2(S[2], t0), S[0]);
        f2 n01 = fma2(g02, t1, fma2(g12, S[2], S[1]));
        f2 n11 = fma2(g12, add2(S[4], t1), S[3]);
        const f2 am = mul2(g02, r[6]), bm = mul2(g12, r[6]);
        const f2 c1 = mul2(cm, r[7]), d1 = mul2(sm, r[7]);
        S[0] = fma2(g02, am, fma2(cm, c1, n00));
        S[1] = fma2(g12, am, fma2(sm, c1, n01));
        S[2] = add2(t0, am);
        S[3] = fma2(g12, bm, fma2(sm, d1, n11));
        S[4] = add2(t1, bm);
        S[5] = add2(S[5], r[9]);
        mu[0] = fma2(u[1], cm, mu[0]);
        mu[1] = fma2(u[1], sm, mu[1]);
        mu[2] = wrap2(add2(add2(mu[2], u[0]), u[2]));
    }
    // sampleOdometry: u~ = u + z * sqrt(variance), variances from the APPLIED command
    f2 un[3];
    {
        const f2 r1 = mul2(u[0], u[0]), tr = mul2(u[1], u[1]), r2 = mul2(u[2], u[2]);
        const f2 v0 = fma2(r1, c_mc.nm.alpha[0], mul2(tr, c_mc.nm.alpha[1]));
        const f2 v1 = fma2(tr, c_mc.nm.alpha[2], mul2(add2(r1, r2), c_mc.nm.alpha[3]));
        const f2 v2 = fma2(r2, c_mc.nm.alpha[0], mul2(tr, c_mc.nm.alpha[1]));
        un[0] = fma2(zn[0], sqrt2(v0), u[0]);
        un[1] = fma2(zn[1], sqrt2(v1), u[1]);
        un[2] = fma2(zn[2], sqrt2(v2), u[2]);
    }
    move2(real, un);
    if (EST == MCSIM_EST_REFERENCE) move2(p, un);
    // 8 sequential scalar range updates (EKFupdate); z_l = |real - m_l| + sqrt(Q) n_l
#pragma unroll
    for (int l = 0; l < NL; ++l) {
        const float lx = c_mc.nm.lx[l], ly = c_mc.nm.ly[l];
        const f2 ex = add2(real[0], -lx), ey = add2(real[1], -ly);
        const f2 eq = fma2(ex, ex, mul2(ey, ey));
        const f2 z = fma2(zn[3 + l], c_mc.sqrtQ, mul2(eq, rsqrt2(eq)));
        const f2 dx = add2(mu[0], -lx), dy = add2(mu[1], -ly);
        const f2 qq = fma2(dx, dx, mul2(dy, dy));
        const f2 rs = rsqrt2(qq);
        const f2 h0v = mul2(dx, rs), h1v = mul2(dy, rs);
        const f2 a0 = fma2(S[0], h0v, mul2(S[1], h1v));
        const f2 a1 = fma2(S[1], h0v, mul2(S[3], h1v));
        const f2 a2 = fma2(S[2], h0v, mul2(S[4], h1v));
        const f2 inv = rcp2(fma2(h0v, a0, fma2(h1v, a1, c_mc.nm.Q)));
        const f2 innov = fma2(neg2(qq), rs, z);
        const f2 k0v = mul2(a0, inv), k1v = mul2(a1, inv), k2v = mul2(a2, inv);
        mu[0] = fma2(k0v, innov, mu[0]);
        mu[1] = fma2(k1v, innov, mu[1]);
        mu[2] = fma2(k2v, innov, mu[2]);
        S[0] = fma2(neg2(k0v), a0, S[0]);
        S[1] = fma2(neg2(k0v), a1, S[1]);
        S[2] = fma2(neg2(k0v), a2, S[2]);
        S[3] = fma2(neg2(k1v), a1, S[3]);
        S[4] = fma2(neg2(k1v), a2, S[4]);
        S[5] = fma2(neg2(k2v), a2, S[5]);
    }
}

template <bool EARLY_EXIT, int EST, int NL>
__global__ void __launch_bounds__(128, MC_P2_CTAS_PER_SM) mc_philox2_kernel(const McLaunch a) {
    extern __shared__ __align__(16) float4 s_tab[];
    __shared__ unsigned s_cnt[2];
    __shared__ unsigned s_hist[MCSIM_MAX_WAYPOINTS];
    const int T = c_mc.T;
    for (int i = threadIdx.x; i < a.sat2_n; i += blockDim.x) s_tab[i] = a.sat2[i];
    for (int i = threadIdx.x; i < T; i += blockDim.x) s_hist[i] = 0;
    if (threadIdx.x < 2) s_cnt[threadIdx.x] = 0;
    __syncthreads();

    unsigned my_hits = 0, my_steps = 0;
    const unsigned n_pairs = c_scene.n_pairs;
    const unsigned long long n_thr = (a.n + 1) >> 1;
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    // block-uniform trip count (every thread of a CTA runs the same number of passes; lanes past the
    // end are masked by valid0 / valid1): all control flow below is warp-uniform, which lets ptxas keep
    // the table reads on the uniform datapath (LDCU) and broadcast them as FFMA2 operands
    for (unsigned long long base = (unsigned long long)blockIdx.x * blockDim.x; base < n_thr; base += stride) {
        const unsigned long long j = base + threadIdx.x;
        const unsigned long long i0 = 2 * j, i1 = 2 * j + 1;
        const bool valid0 = i0 < a.n;
        const bool valid1 = i1 < a.n;
        const unsigned long long id0 = a.offset + i0, id1 = a.offset + i1;
        Rollout2 st;
        init2<EST>(st, c_mc.mu0, id0, id1, a.k0, a.k1);

        unsigned cnt0 = 0, cnt1 = 0;
        int first0 = -1, first1 = -1;
        for (int t = 0; t < T; ++t) {
            // ---- collision test at waypoint t
            const f2* q = (EST == MCSIM_EST_CLOSEDLOOP) ? st.real : st.p;
            f2 sn, cs;
            sincos2(q[2], sn, cs);
            bool h0 = false, h1 = false;
            robot_hits2(s_tab, q[0], q[1], cs, sn, h0, h1);
            if (EARLY_EXIT) {
                my_steps += (valid0 && cnt0 == 0) + (valid1 && cnt1 == 0);
            }
            if (h0) { if (first0 < 0) first0 = t; ++cnt0; }
            if (h1) { if (first1 < 0) first1 = t; ++cnt1; }
            if (EARLY_EXIT && __all_sync(0xffffffffu, (cnt0 || !valid0) && (cnt1 || !valid1))) break;
            if (t == T - 1) break;

            step2<EST, NL>(st, c_plan2 + t * MC_PLAN2_STRIDE, t, id0, id1, a.k0, a.k1);
        }
        if (!EARLY_EXIT) my_steps += (unsigned(valid0) + unsigned(valid1)) * unsigned(T);
        if (cnt0 && valid0) { ++my_hits; atomicAdd(&s_hist[first0], 1u); }
        if (cnt1 && valid1) { ++my_hits; atomicAdd(&s_hist[first1], 1u); }
        if (a.counts) {
            if (valid0) a.counts[i0] = EARLY_EXIT ? (cnt0 != 0) : cnt0;
            if (valid1) a.counts[i1] = EARLY_EXIT ? (cnt1 != 0) : cnt1;
        }
    }

    // ---- aggregation: warp reduce -> shared -> one 64-bit atomic per counter per block
    my_hits = __reduce_add_sync(0xffffffffu, my_hits);
    unsigned long long w_steps = my_steps;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) w_steps += __shfl_xor_sync(0xffffffffu, w_steps, o);
    if ((threadIdx.x & 31) == 0) {
        atomicAdd(&s_cnt[0], my_hits);
        atomicAdd(&a.counters[CNT_STEPS], w_steps);
        atomicAdd(&a.counters[CNT_TESTS], w_steps * n_pairs);  // exhaustive: every live pair at every executed step
    }
    __syncthreads();
    if (threadIdx.x == 0 && s_cnt[0]) atomicAdd(&a.counters[CNT_HITS], (unsigned long long)s_cnt[0]);
    for (int t = threadIdx.x; t < T; t += blockDim.x)
        if (s_hist[t]) atomicAdd(&a.counters[CNT_HIST + t], (unsigned long long)s_hist[t]);
}

#endif
