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
// registers): x*_t (3), u*_t (3), M_t (3) = generateM_EKF(u*_t), M_0 + M_2, angleWrap(theta*_t)
#define MC_PLAN2_STRIDE 12
#ifndef MC_P2_CTAS_PER_SM
#define MC_P2_CTAS_PER_SM 5  // resident 128-thread CTAs per SM the register budget is set for
#endif
#define MC_PLAN2_MAX_T 256
__constant__ float c_plan2[MC_PLAN2_MAX_T * MC_PLAN2_STRIDE];

// two standard normals per lane from two 32-bit words per lane (Box-Muller)
__device__ __forceinline__ void box_muller2(unsigned a0, unsigned b0, unsigned a1, unsigned b1, f2& n0, f2& n1) {  // @phase boxmuller
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

// the same from the 24 radius + 16 angle bits of one step pair per lane (philox.h: step draws).
// MC_ANGLE_PRMT = 1 (default; 3.637 vs 3.628e10 rollout-steps/s): the 16-bit halves become floats on the ALU pipe (one PRMT each: 0x4B000000 | a =
// 2^23 + a, and the 2^23 leaves in a packed add); 0: plain conversion, for which ptxas picks I2F.U16
// (one instruction, but on the XU pipe next to the MUFUs).
#ifndef MC_ANGLE_PRMT
#define MC_ANGLE_PRMT 1
#endif
__device__ __forceinline__ float half16f(unsigned w, int h) {  // @phase draw_bits
#if MC_ANGLE_PRMT
    return __uint_as_float(__byte_perm(w, 0x4B000000u, h ? 0x7632u : 0x7610u));
#else
    return float(h ? (w >> 16) : (w & 0xffffu));
#endif
}
__device__ __forceinline__ void box_muller2_bits(unsigned r0, float a0, unsigned r1, float a1, f2& n0, f2& n1) {  // @phase boxmuller
    const f2 ua = fma2(mk2(float(r0), float(r1)), MCSIM_U24_SCALE, MCSIM_U24_HALF);
#if MC_ANGLE_PRMT
    const f2 th = fma2(add2(mk2(a0, a1), -8388608.0f), MCSIM_A16_SCALE, MCSIM_A16_HALF);
#else
    const f2 th = fma2(mk2(a0, a1), MCSIM_A16_SCALE, MCSIM_A16_HALF);
#endif
    const f2 r = sqrt2(mul2(mk2(mufu_lg2(lo(ua)), mufu_lg2(hi(ua))), -1.3862943611198906f));
    f2 s, c;
    sincos2(th, s, c);
    n0 = mul2(r, c);
    n1 = mul2(r, s);
}

struct Philox2 {
    mcsim_rng::u32x4 a, b;
};
__device__ __forceinline__ Philox2 philox_pair(unsigned long long id0, unsigned long long id1, unsigned blk, unsigned purpose, unsigned k0, unsigned k1) {  // @phase philox
    Philox2 o;
    const mcsim_rng::u32x4 c0 = {unsigned(id0), unsigned(id0 >> 32), blk, purpose};
    const mcsim_rng::u32x4 c1 = {unsigned(id1), unsigned(id1 >> 32), blk, purpose};
    o.a = mcsim_rng::philox4x32_10(c0, k0, k1);
    o.b = mcsim_rng::philox4x32_10(c1, k0, k1);
    return o;
}

// the 3 + NL standard normals of step t -> t+1 for both rollouts of a thread: 12 per two Philox blocks
// (philox.h: step draws); also what mcsim_cuda_step_normals dumps for the distribution tests
template <int NL>
__device__ __forceinline__ void step_normals2(f2* zn, int t, unsigned long long id0, unsigned long long id1, unsigned k0, unsigned k1) {  // @phase draw_bits
#pragma unroll
    for (int g = 0; g < (3 + NL + 11) / 12; ++g) {
        const Philox2 wa = philox_pair(id0, id1, unsigned(t * 8 + 2 * g), unsigned(mcsim_rng::PURPOSE_STEP), k0, k1);
        const Philox2 wb = philox_pair(id0, id1, unsigned(t * 8 + 2 * g + 1), unsigned(mcsim_rng::PURPOSE_STEP), k0, k1);
        unsigned rad0[6], ang0[6], rad1[6], ang1[6];
        mcsim_rng::step_bits(wa.a, wb.a, rad0, ang0);  // (the integer angle halves are dead code here)
        mcsim_rng::step_bits(wa.b, wb.b, rad1, ang1);
        const unsigned aw0[3] = {wb.a.y, wb.a.z, wb.a.w}, aw1[3] = {wb.b.y, wb.b.z, wb.b.w};
#pragma unroll
        for (int j = 0; j < 6; ++j)
            box_muller2_bits(rad0[j], half16f(aw0[j >> 1], j & 1), rad1[j], half16f(aw1[j >> 1], j & 1), zn[12 * g + 2 * j], zn[12 * g + 2 * j + 1]);
    }
}

// ---- packed SAT: both lanes against every live pair ------------------------------------------------
// The same separating-axis test as pair_hit (sat_device.cuh) on the orientation-grouped table
// (mcsim_cuda_set_scene): per class of boxes sharing an axis direction (mod 90 degrees) the relative
// rotation (c, s), the hull centre in the class frame (gx, gy) and the hull's radii on the class
// axes (rbx, rby) are computed once (12 packed instructions); per pair and per two rollouts that
// leaves 8 FFMA2 + 4 FADD2 and 8 FSETP.  Every live pair is still tested at every step.
// Table records (float4, shared memory): class header (cos, sin, count, -), then per pair
// (cx, cy, hx, hy), (-p, -q, -, -) with p = cb cx + sb cy, q = cb cy - sb cx.
struct ClassFrame2 {
    f2 c, s, gx, gy, rbx, rby;
};
// per orientation class: |relative rotation|, hull centre in the class frame, hull radii on the class axes
// AX: every class axis of the scene is (1, 0) and every hull's yaw is 0 (an axis-aligned scene, the usual OpenRAVE env):
// the rotations into the class and hull frames multiply by 1 and add 0 -- dropped at compile time, same bits.
template <bool AX = false>
__device__ __forceinline__ ClassFrame2 class_frame2(const float4 head, f2 ch, f2 sh, f2 hcx, f2 hcy, f2 hhx, f2 hhy) {  // @phase sat_frames
    ClassFrame2 k;
    if (AX) {
        k.c = abs2(ch); k.s = abs2(sh); k.gx = hcx; k.gy = hcy;
        k.rbx = fma2(k.c, hhx, mul2(k.s, hhy));
        k.rby = fma2(k.s, hhx, mul2(k.c, hhy));
        return k;
    }
    k.c = abs2(fma2(ch, head.x, mul2(sh, head.y)));
    k.s = abs2(fma2(sh, head.x, mul2(neg2(ch), head.y)));
    k.gx = fma2(hcx, head.x, mul2(hcy, head.y));
    k.gy = fma2(hcy, head.x, mul2(neg2(hcx), head.y));
    k.rbx = fma2(k.c, hhx, mul2(k.s, hhy));
    k.rby = fma2(k.s, hhx, mul2(k.c, hhy));
    return k;
}
// one pair of a class, both lanes: A = (cx, cy, hx, hy), B = (-p, -q)
__device__ __forceinline__ void class_pair_test2(const ClassFrame2& k, f2 ch, f2 sh, f2 nph, f2 nqh, f2 hhx, f2 hhy, const float4 A, const float2 B, bool& hit0, bool& hit1) {  // @phase sat_pair
    const f2 drx = fma2(ch, A.x, fma2(sh, A.y, nph));
    const f2 dry = fma2(ch, A.y, fma2(neg2(sh), A.x, nqh));
    const f2 r0 = fma2(k.c, A.z, fma2(k.s, A.w, hhx));
    const f2 r1 = fma2(k.s, A.z, fma2(k.c, A.w, hhy));
    const f2 dbx = add2(k.gx, B.x);
    const f2 dby = add2(k.gy, B.y);
    const f2 r2 = add2(k.rbx, A.z);
    const f2 r3 = add2(k.rby, A.w);
    if ((fabsf(lo(drx)) <= lo(r0)) & (fabsf(lo(dry)) <= lo(r1)) & (fabsf(lo(dbx)) <= lo(r2)) & (fabsf(lo(dby)) <= lo(r3))) hit0 = true;
    if ((fabsf(hi(drx)) <= hi(r0)) & (fabsf(hi(dry)) <= hi(r1)) & (fabsf(hi(dbx)) <= hi(r2)) & (fabsf(hi(dby)) <= hi(r3))) hit1 = true;
}

struct HullFrame2 {
    f2 hcx, hcy, ch, sh, nph, nqh, hhx, hhy;
};
template <bool AX = false>
__device__ __forceinline__ HullFrame2 hull_frame2(const HullRec& hr, f2 x, f2 y, f2 ct, f2 st) {  // @phase sat_frames
    HullFrame2 h;
    h.hcx = fma2(ct, hr.ox, fma2(neg2(st), hr.oy, x));
    h.hcy = fma2(st, hr.ox, fma2(ct, hr.oy, y));
    h.ch = AX ? ct : fma2(ct, hr.cp, mul2(neg2(st), hr.sp));
    h.sh = AX ? st : fma2(st, hr.cp, mul2(ct, hr.sp));
    h.nph = neg2(fma2(h.ch, h.hcx, mul2(h.sh, h.hcy)));
    h.nqh = neg2(fma2(h.ch, h.hcy, mul2(neg2(h.sh), h.hcx)));
    h.hhx = bc(hr.hx);
    h.hhy = bc(hr.hy);
    return h;
}

template <bool AX = false>
__device__ __forceinline__ void robot_hits2(const float4* __restrict__ tab, f2 x, f2 y, f2 ct, f2 st, bool& hit0, bool& hit1) {  // @phase sat_exhaustive
    const int nh = c_scene.n_hulls;
    for (int hI = 0; hI < nh; ++hI) {
        const HullRec& hr = c_scene.hulls[hI];
        const HullFrame2 h = hull_frame2<AX>(hr, x, y, ct, st);
        const float4* g = tab + hr.g_first;
        for (int gi = 0; gi < hr.n_groups; ++gi) {
            const float4 head = g[0];
            const int cnt = __float_as_int(head.z);
            const ClassFrame2 k = class_frame2<AX>(head, h.ch, h.sh, h.hcx, h.hcy, h.hhx, h.hhy);
            ++g;
#pragma unroll 4
            for (int b = 0; b < cnt; ++b)
                class_pair_test2(k, h.ch, h.sh, h.nph, h.nqh, h.hhx, h.hhy, g[2 * b], *reinterpret_cast<const float2*>(g + 2 * b + 1), hit0, hit1);
            g += 2 * cnt;
        }
    }
}

struct HullPose2 {
    f2 hcx, hcy, ch, sh, nph, nqh;
    float hhx, hhy;
};

__device__ __forceinline__ HullPose2 place_hull2(const HullRec& hr, f2 x, f2 y, f2 ct, f2 st) {
    HullPose2 h;
    h.hcx = fma2(ct, hr.ox, fma2(neg2(st), hr.oy, x));
    h.hcy = fma2(st, hr.ox, fma2(ct, hr.oy, y));
    h.ch = fma2(ct, hr.cp, mul2(neg2(st), hr.sp));
    h.sh = fma2(st, hr.cp, mul2(ct, hr.sp));
    h.nph = neg2(fma2(h.ch, h.hcx, mul2(h.sh, h.hcy)));
    h.nqh = neg2(fma2(h.ch, h.hcy, mul2(neg2(h.sh), h.hcx)));
    h.hhx = hr.hx;
    h.hhy = hr.hy;
    return h;
}

// pair_hit (sat_device.cuh) for both lanes: 20 packed FP32 instructions + 8 FSETP
__device__ __forceinline__ void pair_test2(const HullPose2& h, const float4 lo4, const float4 hi4, bool& hit0, bool& hit1) {
    const f2 c = abs2(fma2(h.ch, lo4.z, mul2(h.sh, lo4.w)));
    const f2 s = abs2(fma2(h.sh, lo4.z, mul2(neg2(h.ch), lo4.w)));
    const f2 drx = fma2(h.ch, lo4.x, fma2(h.sh, lo4.y, h.nph));
    const f2 dry = fma2(h.ch, lo4.y, fma2(neg2(h.sh), lo4.x, h.nqh));
    const f2 dbx = fma2(h.hcx, lo4.z, fma2(h.hcy, lo4.w, hi4.z));
    const f2 dby = fma2(h.hcy, lo4.z, fma2(neg2(h.hcx), lo4.w, hi4.w));
    const f2 r0 = fma2(c, hi4.x, fma2(s, hi4.y, h.hhx));
    const f2 r1 = fma2(s, hi4.x, fma2(c, hi4.y, h.hhy));
    const f2 r2 = fma2(c, h.hhx, fma2(s, h.hhy, hi4.x));
    const f2 r3 = fma2(s, h.hhx, fma2(c, h.hhy, hi4.y));
    if ((fabsf(lo(drx)) <= lo(r0)) & (fabsf(lo(dry)) <= lo(r1)) & (fabsf(lo(dbx)) <= lo(r2)) & (fabsf(lo(dby)) <= lo(r3))) hit0 = true;
    if ((fabsf(hi(drx)) <= hi(r0)) & (fabsf(hi(dry)) <= hi(r1)) & (fabsf(hi(dbx)) <= hi(r2)) & (fabsf(hi(dby)) <= hi(r3))) hit1 = true;
}

// (exact broad phase: table layout, levels and the exactness argument are in sat_device.cuh)

// prediction (:413-431) on a packed pose
__device__ __forceinline__ void move2(f2 p[3], const f2 u[3]) {  // @phase move
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
__device__ __forceinline__ void init2(Rollout2& st, const float* mu0, unsigned long long id0, unsigned long long id1, unsigned k0, unsigned k1) {  // @phase init
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
    // noise: (rot1, trans, rot2, obs 0 .. NL-1), 12 normals per two Philox blocks (philox.h: step draws)
    f2 zn[12 * ((3 + NL + 11) / 12)];
    step_normals2<NL>(zn, t, id0, id1, k0, k1);
    // applied control (generateL + :719-726): u = u_req = inverseOdometry(mu, x*_{t+1}); u*_0 at t = 0  // @phase controller
    // The commanded translation points from mu at x*_{t+1}: cos / sin of its heading mu_2 + u_0 are (dx, dy) / |d| and
    // the predicted mean g(mu, u) IS x*_{t+1} (that is what the controller asks for), so neither costs a MUFU.SIN / COS
    // or a wrap; at t = 0 (u = u*_0, mu = x*_0) the general expressions are evaluated.
    f2 u[3], cm, sm, gx, gy, tr;   // gx, gy = u_1 (cos, sin)(mu_2 + u_0); tr = u_1^2
    const float* rn = r + MC_PLAN2_STRIDE;
    if (t == 0) {
        u[0] = bc(r[3]); u[1] = bc(r[4]); u[2] = bc(r[5]);
        sincos2(add2(mu[2], u[0]), sm, cm);
        gx = mul2(u[1], cm);
        gy = mul2(u[1], sm);
        tr = mul2(u[1], u[1]);
        mu[0] = add2(mu[0], gx);
        mu[1] = add2(mu[1], gy);
        mu[2] = wrap2(add2(add2(mu[2], u[0]), u[2]));
    } else {
        gx = add2(neg2(mu[0]), rn[0]);
        gy = add2(neg2(mu[1]), rn[1]);
        tr = fma2(gx, gx, mul2(gy, gy));
        const f2 ri = rsqrt2(mk2(fmaxf(lo(tr), 1e-30f), fmaxf(hi(tr), 1e-30f)));
        u[0] = wrap2(sub2(atan2_2(gy, gx), mu[2]));
        u[1] = mul2(tr, ri);
        u[2] = wrap2(sub2(add2(neg2(mu[2]), rn[2]), u[0]));
        cm = mul2(gx, ri);
        sm = mul2(gy, ri);
        mu[0] = bc(rn[0]);
        mu[1] = bc(rn[1]);
        mu[2] = bc(rn[10]);   // angleWrap(theta*_{t+1}) = angleWrap(mu_2 + u_0 + u_2) mod 2 pi
    }
    // EKF predict: Sigma' = G Sigma G^T + V M V^T (M from the nominal command); G_02 = -gy, G_12 = gx  // @phase predict
    {
        const f2 g02 = neg2(gy), g12 = gx;
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
    }
    // sampleOdometry: u~ = u + z * sqrt(variance), variances from the APPLIED command  // @phase odometry
    f2 un[3];
    {
        const f2 r1 = mul2(u[0], u[0]), r2 = mul2(u[2], u[2]);
        const f2 v0 = fma2(r1, c_mc.nm.alpha[0], mul2(tr, c_mc.nm.alpha[1]));
        const f2 v1 = fma2(tr, c_mc.nm.alpha[2], mul2(add2(r1, r2), c_mc.nm.alpha[3]));
        const f2 v2 = fma2(r2, c_mc.nm.alpha[0], mul2(tr, c_mc.nm.alpha[1]));
        un[0] = fma2(zn[0], sqrt2(v0), u[0]);
        un[1] = fma2(zn[1], sqrt2(v1), u[1]);
        un[2] = fma2(zn[2], sqrt2(v2), u[2]);
    }
    move2(real, un);
    if (EST == MCSIM_EST_REFERENCE) move2(p, un);
    // 8 sequential scalar range updates (EKFupdate); z_l = |real - m_l| + sqrt(Q) n_l  // @phase update
    // With d = (mu - m_l)_xy, q = |d|^2 and H = d^T / sqrt(q):  a~ = Sigma d,  D = d . a~ + Q q  (= q S of :905),
    //   K (z - zhat) = a~ (sqrt(q) z - q) / D,   K H Sigma = a~ a~^T / D
    // -- the same update with ONE reciprocal on the dependent chain mu_l -> mu_{l+1}; sqrt(q) and the sampled range
    // run beside it (the 8 updates are a serial chain per rollout: its latency, not its instruction count, is what
    // the kernel waits for)
#pragma unroll
    for (int l = 0; l < NL; ++l) {
        const float lx = c_mc.nm.lx[l], ly = c_mc.nm.ly[l];
        const f2 ex = add2(real[0], -lx), ey = add2(real[1], -ly);
        const f2 z = fma2(zn[3 + l], c_mc.sqrtQ, sqrt2(fma2(ex, ex, mul2(ey, ey))));
        const f2 dx = add2(mu[0], -lx), dy = add2(mu[1], -ly);
        const f2 a0 = fma2(S[0], dx, mul2(S[1], dy));
        const f2 a1 = fma2(S[1], dx, mul2(S[3], dy));
        const f2 a2 = fma2(S[2], dx, mul2(S[4], dy));
        const f2 qq = fma2(dx, dx, mul2(dy, dy));
        const f2 inv = rcp2(fma2(dx, a0, fma2(dy, a1, mul2(qq, c_mc.nm.Q))));
        const f2 w = fma2(sqrt2(qq), z, neg2(qq));
        const f2 b0 = mul2(a0, inv), b1 = mul2(a1, inv), b2 = mul2(a2, inv);
        mu[0] = fma2(b0, w, mu[0]);
        mu[1] = fma2(b1, w, mu[1]);
        mu[2] = fma2(b2, w, mu[2]);
        S[0] = fma2(neg2(b0), a0, S[0]);
        S[1] = fma2(neg2(b0), a1, S[1]);
        S[2] = fma2(neg2(b0), a2, S[2]);
        S[3] = fma2(neg2(b1), a1, S[3]);
        S[4] = fma2(neg2(b1), a2, S[4]);
        S[5] = fma2(neg2(b2), a2, S[5]);
    }
}

// shared-memory tables of the packed kernels: grouped pair table | candidate lists
struct SatTables {
    const float4* tab;
    const unsigned char* ccnt;      // [T][n_hulls][MC_CULL_LEVELS]
    const unsigned short* cidx;     // [T][n_hulls][MC_CULL_CAP] x (pair record, class header)
    int nh;
    unsigned n_pairs;
    bool cull;
};
__device__ __forceinline__ SatTables stage_tables(float4* s_tab, const McLaunch& a, int T) {
    SatTables tb;
    tb.nh = c_scene.n_hulls;
    tb.n_pairs = c_scene.n_pairs;
    tb.cull = a.cull != nullptr;
    unsigned* s_cull = reinterpret_cast<unsigned*>(s_tab + a.sat2_n);
    for (int i = threadIdx.x; i < a.sat2_n; i += blockDim.x) s_tab[i] = a.sat2[i];
    if (tb.cull)
        for (int i = threadIdx.x; i < a.cull_words; i += blockDim.x) s_cull[i] = a.cull[i];
    tb.tab = s_tab;
    tb.ccnt = reinterpret_cast<const unsigned char*>(s_cull);
    tb.cidx = reinterpret_cast<const unsigned short*>(s_cull + (T * tb.nh * MC_CULL_LEVELS + 3) / 4);
    return tb;
}

// collision test of both lanes at waypoint t (q = the tested pose).  in0 / in1: lanes whose result matters (they set
// the warp's list level); returns the pair tests executed per such lane.
template <bool AX = false>
__device__ __forceinline__ unsigned collide2(const SatTables& tb, const f2* q, int t, bool in0, bool in1, bool& h0, bool& h1) {  // @phase sat_select
    f2 sn, cs;
    sincos2(q[2], sn, cs);
    int lev = MC_CULL_LEVELS;
    const int nh = tb.nh;
    if (tb.cull) {
        // squared distance of the tested poses from x*_t, max over the warp's lanes that matter
        const float* r = c_plan2 + t * MC_PLAN2_STRIDE;
        const f2 ex = add2(q[0], -r[0]), ey = add2(q[1], -r[1]);
        const f2 d2 = fma2(ex, ex, mul2(ey, ey));
        const unsigned u0 = in0 ? __float_as_uint(lo(d2)) : 0u, u1 = in1 ? __float_as_uint(hi(d2)) : 0u;
        const float dm = __uint_as_float(__reduce_max_sync(0xffffffffu, u0 > u1 ? u0 : u1));  // NaN orders above every finite value
        lev = cull_level(dm);
        if (lev < MC_CULL_LEVELS && tb.ccnt[t * nh * MC_CULL_LEVELS + lev] == MC_CULL_OVERFLOW) lev = MC_CULL_LEVELS;
    }
    if (lev < MC_CULL_LEVELS) {
        unsigned tested = 0;
        for (int hI = 0; hI < nh; ++hI) {
            const int n = tb.ccnt[(t * nh + hI) * MC_CULL_LEVELS + lev];
            if (n == 0) continue;
            const HullFrame2 h = hull_frame2<AX>(c_scene.hulls[hI], q[0], q[1], cs, sn);
            const ushort2* ci = reinterpret_cast<const ushort2*>(tb.cidx) + (t * nh + hI) * MC_CULL_CAP;
            if (AX) {   // one class, and its frame is the hull's: nothing depends on the class header
                const ClassFrame2 cf = class_frame2<true>(make_float4(1.f, 0.f, 0.f, 0.f), h.ch, h.sh, h.hcx, h.hcy, h.hhx, h.hhy);
                for (int k = 0; k < n; ++k) {
                    const unsigned ox = ci[k].x;
                    class_pair_test2(cf, h.ch, h.sh, h.nph, h.nqh, h.hhx, h.hhy, tb.tab[ox], *reinterpret_cast<const float2*>(tb.tab + ox + 1), h0, h1);
                }
            } else {
                ClassFrame2 cf;
                unsigned cls = 0xffffffffu;
                for (int k = 0; k < n; ++k) {
                    const ushort2 o = ci[k];  // x: pair record, y: class header
                    if (o.y != cls) {         // consecutive candidates mostly share their class
                        cls = o.y;
                        cf = class_frame2<false>(tb.tab[o.y], h.ch, h.sh, h.hcx, h.hcy, h.hhx, h.hhy);
                    }
                    class_pair_test2(cf, h.ch, h.sh, h.nph, h.nqh, h.hhx, h.hhy, tb.tab[o.x], *reinterpret_cast<const float2*>(tb.tab + o.x + 1), h0, h1);
                }
            }
            tested += n;
        }
        return tested;
    }
    robot_hits2<AX>(tb.tab, q[0], q[1], cs, sn, h0, h1);
    return tb.n_pairs;
}

// warp reduce -> shared -> one 64-bit atomic per counter per block
__device__ __forceinline__ void flush_counters(const McLaunch& a, int T, unsigned my_hits, unsigned long long my_steps, unsigned long long my_tests,
                                               unsigned* s_cnt, const unsigned* s_hist) {  // @phase epilogue
    my_hits = __reduce_add_sync(0xffffffffu, my_hits);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        my_steps += __shfl_xor_sync(0xffffffffu, my_steps, o);
        my_tests += __shfl_xor_sync(0xffffffffu, my_tests, o);
    }
    if ((threadIdx.x & 31) == 0) {
        atomicAdd(&s_cnt[0], my_hits);
        atomicAdd(&a.counters[CNT_STEPS], my_steps);
        atomicAdd(&a.counters[CNT_TESTS], my_tests);  // executed pair tests
    }
    __syncthreads();
    if (threadIdx.x == 0 && s_cnt[0]) atomicAdd(&a.counters[CNT_HITS], (unsigned long long)s_cnt[0]);
    for (int t = threadIdx.x; t < T; t += blockDim.x)
        if (s_hist[t]) atomicAdd(&a.counters[CNT_HIST + t], (unsigned long long)s_hist[t]);
}

template <bool EARLY_EXIT, int EST, int NL, bool AX = false>
__global__ void __launch_bounds__(128, MC_P2_CTAS_PER_SM) mc_philox2_kernel(const McLaunch a) {
    extern __shared__ __align__(16) float4 s_tab[];  // @phase prologue
    __shared__ unsigned s_cnt[2];
    __shared__ unsigned s_hist[MCSIM_MAX_WAYPOINTS];
    const int T = c_mc.T;
    const SatTables tb = stage_tables(s_tab, a, T);
    for (int i = threadIdx.x; i < T; i += blockDim.x) s_hist[i] = 0;
    if (threadIdx.x < 2) s_cnt[threadIdx.x] = 0;
    __syncthreads();

    unsigned my_hits = 0;
    unsigned long long my_steps = 0, my_tests = 0;
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

        unsigned cnt0 = 0, cnt1 = 0, pass_steps = 0, pass_tests = 0;
        int first0 = -1, first1 = -1;
        for (int t = 0; t < T; ++t) {
            // ---- collision test at waypoint t
            const bool in0 = EARLY_EXIT ? (valid0 && cnt0 == 0) : valid0, in1 = EARLY_EXIT ? (valid1 && cnt1 == 0) : valid1;
            const unsigned live = unsigned(in0) + unsigned(in1);
            bool h0 = false, h1 = false;
            const unsigned tested = collide2<AX>(tb, (EST == MCSIM_EST_CLOSEDLOOP) ? st.real : st.p, t, in0, in1, h0, h1);
            pass_tests += live * tested;
            pass_steps += live;  // @phase bookkeeping
            if (h0) { if (first0 < 0) first0 = t; ++cnt0; }
            if (h1) { if (first1 < 0) first1 = t; ++cnt1; }
            if (EARLY_EXIT && __all_sync(0xffffffffu, (cnt0 || !valid0) && (cnt1 || !valid1))) break;
            if (t == T - 1) break;

            step2<EST, NL>(st, c_plan2 + t * MC_PLAN2_STRIDE, t, id0, id1, a.k0, a.k1);
        }
        my_steps += pass_steps;
        my_tests += pass_tests;
        if (cnt0 && valid0) { ++my_hits; atomicAdd(&s_hist[first0], 1u); }
        if (cnt1 && valid1) { ++my_hits; atomicAdd(&s_hist[first1], 1u); }
        if (a.counts) {
            if (valid0) a.counts[i0] = EARLY_EXIT ? (cnt0 != 0) : cnt0;
            if (valid1) a.counts[i1] = EARLY_EXIT ? (cnt1 != 0) : cnt1;
        }
    }
    flush_counters(a, T, my_hits, my_steps, my_tests, s_cnt, s_hist);
}

// ---- early exit with survivor compaction -----------------------------------------------------------------------
// "Each rollout exits early on its first hit" (north_star): inside one launch a warp can only leave when all 64 of
// its rollouts have hit, and on the committed plan 76 % of the warps keep a survivor to the last waypoint.  So the
// plan is cut into segments of MC_SEG waypoints; a launch runs ONE segment for the rollouts that are still free at
// its first waypoint and appends the survivors' state (15 floats + the 64-bit id = 68 bytes, structure of arrays)
// to a compacted buffer in HBM for the next launch, which therefore runs full warps again.  Philox counters are
// derived from (id, waypoint), so flags, hit totals and the first-hit histogram are bit-identical to the plain
// kernel's.  Traffic: 136 bytes per surviving rollout and segment boundary -- 2 ms of HBM time per 10^8 rollouts
// against 146 ms of arithmetic.
#ifndef MC_SEG
#define MC_SEG 4
#endif
#define MC_STATE_WORDS 17   // p[3], real[3], mu[3], S[6], id lo, id hi
struct McSegLaunch {
    McLaunch base;              // n = rollouts of the chunk (first segment), offset, keys, counters, tables
    int t0, t1;                 // waypoints [t0, t1) are tested; the state leaves at waypoint t1
    const float* in;            // survivors entering at t0: [MC_STATE_WORDS][cap]; NULL: first segment (init from ids)
    float* out;                 // survivors leaving at t1 (NULL: t1 == T)
    const unsigned* n_in;       // device: rollouts in `in`
    unsigned* n_out;            // device: rollouts appended to `out`
    unsigned cap;
};

template <int EST, int NL, bool AX = false>
__global__ void __launch_bounds__(128, MC_P2_CTAS_PER_SM) mc_philox2_segment_kernel(const McSegLaunch g) {
    extern __shared__ __align__(16) float4 s_tab[];  // @phase prologue
    __shared__ unsigned s_cnt[2];
    __shared__ unsigned s_hist[MCSIM_MAX_WAYPOINTS];
    __shared__ unsigned s_warp[4], s_base;
    const McLaunch& a = g.base;
    const int T = c_mc.T;
    const SatTables tb = stage_tables(s_tab, a, T);
    for (int i = threadIdx.x; i < T; i += blockDim.x) s_hist[i] = 0;
    if (threadIdx.x < 2) s_cnt[threadIdx.x] = 0;
    __syncthreads();

    unsigned my_hits = 0;
    unsigned long long my_steps = 0, my_tests = 0;
    const unsigned long long n = g.in ? (unsigned long long)*g.n_in : a.n;
    const unsigned long long n_thr = (n + 1) >> 1;
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    for (unsigned long long base = (unsigned long long)blockIdx.x * blockDim.x; base < n_thr; base += stride) {
        const unsigned long long j = base + threadIdx.x;
        const unsigned long long i0 = 2 * j, i1 = 2 * j + 1;
        const bool valid0 = i0 < n, valid1 = i1 < n;
        unsigned long long id0, id1;
        Rollout2 st;
        if (g.in == nullptr) {
            id0 = a.offset + i0; id1 = a.offset + i1;
            init2<EST>(st, c_mc.mu0, id0, id1, a.k0, a.k1);
        } else {   // @phase state_io
            const float2* in2 = reinterpret_cast<const float2*>(g.in);   // cap is even: slots 2j, 2j + 1 are one float2
            const size_t row = g.cap / 2;
            float2 v[MC_STATE_WORDS];
#pragma unroll
            for (int k = 0; k < MC_STATE_WORDS; ++k) v[k] = valid0 ? __ldcs(in2 + k * row + j) : make_float2(0.f, 0.f);
#pragma unroll
            for (int k = 0; k < 3; ++k) { st.p[k] = mk2(v[k].x, v[k].y); st.real[k] = mk2(v[3 + k].x, v[3 + k].y); st.mu[k] = mk2(v[6 + k].x, v[6 + k].y); }
#pragma unroll
            for (int k = 0; k < 6; ++k) st.S[k] = mk2(v[9 + k].x, v[9 + k].y);
            id0 = (unsigned long long)__float_as_uint(v[15].x) | ((unsigned long long)__float_as_uint(v[16].x) << 32);
            id1 = (unsigned long long)__float_as_uint(v[15].y) | ((unsigned long long)__float_as_uint(v[16].y) << 32);
        }
        bool free0 = valid0, free1 = valid1;   // not yet collided
        unsigned pass_steps = 0, pass_tests = 0;
        for (int t = g.t0; t < g.t1; ++t) {
            const unsigned live = unsigned(free0) + unsigned(free1);
            bool h0 = false, h1 = false;
            const unsigned tested = collide2<AX>(tb, (EST == MCSIM_EST_CLOSEDLOOP) ? st.real : st.p, t, free0, free1, h0, h1);
            pass_tests += live * tested;
            pass_steps += live;  // @phase bookkeeping
            if (h0 && free0) { free0 = false; ++my_hits; atomicAdd(&s_hist[t], 1u); if (a.counts) a.counts[id0 - a.offset] = 1u; }
            if (h1 && free1) { free1 = false; ++my_hits; atomicAdd(&s_hist[t], 1u); if (a.counts) a.counts[id1 - a.offset] = 1u; }
            if (__all_sync(0xffffffffu, !free0 && !free1)) break;
            if (t == T - 1) break;
            step2<EST, NL>(st, c_plan2 + t * MC_PLAN2_STRIDE, t, id0, id1, a.k0, a.k1);
        }
        my_steps += pass_steps;
        my_tests += pass_tests;
        if (g.out) {   // @phase state_io
            // survivors -> compacted slots: warp ballots, one atomic per CTA and pass
            const unsigned m0 = __ballot_sync(0xffffffffu, free0), m1 = __ballot_sync(0xffffffffu, free1);
            if (lane == 0) s_warp[warp] = __popc(m0) + __popc(m1);
            __syncthreads();
            if (threadIdx.x == 0) {
                const unsigned tot = s_warp[0] + s_warp[1] + s_warp[2] + s_warp[3];
                s_base = tot ? atomicAdd(g.n_out, tot) : 0u;
            }
            __syncthreads();
            unsigned slot = s_base;
            for (unsigned w = 0; w < warp; ++w) slot += s_warp[w];
            const unsigned below = (1u << lane) - 1u;
            const unsigned slot0 = slot + __popc(m0 & below), slot1 = slot + __popc(m0) + __popc(m1 & below);
            float* out = g.out;
            const size_t cap = g.cap;
            if (free0) {
#pragma unroll
                for (int k = 0; k < 3; ++k) { out[k * cap + slot0] = lo(st.p[k]); out[(3 + k) * cap + slot0] = lo(st.real[k]); out[(6 + k) * cap + slot0] = lo(st.mu[k]); }
#pragma unroll
                for (int k = 0; k < 6; ++k) out[(9 + k) * cap + slot0] = lo(st.S[k]);
                out[15 * cap + slot0] = __uint_as_float(unsigned(id0));
                out[16 * cap + slot0] = __uint_as_float(unsigned(id0 >> 32));
            }
            if (free1) {
#pragma unroll
                for (int k = 0; k < 3; ++k) { out[k * cap + slot1] = hi(st.p[k]); out[(3 + k) * cap + slot1] = hi(st.real[k]); out[(6 + k) * cap + slot1] = hi(st.mu[k]); }
#pragma unroll
                for (int k = 0; k < 6; ++k) out[(9 + k) * cap + slot1] = hi(st.S[k]);
                out[15 * cap + slot1] = __uint_as_float(unsigned(id1));
                out[16 * cap + slot1] = __uint_as_float(unsigned(id1 >> 32));
            }
            __syncthreads();   // s_warp / s_base are reused by the next pass
        }
    }
    flush_counters(a, T, my_hits, my_steps, my_tests, s_cnt, s_hist);
}

// test hook (mcsim_cuda_step_normals): the 12 step normals of rollouts id0 .. id0 + n - 1 at step t exactly as
// mc_philox2_kernel draws them (same device function: MUFU lg2 / sqrt / sin / cos on 24 + 16 bits)
__global__ void mc_step_normals_kernel(unsigned long long id0, unsigned long long n, int t, unsigned k0, unsigned k1, float* __restrict__ out) {
    const unsigned long long j = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    const unsigned long long i0 = 2 * j, i1 = 2 * j + 1;
    if (i0 >= n) return;
    f2 zn[12];
    step_normals2<8>(zn, t, id0 + i0, id0 + i1, k0, k1);
#pragma unroll
    for (int k = 0; k < 12; ++k) {
        out[i0 * 12 + k] = lo(zn[k]);
        if (i1 < n) out[i1 * 12 + k] = hi(zn[k]);
    }
}

#endif
