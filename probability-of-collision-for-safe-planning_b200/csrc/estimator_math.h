// estimator_math.h -- scalar arithmetic of the odometry motion model, controller and range-only EKF,
// templated on the scalar type and usable from host (FP64 host thread) and device (FP32 per-rollout
// thread of the Philox kernel, FP64 per-component EKF of the GMM path).
//
// Each function states the reference lines it reproduces (mcsimplugin/MCSimulator.h).
#ifndef MCSIM_ESTIMATOR_MATH_H
#define MCSIM_ESTIMATOR_MATH_H

#include <math.h>

#ifdef __CUDACC__
#define MCSIM_HD __host__ __device__ __forceinline__
#else
#define MCSIM_HD inline
#endif

#ifndef MCSIM_MAX_LANDMARKS
#define MCSIM_MAX_LANDMARKS 16
#endif

namespace mcsim_math {

// #define MPI 3.14159265358979323846 (MCSimulator.h:43)
#define MCSIM_PI 3.14159265358979323846

template <typename T>
struct NoiseModel {
    T alpha[4];  // squared alphas
    T Q;         // range variance
    int L;
    T lx[MCSIM_MAX_LANDMARKS];
    T ly[MCSIM_MAX_LANDMARKS];
};

MCSIM_HD void sincos_t(double a, double& s, double& c) { s = sin(a); c = cos(a); }
MCSIM_HD void sincos_t(float a, float& s, float& c) {
#ifdef __CUDA_ARCH__
    sincosf(a, &s, &c);
#else
    s = sinf(a); c = cosf(a);
#endif
}
MCSIM_HD double sqrt_t(double a) { return sqrt(a); }
MCSIM_HD float sqrt_t(float a) { return sqrtf(a); }
MCSIM_HD double atan2_t(double y, double x) { return atan2(y, x); }
MCSIM_HD float atan2_t(float y, float x) { return atan2f(y, x); }

// angleWrap (:56-65): closed interval [0, 2*pi]; 2*pi itself is kept.
template <typename T>
MCSIM_HD T angle_wrap(T a) {
    const T two_pi = T(2 * MCSIM_PI);
    while (a < T(0)) a += two_pi;
    while (a > two_pi) a -= two_pi;
    return a;
}

// prediction (:413-431) == moveParticles (:300-322) for one column.
template <typename T>
MCSIM_HD void predict_pose(const T s[3], const T u[3], T out[3]) {
    T sn, cs;
    sincos_t(s[2] + u[0], sn, cs);
    out[0] = s[0] + u[1] * cs;
    out[1] = s[1] + u[1] * sn;
    out[2] = angle_wrap(s[2] + u[0] + u[2]);
}

// inverseOdometry (:434-449)
template <typename T>
MCSIM_HD void inverse_odometry(const T p1[3], const T p2[3], T u[3]) {
    const T dx = p2[0] - p1[0], dy = p2[1] - p1[1];
    const T r1 = angle_wrap(atan2_t(dy, dx) - p1[2]);
    u[0] = r1;
    u[1] = sqrt_t(dx * dx + dy * dy);
    u[2] = angle_wrap(p2[2] - p1[2] - r1);
}

// generateL (:532-553) followed by appliedcontrol = u* + L (mu - x*) (:719-726).
// The divide-then-multiply order is kept: it is what the reference computes.
template <typename T>
MCSIM_HD void applied_control(const T xs[3], const T mu[3], const T xg[3], const T ustar[3], T u[3]) {
    T ureq[3];
    inverse_odometry(mu, xg, ureq);
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        const T xhat = mu[i] - xs[i];
        const T ubar = ureq[i] - ustar[i];
        const T gain = ubar / (xhat != T(0) ? xhat : T(0.1));
        u[i] = ustar[i] + gain * xhat;
    }
}

// generateM_EKF (:495-513): diagonal of M, from the NOMINAL command.
template <typename T>
MCSIM_HD void motion_noise(const NoiseModel<T>& nm, const T u[3], T M[3]) {
    const T r1 = u[0] * u[0], tr = u[1] * u[1], r2 = u[2] * u[2];
    M[0] = nm.alpha[0] * r1 + nm.alpha[1] * tr;
    M[1] = nm.alpha[2] * tr + nm.alpha[3] * r1 + nm.alpha[3] * r2;
    M[2] = nm.alpha[0] * r2 + nm.alpha[1] * tr;
}

// sampleOdometry (:391-403): noisy command from the APPLIED command; z = 3 standard normals in the
// order rot1, trans, rot2; sampleNormal = mean + randn * sqrt(variance) (:51-53).
template <typename T>
MCSIM_HD void sample_odometry(const NoiseModel<T>& nm, const T u[3], const T* z, T un[3]) {
    const T r1 = u[0] * u[0], tr = u[1] * u[1], r2 = u[2] * u[2];
    un[0] = u[0] + z[0] * sqrt_t(nm.alpha[0] * r1 + nm.alpha[1] * tr);
    un[1] = u[1] + z[1] * sqrt_t(nm.alpha[2] * tr + nm.alpha[3] * (r1 + r2));
    un[2] = u[2] + z[2] * sqrt_t(nm.alpha[0] * r2 + nm.alpha[1] * tr);
}

// EKFpredict (:868-881) with G (:517-529), V (:453-468), R = V M V^T, Sigma' = G Sigma G^T + R.
// S and pS are full 3x3 row-major: the reference's non-Joseph update leaves Sigma slightly
// asymmetric and its products use every entry.
template <typename T>
MCSIM_HD void ekf_predict(const T mu[3], const T S[9], const T u[3], const T M[3], T pmu[3], T pS[9]) {
    T sn, cs;
    sincos_t(mu[2] + u[0], sn, cs);
    const T g02 = -u[1] * sn, g12 = u[1] * cs;
    // GS = G * S
    T GS[9];
#pragma unroll
    for (int j = 0; j < 3; ++j) {
        GS[0 + j] = S[0 + j] + g02 * S[6 + j];
        GS[3 + j] = S[3 + j] + g12 * S[6 + j];
        GS[6 + j] = S[6 + j];
    }
    // V = [[g02, cs, 0], [g12, sn, 0], [1, 0, 1]];  R_ij = sum_k V_ik M_k V_jk
    const T V[9] = {g02, cs, T(0), g12, sn, T(0), T(1), T(0), T(1)};
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        // (G S G^T)_i. = [GS_i0 + g02 GS_i2, GS_i1 + g12 GS_i2, GS_i2]
        const T a0 = GS[3 * i + 0] + GS[3 * i + 2] * g02;
        const T a1 = GS[3 * i + 1] + GS[3 * i + 2] * g12;
        const T a2 = GS[3 * i + 2];
        const T a[3] = {a0, a1, a2};
#pragma unroll
        for (int j = 0; j < 3; ++j) {
            const T r = V[3 * i + 0] * M[0] * V[3 * j + 0] + V[3 * i + 1] * M[1] * V[3 * j + 1] +
                        V[3 * i + 2] * M[2] * V[3 * j + 2];
            pS[3 * i + j] = a[j] + r;
        }
    }
    predict_pose(mu, u, pmu);
}

// observation (:368-381)
template <typename T>
MCSIM_HD T range_to(const NoiseModel<T>& nm, const T s[3], int l) {
    const T dx = s[0] - nm.lx[l], dy = s[1] - nm.ly[l];
    return sqrt_t(dx * dx + dy * dy);
}

// EKFupdate (:883-929): L sequential scalar range updates, in place; H row from makeHRow (:470-492);
// K = Sigma H^T * (1/S); Sigma <- (I - K H) Sigma (non-Joseph); no angle wrap afterwards.
// One landmark (the loop body); callers with a compile-time landmark count unroll it themselves.
template <typename T>
MCSIM_HD void ekf_update_one(const T lx, const T ly, const T Q, T mu[3], T S[9], const T zl) {
    const T dx = mu[0] - lx, dy = mu[1] - ly;
    const T q = dx * dx + dy * dy;
    const T rq = sqrt_t(q);
    const T h0 = dx / rq, h1 = dy / rq;
    // H Sigma (1x3), then S = H Sigma H^T + Q
    const T hs0 = h0 * S[0] + h1 * S[3];
    const T hs1 = h0 * S[1] + h1 * S[4];
    const T s = hs0 * h0 + hs1 * h1 + Q;
    const T inv = T(1) / s;
    // K = Sigma H^T / S
    const T k0 = (S[0] * h0 + S[1] * h1) * inv;
    const T k1 = (S[3] * h0 + S[4] * h1) * inv;
    const T k2 = (S[6] * h0 + S[7] * h1) * inv;
    const T innov = zl - rq;
    mu[0] += k0 * innov;
    mu[1] += k1 * innov;
    mu[2] += k2 * innov;
    // A = I - K H (third column of H is 0), Sigma <- A Sigma
    const T a00 = T(1) - k0 * h0, a01 = -k0 * h1;
    const T a10 = -k1 * h0, a11 = T(1) - k1 * h1;
    const T a20 = -k2 * h0, a21 = -k2 * h1;
    T N[9];
#pragma unroll
    for (int j = 0; j < 3; ++j) {
        N[0 + j] = a00 * S[0 + j] + a01 * S[3 + j];
        N[3 + j] = a10 * S[0 + j] + a11 * S[3 + j];
        N[6 + j] = a20 * S[0 + j] + a21 * S[3 + j] + S[6 + j];
    }
#pragma unroll
    for (int j = 0; j < 9; ++j) S[j] = N[j];
}

// EKFupdate for one landmark with ONE division on the dependent chain mu_l -> mu_{l+1} (the 8 updates of a step are
// serial per component, and in FP64 a square root and each division cost ~200 dependent cycles).  With d = (mu - m_l)_xy,
// q = |d|^2 and H = d^T / sqrt(q):  c = Sigma_{.,xy} d  (= Sigma H^T sqrt(q)),  r = d^T Sigma_{xy,.}  (= H Sigma sqrt(q)),
// D = d . r_xy + Q q  (= q S of :905),   K (z - zhat) = c (sqrt(q) z - q) / D,   K H Sigma = c r^T / D
// -- the reference's update (ekf_update_one above, non-Joseph, Sigma's asymmetry kept: c and r are computed separately)
// up to rounding; sqrt(q) is off the chain.  Used by the device GMM path in Philox mode, where the bar is statistical;
// replay mode and the real-robot thread keep the reference's order of operations.
template <typename T>
MCSIM_HD void ekf_update_one_chain1(const T lx, const T ly, const T Q, T mu[3], T S[9], const T zl) {
    const T dx = mu[0] - lx, dy = mu[1] - ly;
    const T q = dx * dx + dy * dy;
    const T c0 = S[0] * dx + S[1] * dy, c1 = S[3] * dx + S[4] * dy, c2 = S[6] * dx + S[7] * dy;
    const T r0 = dx * S[0] + dy * S[3], r1 = dx * S[1] + dy * S[4], r2 = dx * S[2] + dy * S[5];
    const T inv = T(1) / (dx * r0 + dy * r1 + Q * q);
    const T w = (sqrt_t(q) * zl - q) * inv;
    mu[0] += c0 * w;
    mu[1] += c1 * w;
    mu[2] += c2 * w;
    const T b0 = c0 * inv, b1 = c1 * inv, b2 = c2 * inv;
    S[0] -= b0 * r0; S[1] -= b0 * r1; S[2] -= b0 * r2;
    S[3] -= b1 * r0; S[4] -= b1 * r1; S[5] -= b1 * r2;
    S[6] -= b2 * r0; S[7] -= b2 * r1; S[8] -= b2 * r2;
}

// ---- deviation-state filter (method variant; the reference has none) ----------------------------
// The paper's first item of future work (finalpaper/ajaay_paper.tex:893, 901): "produce a KF estimate of the deviation of
// a robot's configuration from the nominal configuration", as Patil et al. / van den Berg's LQG-MP do.  State
// delta_t = x_t - x*_t; motion and sensor models are linearised ONCE about the nominal plan (x*_t, u*_t), not about the
// current estimate:
//   predict   delta' = G* delta + V* (u - u*_t),   Sigma' = G* Sigma G*^T + V* M V*^T,   G*, V* at (x*_t, u*_t)
//   update l  nu = z_l - |x*_{t+1} - m_l| - H*_l delta,  K = Sigma H*^T / (H* Sigma H*^T + Q),
//             delta += K nu,  Sigma <- (I - K H*) Sigma,  H*_l at x*_{t+1} for every landmark (no relinearisation)
// The filter is linear in the state, so mean and covariance of the belief do not interact.  Interfaces mirror
// ekf_predict / ekf_update_one with the mean kept as an absolute pose (mu = x* + delta): xs = x*_t, xn = x*_{t+1},
// us = u*_t, u = the command applied by the real robot.
template <typename T>
MCSIM_HD T wrap_pm_pi(T a) {
    while (a > T(MCSIM_PI)) a -= T(2 * MCSIM_PI);
    while (a <= -T(MCSIM_PI)) a += T(2 * MCSIM_PI);
    return a;
}
template <typename T>
MCSIM_HD void lkf_predict(const T mu[3], const T S[9], const T xs[3], const T xn[3], const T us[3], const T u[3], const T M[3], T pmu[3], T pS[9]) {
    T sn, cs;
    sincos_t(xs[2] + us[0], sn, cs);
    const T g02 = -us[1] * sn, g12 = us[1] * cs;
    T GS[9];
#pragma unroll
    for (int j = 0; j < 3; ++j) {
        GS[0 + j] = S[0 + j] + g02 * S[6 + j];
        GS[3 + j] = S[3 + j] + g12 * S[6 + j];
        GS[6 + j] = S[6 + j];
    }
    const T V[9] = {g02, cs, T(0), g12, sn, T(0), T(1), T(0), T(1)};
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        const T a[3] = {GS[3 * i + 0] + GS[3 * i + 2] * g02, GS[3 * i + 1] + GS[3 * i + 2] * g12, GS[3 * i + 2]};
#pragma unroll
        for (int j = 0; j < 3; ++j)
            pS[3 * i + j] = a[j] + V[3 * i + 0] * M[0] * V[3 * j + 0] + V[3 * i + 1] * M[1] * V[3 * j + 1] + V[3 * i + 2] * M[2] * V[3 * j + 2];
    }
    // delta' = G* delta + V* du
    // heading and rotation deviations are differences of angles kept in [0, 2 pi] (angleWrap): taken on (-pi, pi]
    const T d0 = mu[0] - xs[0], d1 = mu[1] - xs[1], d2 = wrap_pm_pi(mu[2] - xs[2]);
    const T du0 = wrap_pm_pi(u[0] - us[0]), du1 = u[1] - us[1], du2 = wrap_pm_pi(u[2] - us[2]);
    pmu[0] = xn[0] + d0 + g02 * d2 + g02 * du0 + cs * du1;
    pmu[1] = xn[1] + d1 + g12 * d2 + g12 * du0 + sn * du1;
    pmu[2] = xn[2] + d2 + du0 + du2;
}
template <typename T>
MCSIM_HD void lkf_update_one(const T lx, const T ly, const T Q, const T xn[3], T mu[3], T S[9], const T zl) {
    const T dx = xn[0] - lx, dy = xn[1] - ly;
    const T rq = sqrt_t(dx * dx + dy * dy);
    const T h0 = dx / rq, h1 = dy / rq;
    const T hs0 = h0 * S[0] + h1 * S[3];
    const T hs1 = h0 * S[1] + h1 * S[4];
    const T inv = T(1) / (hs0 * h0 + hs1 * h1 + Q);
    const T k0 = (S[0] * h0 + S[1] * h1) * inv;
    const T k1 = (S[3] * h0 + S[4] * h1) * inv;
    const T k2 = (S[6] * h0 + S[7] * h1) * inv;
    const T innov = zl - rq - (h0 * (mu[0] - xn[0]) + h1 * (mu[1] - xn[1]));
    mu[0] += k0 * innov;
    mu[1] += k1 * innov;
    mu[2] += k2 * innov;
    const T a00 = T(1) - k0 * h0, a01 = -k0 * h1;
    const T a10 = -k1 * h0, a11 = T(1) - k1 * h1;
    const T a20 = -k2 * h0, a21 = -k2 * h1;
    T N[9];
#pragma unroll
    for (int j = 0; j < 3; ++j) {
        N[0 + j] = a00 * S[0 + j] + a01 * S[3 + j];
        N[3 + j] = a10 * S[0 + j] + a11 * S[3 + j];
        N[6 + j] = a20 * S[0 + j] + a21 * S[3 + j] + S[6 + j];
    }
#pragma unroll
    for (int j = 0; j < 9; ++j) S[j] = N[j];
}

template <typename T>
MCSIM_HD void ekf_update(const NoiseModel<T>& nm, T mu[3], T S[9], const T* z, const int L) {
#pragma unroll
    for (int l = 0; l < L; ++l) ekf_update_one(nm.lx[l], nm.ly[l], nm.Q, mu, S, z[l]);
}

}  // namespace mcsim_math
#endif
