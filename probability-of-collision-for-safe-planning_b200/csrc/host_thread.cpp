// host_thread.cpp -- the shared closed-loop "real robot" of the reference estimator, FP64, host.
//
// In the reference one real robot is simulated per run and ITS noisy command moves every particle
// (mcsimplugin/MCSimulator.h:754,760).  In replay (deterministic) mode that thread is
// O(T) scalar work shared by all rollouts, so it runs here on the host once per run and only the
// resulting command table u~[3 x (T-1)] goes to the GPU.  The same arithmetic, templated on the
// scalar type, is what each rollout carries in registers in Philox mode (mc_rollout_sm100.cu).
//
// Follows, statement by statement (FP64, same evaluation order where it matters):
//   angleWrap            MCSimulator.h:56-65      prediction        :413-431
//   sampleOdometry       :391-410                 inverseOdometry   :434-449
//   generateL + applied  :532-553, :719-726       generateM_EKF     :495-513
//   EKFpredict           :868-881 (G :517-529, V :453-468)
//   observation/sample   :368-387                 EKFupdate         :883-929 (H row :470-492)
#include "mcsim_cuda.h"
#include "estimator_math.h"

#include <cmath>
#include <cstring>
#include <vector>

extern "C" int mcsim_host_thread(const mcsim_mc_params* p, const double* normals, size_t n_normals,
                                 double* applied_noisy, double* applied, double* real_path,
                                 double* mu_path, double* cov_path, double* obs) {
    if (!p || !normals || p->T < 1 || !p->trajectory || !p->odometry) return MCSIM_ERR_ARG;
    const int T = p->T, L = p->n_landmarks;
    if (L < 0 || L > MCSIM_MAX_LANDMARKS || (L > 0 && !p->landmarks)) return MCSIM_ERR_ARG;
    if (n_normals < size_t(T - 1) * size_t(3 + L)) return MCSIM_ERR_DRAWS;

    using namespace mcsim_math;
    NoiseModel<double> nm;
    for (int i = 0; i < 4; ++i) nm.alpha[i] = p->alphas[i];
    nm.Q = p->Q;
    nm.L = L;
    for (int l = 0; l < L; ++l) { nm.lx[l] = p->landmarks[l]; nm.ly[l] = p->landmarks[L + l]; }

    // mu = initialmu = trajectory column 0 (:164, :656); cov = initialcovariance (:657)
    double mu[3] = {p->trajectory[0], p->trajectory[T], p->trajectory[2 * T]};
    double S[9];
    std::memcpy(S, p->cov0, sizeof(S));
    double real[3] = {mu[0], mu[1], mu[2]};  // realstate = mu (:685)

    auto put_state = [&](int t) {
        if (real_path) for (int k = 0; k < 3; ++k) real_path[k * T + t] = real[k];
        if (mu_path) for (int k = 0; k < 3; ++k) mu_path[k * T + t] = mu[k];
        if (cov_path) std::memcpy(cov_path + size_t(t) * 9, S, sizeof(S));
    };
    put_state(0);

    const double* z = normals;
    for (int t = 0; t < T - 1; ++t) {
        const double ustar[3] = {p->odometry[t], p->odometry[(T - 1) + t], p->odometry[2 * (T - 1) + t]};
        const double xs[3] = {p->trajectory[t], p->trajectory[T + t], p->trajectory[2 * T + t]};
        const double xg[3] = {p->trajectory[t + 1], p->trajectory[T + t + 1], p->trajectory[2 * T + t + 1]};
        double M[3];
        motion_noise(nm, ustar, M);              // generateM_EKF on the NOMINAL command (:698-701)
        double u[3];
        applied_control(xs, mu, xg, ustar, u);   // gain + applied control (:706-726)
        double pmu[3], pS[9];
        ekf_predict(mu, S, u, M, pmu, pS);       // :746
        double un[3];
        sample_odometry(nm, u, z, un);           // 3 draws, order rot1, trans, rot2 (:401-403)
        z += 3;
        double nxt[3];
        predict_pose(real, un, nxt);             // :405
        for (int k = 0; k < 3; ++k) real[k] = nxt[k];   // :776
        double zobs[MCSIM_MAX_LANDMARKS];
        for (int l = 0; l < L; ++l) zobs[l] = range_to(nm, real, l) + z[l] * std::sqrt(nm.Q);  // :383-387
        z += L;
        ekf_update(nm, pmu, pS, zobs, L);          // in place, 8 sequential scalar updates (:883-929)
        for (int k = 0; k < 3; ++k) mu[k] = pmu[k];
        std::memcpy(S, pS, sizeof(S));
        if (applied_noisy) for (int k = 0; k < 3; ++k) applied_noisy[k * (T - 1) + t] = un[k];
        if (applied) for (int k = 0; k < 3; ++k) applied[k * (T - 1) + t] = u[k];
        if (obs) for (int l = 0; l < L; ++l) obs[size_t(t) * L + l] = zobs[l];
        put_state(t + 1);
    }
    return MCSIM_OK;
}
