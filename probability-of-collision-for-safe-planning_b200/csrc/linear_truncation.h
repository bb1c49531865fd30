// linear_truncation.h -- analytic truncation of a Gaussian belief against linear constraints, host and
// device.  The reference has only a predicate stub for this (linConstraint.py:15-30, gaussprop.py:58-73, never called
// from C++); the arithmetic follows the method its slides cite (Patil, van den Berg, Alterovitz, "Estimating probability
// of collision for safe motion planning under Gaussian motion and sensing uncertainty", ICRA 2012, Sec. III-C): project
// N(mean, cov) on the constraint normal a, truncate the 1-D marginal y = a^T x to the feasible side y <= b, and propagate
// the truncated moments back through the conditional x | y; the changes of all constraints are computed from the
// UNTRUNCATED belief and summed (order-independent).  PARITY UNPINNED: no reference implementation exists; checked
// against the closed-form truncated normal (scipy), sampled moments (tests/test_linear_truncation.py) and the sampled
// truncation of the GMM mode on a half-plane scene (tests/test_variants_gpu.py).
#ifndef MCSIM_LINEAR_TRUNCATION_H
#define MCSIM_LINEAR_TRUNCATION_H

#include <math.h>

#ifdef __CUDACC__
#define MCSIM_LT_HD __host__ __device__ inline
#else
#define MCSIM_LT_HD inline
#endif

namespace mcsim_lt {

MCSIM_LT_HD double pdf(double z) { return 0.3989422804014327 * exp(-0.5 * z * z); }
MCSIM_LT_HD double cdf(double z) { return 0.5 * erfc(-z * 0.7071067811865476); }

// Returns 0 = ok, 2 = a constraint normal with no variance along it, 5 = no feasible mass left.  free_mass (optional):
// product over the constraints of the feasible mass Phi(alpha_k) of each 1-D marginal (the belief's collision-free
// fraction under the same independence the summed updates assume).
MCSIM_LT_HD int truncate(const double mean[3], const double cov[9], int n, const double* A, const double* b,
                         double mean_out[3], double cov_out[9], double* p_violation, double* free_mass) {
    double dm[3] = {0, 0, 0}, dC[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0}, fm = 1.0;
    for (int k = 0; k < n; ++k) {
        const double* a = A + 3 * k;
        double Sa[3], mu_y = 0, var_y = 0;
        for (int i = 0; i < 3; ++i) {
            Sa[i] = cov[3 * i] * a[0] + cov[3 * i + 1] * a[1] + cov[3 * i + 2] * a[2];
            mu_y += a[i] * mean[i];
        }
        for (int i = 0; i < 3; ++i) var_y += a[i] * Sa[i];
        if (!(var_y > 0)) return 2;
        const double sd = sqrt(var_y), alpha = (b[k] - mu_y) / sd;
        const double Phi = cdf(alpha);
        if (p_violation) p_violation[k] = 1.0 - Phi;
        fm *= Phi;
        if (!(Phi > 1e-300)) return 5;
        const double lambda = pdf(alpha) / Phi;
        const double m = mu_y - lambda * sd;                                  // E[y | y <= b]
        const double s2 = var_y * (1.0 - alpha * lambda - lambda * lambda);   // Var[y | y <= b]
        const double gm = (mu_y - m) / var_y, gc = (var_y - s2) / (var_y * var_y);
        for (int i = 0; i < 3; ++i) {
            dm[i] += Sa[i] * gm;
            for (int j = 0; j < 3; ++j) dC[3 * i + j] += Sa[i] * Sa[j] * gc;
        }
    }
    for (int i = 0; i < 3; ++i) mean_out[i] = mean[i] - dm[i];
    for (int i = 0; i < 9; ++i) cov_out[i] = cov[i] - dC[i];
    if (free_mass) *free_mass = fm;
    return 0;
}

}  // namespace mcsim_lt
#endif
