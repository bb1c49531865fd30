// linear_truncation.cpp -- analytic truncation of a Gaussian belief against linear constraints
// (SURVEY.md A12).  The reference has only a predicate stub for this (linConstraint.py:15-30,
// gaussprop.py:58-73, never called from C++); the arithmetic follows the method its slides cite
// (Patil, van den Berg, Alterovitz, "Estimating probability of collision for safe motion planning
// under Gaussian motion and sensing uncertainty", ICRA 2012, Sec. III-C): project N(mean, cov) on the
// constraint normal a, truncate the 1-D marginal y = a^T x to the feasible side y <= b, and propagate
// the truncated moments back through the conditional x | y.  PARITY UNPINNED: no reference
// implementation exists; tests/test_linear_truncation.py checks it against the closed-form truncated
// normal (scipy) and against sampled moments.
#include <math.h>

#include "mcsim_cuda.h"

namespace {
const double kInvSqrt2Pi = 0.3989422804014327;
inline double pdf(double z) { return kInvSqrt2Pi * exp(-0.5 * z * z); }
inline double cdf(double z) { return 0.5 * erfc(-z * 0.7071067811865476); }
}  // namespace

extern "C" int mcsim_truncate_linear(const double mean[3], const double cov[9], int n, const double* A, const double* b,
                                     double mean_out[3], double cov_out[9], double* p_violation) {
    if (!mean || !cov || !mean_out || !cov_out || n < 0 || (n && (!A || !b))) return MCSIM_ERR_ARG;
    double dm[3] = {0, 0, 0}, dC[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    for (int k = 0; k < n; ++k) {
        const double* a = A + 3 * k;
        double Sa[3], mu_y = 0, var_y = 0;
        for (int i = 0; i < 3; ++i) {
            Sa[i] = cov[3 * i] * a[0] + cov[3 * i + 1] * a[1] + cov[3 * i + 2] * a[2];
            mu_y += a[i] * mean[i];
        }
        for (int i = 0; i < 3; ++i) var_y += a[i] * Sa[i];
        if (!(var_y > 0)) return MCSIM_ERR_ARG;
        const double sd = sqrt(var_y), alpha = (b[k] - mu_y) / sd;
        const double Phi = cdf(alpha);
        if (p_violation) p_violation[k] = 1.0 - Phi;
        if (!(Phi > 1e-300)) return MCSIM_ERR_DEGENERATE;  // no feasible mass left
        const double lambda = pdf(alpha) / Phi;
        const double m = mu_y - lambda * sd;                                  // E[y | y <= b]
        const double s2 = var_y * (1.0 - alpha * lambda - lambda * lambda);   // Var[y | y <= b]
        // deltas of this constraint, each computed from the UNTRUNCATED belief and summed (order-independent)
        const double gm = (mu_y - m) / var_y, gc = (var_y - s2) / (var_y * var_y);
        for (int i = 0; i < 3; ++i) {
            dm[i] += Sa[i] * gm;
            for (int j = 0; j < 3; ++j) dC[3 * i + j] += Sa[i] * Sa[j] * gc;
        }
    }
    for (int i = 0; i < 3; ++i) mean_out[i] = mean[i] - dm[i];
    for (int i = 0; i < 9; ++i) cov_out[i] = cov[i] - dC[i];
    return MCSIM_OK;
}
