// linear_truncation.cpp -- host entry point of the analytic linear-constraint truncation (linear_truncation.h).
#include "linear_truncation.h"

#include "mcsim_cuda.h"

extern "C" int mcsim_truncate_linear(const double mean[3], const double cov[9], int n, const double* A, const double* b,
                                     double mean_out[3], double cov_out[9], double* p_violation) {
    if (!mean || !cov || !mean_out || !cov_out || n < 0 || (n && (!A || !b))) return MCSIM_ERR_ARG;
    const int rc = mcsim_lt::truncate(mean, cov, n, A, b, mean_out, cov_out, p_violation, nullptr);
    return rc == 0 ? MCSIM_OK : (rc == 5 ? MCSIM_ERR_DEGENERATE : MCSIM_ERR_ARG);
}
