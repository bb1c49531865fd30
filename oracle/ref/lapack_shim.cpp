// TEST INFRASTRUCTURE ONLY (oracle/): the handful of LAPACK entry points the vendored Armadillo
// references when the reference estimator is compiled (no system LAPACK exists in this image).
// Own code, written for the tiny (n <= 4) matrices on this path:
//   dpotrf_  <- arma::chol inside mvnrnd        (armadillo_bits/glue_mvnrnd_meat.hpp:96)
//   dpotrf_/dpotri_ <- inv_sympd, dgetrf_ <- log_det   (gmm_full_meat.hpp:980-1046 init_constants)
//   dsyevd_  <- eig_sym fallback of mvnrnd for a non-SPD covariance (glue_mvnrnd_meat.hpp:98-132)
//   dgetri_, dgesdd_, dpbtrf_ : referenced by header instantiations; dgetri_ implemented, the
//   other two abort loudly because nothing on the path reaches them.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <algorithm>

extern "C" {

void dpotrf_(const char* uplo, const int* n_, double* a, const int* lda_, int* info) {
    const int n = *n_, lda = *lda_;
    *info = 0;
    const bool lower = (*uplo == 'L' || *uplo == 'l');
#define A_(i, j) a[(i) + (j) * lda]
    if (lower) {
        for (int j = 0; j < n; ++j) {
            double d = A_(j, j);
            for (int k = 0; k < j; ++k) d -= A_(j, k) * A_(j, k);
            if (!(d > 0.0)) { *info = j + 1; return; }
            d = std::sqrt(d);
            A_(j, j) = d;
            for (int i = j + 1; i < n; ++i) {
                double s = A_(i, j);
                for (int k = 0; k < j; ++k) s -= A_(i, k) * A_(j, k);
                A_(i, j) = s / d;
            }
        }
    } else {
        for (int j = 0; j < n; ++j) {
            double d = A_(j, j);
            for (int k = 0; k < j; ++k) d -= A_(k, j) * A_(k, j);
            if (!(d > 0.0)) { *info = j + 1; return; }
            d = std::sqrt(d);
            A_(j, j) = d;
            for (int i = j + 1; i < n; ++i) {
                double s = A_(j, i);
                for (int k = 0; k < j; ++k) s -= A_(k, j) * A_(k, i);
                A_(j, i) = s / d;
            }
        }
    }
#undef A_
}

// inverse of a SPD matrix from its Cholesky factor (either triangle); both triangles are written
void dpotri_(const char* uplo, const int* n_, double* a, const int* lda_, int* info) {
    const int n = *n_, lda = *lda_;
    *info = 0;
    const bool lower = (*uplo == 'L' || *uplo == 'l');
    std::vector<double> L(n * n, 0.0), Li(n * n, 0.0);
    for (int i = 0; i < n; ++i)
        for (int j = 0; j <= i; ++j) L[i + j * n] = lower ? a[i + j * lda] : a[j + i * lda];
    // invert lower-triangular L by forward substitution, column by column
    for (int c = 0; c < n; ++c) {
        for (int i = c; i < n; ++i) {
            double s = (i == c) ? 1.0 : 0.0;
            for (int k = c; k < i; ++k) s -= L[i + k * n] * Li[k + c * n];
            if (L[i + i * n] == 0.0) { *info = i + 1; return; }
            Li[i + c * n] = s / L[i + i * n];
        }
    }
    // A^-1 = Li^T Li
    for (int i = 0; i < n; ++i)
        for (int j = 0; j < n; ++j) {
            double s = 0.0;
            for (int k = std::max(i, j); k < n; ++k) s += Li[k + i * n] * Li[k + j * n];
            a[i + j * lda] = s;
        }
}

void dgetrf_(const int* m_, const int* n_, double* a, const int* lda_, int* ipiv, int* info) {
    const int m = *m_, n = *n_, lda = *lda_;
    *info = 0;
    const int mn = std::min(m, n);
    for (int k = 0; k < mn; ++k) {
        int p = k;
        double best = std::fabs(a[k + k * lda]);
        for (int i = k + 1; i < m; ++i)
            if (std::fabs(a[i + k * lda]) > best) { best = std::fabs(a[i + k * lda]); p = i; }
        ipiv[k] = p + 1;
        if (best == 0.0) { if (*info == 0) *info = k + 1; continue; }
        if (p != k) for (int j = 0; j < n; ++j) std::swap(a[k + j * lda], a[p + j * lda]);
        for (int i = k + 1; i < m; ++i) {
            a[i + k * lda] /= a[k + k * lda];
            for (int j = k + 1; j < n; ++j) a[i + j * lda] -= a[i + k * lda] * a[k + j * lda];
        }
    }
}

void dgetri_(const int* n_, double* a, const int* lda_, const int* ipiv, double* work, const int* lwork, int* info) {
    const int n = *n_, lda = *lda_;
    *info = 0;
    if (*lwork == -1) { work[0] = double(std::max(1, n)); return; }
    // solve A X = I using the LU factors: P A = L U
    std::vector<double> X(n * n, 0.0);
    for (int c = 0; c < n; ++c) {
        std::vector<double> b(n, 0.0);
        b[c] = 1.0;
        for (int k = 0; k < n; ++k) std::swap(b[k], b[ipiv[k] - 1]);
        for (int i = 0; i < n; ++i) for (int k = 0; k < i; ++k) b[i] -= a[i + k * lda] * b[k];
        for (int i = n - 1; i >= 0; --i) {
            for (int k = i + 1; k < n; ++k) b[i] -= a[i + k * lda] * b[k];
            if (a[i + i * lda] == 0.0) { *info = i + 1; return; }
            b[i] /= a[i + i * lda];
        }
        for (int i = 0; i < n; ++i) X[i + c * n] = b[i];
    }
    for (int i = 0; i < n; ++i) for (int j = 0; j < n; ++j) a[i + j * lda] = X[i + j * n];
}

// symmetric eigen-decomposition by cyclic Jacobi; eigenvalues ascending, eigenvectors in columns of a
void dsyevd_(const char* jobz, const char* uplo, const int* n_, double* a, const int* lda_, double* w,
             double* work, const int* lwork, int* iwork, const int* liwork, int* info) {
    const int n = *n_, lda = *lda_;
    *info = 0;
    if (*lwork == -1 || *liwork == -1) { work[0] = double(std::max(1, 1 + 6 * n + 2 * n * n)); iwork[0] = std::max(1, 3 + 5 * n); return; }
    const bool lower = (*uplo == 'L' || *uplo == 'l');
    std::vector<double> S(n * n), V(n * n, 0.0);
    for (int i = 0; i < n; ++i)
        for (int j = 0; j < n; ++j) {
            const int r = lower ? std::max(i, j) : std::min(i, j);
            const int c = lower ? std::min(i, j) : std::max(i, j);
            S[i + j * n] = a[r + c * lda];
        }
    for (int i = 0; i < n; ++i) V[i + i * n] = 1.0;
    for (int sweep = 0; sweep < 100; ++sweep) {
        double off = 0.0;
        for (int i = 0; i < n; ++i) for (int j = i + 1; j < n; ++j) off += S[i + j * n] * S[i + j * n];
        if (off < 1e-300) break;
        for (int p = 0; p < n; ++p)
            for (int q = p + 1; q < n; ++q) {
                if (S[p + q * n] == 0.0) continue;
                const double theta = (S[q + q * n] - S[p + p * n]) / (2.0 * S[p + q * n]);
                const double t = (theta >= 0 ? 1.0 : -1.0) / (std::fabs(theta) + std::sqrt(theta * theta + 1.0));
                const double c = 1.0 / std::sqrt(t * t + 1.0), s = t * c;
                for (int k = 0; k < n; ++k) {
                    const double skp = S[k + p * n], skq = S[k + q * n];
                    S[k + p * n] = c * skp - s * skq;
                    S[k + q * n] = s * skp + c * skq;
                }
                for (int k = 0; k < n; ++k) {
                    const double spk = S[p + k * n], sqk = S[q + k * n];
                    S[p + k * n] = c * spk - s * sqk;
                    S[q + k * n] = s * spk + c * sqk;
                }
                for (int k = 0; k < n; ++k) {
                    const double vkp = V[k + p * n], vkq = V[k + q * n];
                    V[k + p * n] = c * vkp - s * vkq;
                    V[k + q * n] = s * vkp + c * vkq;
                }
            }
    }
    std::vector<int> order(n);
    for (int i = 0; i < n; ++i) order[i] = i;
    std::sort(order.begin(), order.end(), [&](int x, int y) { return S[x + x * n] < S[y + y * n]; });
    for (int j = 0; j < n; ++j) {
        w[j] = S[order[j] + order[j] * n];
        if (*jobz == 'V' || *jobz == 'v')
            for (int i = 0; i < n; ++i) a[i + j * lda] = V[i + order[j] * n];
    }
}

void dgesdd_(...) {
    std::fprintf(stderr, "oracle lapack shim: dgesdd_ is not on the collision-probability path\n");
    std::abort();
}

void dpbtrf_(...) {
    std::fprintf(stderr, "oracle lapack shim: dpbtrf_ is not on the collision-probability path\n");
    std::abort();
}

}  // extern "C"
