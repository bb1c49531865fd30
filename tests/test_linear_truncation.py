"""Analytic linear-constraint truncation, host FP64, CPU test.  PARITY UNPINNED: the
reference has no implementation (linConstraint.py:15-30 is a predicate stub); the checks are the closed
form of the truncated normal (scipy) and sampled moments."""
import numpy as np
from scipy import stats

import pcoll_b200

abi = pcoll_b200.cuda_abi


def test_axis_constraint_equals_truncated_normal_closed_form():
    mean, cov = np.array([1.0, -2.0, 0.5]), np.diag([0.04, 0.09, 0.01])
    m, c, pv = abi.truncate_linear(mean, cov, [[1.0, 0.0, 0.0]], [1.1])
    tn = stats.truncnorm(-np.inf, (1.1 - 1.0) / 0.2, loc=1.0, scale=0.2)
    np.testing.assert_allclose(m[0], tn.mean(), rtol=1e-12)
    np.testing.assert_allclose(c[0, 0], tn.var(), rtol=1e-10)
    np.testing.assert_allclose(m[1:], mean[1:])            # independent coordinates untouched
    np.testing.assert_allclose(c[1:, 1:], cov[1:, 1:])
    np.testing.assert_allclose(pv[0], stats.norm.sf(0.5), rtol=1e-12)


def test_oblique_constraint_matches_sampled_moments():
    rng = np.random.default_rng(3)
    a = rng.normal(size=(3, 3))
    cov = a @ a.T * 0.01 + 0.002 * np.eye(3)
    mean = np.array([3.0, 1.0, 1.5])
    n = np.array([0.6, -0.8, 0.3])
    b = float(n @ mean + 0.05)
    m, c, pv = abi.truncate_linear(mean, cov, [n], [b])
    x = rng.multivariate_normal(mean, cov, size=2_000_000)
    keep = x @ n <= b
    xs = x[keep]
    np.testing.assert_allclose(pv[0], 1.0 - keep.mean(), atol=2e-3)
    np.testing.assert_allclose(m, xs.mean(0), atol=5e-4)
    np.testing.assert_allclose(c, np.cov(xs.T), atol=2e-4)
    assert np.all(np.linalg.eigvalsh(c) > 0) and np.allclose(c, c.T)


def test_far_constraint_is_identity_and_sum_is_order_independent():
    mean, cov = np.zeros(3), np.eye(3) * 0.01
    m, c, pv = abi.truncate_linear(mean, cov, [[1.0, 0.0, 0.0]], [10.0])
    np.testing.assert_allclose(m, mean, atol=1e-15)
    np.testing.assert_allclose(c, cov, atol=1e-15)
    assert pv[0] == 0.0
    A, b = np.array([[1.0, 0.2, 0.0], [0.0, 1.0, 0.0], [-1.0, 0.0, 0.1]]), np.array([0.05, 0.1, 0.08])
    m1, c1, p1 = abi.truncate_linear(mean, cov, A, b)
    m2, c2, p2 = abi.truncate_linear(mean, cov, A[::-1], b[::-1])
    np.testing.assert_allclose(m1, m2, rtol=1e-12, atol=1e-15)
    np.testing.assert_allclose(c1, c2, rtol=1e-12, atol=1e-15)
    np.testing.assert_allclose(p1, p2[::-1])
