"""CPU tests: the FP64 host real-robot thread (csrc/host_thread.cpp) against the reference's own
EKF_GaussProp state (mu, Sigma per waypoint) recorded by oracle/_ref on the same injected draws
(tests/golden/mc_replay_*.npz), and the C-ABI export list."""
import ctypes
import os
import re

import numpy as np
import pytest

from conftest import ROOT, golden


def test_cuda_abi_exports_every_declared_symbol():
    import pcoll_b200
    lib = pcoll_b200.cuda_abi.lib()
    hdr = open(os.path.join(ROOT, "include", "mcsim_cuda.h")).read()
    names = set(re.findall(r"\b(mcsim_(?:cuda_|host_)\w+)\s*\(", hdr))
    assert len(names) >= 10
    for n in sorted(names):
        assert hasattr(lib, n), "libmcsim_cuda.so does not export %s" % n


def test_no_device_is_a_loud_error():
    import pcoll_b200
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(pcoll_b200.cuda_abi.McsimCudaError, match="no CPU fallback"):
        pcoll_b200.cuda_abi.Context(0)


@pytest.mark.parametrize("scene", ["pr2test2", "pr2custom"])
def test_host_thread_matches_reference_ekf(plan, scene):
    import pcoll_b200
    g = golden("mc_replay_%s.npz" % scene)
    N = g["counts"].shape[0]
    normals = g["draws"][3 * N:]
    out = pcoll_b200.cuda_abi.host_thread(plan, normals)
    # mu / Sigma after every step: reference values recorded through the CheckCollision hook
    # (the hook fires before the EKF update of the step, so golden row t holds mu_{t-1} for
    # 1 <= t <= T-2, and row T-1 holds the final mu_{T-1}: oracle/ref/ref_oracle.cpp check_hook)
    T = g["mu_path"].shape[0]
    ours_mu = np.vstack([out["mu_path"].T[:1], out["mu_path"].T[:T - 2], out["mu_path"].T[T - 1:]])
    ours_cov = np.vstack([out["cov_path"][:1], out["cov_path"][:T - 2], out["cov_path"][T - 1:]])
    np.testing.assert_allclose(ours_mu, g["mu_path"], rtol=1e-12, atol=1e-13)
    np.testing.assert_allclose(ours_cov, g["cov_path"], rtol=1e-9, atol=1e-16)
    # particle 0 of the reference moved by the same noisy commands: pose differences telescope
    poses = g["poses_first64"]  # [T, 64, 3]
    p = poses[0, 0].copy()
    for t in range(out["applied_noisy"].shape[1]):
        u = out["applied_noisy"][:, t]
        p = np.array([p[0] + u[1] * np.cos(p[2] + u[0]), p[1] + u[1] * np.sin(p[2] + u[0]), p[2] + u[0] + u[2]])
        while p[2] < 0:
            p[2] += 2 * 3.14159265358979323846
        while p[2] > 2 * 3.14159265358979323846:
            p[2] -= 2 * 3.14159265358979323846
        np.testing.assert_allclose(p, poses[t + 1, 0], rtol=1e-12, atol=1e-12)


def test_host_thread_rejects_short_draws(plan):
    import pcoll_b200
    with pytest.raises(pcoll_b200.cuda_abi.McsimCudaError):
        pcoll_b200.cuda_abi.host_thread(plan, np.zeros(10))


def test_py2_str_marshalling():
    import pcoll_b200
    assert pcoll_b200.py2_str(0.15) == "0.15"
    assert pcoll_b200.py2_str(-3.4) == "-3.4"
    assert pcoll_b200.py2_str(2.0) == "2.0"
    assert pcoll_b200.py2_str(1.5707963267948966) == "1.57079632679"
    assert pcoll_b200.py2_str(6.25e-08) == "6.25e-08"
