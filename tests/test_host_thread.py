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
    names = set(re.findall(r"\b(mcsim_(?:cuda_|host_|step_|truncate_)\w+)\s*\(", hdr))
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


def _philox4x32_10(c, k0, k1):
    """Philox4x32-10 (Salmon et al., SC'11) on uint32 arrays c[4]; an independent numpy restatement."""
    c = [np.asarray(x, dtype=np.uint64) for x in c]
    k0 = np.uint64(k0); k1 = np.uint64(k1)
    M = np.uint64(0xFFFFFFFF)
    for _ in range(10):
        p0 = np.uint64(0xD2511F53) * c[0]
        p1 = np.uint64(0xCD9E8D57) * c[2]
        c = [((p1 >> np.uint64(32)) ^ c[1] ^ k0) & M, p1 & M, ((p0 >> np.uint64(32)) ^ c[3] ^ k1) & M, p0 & M]
        k0 = (k0 + np.uint64(0x9E3779B9)) & M
        k1 = (k1 + np.uint64(0xBB67AE85)) & M
    return c


def _step_normals_numpy(seed, ids, t):
    """The documented layout of the Philox-mode step draws (csrc/philox.h): two blocks, six Box-Muller pairs at
    24 radius bits + 16 angle bits."""
    ids = np.asarray(ids, dtype=np.uint64)
    lo, hi = ids & np.uint64(0xFFFFFFFF), ids >> np.uint64(32)
    z = np.zeros_like(lo)
    k0, k1 = seed & 0xFFFFFFFF, seed >> 32
    a = _philox4x32_10([lo, hi, z + np.uint64(8 * t), z], k0, k1)       # PURPOSE_STEP = 0
    b = _philox4x32_10([lo, hi, z + np.uint64(8 * t + 1), z], k0, k1)
    s8, s16, m8, m16 = np.uint64(8), np.uint64(16), np.uint64(0xFF), np.uint64(0xFFFF)
    rad = [a[0] >> s8, a[1] >> s8, a[2] >> s8, a[3] >> s8, b[0] >> s8, (a[0] & m8) | ((a[1] & m8) << s8) | ((a[2] & m8) << s16)]
    ang = [b[1] & m16, b[1] >> s16, b[2] & m16, b[2] >> s16, b[3] & m16, b[3] >> s16]
    out = np.empty((len(ids), 12))
    for j in range(6):
        r = np.sqrt(-2.0 * np.log((rad[j].astype(np.float64) + 0.5) / 2.0 ** 24))
        th = 2 * np.pi * (ang[j].astype(np.float64) + 0.5) / 65536.0
        out[:, 2 * j] = r * np.cos(th)
        out[:, 2 * j + 1] = r * np.sin(th)
    return out


def test_step_draws_layout_and_moments():
    """mcsim_step_normals (host FP64 view of the bits the Philox-mode kernels consume) against the numpy restatement
    of the documented layout, and the law of the 12 normals of a step: mean 0, variance 1, kurtosis 3, no
    correlation between any two of them (240 of the 256 bits, no bit used twice)."""
    import pcoll_b200
    lib = pcoll_b200.cuda_abi.lib()
    seed = 0x5AFEC0DE
    buf = (ctypes.c_double * 12)()
    ids = [0, 1, 12345, (1 << 32) + 7, (1 << 40) + 99]
    for t in (0, 3, 54):
        want = _step_normals_numpy(seed, ids, t)
        for i, rid in enumerate(ids):
            assert lib.mcsim_step_normals(seed, rid, t, buf) == 0
            np.testing.assert_allclose(np.array(buf[:]), want[i], rtol=1e-12, atol=1e-14)
    z = np.vstack([_step_normals_numpy(seed, np.arange(40000), t) for t in range(5)])   # 200,000 x 12
    n = z.shape[0]
    assert np.all(np.abs(z.mean(0)) < 5 / np.sqrt(n))
    assert np.all(np.abs(z.var(0) - 1) < 5 * np.sqrt(2.0 / n))
    assert np.all(np.abs((z ** 4).mean(0) - 3) < 5 * np.sqrt(96.0 / n))
    c = np.corrcoef(z.T) - np.eye(12)
    assert np.abs(c).max() < 5 / np.sqrt(n)
    c2 = np.corrcoef((z ** 2).T) - np.eye(12)     # Box-Muller partners share a radius: uncorrelated squares all the same
    assert np.abs(c2).max() < 5 / np.sqrt(n)
