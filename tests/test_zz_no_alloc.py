"""mcsim_cuda_em_fit_static (include/mcsim_cuda.h): the EM fit with its scratch in a file-scope __device__ array of
the library -- no cudaMalloc inside the call.  No counterpart in the reference (added for callers that cannot allocate);
checked against the existing entry point mcsim_cuda_em_fit on the same inputs, with the tolerances of
tests/test_em_gpu.py::test_em_batched_fits_equal_single_fits (the FP64 atomics of the E-step are unordered, so two
runs of EITHER entry point agree to rounding, not bitwise).

Named test_zz_* so that pytest collects it last: the shared scratch array keeps the previous call's contents, and this
file has to pass both alone in a fresh process and after every other GPU test of the session."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _device_buffers(torch, x, h, m, c, iters):
    F = x.shape[0]
    dev = "cuda:0"
    return dict(
        x=torch.from_numpy(np.ascontiguousarray(x, dtype=np.float32)).to(dev),
        h=torch.from_numpy(np.ascontiguousarray(h, dtype=np.float64)).to(dev),
        m=torch.from_numpy(np.ascontiguousarray(m, dtype=np.float64)).to(dev),
        c=torch.from_numpy(np.ascontiguousarray(c, dtype=np.float64)).to(dev),
        lp=torch.full((F, max(iters, 1)), 7.0, dtype=torch.float64, device=dev),       # poisoned: the call must overwrite
        it=torch.full((F,), -5, dtype=torch.int32, device=dev))


def _static_fit(torch, abi, x, h, m, c, iters, buf=None, check_no_alloc=False):
    """x [F, 3, n]; returns the fitted parameters read back from the caller-owned device buffers."""
    F, _, n = x.shape
    K = h.shape[1]
    if buf is None:
        buf = _device_buffers(torch, x, h, m, c, iters)
    else:   # reuse the caller's buffers: parameters are in/out, so load the initial ones again
        buf["h"].copy_(torch.from_numpy(np.ascontiguousarray(h)))
        buf["m"].copy_(torch.from_numpy(np.ascontiguousarray(m)))
        buf["c"].copy_(torch.from_numpy(np.ascontiguousarray(c)))
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info(0)[0]
    abi.em_fit_static(0, buf["x"].data_ptr(), n, K, F, iters, buf["h"].data_ptr(), buf["m"].data_ptr(), buf["c"].data_ptr(),
                      buf["lp"].data_ptr(), buf["it"].data_ptr(), stream=None)
    torch.cuda.synchronize()
    if check_no_alloc:
        assert torch.cuda.mem_get_info(0)[0] == free0, "device memory in use changed across mcsim_cuda_em_fit_static"
    out = dict(hefts=buf["h"].cpu().numpy(), means=buf["m"].cpu().numpy(), fcovs=buf["c"].cpu().numpy().reshape(F, K, 3, 3),
               avg_log_p=buf["lp"].cpu().numpy()[:, :iters], iters_run=buf["it"].cpu().numpy())
    return out, buf


def _problem(seed, n, K, F):
    from pcoll_b200 import inputs
    xs, hs, ms, cs = [], [], [], []
    for f in range(F):
        x, h, m, c = inputs.synthetic_em_problem(seed + f, n, K)
        xs.append(x); hs.append(h); ms.append(m); cs.append(c)
    return np.stack(xs), np.stack(hs), np.stack(ms), np.stack(cs)


def _same(a, b):
    np.testing.assert_allclose(a["means"], b["means"], rtol=1e-9, atol=1e-12)
    np.testing.assert_allclose(a["fcovs"], np.asarray(b["fcovs"]).reshape(a["fcovs"].shape), rtol=1e-6, atol=1e-12)
    np.testing.assert_allclose(a["hefts"], b["hefts"], rtol=1e-9)
    np.testing.assert_array_equal(a["iters_run"], b["iters_run"])
    np.testing.assert_allclose(a["avg_log_p"], np.asarray(b["avg_log_p"]).reshape(a["avg_log_p"].shape), rtol=1e-9, equal_nan=True)


# (K, n, fits, iterations): the bulk-copy kernel (K <= 4, n % 4 == 0), the split kernel (K = 16), the one-warp kernel on ragged rows
CASES = [(16, 200_000, 1, 6), (3, 400_000, 2, 5), (6, 100_003, 3, 4), (1, 65_536, 8, 3)]


def test_em_fit_static_equals_em_fit_on_two_calls_in_a_row(ctx):
    import torch
    import pcoll_b200
    abi = pcoll_b200.cuda_abi
    for K, n, F, iters in CASES:
        x, h, m, c = _problem(7100 + K, n, K, F)
        want = ctx.em_fit(x, h, m, c, iters)
        first, buf = _static_fit(torch, abi, x, h, m, c, iters)
        # second call in a row on the same buffers: the scratch array now holds the first call's state (its
        # statistics are zeroed by the last M-step, its convergence flags and trace are not)
        second, _ = _static_fit(torch, abi, x, h, m, c, iters, buf=buf, check_no_alloc=True)
        _same(first, want)
        _same(second, want)
        assert np.all(np.isfinite(first["avg_log_p"][:, 0]))
    # a call of another shape in between leaves nothing behind: first case again, after the last one
    K, n, F, iters = CASES[0]
    x, h, m, c = _problem(7100 + K, n, K, F)
    again, _ = _static_fit(torch, abi, x, h, m, c, iters)
    _same(again, ctx.em_fit(x, h, m, c, iters))


def test_em_fit_static_stopped_fit_leaves_no_stale_flags(ctx):
    """A fit that stops early leaves done != 0 in the shared scratch; the next call must start from done = 0.  The stop
    is forced deterministically: one NaN sample makes the average log-likelihood non-finite, which ends the fit after
    its first iteration with the parameters untouched (learn() fails there, gmm_full_meat.hpp:2447-2465)."""
    import torch
    import pcoll_b200
    abi = pcoll_b200.cuda_abi
    x, h, m, c = _problem(7300, 50_000, 2, 2)
    x[1, 0, 123] = np.nan                       # second fit only
    a, _ = _static_fit(torch, abi, x, h, m, c, 5)
    want = ctx.em_fit(x, h, m, c, 5)
    np.testing.assert_array_equal(a["iters_run"], [5, 1])
    np.testing.assert_array_equal(want["iters_run"], [5, 1])
    np.testing.assert_array_equal(a["means"][1], m[1])
    np.testing.assert_allclose(a["means"][0], want["means"][0], rtol=1e-9, atol=1e-12)
    assert np.all(np.isnan(a["avg_log_p"][1]))
    x2, h2, m2, c2 = _problem(7400, 80_000, 4, 2)
    b, _ = _static_fit(torch, abi, x2, h2, m2, c2, 5)
    _same(b, ctx.em_fit(x2, h2, m2, c2, 5))
    np.testing.assert_array_equal(b["iters_run"], [5, 5])


def test_em_fit_static_rejects_sizes_beyond_the_scratch():
    import torch
    import pcoll_b200
    abi = pcoll_b200.cuda_abi
    t = torch.zeros(1024, dtype=torch.float64, device="cuda:0")
    ok = dict(device=0, x_ptr=t.data_ptr(), n=16, K=2, n_fits=1, iters=1, hefts_ptr=t.data_ptr(), means_ptr=t.data_ptr(), fcovs_ptr=t.data_ptr())
    for bad in (dict(n_fits=abi.EM_STATIC_MAX_FITS + 1), dict(iters=abi.EM_STATIC_MAX_ITERS + 1), dict(K=17), dict(K=0),
                dict(n=1 << 31), dict(x_ptr=None), dict(device=-1)):
        with pytest.raises(abi.McsimCudaError) as e:
            abi.em_fit_static(**dict(ok, **bad))
        assert "em_fit_static" in str(e.value)
    assert "beyond the static scratch" in _message(abi, dict(ok, n_fits=9))


def _message(abi, kw):
    try:
        abi.em_fit_static(**kw)
    except abi.McsimCudaError as e:
        return str(e)
    return ""
