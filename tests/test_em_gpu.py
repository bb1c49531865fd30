"""GPU parity tests of the EM operator (BASELINE config 5) through the C ABI (mcsim_cuda_em_*).

Oracle: the reference repository's vendored Armadillo 8.400.0 gmm_full::learn(keep_existing, km_iter = 0)
(SURVEY.md A13): committed known-answer vectors tests/golden/em_K*.npz (made by tests/golden/make_em.py)
and, where oracle/_ref is present, live runs on fresh seeds.  Bar (BASELINE.json north_star): hefts,
means and covariances within 1e-4 relative (covariances relative to the largest entry of the
component's matrix; means relative to max(|mean|, 1))."""
import numpy as np
import pytest

from conftest import golden

pytestmark = pytest.mark.gpu

RTOL = 1e-4


def _errors(out, hefts, means, fcovs, f=0):
    eh = np.max(np.abs(out["hefts"][f] - hefts) / np.maximum(np.abs(hefts), 1e-3))
    em = np.max(np.abs(out["means"][f] - means) / np.maximum(np.abs(means), 1.0))
    scale = np.max(np.abs(fcovs), axis=(1, 2), keepdims=True)
    ec = np.max(np.abs(out["fcovs"][f] - fcovs) / scale)
    return eh, em, ec


@pytest.mark.parametrize("K", [1, 3, 8, 16])
def test_em_matches_armadillo_golden(ctx, K):
    g = golden("em_K%d.npz" % K)
    for it in (1, 5, 20):
        out = ctx.em_fit(g["x"], g["hefts0"], g["means0"], g["fcovs0"], it)
        eh, em, ec = _errors(out, g["hefts_%d" % it], g["means_%d" % it], g["fcovs_%d" % it])
        print("K=%d iters=%d: heft err %.2e mean err %.2e cov err %.2e" % (K, it, eh, em, ec))
        assert out["iters_run"][0] == it
        assert max(eh, em, ec) < RTOL
        np.testing.assert_allclose(out["hefts"].sum(), 1.0, atol=1e-12)


def test_em_batched_fits_equal_single_fits(ctx):
    """n_fits > 1 in one call == the same fits one by one (independent accumulators per fit)."""
    import pcoll_b200
    probs = [pcoll_b200.inputs.synthetic_em_problem(100 + f, 5001, 3, centre=(f * 0.5, 1.0, 0.3 * f)) for f in range(5)]
    x = np.stack([p[0] for p in probs]); h = np.stack([p[1] for p in probs])
    m = np.stack([p[2] for p in probs]); c = np.stack([p[3] for p in probs])
    both = ctx.em_fit(x, h, m, c, 4)
    for f in range(5):
        one = ctx.em_fit(x[f], h[f], m[f], c[f], 4)
        np.testing.assert_allclose(both["means"][f], one["means"][0], rtol=1e-9, atol=1e-12)
        np.testing.assert_allclose(both["fcovs"][f], one["fcovs"][0], rtol=1e-6, atol=1e-12)
        np.testing.assert_allclose(both["hefts"][f], one["hefts"][0], rtol=1e-9)


def test_em_vs_live_reference_ragged_sizes(ctx):
    import oracle_ref
    import pcoll_b200
    if not oracle_ref.available():
        pytest.skip("oracle/_ref not built")
    for seed, n, K in ((1, 257, 2), (2, 1000, 5), (3, 33333, 12)):
        x, h, m, c = pcoll_b200.inputs.synthetic_em_problem(seed, n, K)
        rm, rc, rh, rl = oracle_ref.em_fit(x.T.astype(np.float64), m, c.reshape(K, 9), h, 3)
        out = ctx.em_fit(x, h, m, c, 3)
        eh, em, ec = _errors(out, rh, rm, rc.reshape(K, 3, 3))
        print("n=%d K=%d: heft err %.2e mean err %.2e cov err %.2e" % (n, K, eh, em, ec))
        assert max(eh, em, ec) < RTOL


@pytest.mark.parametrize("K", [1, 3, 16])
def test_em_config5_size_vs_armadillo(ctx, K):
    """BASELINE config 5 at its real size: n = 10^6 samples, 3 EM iterations, against arma::gmm_full::learn run on the
    same samples (gmm_full_meat.hpp:2646-2734; committed parameters tests/golden/em_1e6_K*.npz made by
    tests/golden/make_round2.py -- the samples are regenerated from the seed).  FP32 per-thread partial sums are what
    grows with n: this is the size at which they have to hold 1e-4.  All three E-step variants: n % 4 == 0 takes the
    bulk-copy kernel for K <= 4 (em_estep_tma_kernel) and the scalar two-launch kernel for K = 16; n - 1 samples
    (rows no longer 16-byte aligned) takes the packed-pair kernel (em_estep2_kernel) -- that one against Armadillo run
    live on the shortened sample."""
    import pcoll_b200
    g = golden("em_1e6_K%d.npz" % K)
    n = int(g["n"])
    x, h, m, c = pcoll_b200.inputs.synthetic_em_problem(int(g["seed"]), n, K)
    out = ctx.em_fit(x, h, m, c, int(g["iters"]))
    eh, em, ec = _errors(out, g["hefts"], g["means"], g["fcovs"])
    print("n=1e6 K=%d: heft err %.2e mean err %.2e cov err %.2e; avg log p %.6f (Armadillo after update %.6f)" %
          (K, eh, em, ec, out["avg_log_p"][0, -1], float(g["avg_log_p"])))
    assert out["iters_run"][0] == int(g["iters"])
    assert max(eh, em, ec) < RTOL
    import oracle_ref
    if K <= 4 and oracle_ref.available():
        xs = np.ascontiguousarray(x[:, :n - 1])
        rm, rc, rh, rl = oracle_ref.em_fit(xs.T.astype(np.float64), m, c.reshape(K, 9), h, 3)
        o2 = ctx.em_fit(xs, h, m, c, 3)
        eh, em, ec = _errors(o2, rh, rm, rc.reshape(K, 3, 3))
        print("n=1e6-1 K=%d (packed-pair kernel): heft err %.2e mean err %.2e cov err %.2e" % (K, eh, em, ec))
        assert max(eh, em, ec) < RTOL


def test_em_log_likelihood_is_monotone_and_resident_reuse(ctx):
    """EM never decreases the average log-likelihood (size-independent property, full config-5 size);
    second call reuses the resident samples (x = None)."""
    import pcoll_b200
    x, h, m, c = pcoll_b200.inputs.synthetic_em_problem(7, 1_000_000, 6)
    out = ctx.em_fit(x, h, m, c, 10)
    tr = out["avg_log_p"][0]
    assert np.all(np.isfinite(tr)) and np.all(np.diff(tr) > -1e-5), tr
    again = ctx.em_fit(None, h, m, c, 10)
    np.testing.assert_allclose(again["means"], out["means"], rtol=1e-6)


def test_em_split_api_two_shards_equal_whole(ctx):
    """estep on two half shards + summed statistics + mstep == the unsharded fit (what the NCCL
    all-reduce does across GPUs), run on one device with two contexts."""
    import torch
    import pcoll_b200
    abi = pcoll_b200.cuda_abi
    x, h, m, c = pcoll_b200.inputs.synthetic_em_problem(11, 20000, 4)
    whole = ctx.em_fit(x, h, m, c, 5)
    a, b = abi.Context(0), abi.Context(0)
    try:
        a.em_prepare(x[:, :10000], h, m, c, 5, n_total=20000)
        b.em_prepare(x[:, 10000:], h, m, c, 5, n_total=20000)
        for _ in range(5):
            a.em_estep(); b.em_estep()
            torch.cuda.synchronize()
            sa, sb = a.em_stats_tensor(), b.em_stats_tensor()
            tot = sa + sb
            sa.copy_(tot); sb.copy_(tot)
            torch.cuda.synchronize()
            a.em_mstep(); b.em_mstep()
        ra, rb = a.em_collect(), b.em_collect()
    finally:
        a.close(); b.close()
    np.testing.assert_array_equal(ra["means"], rb["means"])
    np.testing.assert_allclose(ra["means"], whole["means"], rtol=1e-6)
    np.testing.assert_allclose(ra["fcovs"], whole["fcovs"], rtol=1e-4, atol=1e-9)
    np.testing.assert_allclose(ra["hefts"], whole["hefts"], rtol=1e-5)


def test_em_bulk_copy_path_vs_live_reference(ctx):
    """K <= 4 with n % 4 == 0 streams through em_estep_tma_kernel (cp.async.bulk + mbarrier pipeline, 2,048-sample
    tiles, 3 stages): many tiles per CTA, a ragged last tile, several fits per launch, against Armadillo run here;
    and the same fit one sample shorter (n % 4 != 0: the register-prefetch kernel) must agree with it closely."""
    import oracle_ref
    import pcoll_b200
    if not oracle_ref.available():
        pytest.skip("oracle/_ref not built")
    for K, n in ((1, 40004), (2, 123460), (2, 2048 * 9), (4, 50000)):
        probs = [pcoll_b200.inputs.synthetic_em_problem(50 + f, n, K, centre=(0.3 * f, 1.0, 0.2)) for f in range(3)]
        x = np.stack([p[0] for p in probs]); h = np.stack([p[1] for p in probs])
        m = np.stack([p[2] for p in probs]); c = np.stack([p[3] for p in probs])
        out = ctx.em_fit(x, h, m, c, 4)
        for f in range(3):
            rm, rc, rh, rl = oracle_ref.em_fit(x[f].T.astype(np.float64), m[f], c[f].reshape(K, 9), h[f], 4)
            eh, em, ec = _errors(out, rh, rm, rc.reshape(K, 3, 3), f)
            assert max(eh, em, ec) < RTOL, (K, n, f, eh, em, ec)
        shorter = ctx.em_fit(np.ascontiguousarray(x[:, :, :n - 1]), h, m, c, 4)
        np.testing.assert_allclose(shorter["means"], out["means"], rtol=1e-3, atol=1e-4)


@pytest.mark.parametrize("K,n", [(2, 40000), (3, 39999), (4, 40000), (5, 30001), (8, 30000), (9, 30000), (10, 30000), (12, 30000), (16, 30000)])
def test_em_moment_forms_agree(ctx, K, n):
    """The statistics about one common point per fit (MCSIM_EM_MOMENTS_SHIFTED: x - c and its products once per sample)
    and about each component's own mean (_CENTRED) are the same sufficient statistics; on mixtures whose components lie
    within a few sigma of each other -- the propagated beliefs this operator is for -- AUTO picks SHIFTED, and both forms
    meet the 1e-4 bar against Armadillo (run here when oracle/_ref is built) and agree far more closely with each other.
    The shapes cover every E-step kernel: bulk-copy packed (K <= 4, n % 4 == 0), register-prefetch packed (K <= 5),
    one warp per sample tile (K = 6..9), components split over two warps (K > 9)."""
    import oracle_ref
    import pcoll_b200
    abi = pcoll_b200.cuda_abi
    x, h, m, c = pcoll_b200.inputs.synthetic_em_problem(300 + K, n, K)
    outs = {}
    try:
        for form in (abi.EM_MOMENTS_CENTRED, abi.EM_MOMENTS_SHIFTED, abi.EM_MOMENTS_AUTO):
            ctx.em_set_moments(form)
            outs[form] = ctx.em_fit(x, h, m, c, 6)
            used = ctx.em_moments_in_use(0)
            if form != abi.EM_MOMENTS_AUTO:
                assert used == form
            else:   # components a few sigma apart: AUTO takes the shifted form where it pays (the scalar kernels, K >= 6)
                auto_form = abi.EM_MOMENTS_SHIFTED if K >= 6 else abi.EM_MOMENTS_CENTRED
                assert used == auto_form
    finally:
        ctx.em_set_moments(abi.EM_MOMENTS_AUTO)
    cen, shf, aut = (outs[f] for f in (abi.EM_MOMENTS_CENTRED, abi.EM_MOMENTS_SHIFTED, abi.EM_MOMENTS_AUTO))
    eh, em, ec = _errors(shf, cen["hefts"][0], cen["means"][0], cen["fcovs"][0])
    print("K=%d n=%d shifted vs centred: heft %.2e mean %.2e cov %.2e" % (K, n, eh, em, ec))
    assert max(eh, em, ec) < 2e-5
    same = outs[auto_form]
    eh, em, ec = _errors(aut, same["hefts"][0], same["means"][0], same["fcovs"][0])
    assert max(eh, em, ec) < 1e-9      # the same kernels and form: only the order of the FP64 atomics differs
    if oracle_ref.available():
        rm, rc, rh, rl = oracle_ref.em_fit(x.T.astype(np.float64), m, c.reshape(K, 9), h, 6)
        for name, o in (("centred", cen), ("shifted", shf)):
            eh, em, ec = _errors(o, rh, rm, rc.reshape(K, 3, 3))
            print("   %s vs Armadillo: heft %.2e mean %.2e cov %.2e" % (name, eh, em, ec))
            assert max(eh, em, ec) < RTOL


def test_em_far_apart_components_keep_the_centred_form(ctx):
    """Clusters 40 m apart with sigma = 0.03 m: about a common point the FP32 second moments would carry
    (40 / 0.03)^2 = 2e6 times the rounding of the centred ones, so AUTO must stay with the centred form -- and the
    forced shifted form visibly loses the covariances (that is what the guard is for)."""
    import pcoll_b200
    abi = pcoll_b200.cuda_abi
    rng = np.random.Generator(np.random.PCG64(77))
    n, K = 60000, 3
    centres = np.array([[0.0, 0.0, 0.0], [40.0, 0.0, 1.0], [0.0, 40.0, 2.0]])
    lab = rng.integers(0, K, size=n)
    x = (centres[lab] + rng.normal(0.0, 0.03, size=(n, 3))).T.astype(np.float32)
    h = np.full(K, 1.0 / K); m = centres + 0.01; c = np.tile(np.eye(3) * 0.01, (K, 1, 1))
    try:
        ctx.em_set_moments(abi.EM_MOMENTS_AUTO)
        aut = ctx.em_fit(x, h, m, c, 4)
        assert ctx.em_moments_in_use(0) == abi.EM_MOMENTS_CENTRED
        ctx.em_set_moments(abi.EM_MOMENTS_CENTRED)
        cen = ctx.em_fit(x, h, m, c, 4)
        ctx.em_set_moments(abi.EM_MOMENTS_SHIFTED)
        shf = ctx.em_fit(x, h, m, c, 4)
    finally:
        ctx.em_set_moments(abi.EM_MOMENTS_AUTO)
    eh, em, ec = _errors(aut, cen["hefts"][0], cen["means"][0], cen["fcovs"][0])
    assert max(eh, em, ec) < 1e-9
    x64 = x.astype(np.float64)
    for g in range(K):
        pts = x64[:, lab == g]
        np.testing.assert_allclose(aut["means"][0][g], pts.mean(axis=1), atol=1e-5)
        np.testing.assert_allclose(aut["fcovs"][0][g], np.cov(pts, bias=True), rtol=1e-3, atol=1e-7)
    worst = max(np.max(np.abs(shf["fcovs"][0][g] - cen["fcovs"][0][g])) / np.max(np.abs(cen["fcovs"][0][g])) for g in range(K))
    print("forced shifted form on far-apart clusters: covariance error %.2e relative" % worst)
    assert worst > 1e-4
