"""GPU parity tests of the GMM mode (sample -> collide -> moment refit per waypoint), through the C ABI.

Deterministic (replay) mode: the reference's own runGMMEstimation() (oracle/_ref, golden vectors in
tests/golden/gmm_replay_*.npz: injected normals + the multinomial counts it drew) must be reproduced:
per-waypoint colliding fractions exactly (the replay-mode SAT is FP64, so the flags are the reference's),
mixture weights / means / covariances and P(collision) within 1e-4 relative (BASELINE.json north_star).
Philox mode: P inside the spread of the reference's own run-to-run distribution; statistical only."""
import numpy as np
import pytest

from conftest import golden

pytestmark = pytest.mark.gpu
RTOL = 1e-4  # north_star: GMM weights/means/covariances and P(collision) within 1e-4 relative


def _assert_mixture_close(ours, ref, what):
    # weights and means: relative to max(|ref|, 1) (angles and coordinates near 0 have no relative scale);
    # covariances: relative to the component's largest variance
    np.testing.assert_allclose(ours[..., 0], ref[..., 0], rtol=RTOL, atol=RTOL, err_msg=what + " weights")
    np.testing.assert_allclose(ours[..., 1:4], ref[..., 1:4], rtol=RTOL, atol=RTOL, err_msg=what + " means")
    scale = np.abs(ref[..., 4:]).max(axis=-1, keepdims=True)
    assert np.max(np.abs(ours[..., 4:] - ref[..., 4:]) / scale) < RTOL, what + " covariances"


@pytest.mark.parametrize("K", [1, 3])
def test_gmm_replay_matches_reference_golden(ctx, plan, scenes, K):
    import pcoll_b200
    abi = pcoll_b200.cuda_abi
    g = golden("gmm_replay_pr2test2_K%d.npz" % K)
    ctx.set_scene(scenes["pr2test2"]["boxes"], scenes["pr2test2"]["hulls"])
    out = ctx.gmm_run(plan, K, int(g["n_samples"]), mode=abi.MODE_REPLAY, draws=g["draws"], counts=g["gmm_counts"], record=True)
    assert out["degenerate"] == 0
    # colliding fraction per waypoint: exact (integer count / N)
    np.testing.assert_array_equal(out["probabilities"], g["probabilities"])
    assert abs(out["P"] - float(g["P"])) <= RTOL * float(g["P"])
    _assert_mixture_close(out["gmm_pre"], g["gmm_pre"], "pre-truncation mixtures")
    _assert_mixture_close(out["gmm_post"], g["gmm_post"], "final mixture")


def test_gmm_replay_vs_live_reference(ctx, plan, scenes):
    """Fresh draws, other K / N (incl. a component count that is not a warp multiple), reference run live."""
    import oracle_ref
    import pcoll_b200
    abi = pcoll_b200.cuda_abi
    if not oracle_ref.available():
        pytest.skip("oracle/_ref not built")
    T = plan["trajectory"].shape[1]
    sc = scenes["pr2test2"]
    ctx.set_scene(sc["boxes"], sc["hulls"])
    for seed, K, N in ((11, 2, 777), (12, 5, 1501)):
        draws = np.random.Generator(np.random.PCG64(seed)).standard_normal(3 * N * T + 11 * (T - 1))
        try:
            ref = oracle_ref.gmm_run(plan, sc["boxes"], sc["hulls"], K, N, draws=draws, multinomial_seed=seed)
        except RuntimeError:
            continue  # the reference throws when a component loses all its samples
        out = ctx.gmm_run(plan, K, N, mode=abi.MODE_REPLAY, draws=draws, counts=ref["gmm_counts"], record=True)
        np.testing.assert_array_equal(out["probabilities"], ref["probabilities"])
        assert abs(out["P"] - ref["P"]) <= RTOL * ref["P"]
        _assert_mixture_close(out["gmm_pre"], ref["gmm_pre"], "pre-truncation mixtures K=%d" % K)


def test_gmm_replay_config2_size_vs_live_reference(ctx, plan, scenes):
    """BASELINE config 2 at its real size: K = 3 components, N = 10,000 samples per waypoint (MCSimulation.py:204-205),
    all 56 waypoints, deterministic mode against the reference's own runGMMEstimation() run live on the same 1.68 M
    injected normals and its multinomial counts (MCSimulator.h:570-642): per-waypoint colliding fractions exact,
    every pre-truncation mixture, the final mixture and P within 1e-4 relative."""
    import oracle_ref
    import pcoll_b200
    abi = pcoll_b200.cuda_abi
    if not oracle_ref.available():
        pytest.skip("oracle/_ref not built")
    T = plan["trajectory"].shape[1]
    sc = scenes["pr2test2"]
    ctx.set_scene(sc["boxes"], sc["hulls"])
    K, N = 3, 10000
    for seed in (2018, 2019):
        draws = np.random.Generator(np.random.PCG64(seed)).standard_normal(3 * N * T + 11 * (T - 1))
        ref = oracle_ref.gmm_run(plan, sc["boxes"], sc["hulls"], K, N, draws=draws, multinomial_seed=seed)
        out = ctx.gmm_run(plan, K, N, mode=abi.MODE_REPLAY, draws=draws, counts=ref["gmm_counts"], record=True)
        assert out["degenerate"] == 0
        np.testing.assert_array_equal(out["probabilities"], ref["probabilities"])
        assert abs(out["P"] - ref["P"]) <= RTOL * ref["P"]
        _assert_mixture_close(out["gmm_pre"], ref["gmm_pre"], "pre-truncation mixtures, N = 10,000")
        _assert_mixture_close(out["gmm_post"], ref["gmm_post"], "final mixture, N = 10,000")
        print("config 2 (K=3, N=10,000) seed %d: P gpu %.6f reference %.6f, colliding samples %d" %
              (seed, out["P"], ref["P"], int(round(ref["probabilities"].sum() * N))))


def test_gmm_truncate_single_shot_matches_first_waypoint(ctx, plan, scenes):
    import pcoll_b200
    abi = pcoll_b200.cuda_abi
    g = golden("gmm_replay_pr2test2_K3.npz")
    N, K = int(g["n_samples"]), 3
    ctx.set_scene(scenes["pr2test2"]["boxes"], scenes["pr2test2"]["hulls"])
    pre = g["gmm_pre"][0]
    out = ctx.gmm_truncate(pre[:, 0], pre[:, 1:4], pre[:, 4:].reshape(K, 3, 3), N, mode=abi.MODE_REPLAY,
                           counts=g["gmm_counts"][0], normals=g["draws"][:3 * N])
    assert out["p_collision"] == g["probabilities"][0]
    assert int(out["n_coll"].sum() + out["n_free"].sum()) == N
    np.testing.assert_array_equal(out["n_coll"] + out["n_free"], g["gmm_counts"][0].astype(np.uint64))
    np.testing.assert_allclose(out["weights"], out["n_free"] / out["n_free"].sum(), rtol=1e-15)


def test_gmm_truncate_rejects_bad_counts(ctx, scenes):
    import pcoll_b200
    abi = pcoll_b200.cuda_abi
    ctx.set_scene(scenes["pr2test2"]["boxes"], scenes["pr2test2"]["hulls"])
    with pytest.raises(abi.McsimCudaError):
        ctx.gmm_truncate([1.0], [[0, 0, 0]], [np.eye(3) * 1e-3], 10, mode=abi.MODE_REPLAY, counts=[9], normals=np.zeros(30))
    with pytest.raises(abi.McsimCudaError):
        ctx.gmm_truncate([1.0], [[0, 0, 0]], [-np.eye(3)], 10)


def test_gmm_philox_statistics(ctx, plan, scenes):
    """Philox mode: the estimate is a random variable (one shared real-robot thread per run, exactly as in the
    reference); its mean over trajectories must agree with the reference's own run-to-run mean."""
    import oracle_ref
    import pcoll_b200
    abi = pcoll_b200.cuda_abi
    sc = scenes["pr2test2"]
    ctx.set_scene(sc["boxes"], sc["hulls"])
    K, N, R = 3, 4000, 48
    ours = np.array([ctx.gmm_run(plan, K, N, mode=abi.MODE_PHILOX, stream_id=i)["P"] for i in range(R)])
    # same call twice: bit-identical counts -> identical P up to FP64 atomics order
    again = ctx.gmm_run(plan, K, N, mode=abi.MODE_PHILOX, stream_id=0)["P"]
    assert abs(again - ours[0]) < 1e-9
    if not oracle_ref.available():
        pytest.skip("oracle/_ref not built")
    ref = []
    for s in range(R):
        try:
            ref.append(oracle_ref.gmm_run(plan, sc["boxes"], sc["hulls"], K, N, mt_seed=s, multinomial_seed=s, record=False)["P"])
        except RuntimeError:
            pass
    ref = np.array(ref)
    half = 2.576 * np.sqrt(ours.var(ddof=1) / len(ours) + ref.var(ddof=1) / len(ref))
    print("GMM philox: ours mean %.4f (std %.4f), reference mean %.4f (std %.4f), 99%% half-width %.4f" %
          (ours.mean(), ours.std(ddof=1), ref.mean(), ref.std(ddof=1), half))
    assert abs(ours.mean() - ref.mean()) < half


def test_gmm_batch_members_equal_single_runs(ctx, plan, scenes):
    """mcsim_cuda_gmm_run_batch: member b == mcsim_cuda_gmm_run on Philox subsequence stream0 + b (the FP64
    sums are accumulated in a different order: agreement to 1e-9, not bit-exact)."""
    import pcoll_b200
    abi = pcoll_b200.cuda_abi
    ctx.set_scene(scenes["pr2test2"]["boxes"], scenes["pr2test2"]["hulls"])
    batch = ctx.gmm_run_batch(plan, 3, 5000, 7, stream0=11)
    for b in range(7):
        one = ctx.gmm_run(plan, 3, 5000, mode=abi.MODE_PHILOX, stream_id=11 + b)
        np.testing.assert_allclose(batch["probabilities"][b], one["probabilities"], rtol=1e-9, atol=1e-12)
        np.testing.assert_allclose(batch["P"][b], one["P"], rtol=1e-9)
    assert len(set(np.round(batch["P"], 6))) > 1     # different subsequences, different runs
    # different plans in one batch: plan 0 = the committed plan, plan 1 = the same plan driven backwards in time is
    # not a valid plan, so use a shifted copy (same odometry, start moved 5 cm)
    tr = np.stack([plan["trajectory"], plan["trajectory"] + np.array([[0.05], [0.0], [0.0]])])
    od = np.stack([plan["odometry"], plan["odometry"]])
    two = ctx.gmm_run_batch(plan, 3, 5000, 2, trajectories=tr, odometry=od, stream0=11)
    np.testing.assert_allclose(two["P"][0], batch["P"][0], rtol=1e-9)
    assert two["P"][1] != two["P"][0]


def test_gmm_broad_phase_is_exact(ctx, plan, scenes):
    """Philox-mode GMM sampling with the candidate lists (default) against testing every live pair: the collision
    flag of every sample is the same (same SAT arithmetic on the listed pairs), so the per-waypoint colliding
    fractions -- ratios of integer counts -- and P are identical, the refitted mixtures agree to the order of the
    FP64 atomics.  Single runs on both scenes, and a batch whose members have different plans
    (their own lists)."""
    import pcoll_b200
    abi = pcoll_b200.cuda_abi
    T = plan["trajectory"].shape[1]
    for name in ("pr2test2", "pr2custom"):
        ctx.set_scene(scenes[name]["boxes"], scenes[name]["hulls"])
        a = ctx.gmm_run(plan, 3, 20000, mode=abi.MODE_PHILOX, stream_id=5, record=True)
        b = ctx.gmm_run(plan, 3, 20000, mode=abi.MODE_PHILOX, stream_id=5, record=True, exhaustive=True)
        np.testing.assert_array_equal(a["probabilities"], b["probabilities"])
        np.testing.assert_allclose(a["gmm_post"], b["gmm_post"], rtol=1e-9, atol=1e-15)   # FP64 atomics: order varies
        assert abs(a["P"] - b["P"]) <= 1e-12
        assert b["pair_tests"] == 20000 * T * ctx.n_pairs
        print("%s: P = %.5f, pair tests per sample %.2f (every pair: %d)" % (name, a["P"], a["pair_tests"] / (20000 * T), ctx.n_pairs))
        assert a["pair_tests"] < 0.3 * b["pair_tests"]
    tr = np.stack([plan["trajectory"], plan["trajectory"] + np.array([[0.05], [0.0], [0.0]]), plan["trajectory"] + np.array([[0.0], [-0.08], [0.0]])])
    od = np.stack([plan["odometry"]] * 3)
    a = ctx.gmm_run_batch(plan, 3, 5000, 3, trajectories=tr, odometry=od, stream0=11)
    b = ctx.gmm_run_batch(plan, 3, 5000, 3, trajectories=tr, odometry=od, stream0=11, exhaustive=True)
    np.testing.assert_array_equal(a["probabilities"], b["probabilities"])
    assert a["pair_tests"] < b["pair_tests"] == 3 * 5000 * T * ctx.n_pairs


def test_gmm_batch_fused_kernel_members_equal_single_runs(ctx, plan, scenes):
    """Batches of >= 64 members run as one CTA per member for the whole plan (gmm_member_run_kernel): a member must
    still equal the single run on its Philox subsequence (same draws, same refit code; FP64 sums in another order),
    with and without the broad phase, and with members on different plans."""
    import pcoll_b200
    abi = pcoll_b200.cuda_abi
    T = plan["trajectory"].shape[1]
    ctx.set_scene(scenes["pr2test2"]["boxes"], scenes["pr2test2"]["hulls"])
    B = 80
    batch = ctx.gmm_run_batch(plan, 3, 5000, B, stream0=300)
    every = ctx.gmm_run_batch(plan, 3, 5000, B, stream0=300, exhaustive=True)
    np.testing.assert_array_equal(batch["probabilities"], every["probabilities"])
    assert every["pair_tests"] == B * 5000 * T * ctx.n_pairs and batch["pair_tests"] < 0.3 * every["pair_tests"]
    for b in (0, 1, 41, 79):
        one = ctx.gmm_run(plan, 3, 5000, mode=abi.MODE_PHILOX, stream_id=300 + b)
        np.testing.assert_allclose(batch["probabilities"][b], one["probabilities"], rtol=1e-9, atol=1e-12)
        np.testing.assert_allclose(batch["P"][b], one["P"], rtol=1e-9)
    shift = np.zeros((B, 3, 1))
    shift[:, 0, 0] = np.linspace(-0.04, 0.04, B)
    tr = plan["trajectory"][None] + shift
    od = np.repeat(plan["odometry"][None], B, axis=0)
    many = ctx.gmm_run_batch(plan, 2, 4000, B, trajectories=tr, odometry=od, stream0=7)
    few = ctx.gmm_run_batch(plan, 2, 4000, 3, trajectories=tr[[0, 17, 79]], odometry=od[:3], stream0=7)
    # members 17 and 79 of `many` use subsequences 24 and 86; re-run them alone through the per-waypoint kernels
    np.testing.assert_allclose(many["P"][0], few["P"][0], rtol=1e-9)
    for b in (17, 79):
        alone = ctx.gmm_run_batch(plan, 2, 4000, 1, trajectories=tr[[b]], odometry=od[:1], stream0=7 + b)
        np.testing.assert_allclose(many["probabilities"][b], alone["probabilities"][0], rtol=1e-9, atol=1e-12)


@pytest.mark.parametrize("K,N", [(1, 3000), (16, 4000), (3, 40)])
def test_gmm_batch_fused_kernel_edge_cases(ctx, plan, scenes, K, N):
    """One-CTA-per-run kernel against the per-waypoint kernels (the same members run as a batch of 2, below the
    64-run threshold): one component, the maximum component count, and so few samples that components starve
    (degenerate refits are counted, not fatal); a two-waypoint plan."""
    ctx.set_scene(scenes["pr2custom"]["boxes"], scenes["pr2custom"]["hulls"])
    B = 64
    fused = ctx.gmm_run_batch(plan, K, N, B, stream0=1000)
    for b0 in (0, 62):
        pair = ctx.gmm_run_batch(plan, K, N, 2, stream0=1000 + b0)
        np.testing.assert_allclose(fused["probabilities"][b0:b0 + 2], pair["probabilities"], rtol=1e-9, atol=1e-12)
    assert np.all((fused["P"] >= 0) & (fused["P"] <= 1))
    short = dict(plan)
    short["trajectory"] = plan["trajectory"][:, :2]
    short["odometry"] = plan["odometry"][:, :1]
    f2 = ctx.gmm_run_batch(short, K, N, B, stream0=5)
    p2 = ctx.gmm_run_batch(short, K, N, 2, stream0=5)
    np.testing.assert_allclose(f2["probabilities"][:2], p2["probabilities"], rtol=1e-9, atol=1e-12)


@pytest.mark.parametrize("B,K,N", [(256, 3, 2000), (592, 5, 1500), (592, 2, 2049)])
def test_gmm_batch_fused_kernel_cta_shapes(ctx, plan, scenes, B, K, N):
    """The one-CTA-per-run kernel is launched with 512, 256 or 128 threads per run depending on the batch size
    (gmm_member_threads: 148-SM B200 -> 256 runs take 256 threads, 592 take 128); the sample -> thread mapping changes with
    the shape, the draws per sample index do not: members equal the per-waypoint kernels' runs whatever the shape.
    K = 5 takes the count pass's generic path (more than 4 components), 2,049 samples leave a one-sample tail window."""
    ctx.set_scene(scenes["pr2test2"]["boxes"], scenes["pr2test2"]["hulls"])
    fused = ctx.gmm_run_batch(plan, K, N, B, stream0=4000)
    for b0 in (0, B // 2 - 1, B - 2):
        pair = ctx.gmm_run_batch(plan, K, N, 2, stream0=4000 + b0)
        np.testing.assert_allclose(fused["probabilities"][b0:b0 + 2], pair["probabilities"], rtol=1e-9, atol=1e-12)
    assert np.all((fused["P"] >= 0) & (fused["P"] <= 1)) and fused["P"].std() > 0


@pytest.mark.parametrize("K,N", [(3, 10000), (1, 10001), (16, 3000), (2, 100), (3, 32768)])
def test_gmm_cluster_kernel_equals_per_waypoint_kernels(ctx, plan, scenes, K, N, monkeypatch):
    """One-cluster-per-run kernel (single runs and batches below 64 in Philox mode, N <= 32,768) against the
    per-waypoint kernels (mcsim_gmm_run_params.kernel = MCSIM_GMM_KERNEL_PER_WAYPOINT keeps them): same Philox draws per sample index, FP64 sums in a
    different order -> per-waypoint fractions, recorded mixtures and P to 1e-9.  Cases: the reference's default, a
    sample count that is no multiple of 4, the maximum component count, so few samples that most CTAs of the cluster
    idle and components starve, the largest routed sample count; a two-waypoint plan; a batch on 8-CTA clusters."""
    import pcoll_b200
    abi = pcoll_b200.cuda_abi
    ctx.set_scene(scenes["pr2test2"]["boxes"], scenes["pr2test2"]["hulls"])
    short = dict(plan)
    short["trajectory"] = plan["trajectory"][:, :2]
    short["odometry"] = plan["odometry"][:, :1]
    for pl in (plan, short):
        a = ctx.gmm_run(pl, K, N, mode=abi.MODE_PHILOX, stream_id=21, record=True)
        b = ctx.gmm_run(pl, K, N, mode=abi.MODE_PHILOX, stream_id=21, record=True, kernel=abi.GMM_KERNEL_PER_WAYPOINT)
        np.testing.assert_allclose(a["probabilities"], b["probabilities"], rtol=1e-9, atol=1e-12)
        np.testing.assert_allclose(a["P"], b["P"], rtol=1e-9)
        np.testing.assert_allclose(a["gmm_pre"], b["gmm_pre"], rtol=1e-7, atol=1e-12)
        np.testing.assert_allclose(a["gmm_post"], b["gmm_post"], rtol=1e-7, atol=1e-12)
        assert a["degenerate"] == b["degenerate"]
    batch = ctx.gmm_run_batch(plan, K, N, 12, stream0=40)       # 12 runs: 8 CTAs per cluster
    for b in (0, 11):
        one = ctx.gmm_run(plan, K, N, mode=abi.MODE_PHILOX, stream_id=40 + b, kernel=abi.GMM_KERNEL_PER_WAYPOINT)
        np.testing.assert_allclose(batch["probabilities"][b], one["probabilities"], rtol=1e-9, atol=1e-12)


@pytest.mark.parametrize("T", [120, 250])
def test_gmm_cluster_kernel_long_plans(ctx, plan, scenes, T, monkeypatch):
    """Plans whose candidate lists make static + dynamic shared memory of the cluster kernel pass 48 KB (T = 120: 23 KB of
    lists beside 30 KB of static slots; T = 250: 48 KB of lists): same results as the per-waypoint kernels."""
    import pcoll_b200
    abi, inp = pcoll_b200.cuda_abi, pcoll_b200.inputs
    boxes = inp.synthetic_obstacle_field(5, 64)
    ctx.set_scene(boxes, scenes["pr2test2"]["hulls"])
    tr, od = inp.synthetic_plans(5, 1, T, boxes, clear=-0.1)
    long_plan = dict(plan)
    long_plan["trajectory"], long_plan["odometry"] = tr[0], od[0]
    a = ctx.gmm_run(long_plan, 3, 4000, mode=abi.MODE_PHILOX, stream_id=3)
    b = ctx.gmm_run(long_plan, 3, 4000, mode=abi.MODE_PHILOX, stream_id=3, kernel=abi.GMM_KERNEL_PER_WAYPOINT)
    np.testing.assert_allclose(a["probabilities"], b["probabilities"], rtol=1e-9, atol=1e-12)
    assert a["kernel_ms"] < b["kernel_ms"]
