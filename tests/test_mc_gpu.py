"""GPU parity tests of the MC rollout kernel, through the C ABI (include/mcsim_cuda.h).

Deterministic (replay) mode: injected draws; per-rollout collision counts / flags must be bit-exact
with the reference estimator (oracle/_ref, golden vectors in tests/golden/mc_replay_*.npz) except
for rollouts that have a pose within NEAR_EPS of contact, which are counted and reported
separately (BASELINE.json north_star); poses within 1e-5 relative.
Philox mode: P inside the 99 % confidence interval of the reference's run-to-run mean."""
import json
import os

import numpy as np
import pytest

from conftest import GOLD, golden

pytestmark = pytest.mark.gpu

NEAR_EPS = 1e-4     # metres: |clearance| below this marks a near-contact pose
POSE_RTOL = 1e-5    # north_star: propagated poses within 1e-5 relative


def _replay(ctx, plan, scene, draws, N, **kw):
    import pcoll_b200
    abi = pcoll_b200.cuda_abi
    ctx.set_scene(scene["boxes"], scene["hulls"])
    ht = abi.host_thread(plan, draws[3 * N:])
    p, keep = abi.make_mc_params(plan, n_rollouts=N, mode=abi.MODE_REPLAY, applied=ht["applied_noisy"],
                                 z0=draws[:3 * N], near_eps=NEAR_EPS, early_exit=kw.pop("early_exit", False))
    return ctx.mc_run(p, **kw)


@pytest.mark.parametrize("name", ["pr2test2", "pr2custom"])
def test_replay_counts_bit_exact_vs_reference_golden(ctx, plan, scenes, name):
    g = golden("mc_replay_%s.npz" % name)
    N = g["counts"].shape[0]
    out = _replay(ctx, plan, scenes[name], g["draws"], N, want_counts=True, want_hist=True)
    near = (np.abs(g["clearance"].astype(np.float64)) < NEAR_EPS).any(axis=0)  # rollouts with a near-contact pose
    clean = ~near
    assert clean.sum() >= 0.95 * N
    np.testing.assert_array_equal(out["counts"][clean], g["counts"][clean])
    # hit count: exact on the clean set; the near-contact rollouts are reported, not asserted
    assert int((out["counts"][clean] > 0).sum()) == int((g["counts"][clean] > 0).sum())
    n_diff = int((out["counts"][near] != g["counts"][near]).sum())
    print("%s: N=%d hits gpu=%d ref=%d; near-contact rollouts=%d (of which %d differ)" %
          (name, N, out["hits"], int((g["counts"] > 0).sum()), int(near.sum()), n_diff))
    assert abs(out["hits"] - int((g["counts"] > 0).sum())) <= int(near.sum())
    assert out["executed_steps"] == N * plan["trajectory"].shape[1]
    assert int(out["first_hit_hist"].sum()) == out["hits"]


@pytest.mark.parametrize("name", ["pr2test2", "pr2custom"])
def test_replay_poses_and_clearance(ctx, plan, scenes, name):
    g = golden("mc_replay_%s.npz" % name)
    N = g["counts"].shape[0]
    out = _replay(ctx, plan, scenes[name], g["draws"], N, want_counts=True, audit=True)
    poses = np.transpose(out["poses"], (0, 2, 1))[:, :64, :].astype(np.float64)   # [T, 64, 3]
    ref = g["poses_first64"]
    # angles live on [0, 2 pi]: compare on the circle
    d = poses - ref
    d[..., 2] = (d[..., 2] + np.pi) % (2 * np.pi) - np.pi
    scale = np.maximum(np.abs(ref), 1.0)
    assert np.max(np.abs(d) / scale) < POSE_RTOL
    np.testing.assert_allclose(out["clearance"], g["clearance"], atol=2e-5)
    gpu_near = int((np.abs(out["clearance"]) < NEAR_EPS).sum())
    assert out["near_contact"] == gpu_near
    # flags from the audit launch agree with the exhaustive launch wherever the pose is not near contact
    far = np.abs(g["clearance"]) >= NEAR_EPS
    np.testing.assert_array_equal((out["clearance"] <= 0)[far], (g["clearance"] <= 0)[far])


def test_replay_early_exit_flags(ctx, plan, scenes):
    g = golden("mc_replay_pr2test2.npz")
    N = g["counts"].shape[0]
    full = _replay(ctx, plan, scenes["pr2test2"], g["draws"], N, want_counts=True)
    ee = _replay(ctx, plan, scenes["pr2test2"], g["draws"], N, want_counts=True, want_hist=True, early_exit=True)
    np.testing.assert_array_equal(ee["counts"], (full["counts"] > 0).astype(np.uint32))
    assert ee["hits"] == full["hits"]
    assert ee["executed_steps"] < full["executed_steps"]


@pytest.mark.parametrize("name", ["pr2test2", "pr2custom"])
def test_replay_broad_phase_equals_every_pair(ctx, plan, scenes, name):
    """The replay kernel's broad phase (the plan's candidate lists, level chosen per warp: the reference driver's default
    `shared` mode runs through it) against testing every live pair at every pose: identical per-rollout counts wherever
    no pose is within 1e-4 m of contact (the listed pairs go through the grouped-table arithmetic, the exhaustive loop
    through the pair table: they may round differently AT contact), with and without early exit; fewer pair tests."""
    import pcoll_b200
    abi = pcoll_b200.cuda_abi
    ctx.set_scene(scenes[name]["boxes"], scenes[name]["hulls"])
    T = plan["trajectory"].shape[1]
    N = 20011
    draws = np.random.Generator(np.random.PCG64(77)).standard_normal(3 * N + 11 * (T - 1))
    ht = abi.host_thread(plan, draws[3 * N:])
    kw = dict(n_rollouts=N, mode=abi.MODE_REPLAY, applied=ht["applied_noisy"], z0=draws[:3 * N])
    aud = ctx.mc_run(abi.make_mc_params(plan, early_exit=False, **kw)[0], audit=True)
    near = (np.abs(aud["clearance"]) < NEAR_EPS).any(axis=0)
    for ee in (False, True):
        a = ctx.mc_run(abi.make_mc_params(plan, early_exit=ee, **kw)[0], want_counts=True, want_hist=True)
        b = ctx.mc_run(abi.make_mc_params(plan, early_exit=ee, exhaustive=True, **kw)[0], want_counts=True, want_hist=True)
        np.testing.assert_array_equal(a["counts"][~near], b["counts"][~near])
        assert abs(a["hits"] - b["hits"]) <= int(near.sum())
        assert b["pair_tests"] == b["executed_steps"] * ctx.n_pairs and a["pair_tests"] < 0.3 * b["pair_tests"]
        print("%s ee=%d: replay kernel %.1f us with the lists (%.2f pair tests per pose), %.1f us testing all %d pairs" % (
            name, ee, a["kernel_ms"] * 1e3, a["pair_tests"] / a["executed_steps"], b["kernel_ms"] * 1e3, ctx.n_pairs))


def test_replay_vs_live_reference_random_draws(ctx, plan, scenes):
    """Fresh seeds and sizes (incl. ragged, non-multiple-of-warp N) against the reference run live."""
    import oracle_ref
    if not oracle_ref.available():
        pytest.skip("oracle/_ref not built")
    T = plan["trajectory"].shape[1]
    for seed, N in ((1, 1), (2, 33), (3, 257), (4, 1500)):
        draws = np.random.Generator(np.random.PCG64(seed)).standard_normal(3 * N + 11 * (T - 1))
        ref = oracle_ref.mc_run(plan, scenes["pr2test2"]["boxes"], scenes["pr2test2"]["hulls"], N, draws=draws)
        out = _replay(ctx, plan, scenes["pr2test2"], draws, N, want_counts=True)
        near = (np.abs(ref["clearance"]) < NEAR_EPS).any(axis=0)
        np.testing.assert_array_equal(out["counts"][~near], ref["counts"][~near])


def test_zero_rollouts(ctx, plan, scenes):
    import pcoll_b200
    abi = pcoll_b200.cuda_abi
    ctx.set_scene(scenes["pr2test2"]["boxes"], scenes["pr2test2"]["hulls"])
    p, keep = abi.make_mc_params(plan, n_rollouts=0, mode=abi.MODE_PHILOX)
    out = ctx.mc_run(p)
    assert out["hits"] == 0 and out["executed_steps"] == 0


def test_philox_inside_reference_confidence_interval(ctx, plan, scenes):
    import pcoll_b200
    abi = pcoll_b200.cuda_abi
    dist = json.load(open(os.path.join(GOLD, "mc_dist_pr2test2.json")))
    half = 2.576 * dist["std"] / np.sqrt(dist["runs"])
    ctx.set_scene(scenes["pr2test2"]["boxes"], scenes["pr2test2"]["hulls"])
    N = 1 << 20
    p, keep = abi.make_mc_params(plan, n_rollouts=N, mode=abi.MODE_PHILOX, early_exit=True)
    out = ctx.mc_run(p)
    print("philox P=%.5f reference mean=%.5f +- %.5f (99%% CI, %d reference runs)" % (out["P"], dist["mean"], half, dist["runs"]))
    assert half < 0.003      # 2,000 reference runs (tests/golden/make_round2.py)
    assert abs(out["P"] - dist["mean"]) < half
    # early exit does not change the estimate
    p2, keep2 = abi.make_mc_params(plan, n_rollouts=N, mode=abi.MODE_PHILOX, early_exit=False)
    out2 = ctx.mc_run(p2, want_counts=True)
    assert out2["hits"] == out["hits"]
    assert out2["executed_steps"] == N * plan["trajectory"].shape[1]


def test_philox_step_normals_distribution(ctx):
    """The packed kernel's step draws (24 radius + 16 angle bits per Box-Muller pair, MUFU lg2 / sqrt / sin / cos,
    csrc/philox.h) dumped from the device function the kernel itself calls: 10^7 draws against N(0, 1) --
    Kolmogorov-Smirnov distance below the 1 % critical value 1.628 / sqrt(n), tail masses beyond 3, 4 and 5 sigma
    inside 4.5-sigma Poisson bounds, moments and pair correlation per draw slot -- and against the FP64 host
    recomputation of the same bits (mcsim_step_normals)."""
    import ctypes as C
    from scipy import stats
    import pcoll_b200
    abi = pcoll_b200.cuda_abi
    n_roll = 833_334                       # x 12 = 10,000,008 draws
    z = ctx.step_normals(n_roll, t=17, id0=5_000_000_000)   # ids beyond 2^32: the high counter word is in play
    assert z.shape == (n_roll, 12) and np.isfinite(z).all()
    flat = z.astype(np.float64).ravel()
    n = flat.size
    d, pval = stats.kstest(flat, "norm")
    print("KS distance %.3e (1%% critical %.3e), max |z| %.3f" % (d, 1.628 / np.sqrt(n), np.abs(flat).max()))
    assert d < 1.628 / np.sqrt(n)
    for k, p_tail in ((3.0, 2.6997960632601866e-3), (4.0, 6.334248366623973e-5), (5.0, 5.733031437583878e-7)):
        got, want = int((np.abs(flat) > k).sum()), n * p_tail
        print("|z| > %.0f: %d draws, expected %.1f" % (k, got, want))
        assert abs(got - want) < 4.5 * np.sqrt(want) + 1
    assert np.abs(flat).max() < 5.9       # sqrt(2 * 25 ln 2) = 5.887: the 24-bit radius cap
    # per slot: mean, variance, kurtosis, correlation with the partner of its Box-Muller pair
    for j in range(12):
        col = z[:, j].astype(np.float64)
        assert abs(col.mean()) < 4.5 / np.sqrt(n_roll)
        assert abs(col.var() - 1.0) < 4.5 * np.sqrt(2.0 / n_roll)
        assert abs((col ** 4).mean() - 3.0) < 4.5 * np.sqrt(96.0 / n_roll)
    for j in range(0, 12, 2):
        r = np.corrcoef(z[:, j], z[:, j + 1])[0, 1]
        assert abs(r) < 4.5 / np.sqrt(n_roll)
    r = np.corrcoef(z[:, 0], z[:, 2])[0, 1]
    assert abs(r) < 4.5 / np.sqrt(n_roll)
    # FP64 recomputation from the same Philox bits on the host
    host = np.zeros(12)
    worst = 0.0
    for i in (0, 1, 2, 3, 77, 4096, n_roll - 1):
        assert abi.lib().mcsim_step_normals(0x5AFEC0DE, 5_000_000_000 + i, 17, host.ctypes.data_as(C.POINTER(C.c_double))) == 0
        worst = max(worst, float(np.abs(z[i] - host).max()))
    print("device FP32 / MUFU draws vs FP64 host recomputation: max |diff| %.2e" % worst)
    assert worst < 2e-5


def test_philox_shard_invariance(ctx, plan, scenes):
    """Rollout-id-derived counters: the hit total does not depend on how ids are split."""
    import pcoll_b200
    abi = pcoll_b200.cuda_abi
    ctx.set_scene(scenes["pr2test2"]["boxes"], scenes["pr2test2"]["hulls"])
    N = 100003
    p, keep = abi.make_mc_params(plan, n_rollouts=N, mode=abi.MODE_PHILOX)
    whole = ctx.mc_run(p, want_hist=True)
    for G in (2, 4, 8):
        hits, hist = 0, np.zeros_like(whole["first_hit_hist"])
        for g in range(G):
            lo, hi = g * N // G, (g + 1) * N // G
            pg, kg = abi.make_mc_params(plan, n_rollouts=hi - lo, rollout_offset=lo, mode=abi.MODE_PHILOX)
            o = ctx.mc_run(pg, want_hist=True)
            hits += o["hits"]
            hist += o["first_hit_hist"]
        assert hits == whole["hits"]
        np.testing.assert_array_equal(hist, whole["first_hit_hist"])


@pytest.mark.parametrize("est", ["reference", "closedloop"])
@pytest.mark.parametrize("name", ["pr2test2", "pr2custom"])
def test_philox_packed_kernel_matches_precise_kernel(ctx, plan, scenes, name, est):
    """mc_philox2_kernel (two rollouts per thread, packed FP32 pairs, MUFU approximations) against
    mc_rollout_kernel<PHILOX> (scalar, IEEE division / sqrt, libm trigonometry) on the same Philox
    counters: the rollouts are the same random experiments, so the per-rollout collision counts
    may differ only where a pose is within rounding of contact.  Bars: >= 99 % of the rollouts have
    identical counts, |dP| <= 2e-3; odd N exercises the masked second lane."""
    import pcoll_b200
    abi = pcoll_b200.cuda_abi
    ctx.set_scene(scenes[name]["boxes"], scenes[name]["hulls"])
    e = abi.EST_REFERENCE if est == "reference" else abi.EST_CLOSEDLOOP
    N = 200001
    outs = {}
    for precise in (False, True):
        p, keep = abi.make_mc_params(plan, n_rollouts=N, mode=abi.MODE_PHILOX, estimator=e, early_exit=False,
                                     precise=precise, rollout_offset=12345)
        outs[precise] = ctx.mc_run(p, want_counts=True, want_hist=True)
    a, b = outs[False], outs[True]
    same = float((a["counts"] == b["counts"]).mean())
    flags = float(((a["counts"] > 0) == (b["counts"] > 0)).mean())
    print("%s/%s: packed P=%.5f precise P=%.5f identical counts %.4f identical flags %.5f" % (name, est, a["P"], b["P"], same, flags))
    assert same >= 0.99 and flags >= 0.999
    assert abs(a["P"] - b["P"]) <= 2e-3
    assert a["executed_steps"] == b["executed_steps"] == N * plan["trajectory"].shape[1]
    assert a["pair_tests"] <= b["pair_tests"]     # packed kernel: broad phase on by default
    assert int(a["first_hit_hist"].sum()) == a["hits"] == int((a["counts"] > 0).sum())
    # early exit on the packed kernel: same flags as its own exhaustive run
    p, keep = abi.make_mc_params(plan, n_rollouts=N, mode=abi.MODE_PHILOX, estimator=e, early_exit=True, rollout_offset=12345)
    ee = ctx.mc_run(p, want_counts=True, want_hist=True)
    np.testing.assert_array_equal(ee["counts"], (a["counts"] > 0).astype(np.uint32))
    assert ee["hits"] == a["hits"] and ee["executed_steps"] < a["executed_steps"]
    np.testing.assert_array_equal(ee["first_hit_hist"], a["first_hit_hist"])


@pytest.mark.parametrize("est", ["reference", "closedloop"])
@pytest.mark.parametrize("name", ["pr2test2", "pr2custom"])
def test_philox_broad_phase_is_exact(ctx, plan, scenes, name, est):
    """The packed kernel's broad phase (per-waypoint candidate lists chosen by the warp's largest deviation from
    x*_t, exhaustive fallback) must give the SAME per-rollout collision counts, first-hit histogram and executed
    steps as testing every live pair at every step -- bit-exact, with and without early exit; only the executed
    pair tests shrink.  A second plan with a 36 x larger initial covariance (sigma = 0.19 m) drives warps through every level
    and into the exhaustive fallback."""
    import pcoll_b200
    abi = pcoll_b200.cuda_abi
    ctx.set_scene(scenes[name]["boxes"], scenes[name]["hulls"])
    e = abi.EST_REFERENCE if est == "reference" else abi.EST_CLOSEDLOOP
    N = 150001
    wide = dict(plan)
    wide["cov0"] = np.asarray(plan["cov0"]) * np.array([[36.0], [36.0], [25.0]])
    for label, pl in (("committed", plan), ("wide start", wide)):
        for ee in (False, True):
            outs = {}
            for exhaustive in (False, True):
                p, keep = abi.make_mc_params(pl, n_rollouts=N, mode=abi.MODE_PHILOX, estimator=e, early_exit=ee,
                                             exhaustive=exhaustive, rollout_offset=777)
                outs[exhaustive] = ctx.mc_run(p, want_counts=True, want_hist=True)
            a, b = outs[False], outs[True]
            np.testing.assert_array_equal(a["counts"], b["counts"])
            np.testing.assert_array_equal(a["first_hit_hist"], b["first_hit_hist"])
            assert a["hits"] == b["hits"] and a["executed_steps"] == b["executed_steps"]
            assert b["pair_tests"] == b["executed_steps"] * ctx.n_pairs
            print("%s/%s/%s ee=%d: P=%.4f pair tests per step %.2f (exhaustive %d)" % (
                name, est, label, ee, a["P"], a["pair_tests"] / a["executed_steps"], ctx.n_pairs))
            if label == "committed":
                assert a["pair_tests"] < 0.25 * b["pair_tests"]
            else:
                assert a["pair_tests"] < b["pair_tests"]


@pytest.mark.parametrize("est", ["reference", "closedloop"])
@pytest.mark.parametrize("name", ["pr2test2", "pr2custom"])
def test_early_exit_survivor_compaction_is_exact(ctx, plan, scenes, name, est):
    """Early exit with survivor compaction (runs of >= 2^18 rollouts: one launch per 4 waypoints, the survivors' state
    goes through a compacted buffer in device memory) against the one-launch early-exit kernel and against the full
    per-rollout counts: identical flags for every rollout, identical hit total and first-hit histogram; executed
    rollout-steps = the free ones (sum over waypoints of rollouts without a hit so far) in both early-exit forms -- the
    compacted launches just no longer carry the dead lanes along.  Odd N: masked lane."""
    import pcoll_b200
    abi = pcoll_b200.cuda_abi
    ctx.set_scene(scenes[name]["boxes"], scenes[name]["hulls"])
    e = abi.EST_REFERENCE if est == "reference" else abi.EST_CLOSEDLOOP
    N = (1 << 18) + 4097
    T = plan["trajectory"].shape[1]
    outs = {}
    for key, kw in (("full", dict(early_exit=False)), ("ee", dict(early_exit=True, no_compaction=True)), ("seg", dict(early_exit=True))):
        p, keep = abi.make_mc_params(plan, n_rollouts=N, mode=abi.MODE_PHILOX, estimator=e, rollout_offset=(1 << 33) + 5, **kw)
        outs[key] = ctx.mc_run(p, want_counts=True, want_hist=True)
    flags = (outs["full"]["counts"] > 0).astype(np.uint32)
    for key in ("ee", "seg"):
        np.testing.assert_array_equal(outs[key]["counts"], flags)
        np.testing.assert_array_equal(outs[key]["first_hit_hist"], outs["full"]["first_hit_hist"])
        assert outs[key]["hits"] == outs["full"]["hits"] == int(flags.sum())
    # free rollout-steps: a rollout is tested at waypoints 0 .. first hit
    hist = outs["full"]["first_hit_hist"].astype(np.int64)
    free_steps = int((hist * (np.arange(T) + 1)).sum() + (N - hist.sum()) * T)
    assert outs["seg"]["executed_steps"] == outs["ee"]["executed_steps"] == free_steps
    # (pair tests differ a little either way: compacted warps hold other rollouts, so the warp-wide list level differs)
    print("%s/%s: P=%.4f free steps %.3f of nominal; kernel ms: full %.3f, one-launch early exit %.3f, compaction %.3f" % (
        name, est, outs["full"]["P"], free_steps / (N * T), outs["full"]["kernel_ms"], outs["ee"]["kernel_ms"], outs["seg"]["kernel_ms"]))


def test_early_exit_compaction_across_chunks(ctx, plan, scenes):
    """More than one chunk of 2^24 rollouts (the survivor buffers are reused chunk after chunk) and a plan whose length is
    no multiple of the segment length: hit total and histogram equal to the one-launch kernel's."""
    import pcoll_b200
    abi = pcoll_b200.cuda_abi
    ctx.set_scene(scenes["pr2test2"]["boxes"], scenes["pr2test2"]["hulls"])
    short = dict(plan)
    short["trajectory"] = np.ascontiguousarray(plan["trajectory"][:, :30])
    short["odometry"] = np.ascontiguousarray(plan["odometry"][:, :29])
    N = (1 << 24) + 70001
    outs = []
    for nc in (True, False):
        p, keep = abi.make_mc_params(short, n_rollouts=N, mode=abi.MODE_PHILOX, early_exit=True, no_compaction=nc)
        outs.append(ctx.mc_run(p, want_hist=True))
    assert outs[0]["hits"] == outs[1]["hits"] and outs[0]["executed_steps"] == outs[1]["executed_steps"]
    np.testing.assert_array_equal(outs[0]["first_hit_hist"], outs[1]["first_hit_hist"])


def test_philox_edge_cases_empty_scene_single_waypoint_other_landmark_counts(ctx, plan, scenes):
    """Empty obstacle set (no live pair): no rollout ever hits; T = 1: only the start poses are tested;
    a landmark count other than 8 or a plan longer than the packed kernel's table falls back to the scalar
    kernel and must agree with the packed one statistically."""
    import pcoll_b200
    abi = pcoll_b200.cuda_abi
    hulls = scenes["pr2test2"]["hulls"]
    ctx.set_scene(np.zeros((0, 8)), hulls)
    p, keep = abi.make_mc_params(plan, n_rollouts=5001, mode=abi.MODE_PHILOX, early_exit=False)
    out = ctx.mc_run(p, want_counts=True)
    assert out["hits"] == 0 and out["pair_tests"] == 0 and out["executed_steps"] == 5001 * plan["trajectory"].shape[1]
    ctx.set_scene(scenes["pr2test2"]["boxes"], hulls)
    one = dict(plan)
    one["trajectory"] = np.ascontiguousarray(plan["trajectory"][:, :1])
    one["odometry"] = np.zeros((3, 0))
    p, keep = abi.make_mc_params(one, n_rollouts=4097, mode=abi.MODE_PHILOX, early_exit=False)
    out = ctx.mc_run(p)
    assert out["executed_steps"] == 4097 and out["hits"] <= 4097
    # 6 landmarks -> scalar kernel (NL generic); compare P with the 8-landmark packed run loosely (different filter)
    six = dict(plan)
    six["landmarks"] = np.ascontiguousarray(plan["landmarks"][:, :6])
    p6, keep6 = abi.make_mc_params(six, n_rollouts=100000, mode=abi.MODE_PHILOX)
    p8, keep8 = abi.make_mc_params(plan, n_rollouts=100000, mode=abi.MODE_PHILOX)
    P6, P8 = ctx.mc_run(p6)["P"], ctx.mc_run(p8)["P"]
    assert 0.5 < P6 <= 1.0 and abs(P6 - P8) < 0.1
    # T = 300 > 256 waypoints: plan repeated back and forth is not needed, a long straight plan will do
    T = 300
    long_plan = dict(plan)
    tr = np.zeros((3, T)); tr[0] = -3.0 + 0.01 * np.arange(T); tr[1] = 1.3
    od = np.zeros((3, T - 1)); od[1] = 0.01
    long_plan["trajectory"], long_plan["odometry"] = tr, od
    pl, keepl = abi.make_mc_params(long_plan, n_rollouts=20000, mode=abi.MODE_PHILOX, early_exit=False)
    out = ctx.mc_run(pl)
    assert out["executed_steps"] == 20000 * T
    sc = ctx.score_plans(long_plan, tr[None], od[None], 20000, margin=-1.0)   # packed law, plan table in shared memory
    assert abs(sc["P"][0] - out["P"]) < 0.02, (sc["P"][0], out["P"])


@pytest.mark.parametrize("est", ["reference", "closedloop"])
@pytest.mark.parametrize("name", ["pr2test2", "pr2custom"])
def test_axis_aligned_instantiations_equal_the_generic_kernel(ctx, plan, scenes, name, est):
    """On an axis-aligned scene (every box class axis (1, 0), every hull's yaw 0 -- both committed scenes) the packed
    kernels run instantiations that drop the rotations into the class and hull frames (multiplications by 1, additions
    of 0: the same bits).  One more box, rotated by 30 degrees and 100 m away, makes the scene generic without touching a
    single collision: per-rollout counts, first-hit histogram and executed steps must be identical -- plain launch,
    one-launch early exit and the survivor-compaction path (N >= 2^18)."""
    import pcoll_b200
    abi = pcoll_b200.cuda_abi
    e = abi.EST_REFERENCE if est == "reference" else abi.EST_CLOSEDLOOP
    boxes = np.asarray(scenes[name]["boxes"], dtype=np.float64)
    far = np.array([[100.0, 100.0, np.cos(np.pi / 6), np.sin(np.pi / 6), 0.3, 0.2, boxes[:, 6].min(), boxes[:, 7].max()]])
    outs = {}
    for label, bx in (("axis-aligned", boxes), ("generic", np.vstack([boxes, far]))):
        ctx.set_scene(bx, scenes[name]["hulls"])
        for ee, N in ((False, 150001), (True, 150001), (True, (1 << 18) + 77)):
            p, keep = abi.make_mc_params(plan, n_rollouts=N, mode=abi.MODE_PHILOX, estimator=e, early_exit=ee, rollout_offset=4242)
            outs[(label, ee, N)] = ctx.mc_run(p, want_counts=True, want_hist=True)
    for (label, ee, N), a in outs.items():
        if label != "axis-aligned":
            continue
        b = outs[("generic", ee, N)]
        np.testing.assert_array_equal(a["counts"], b["counts"])
        np.testing.assert_array_equal(a["first_hit_hist"], b["first_hit_hist"])
        assert a["hits"] == b["hits"] and a["executed_steps"] == b["executed_steps"]
    ctx.set_scene(scenes[name]["boxes"], scenes[name]["hulls"])
