"""GPU tests of batched plan scoring (BASELINE config 4), through the C ABI (mcsim_cuda_score_*).

The per-rollout law is the MC Philox law, whose parity with the reference is established in
tests/test_mc_gpu.py; here: (1) a batch of one plan reproduces mcsim_cuda_mc_run on the same Philox
counters, (2) the broad phase is exact (same hits as testing every live pair at every step),
(3) results do not depend on how plans are split over calls (the multi-GPU sharding)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _hulls():
    from conftest import golden
    return golden("scene_pr2test2.npz")["hulls"]


def test_single_plan_batch_equals_mc_run(ctx, plan, scenes):
    import pcoll_b200
    abi = pcoll_b200.cuda_abi
    for name in ("pr2test2", "pr2custom"):
        ctx.set_scene(scenes[name]["boxes"], scenes[name]["hulls"])
        N = 100001
        p, keep = abi.make_mc_params(plan, n_rollouts=N, mode=abi.MODE_PHILOX, early_exit=False)
        mc = ctx.mc_run(p)
        for margin in (-1.0, 0.3):
            sc = ctx.score_plans(plan, plan["trajectory"][None], plan["odometry"][None], N, margin=margin)
            # same counters and arithmetic except the SAT grouping (rounding at contact): |d hits| tiny
            assert abs(int(sc["hits"][0]) - mc["hits"]) <= 5e-4 * N, (name, margin, sc["hits"], mc["hits"])
            assert sc["executed_steps"] == N * plan["trajectory"].shape[1]
        ee = ctx.score_plans(plan, plan["trajectory"][None], plan["odometry"][None], N, margin=0.3, early_exit=True)
        assert ee["hits"][0] == sc["hits"][0] and ee["executed_steps"] < sc["executed_steps"]


def test_scoring_inside_reference_confidence_interval(ctx, plan):
    """Batched scoring against the ORACLE, not against another CUDA path: 3 plans of the config-4 generator on its
    256-box field; the reference's own runSimulation() (oracle/_ref, MCSimulator.h:649-864) was run 256 x 1,000
    particles per plan with the same boxes (tests/golden/score_oracle.json, made by tests/golden/make_round2.py).
    score_plans' P (65,536 rollouts per plan, Philox law = the expectation of the reference's per-run estimate)
    must fall inside the 99 % confidence interval of the reference's run-to-run mean; with the broad phase and
    with every pair tested."""
    import json
    import os
    import pcoll_b200
    from conftest import GOLD
    inp = pcoll_b200.inputs
    ref = json.load(open(os.path.join(GOLD, "score_oracle.json")))
    boxes = inp.synthetic_obstacle_field()
    tr, od = inp.synthetic_plans(ref["seed"], len(ref["plans"]), ref["T"], boxes, clear=ref["clear"])
    for q, r in enumerate(ref["plans"]):
        np.testing.assert_allclose(tr[q, :, 0], r["start"], rtol=0, atol=1e-12)   # same plans as the oracle saw
    ctx.set_scene(boxes, _hulls())
    R = 65536
    for margin in (3.0, -1.0):
        out = ctx.score_plans(plan, tr, od, R, margin=margin)
        for q, r in enumerate(ref["plans"]):
            half = 2.576 * np.sqrt(r["std"] ** 2 / ref["runs"] + out["P"][q] * (1 - out["P"][q]) / R)
            print("plan %d margin %.0f: P %.4f, reference %.4f +- %.4f" % (q, margin, out["P"][q], r["mean"], half))
            assert abs(out["P"][q] - r["mean"]) < half, (q, margin, out["P"][q], r["mean"], half)
    assert min(r["mean"] for r in ref["plans"]) < 0.9     # the anchor is informative (not all plans at P = 1)


def test_broad_phase_is_exact_and_split_invariant(ctx, plan):
    import pcoll_b200
    inp = pcoll_b200.inputs
    boxes = inp.synthetic_obstacle_field()
    ctx.set_scene(boxes, _hulls())
    # clear = -0.15: nominal paths graze the obstacles, so P spans (0, 1)
    tr, od = inp.synthetic_plans(7, 24, 120, boxes, clear=-0.15)
    R = 4096 + 37
    full = ctx.score_plans(plan, tr, od, R, margin=-1.0)
    cull = ctx.score_plans(plan, tr, od, R, margin=1.5)
    tight = ctx.score_plans(plan, tr, od, R, margin=0.02)   # most warps fall back to the exhaustive test
    np.testing.assert_array_equal(cull["hits"], full["hits"])
    np.testing.assert_array_equal(tight["hits"], full["hits"])
    print("pair tests per rollout-step: exhaustive %.1f, margin 1.5 m %.2f, margin 0.02 m %.1f" % tuple(o["pair_tests"] / o["executed_steps"] for o in (full, cull, tight)))
    assert cull["pair_tests"] < 0.1 * full["pair_tests"] and full["pair_tests"] == full["executed_steps"] * 256
    assert cull["pair_tests"] < tight["pair_tests"] <= full["pair_tests"]
    assert 0 < (full["P"] > 0.01).sum() and (full["P"] < 0.99).sum() > 0
    print("P per plan:", np.round(full["P"], 3))
    # plans split over calls (what the multi-GPU sharding does)
    parts = [ctx.score_plans(plan, tr[a:b], od[a:b], R, plan_offset=a, margin=1.5)["hits"] for a, b in ((0, 5), (5, 24))]
    np.testing.assert_array_equal(np.concatenate(parts), full["hits"])
    # device-resident form
    ctx.score_prepare(plan, tr, od, R, margin=1.5)
    ctx.score_launch()
    np.testing.assert_array_equal(ctx.score_collect()["hits"], full["hits"])


def test_planner_to_scorer_loop(tmp_path):
    """tools/plan_and_score.py: candidates planned on the lattice, scored in one launch, safest one written as
    trajectory.dat / odometry.dat; a wider berth must not be less safe than grazing the obstacles."""
    import importlib.util
    import os
    from conftest import ROOT
    spec = importlib.util.spec_from_file_location("plan_and_score", os.path.join(ROOT, "tools", "plan_and_score.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    out = mod.main(["--rollouts", "16384", "--outdir", str(tmp_path)])
    P = [c["P_collision"] for c in out["candidates"]]
    assert len(P) >= 3 and all(0.0 <= p <= 1.0 for p in P)
    assert P[-1] <= P[0] + 0.02 and out["chosen"] == int(np.argmin(P))
    assert os.path.isfile(out["files"][0]) and os.path.isfile(out["files"][1])


def test_broad_phase_levels_with_overflowing_lists(ctx, plan):
    """1,000 boxes: at wide margins the 16 list entries of a waypoint do not hold every pair, so the upper levels
    fall back to the exhaustive test while the lower ones keep their lists -- hits must not change."""
    import pcoll_b200
    inp = pcoll_b200.inputs
    boxes = inp.synthetic_obstacle_field(11, 1000)
    ctx.set_scene(boxes, _hulls())
    tr, od = inp.synthetic_plans(11, 6, 80, boxes, clear=-0.25)
    R = 2048 + 5
    full = ctx.score_plans(plan, tr, od, R, margin=-1.0)
    assert full["pair_tests"] == full["executed_steps"] * 1000
    tests = []
    for m in (0.5, 3.0, 8.0, 40.0):
        o = ctx.score_plans(plan, tr, od, R, margin=m)
        np.testing.assert_array_equal(o["hits"], full["hits"])
        tests.append(o["pair_tests"] / o["executed_steps"])
    print("pair tests per rollout-step at margins 0.5 / 3 / 8 / 40 m: %.1f %.1f %.1f %.1f (exhaustive 1000)" % tuple(tests))
    assert tests[1] < 100
    print("P per plan:", np.round(full["P"], 3))
    assert full["hits"].sum() > 0
