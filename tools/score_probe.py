"""Developer probe (not the bench): the config-4 subset of bench.py (128 plans x 65,536 rollouts x 500 waypoints,
256 boxes) through the scoring kernel at a few broad-phase margins; prints nominal rollout-steps/s and executed pair
tests per step, and checks the hits of a small batch against the exhaustive test."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pcoll_b200  # noqa: E402

abi, inp = pcoll_b200.cuda_abi, pcoll_b200.inputs
gold = os.path.join(ROOT, "tests", "golden")
plan = pcoll_b200.load_plan_json(os.path.join(gold, "plan_committed.json"))
sc = np.load(os.path.join(gold, "scene_pr2test2.npz"))
ctx = abi.Context(0)
boxes = inp.synthetic_obstacle_field(4096, 256)
n_plans, T4, R = 128, 500, 65536
tr, od = inp.synthetic_plans(4096, n_plans, T4, boxes)
ctx.set_scene(boxes, sc["hulls"])
full = ctx.score_plans(plan, tr[:8], od[:8], 4096 + 37, margin=-1.0)
for m in (1.5, 0.7, 0.05):
    np.testing.assert_array_equal(ctx.score_plans(plan, tr[:8], od[:8], 4096 + 37, margin=m)["hits"], full["hits"])
print("hits equal to the exhaustive test at margins 1.5, 0.7, 0.05")
margins = [float(x) for x in sys.argv[1:]] or [3.0, 1.5]
for m in margins:
    ctx.score_prepare(plan, tr, od, R, margin=m)
    ctx.score_launch()
    ctx.score_collect()
    best = 1e30
    for _ in range(2):
        r = ctx.score_plans(plan, tr, od, R, margin=m)
        best = min(best, r["kernel_ms"])
    print("margin %.2f: %.3e nominal rollout-steps/s, %.2f pair tests per step, kernel %.1f ms, P mean %.4f" %
          (m, n_plans * R * T4 / (best * 1e-3), r["pair_tests"] / r["executed_steps"], best, r["P"].mean()))
ctx.close()
