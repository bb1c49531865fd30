"""Target program of the ncu capture of the scoring kernel (profiles/): BASELINE config-4 subset, 32 plans x 16,384
rollouts x 500 waypoints against the 256-box field, 8-level broad phase at margin 3 m.  Usage: ncu_score_target.py [margin]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pcoll_b200  # noqa: E402

abi, inp = pcoll_b200.cuda_abi, pcoll_b200.inputs
margin = float(sys.argv[1]) if len(sys.argv) > 1 else 3.0
gold = os.path.join(ROOT, "tests", "golden")
plan = pcoll_b200.load_plan_json(os.path.join(gold, "plan_committed.json"))
sc = np.load(os.path.join(gold, "scene_pr2test2.npz"))
ctx = abi.Context(0)
boxes = inp.synthetic_obstacle_field(4096, 256)
tr, od = inp.synthetic_plans(4096, 32, 500, boxes)
ctx.set_scene(boxes, sc["hulls"])
for _ in range(2):
    r = ctx.score_plans(plan, tr, od, 16384, margin=margin)
print("margin %s: kernel %.3f ms, %.3e nominal rollout-steps/s, %.2f pair tests per step" %
      (margin, r["kernel_ms"], 32 * 16384 * 500 / (r["kernel_ms"] * 1e-3), r["pair_tests"] / r["executed_steps"]))
ctx.close()
