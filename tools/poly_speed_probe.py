"""Developer probe: Philox-mode MC rate with the box fixture (packed kernel) and with a polygon-hull robot (scalar kernel +
exact broad phase), committed plan, pr2test2 scene."""
import sys, time
import os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, pcoll_b200
from test_polygon_hulls_gpu import _robot
abi = pcoll_b200.cuda_abi
plan = pcoll_b200.load_plan_json(os.path.join(ROOT, "tests", "golden", "plan_committed.json"))
sc = np.load(os.path.join(ROOT, "tests", "golden", "scene_pr2test2.npz"))
ctx = abi.Context(0)
hulls, polys = _robot()
T = plan["trajectory"].shape[1]
for name, args in (("box fixture", (sc["boxes"], sc["hulls"])), ("octagon base + box body", (sc["boxes"], hulls, polys))):
    ctx.set_scene(*args)
    p, keep = abi.make_mc_params(plan, n_rollouts=4_000_000, mode=abi.MODE_PHILOX, early_exit=False)
    ctx.mc_run(p); r = ctx.mc_run(p)
    print("%s: %.3e rollout-steps/s, P %.4f, pair tests per step %.2f" % (name, 4e6 * T / (r["kernel_ms"] * 1e-3), r["P"], r["pair_tests"] / max(r["executed_steps"], 1)))
