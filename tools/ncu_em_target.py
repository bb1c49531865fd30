"""ncu target: EM E-step (K = 3 and K = 16) on 8 fits x 10^6 samples, and the scoring kernel."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pcoll_b200  # noqa: E402

abi = pcoll_b200.cuda_abi
ctx = abi.Context(0)
probs = [pcoll_b200.inputs.synthetic_em_problem(9000 + f, 10 ** 6, 16) for f in range(8)]
x = np.stack([q[0] for q in probs])
for K in (1, 3, 16):
    h = np.stack([np.full(K, 1.0 / K) for q in probs]); m = np.stack([q[2][:K] for q in probs]); c = np.stack([q[3][:K] for q in probs])
    r = ctx.em_fit(x if K == 1 else None, h, m, c, 3)
    print("K=%d kernel %.3f ms" % (K, r["kernel_ms"]))
gold = os.path.join(ROOT, "tests", "golden")
plan = pcoll_b200.load_plan_json(os.path.join(gold, "plan_committed.json"))
sc = np.load(os.path.join(gold, "scene_pr2test2.npz"))
boxes = pcoll_b200.inputs.synthetic_obstacle_field()
tr, od = pcoll_b200.inputs.synthetic_plans(1, 32, 200, boxes)
ctx.set_scene(boxes, sc["hulls"])
for margin in (1.5, -1.0):
    r = ctx.score_plans(plan, tr, od, 16384, margin=margin)
    print("score margin %s kernel %.3f ms" % (margin, r["kernel_ms"]))
ctx.close()
