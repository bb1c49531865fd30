"""Target program of the ncu captures (profiles/): one warm-up and one measured launch of the packed
Philox rollout kernel, committed plan, pr2test2 scene, no early exit.  Usage: ncu_target.py [N] [precise]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pcoll_b200  # noqa: E402

abi = pcoll_b200.cuda_abi
N = int(sys.argv[1]) if len(sys.argv) > 1 else 4_000_000
precise = len(sys.argv) > 2 and sys.argv[2] == "precise"
gold = os.path.join(ROOT, "tests", "golden")
plan = pcoll_b200.load_plan_json(os.path.join(gold, "plan_committed.json"))
sc = np.load(os.path.join(gold, "scene_pr2test2.npz"))
ctx = abi.Context(0)
ctx.set_scene(sc["boxes"], sc["hulls"])
p, keep = abi.make_mc_params(plan, n_rollouts=N, mode=abi.MODE_PHILOX, early_exit=False, precise=precise)
for _ in range(2):
    out = ctx.mc_run(p)
print("N=%d P=%.5f hits=%d kernel %.3f ms  %.3e rollout-steps/s  %.3f pair tests per step" %
      (N, out["P"], out["hits"], out["kernel_ms"], out["executed_steps"] / out["kernel_ms"] * 1e3, out["pair_tests"] / max(out["executed_steps"], 1)))
ctx.close()
