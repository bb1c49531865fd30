"""Target program of the GMM captures (profiles/): one warm-up and one measured batch of runGMMEstimation runs,
committed plan, pr2test2 scene, K = 3, 10^4 samples per waypoint.  Usage: ncu_gmm_target.py [n_traj] [exhaustive]"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pcoll_b200  # noqa: E402

abi = pcoll_b200.cuda_abi
if "--lib" in sys.argv:     # a variant build (tools/build_variant.sh)
    i = sys.argv.index("--lib")
    abi.LIB_PATH = os.path.abspath(sys.argv[i + 1])
    del sys.argv[i:i + 2]
B = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
exhaustive = len(sys.argv) > 2 and sys.argv[2] == "exhaustive"
gold = os.path.join(ROOT, "tests", "golden")
plan = pcoll_b200.load_plan_json(os.path.join(gold, "plan_committed.json"))
sc = np.load(os.path.join(gold, "scene_pr2test2.npz"))
ctx = abi.Context(0)
ctx.set_scene(sc["boxes"], sc["hulls"])
for _ in range(2):
    t0 = time.perf_counter()
    out = ctx.gmm_run_batch(plan, 3, 10000, B, stream0=100, exhaustive=exhaustive)
    dt = time.perf_counter() - t0
print("B=%d P_mean=%.5f kernel %.3f ms (%.0f trajectories/s), call %.3f ms (%.0f trajectories/s), pair tests per sample %.2f" % (
    B, float(np.mean(out["P"])), out["kernel_ms"], B / out["kernel_ms"] * 1e3, dt * 1e3, B / dt, out["pair_tests"] / (B * 10000 * 56)))
ctx.close()
