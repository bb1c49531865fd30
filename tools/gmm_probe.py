"""Developer probe (not the bench): one runGMMEstimation at a time (K = 3, N = 10^4 and 10^6, committed plan,
pr2test2, Philox mode) -- wall time per run and the device time of its launch chain."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pcoll_b200  # noqa: E402

abi = pcoll_b200.cuda_abi
gold = os.path.join(ROOT, "tests", "golden")
plan = pcoll_b200.load_plan_json(os.path.join(gold, "plan_committed.json"))
sc = np.load(os.path.join(gold, "scene_pr2test2.npz"))
ctx = abi.Context(0)
ctx.set_scene(sc["boxes"], sc["hulls"])
for K, ns in ((3, 10000), (1, 10000), (3, 1000000)):
    ctx.gmm_run(plan, K, ns, mode=abi.MODE_PHILOX)
    t0 = time.perf_counter()
    reps = 20
    ks = []
    for i in range(reps):
        g = ctx.gmm_run(plan, K, ns, mode=abi.MODE_PHILOX, seed=1 + i)
        ks.append(g["kernel_ms"])
    dt = (time.perf_counter() - t0) / reps
    print("K=%d N=%d: %.3f ms per run wall (%.0f trajectories/s), launch chain %.3f ms, P=%.5f" % (K, ns, dt * 1e3, 1 / dt, np.median(ks), g["P"]))
ctx.close()
