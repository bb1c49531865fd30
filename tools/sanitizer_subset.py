"""Small invocations of the kernels with the most intricate synchronisation, for compute-sanitizer (memcheck / racecheck /
synccheck): survivor compaction (shared-memory prefix + global appends), the GMM cluster kernel (remote shared-memory
stores + cluster barriers), the one-CTA-per-run GMM kernel, the EM kernels (cp.async.bulk + mbarrier pipeline; named
barriers between the warps of a group), the packed MC kernel.  Results are checked against the plain paths."""
import os
import sys

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
N = (1 << 18) + 77
outs = [ctx.mc_run(abi.make_mc_params(plan, n_rollouts=N, mode=abi.MODE_PHILOX, early_exit=ee, no_compaction=nc)[0], want_hist=True)
        for ee, nc in ((False, False), (True, True), (True, False))]
assert outs[0]["hits"] == outs[1]["hits"] == outs[2]["hits"], [o["hits"] for o in outs]
print("mc_philox2_kernel / mc_philox2_segment_kernel: hits", outs[0]["hits"], "of", N)
a = ctx.gmm_run(plan, 3, 10000, stream_id=3)
b = ctx.gmm_run(plan, 3, 10000, stream_id=3, kernel=abi.GMM_KERNEL_PER_WAYPOINT)
assert abs(a["P"] - b["P"]) < 1e-9
print("gmm_cluster_run_kernel vs per-waypoint kernels: P", a["P"])
g = ctx.gmm_run_batch(plan, 3, 3000, 64, stream0=5)
print("gmm_member_run_kernel (64 runs): mean P", float(g["P"].mean()))
for K, n in ((2, 40000), (5, 30001), (12, 30000), (16, 20004)):
    x, h, m, c = pcoll_b200.inputs.synthetic_em_problem(70 + K, n, K)
    r = ctx.em_fit(x, h, m, c, 3)
    assert np.all(np.isfinite(r["means"]))
    print("EM K=%d n=%d: avg log p %.6f" % (K, n, r["avg_log_p"][0, -1]))
ctx.close()
print("sanitizer subset done")
