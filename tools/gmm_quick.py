"""Developer probe: GMM mode timings (single run through the cluster kernel, batches through the one-CTA-per-run
kernel) and the batch kernel's fraction of the FP32 peak."""
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
gold = os.path.join(ROOT, "tests", "golden")
plan = pcoll_b200.load_plan_json(os.path.join(gold, "plan_committed.json"))
sc = np.load(os.path.join(gold, "scene_pr2test2.npz"))
if len(sys.argv) > 1 and sys.argv[1] == "sweep":
    K_LIST, B_LIST = (), ()
else:
    K_LIST, B_LIST = ((3, 10000), (1, 10000), (16, 10000), (3, 1000000)), (64, 256, 1184, 1332, 2368, 2664)
ctx = abi.Context(0)
fp32 = ctx.measure_peak(0, 100.0)
ctx.set_scene(sc["boxes"], sc["hulls"])
T = plan["trajectory"].shape[1]
for K, ns in K_LIST:
    ctx.gmm_run(plan, K, ns)
    t0 = time.perf_counter()
    for i in range(10):
        g = ctx.gmm_run(plan, K, ns, seed=1 + i)
    dt = (time.perf_counter() - t0) / 10
    print("single run K=%d N=%d: %.1f trajectories/s, kernel %.3f ms, P=%.4f" % (K, ns, 1 / dt, g["kernel_ms"], g["P"]))
for B in B_LIST:
    ctx.gmm_run_batch(plan, 3, 10000, B)
    best = None
    for rep in range(3):
        t0 = time.perf_counter()
        g = ctx.gmm_run_batch(plan, 3, 10000, B, stream0=B * (rep + 1))
        dt = time.perf_counter() - t0
        best = dt if best is None or dt < best else best
    checks = B * T * 10000
    flops = checks * 35 + g["pair_tests"] * 40
    print("batch %5d: %.0f trajectories/s (kernel only %.0f), %.3f ms, %.2f tests/sample, frac of FP32 peak %.4f" % (
        B, B / best, B / (g["kernel_ms"] * 1e-3), g["kernel_ms"], g["pair_tests"] / checks, flops / (g["kernel_ms"] * 1e-3) / 1e12 / fp32))
ctx.close()


def sweep():
    """cluster kernel (one launch per run) against the three-kernels-per-waypoint chain over the sample count"""
    ctx = abi.Context(0)
    ctx.set_scene(sc["boxes"], sc["hulls"])
    for ns in (4096, 8192, 16384, 32768, 65536, 131072, 262144, 524288, 1048576):
        row = []
        for kern in (abi.GMM_KERNEL_CLUSTER, abi.GMM_KERNEL_PER_WAYPOINT):
            ctx.gmm_run(plan, 3, ns, kernel=kern)
            best = min(ctx.gmm_run(plan, 3, ns, seed=2 + i, kernel=kern)["kernel_ms"] for i in range(5))
            row.append(best)
        print("N=%8d: cluster %.3f ms, per-waypoint %.3f ms -> %s" % (ns, row[0], row[1], "cluster" if row[0] < row[1] else "per-waypoint"))
    ctx.close()


if len(sys.argv) > 1 and sys.argv[1] == "sweep":
    sweep()
