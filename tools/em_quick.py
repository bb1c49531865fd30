"""Developer probe: E-step rate over K (16 fits x 10^6 samples resident, 10 iterations): sample-iterations/s, fraction of
the FP32 peak at 61 flops per (sample, component, iteration), fraction of the HBM peak at 12 B per sample-iteration.
`[EM_KS="3 8"] [EM_MOMENTS=0|1|2] python tools/em_quick.py [lib.so ...]`"""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def one(lib_path):
    import numpy as np
    import pcoll_b200
    abi = pcoll_b200.cuda_abi
    if lib_path:
        abi.LIB_PATH = lib_path
    ctx = abi.Context(0)
    fp32 = ctx.measure_peak(0, 100.0)
    ctx.em_set_moments(int(os.environ.get("EM_MOMENTS", "0")))   # 0 auto, 1 centred, 2 shifted
    F, n, iters = int(os.environ.get("EM_F", "16")), 10 ** 6, 10
    if F == 56:   # BASELINE config 5 exactly as bench.py builds it (56 waypoint buffers, 672 MB)
        import bench
        plan = pcoll_b200.load_plan_json(os.path.join(ROOT, "tests", "golden", "plan_committed.json"))
        x, m16, c16 = bench.config5_buffers(plan, n, F)
        probs = [(None, None, m16[f], c16[f]) for f in range(F)]
    else:
        probs = [pcoll_b200.inputs.synthetic_em_problem(9000 + f % 4, n, 16, centre=(0.3 * f, 0.0, 0.0)) for f in range(F)]
        x = np.stack([q[0] for q in probs])
    first = True
    for K in (int(k) for k in os.environ.get("EM_KS", "1 2 3 4 5 6 8 9 12 13 16").split()):
        h = np.full((F, K), 1.0 / K); m = np.stack([q[2][:K] for q in probs]); c = np.stack([q[3][:K] for q in probs])
        ctx.em_fit(x if first else None, h, m, c, 2)
        first = False
        r = ctx.em_fit(None, h, m, c, iters)
        done = int(r["iters_run"].astype(np.int64).sum())
        sec = r["kernel_ms"] * 1e-3
        print("%s K=%2d: %.3f ms, %.3e sample-iterations/s, FP32 frac %.3f, HBM %.0f GB/s, iters_run %d..%d, avg log p %.6f" % (
            os.path.basename(lib_path or "in-tree"), K, r["kernel_ms"], n * done / sec, 61.0 * K * n * done / sec / 1e12 / fp32, 12.0 * n * done / sec / 1e9,
            r["iters_run"].min(), r["iters_run"].max(), float(np.nanmean(r["avg_log_p"][:, -1]))), flush=True)
    ctx.close()


if __name__ == "__main__":
    if len(sys.argv) > 2 and sys.argv[1] == "--one":
        one(sys.argv[2] if sys.argv[2] != "-" else None)
    else:
        for lib in (sys.argv[1:] or ["-"]):
            subprocess.run([sys.executable, os.path.abspath(__file__), "--one", lib], check=False)
