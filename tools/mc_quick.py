"""Developer probe: times the packed MC rollout kernel of one or more builds of libmcsim_cuda.so
(`python tools/mc_quick.py [lib.so ...]`; default: the in-tree library) at the bench size and prints
rollout-steps/s, hits, executed pair tests per step and the fraction of the measured FP32 peak."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def one(lib_path, scene="pr2test2", N=10 ** 8):
    import numpy as np
    import pcoll_b200
    abi = pcoll_b200.cuda_abi
    if lib_path:
        abi.LIB_PATH = lib_path
    gold = os.path.join(ROOT, "tests", "golden")
    plan = pcoll_b200.load_plan_json(os.path.join(gold, "plan_committed.json"))
    sc = np.load(os.path.join(gold, "scene_%s.npz" % scene))
    ctx = abi.Context(0)
    fp32 = ctx.measure_peak(0, 100.0)
    ctx.set_scene(sc["boxes"], sc["hulls"])
    T = plan["trajectory"].shape[1]
    for ee, ex, nc in ((0, 0, 0), (1, 0, 0), (1, 0, 1), (0, 1, 0)):
        n = N if not ex else N // 4
        p, keep = abi.make_mc_params(plan, n_rollouts=n, mode=abi.MODE_PHILOX, early_exit=ee, exhaustive=ex, no_compaction=nc)
        ctx.mc_run(p)
        best = None
        for _ in range(2):
            out = ctx.mc_run(p)
            if best is None or out["kernel_ms"] < best["kernel_ms"]:
                best = out
        ms = best["kernel_ms"]
        flops = best["executed_steps"] * 647 + best["pair_tests"] * 40
        print("%s %s ee=%d ex=%d nc=%d: %.3f ms  %.4e nominal steps/s  hits %d P %.6f  tests/step %.3f  executed %.3f  frac %.4f (peak %.1f)" % (
            os.path.basename(lib_path or "in-tree"), scene, ee, ex, nc, ms, n * T / ms * 1e3, best["hits"], best["P"],
            best["pair_tests"] / max(best["executed_steps"], 1), best["executed_steps"] / (n * T), flops / ms * 1e-9 / fp32, fp32), flush=True)
    ctx.close()


if __name__ == "__main__":
    if len(sys.argv) > 2 and sys.argv[1] == "--one":
        one(sys.argv[2] if sys.argv[2] != "-" else None, *(sys.argv[3:4]))
    else:
        libs = sys.argv[1:] or ["-"]
        for lib in libs:
            subprocess.run([sys.executable, os.path.abspath(__file__), "--one", lib], check=False)
