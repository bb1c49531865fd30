"""Developer probe (not the bench): times the MC rollout kernel device-resident at a few sizes and
prints rollout-steps/s, SAT pair tests/s and the fraction of the measured FP32 peak."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pcoll_b200  # noqa: E402

abi = pcoll_b200.cuda_abi
F_SAT, F_MOVE, F_PHILOX_STEP = 40, 10, 647  # SURVEY.md 8(d) algorithmic flops


def main():
    gold = os.path.join(ROOT, "tests", "golden")
    plan = pcoll_b200.load_plan_json(os.path.join(gold, "plan_committed.json"))
    scene = sys.argv[1] if len(sys.argv) > 1 else "pr2test2"
    sc = np.load(os.path.join(gold, "scene_%s.npz" % scene))
    ctx = abi.Context(0)
    fp32, mufu = ctx.measure_peaks(100.0)
    print("measured peaks: FP32 %.1f TFLOP/s, SFU %.2f Top/s" % (fp32, mufu))
    ctx.set_scene(sc["boxes"], sc["hulls"])
    T = plan["trajectory"].shape[1]
    for ee in (1, 0):
        for N in (1 << 16, 1 << 20, 1 << 24, 10 ** 8):
            p, keep = abi.make_mc_params(plan, n_rollouts=N, mode=abi.MODE_PHILOX, early_exit=ee)
            ctx.mc_run(p)  # warm-up
            t0 = time.time()
            out = ctx.mc_run(p)
            wall = time.time() - t0
            ms = out["kernel_ms"]
            flops = out["executed_steps"] * F_PHILOX_STEP + out["pair_tests"] * F_SAT
            print("early_exit=%d N=%9d kernel %.3f ms (wall %.3f ms) P=%.5f nominal %.3e steps/s executed %.3e steps/s "
                  "%.3e pair tests/s  %.2f TFLOP/s algorithmic = %.1f%% of FP32 peak" %
                  (ee, N, ms, wall * 1e3, out["P"], N * T / ms * 1e3, out["executed_steps"] / ms * 1e3,
                   out["pair_tests"] / ms * 1e3, flops / ms * 1e-9, 100 * flops / ms * 1e-9 / fp32))
    ctx.close()


if __name__ == "__main__":
    main()
