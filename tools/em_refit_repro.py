"""Developer probe: run-to-run reproducibility of the EM-refit GMM run (free samples reach the EM fit in the order the
warps' atomics appended them).  `python tools/em_refit_repro.py [--lib lib.so] [stream_id ...]`"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pcoll_b200  # noqa: E402

abi = pcoll_b200.cuda_abi
if "--lib" in sys.argv:
    i = sys.argv.index("--lib")
    abi.LIB_PATH = os.path.abspath(sys.argv[i + 1])
    del sys.argv[i:i + 2]
gold = os.path.join(ROOT, "tests", "golden")
plan = pcoll_b200.load_plan_json(os.path.join(gold, "plan_committed.json"))
sc = np.load(os.path.join(gold, "scene_pr2test2.npz"))
ctx = abi.Context(0)
ctx.set_scene(sc["boxes"], sc["hulls"])
for sid in [int(s) for s in sys.argv[1:]] or [4, 5, 6, 7]:
    runs = [ctx.gmm_run(plan, 3, 10000, mode=abi.MODE_PHILOX, stream_id=sid, refit=abi.GMM_REFIT_EM, em_iters=3, record=True) for _ in range(4)]
    base = runs[0]
    for r in runs[1:]:
        flips = np.abs(r["probabilities"] - base["probabilities"]) * 10000
        dpost = np.max(np.abs(r["gmm_pre"] - base["gmm_pre"]), axis=(1, 2))   # gmm_pre[t]: the mixture waypoint t samples from
        first = int(np.argmax(dpost > 0)) if dpost is not None and np.any(dpost > 0) else -1
        print("stream %d: flips sum %.0f max %.0f at waypoints %s; first waypoint whose mixture differs: %d (max |d| %.3e)" % (
            sid, flips.sum(), flips.max(), np.nonzero(flips > 0.5)[0].tolist(), first, float(dpost.max()) if dpost is not None else -1))
    if dpost is not None:
        print("   per-waypoint max |d mixture| of the last pair:", np.array2string(dpost, precision=1, max_line_width=200))
ctx.close()
