#!/usr/bin/env python3
"""Planner -> scorer loop (the use the paper motivates): plan several candidate paths
between two poses with different clearances, score all of them in ONE batched launch
(mcsim_cuda_score_plans), pick the safest and write it as trajectory.dat / odometry.dat.

  python tools/plan_and_score.py [--scene pr2test2 | --field SEED] [--rollouts 65536] [--outdir plans_out]
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pcoll_b200  # noqa: E402
from pcoll_b200 import planner  # noqa: E402


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--scene", default=None, choices=["pr2test2", "pr2custom"])
    ap.add_argument("--field", type=int, default=11, help="seed of a synthetic 25-box field in a 12 m room (default)")
    ap.add_argument("--rollouts", type=int, default=65536)
    ap.add_argument("--outdir", default="plans_out")
    ap.add_argument("--device", type=int, default=0)
    a = ap.parse_args(argv)
    gold = os.path.join(ROOT, "tests", "golden")
    model = pcoll_b200.load_plan_json(os.path.join(gold, "plan_committed.json"))   # noise model, landmarks, Sigma_0
    hulls = np.load(os.path.join(gold, "scene_pr2test2.npz"))["hulls"]
    if a.scene:
        boxes = np.load(os.path.join(gold, "scene_%s.npz" % a.scene))["boxes"]
        start, goal = model["trajectory"][:, 0], [2.6, -1.3, 3 * np.pi / 2]
        clearances = (0.0, 0.005, 0.01, 0.02)
    else:
        boxes = pcoll_b200.inputs.synthetic_obstacle_field(a.field, 25, room=12.0)
        lb = planner.live_boxes(boxes, hulls)
        grid = [[x, y, 0.0] for x in np.arange(-5.0, 5.01, 0.5) for y in np.arange(-5.0, 5.01, 0.5)]
        free = sorted((p for p in grid if planner.footprint_gap(p, hulls, lb) > 0.35), key=lambda p: p[0] + p[1])
        start, goal = free[0], free[-1]
        clearances = (0.0, 0.02, 0.05, 0.1, 0.2, 0.3)
    cands = planner.candidate_plans(start, goal, boxes, hulls, clearances=clearances)
    if not cands:
        raise SystemExit("no candidate plan found")
    tr, od = planner.stack_plans([c["path"] for c in cands])
    ctx = pcoll_b200.cuda_abi.Context(a.device)
    ctx.set_scene(boxes, hulls)
    r = ctx.score_plans(model, tr, od, a.rollouts)
    ctx.close()
    table = [dict(clearance=c["clearance"], waypoints=len(c["path"]), P_collision=float(p)) for c, p in zip(cands, r["P"])]
    best = int(np.argmin(r["P"]))
    tp, op = planner.write_plan_files(cands[best]["path"], a.outdir)
    out = dict(start=list(map(float, start)), goal=list(map(float, goal)), rollouts_per_plan=a.rollouts, candidates=table,
               chosen=best, kernel_ms=r["kernel_ms"], files=[tp, op])
    print(json.dumps(out))
    return out


if __name__ == "__main__":
    main()
