"""Host-side planner and plan files, CPU tests."""
import pickle

import numpy as np

import pcoll_b200
from conftest import golden
from pcoll_b200 import planner


def test_inverse_odometry_reproduces_the_committed_odometry(plan):
    """path_odometry (gaussprop.py:147-172) on the committed trajectory == the committed odometry.dat."""
    tr, od = plan["trajectory"], plan["odometry"]
    got = np.array(planner.path_odometry([tr[:, t] for t in range(tr.shape[1])])).T
    d = got - od
    d[[0, 2]] = (d[[0, 2]] + np.pi) % (2 * np.pi) - np.pi       # rotations live on the circle ([0, 2 pi] closed)
    assert np.max(np.abs(d)) < 1e-9


def test_astar_plans_are_lattice_paths_clear_of_the_scene(tmp_path):
    sc = golden("scene_pr2test2.npz")
    boxes, hulls = sc["boxes"], sc["hulls"]
    start, goal = [-3.4, -1.4, 0.0], [2.6, -1.3, 3 * np.pi / 2]
    cands = planner.candidate_plans(start, goal, boxes, hulls, clearances=(0.0, 0.01, 0.2))
    assert len(cands) == 2          # the doorway of this scene leaves < 3 cm on each side: 0.2 m is infeasible
    assert abs(len(cands[0]["path"]) - 56) <= 2   # the committed plan has 56 waypoints on the same lattice
    lb = planner.live_boxes(boxes, hulls)
    for c in cands:
        p = np.array(c["path"])
        assert np.allclose(p[0], start) and np.hypot(*(p[-1, :2] - goal[:2])) <= 0.15 * 0.75
        steps = np.hypot(*np.diff(p[:, :2], axis=0).T)
        assert np.all((steps < 1e-9) | np.isclose(steps, 0.15) | np.isclose(steps, 0.15 * np.sqrt(2)))
        assert min(planner.footprint_gap(q, hulls, lb) for q in p) > c["clearance"]
    tp, op = planner.write_plan_files(cands[0]["path"], str(tmp_path))
    back = pcoll_b200.plan_from_pickles(tp, op, as_sent=False)
    assert back["trajectory"].shape == (3, len(cands[0]["path"])) and back["odometry"].shape == (3, len(cands[0]["path"]) - 1)
    with open(op, "rb") as f:
        assert len(pickle.load(f)) == len(cands[0]["path"]) - 1
    tr, od = planner.stack_plans([c["path"] for c in cands])
    assert tr.shape[0] == 2 and od.shape[2] == tr.shape[2] - 1 and np.all(od[:, 1] >= 0)


def test_wider_clearance_gives_longer_plans_in_an_open_field():
    boxes = pcoll_b200.inputs.synthetic_obstacle_field(11, 25, room=12.0)
    hulls = golden("scene_pr2test2.npz")["hulls"]
    lb = planner.live_boxes(boxes, hulls)
    grid = [[x, y, 0.0] for x in np.arange(-5.0, 5.01, 0.5) for y in np.arange(-5.0, 5.01, 0.5)]
    free = sorted((p for p in grid if planner.footprint_gap(p, hulls, lb) > 0.35), key=lambda p: p[0] + p[1])
    assert len(free) >= 2
    cands = planner.candidate_plans(free[0], free[-1], boxes, hulls, clearances=(0.0, 0.1, 0.3))
    assert len(cands) >= 2
    cost = [sum(np.hypot(*np.diff(np.array(c["path"])[:, :2], axis=0).T)) for c in cands]
    assert all(b >= a - 1e-9 for a, b in zip(cost, cost[1:]))
    for c in cands:
        assert min(planner.footprint_gap(q, hulls, lb) for q in c["path"]) > c["clearance"]
