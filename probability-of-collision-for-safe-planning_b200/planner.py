"""Input side of the hot path: a small A* planner on the reference's state lattice and
the plan files the estimator reads.

The reference plans with `astar_planner.A_Star_Planner_8_euclidean(goal, distanceWeight = 100, angleWeight = 5,
distanceDisc = 0.15, angleDisc = pi / 2, ...)` (hw2_astar.py:117-131); that module is missing from the
repository, so this is a from-scratch planner on the same discretisation, not a restatement: states
(x, y, theta) on a 0.15 m x 0.15 m x 90 degree lattice anchored at the start, 8-connected translations
plus in-place quarter turns, cost = 100 * distance + 5 * |turn|, Euclidean heuristic.  Nominal poses are
kept `clearance` metres away from every obstacle prism in the separating-axis sense (`footprint_gap`, a lower
bound of the true distance, on the same prisms the estimator's SAT uses).

`path_odometry` / `inverse_odometry` follow gaussprop.py:147-172 (== MCSimulator.h:434-449),
`write_plan_files` writes trajectory.dat / odometry.dat as hw2_astar.py:195-204 does (pickles of a list
of [x, y, theta] and a list of [rot1, trans, rot2]).  Planning is host-side, O(lattice); scoring the
candidates is the GPU path (cuda_abi.Context.score_plans)."""
import heapq
import os
import pickle

import numpy as np

TWO_PI = 2.0 * np.pi


def round_angle(a):
    """roundAngle / angleWrap (gaussprop.py, MCSimulator.h:56-74): into [0, 2 pi], both ends kept."""
    while a < 0:
        a += TWO_PI
    while a > TWO_PI:
        a -= TWO_PI
    return a


def inverse_odometry(p1, p2):
    rot1 = round_angle(np.arctan2(p2[1] - p1[1], p2[0] - p1[0]) - p1[2])
    trans = float(np.hypot(p2[0] - p1[0], p2[1] - p1[1]))
    rot2 = round_angle(p2[2] - p1[2] - rot1)
    return np.array([rot1, trans, rot2])


def path_odometry(path):
    return [inverse_odometry(path[t], path[t + 1]) for t in range(len(path) - 1)]


def write_plan_files(path, directory="."):
    """trajectory.dat / odometry.dat of hw2_astar.py:195-204.  Returns the two paths."""
    os.makedirs(directory, exist_ok=True)
    tp, op = os.path.join(directory, "trajectory.dat"), os.path.join(directory, "odometry.dat")
    with open(tp, "wb") as f:
        pickle.dump([[float(v) for v in p] for p in path], f, protocol=2)
    with open(op, "wb") as f:
        pickle.dump([[float(v) for v in u] for u in path_odometry(path)], f, protocol=2)
    return tp, op


def box_distance(points, boxes):
    """Distance from points [n, 2] to every prism's footprint [m]: [n, m] (0 inside)."""
    d = points[:, None, :] - boxes[None, :, 0:2]
    c, s = boxes[None, :, 2], boxes[None, :, 3]
    lx = np.abs(c * d[..., 0] + s * d[..., 1]) - boxes[None, :, 4]
    ly = np.abs(c * d[..., 1] - s * d[..., 0]) - boxes[None, :, 5]
    return np.hypot(np.maximum(lx, 0.0), np.maximum(ly, 0.0))


def live_boxes(boxes, hulls):
    """Prisms whose z-interval overlaps some robot hull's (the rest can never collide: floor, shelves)."""
    keep = np.zeros(len(boxes), dtype=bool)
    for h in hulls:
        keep |= (h[6] < boxes[:, 7]) & (boxes[:, 6] < h[7])
    return boxes[keep]


def robot_reach(hulls):
    return max(float(np.hypot(h[0], h[1]) + np.hypot(h[4], h[5])) for h in hulls)


def footprint_gap(pose, hulls, boxes):
    """Separating-axis gap (a lower bound of the distance) between the robot's hull footprints at `pose` and
    the box footprints, min over the (hull, box) pairs whose z-intervals overlap; +inf without such a pair.
    Planner-side validity check of a NOMINAL pose (what CheckCollision does inside the reference's A*); the
    estimator's per-rollout collision test is the CUDA SAT, not this."""
    x, y, th = pose
    ct, st = np.cos(th), np.sin(th)
    best = np.inf
    for h in hulls:
        b = boxes[(h[6] < boxes[:, 7]) & (boxes[:, 6] < h[7])]
        if len(b) == 0:
            continue
        hcx, hcy = x + ct * h[0] - st * h[1], y + st * h[0] + ct * h[1]
        ch, sh = ct * h[2] - st * h[3], st * h[2] + ct * h[3]
        dx, dy = b[:, 0] - hcx, b[:, 1] - hcy
        c, s = np.abs(ch * b[:, 2] + sh * b[:, 3]), np.abs(sh * b[:, 2] - ch * b[:, 3])
        g = np.maximum.reduce([
            np.abs(ch * dx + sh * dy) - (h[4] + b[:, 4] * c + b[:, 5] * s),
            np.abs(ch * dy - sh * dx) - (h[5] + b[:, 4] * s + b[:, 5] * c),
            np.abs(b[:, 2] * dx + b[:, 3] * dy) - (b[:, 4] + h[4] * c + h[5] * s),
            np.abs(b[:, 2] * dy - b[:, 3] * dx) - (b[:, 5] + h[4] * s + h[5] * c)])
        best = min(best, float(g.min()))
    return best


def astar_lattice(start, goal, boxes, hulls, clearance=0.05, step=0.15, w_dist=100.0, w_angle=5.0, max_expand=400000):
    """A* from start (x, y, theta) to the lattice cell nearest goal.  Returns the list of [x, y, theta] or None."""
    start = np.asarray(start, dtype=np.float64)
    goal = np.asarray(goal, dtype=np.float64)
    h0 = int(np.round(start[2] / (np.pi / 2))) % 4
    gi = (int(np.round((goal[0] - start[0]) / step)), int(np.round((goal[1] - start[1]) / step)),
          int(np.round(goal[2] / (np.pi / 2))) % 4)
    moves = [(dx, dy) for dx in (-1, 0, 1) for dy in (-1, 0, 1) if dx or dy]
    free = {}

    def is_free(ix, iy, ih):
        k = (ix, iy, ih)
        if k not in free:
            free[k] = footprint_gap((start[0] + ix * step, start[1] + iy * step, ih * (np.pi / 2)), hulls, boxes) > clearance
        return free[k]

    def heur(ix, iy, ih):
        dh = abs(ih - gi[2]) % 4
        return w_dist * step * np.hypot(ix - gi[0], iy - gi[1]) + w_angle * (np.pi / 2) * min(dh, 4 - dh)

    s0 = (0, 0, h0)
    if not is_free(0, 0, h0) or not is_free(*gi):
        return None
    best = {s0: 0.0}
    parent = {}
    heap = [(heur(*s0), 0.0, s0)]
    n = 0
    while heap and n < max_expand:
        f, g, s = heapq.heappop(heap)
        if g > best.get(s, np.inf):
            continue
        if s == gi:
            cells = [s]
            while cells[-1] in parent:
                cells.append(parent[cells[-1]])
            return [[start[0] + ix * step, start[1] + iy * step, ih * (np.pi / 2)] for ix, iy, ih in reversed(cells)]
        n += 1
        ix, iy, ih = s
        nxt = [((ix + dx, iy + dy, ih), w_dist * step * np.hypot(dx, dy)) for dx, dy in moves if is_free(ix + dx, iy + dy, ih)]
        nxt += [((ix, iy, (ih + d) % 4), w_angle * (np.pi / 2)) for d in (1, 3) if is_free(ix, iy, (ih + d) % 4)]
        for t, c in nxt:
            ng = g + c
            if ng < best.get(t, np.inf):
                best[t] = ng
                parent[t] = s
                heapq.heappush(heap, (ng + heur(*t), ng, t))
    return None


def candidate_plans(start, goal, boxes, hulls, clearances=(0.0, 0.02, 0.05, 0.1, 0.2, 0.3), step=0.15):
    """One A* plan per clearance (larger clearance = wider berth, possibly a different corridor).
    Returns a list of dict(clearance, path)."""
    lb = live_boxes(np.asarray(boxes, dtype=np.float64), hulls)
    out = []
    for c in clearances:
        path = astar_lattice(start, goal, lb, hulls, clearance=c, step=step)
        if path is not None and len(path) >= 2:
            out.append(dict(clearance=c, path=path))
    return out


def stack_plans(paths):
    """Plans of different lengths -> [n, 3, T] / [n, 3, T-1] arrays for score_plans: shorter plans wait at their
    goal (zero translation: inverse odometry of two identical poses)."""
    T = max(len(p) for p in paths)
    tr = np.zeros((len(paths), 3, T))
    od = np.zeros((len(paths), 3, T - 1))
    for k, p in enumerate(paths):
        q = list(p) + [p[-1]] * (T - len(p))
        tr[k] = np.array(q).T
        od[k] = np.array(path_odometry(q)).T
    return tr, od
