"""Input side of the hot path: the two plan pickles and the parameter defaults of the reference driver.

* `trajectory.dat` / `odometry.dat` are Python-2 protocol-0 pickles (hw2_astar.py:198-204): a list of
  T `[x, y, theta]` lists and a list of T-1 `ndarray(3,)` `[rot1, trans, rot2]`.
* The reference driver sends numbers as Python-2 `str(float)` joined by spaces
  (MCSimulation.py:81-85, :195-198), i.e. rounded to 12 significant digits: `py2_str`.
* Parameter defaults: gaussprop.py:36 (alphas), :39 (Q), :45-46 (landmarks), :56 (Sigma0);
  MCSimulation.py:164 (numParticles), :204-205 (numGaussians, numGMMSamples).
"""
import json
import pickle

import numpy as np


def py2_str(x):
    """Python-2 str(float): repr with 12 significant digits ('%.12g'), '.0' appended to integers."""
    s = "%.12g" % float(x)
    if "." not in s and "e" not in s and "n" not in s and "i" not in s:
        s += ".0"
    return s


def default_params():
    return dict(
        alphas=np.square(np.array([0.00025, 0.0025, 0.0025, 0.0025])),
        Q=float(np.square(0.2)),
        landmarks=np.array([[3, -3, 0, 0, -3, 3, -3, 3], [0, 0, 2, -2, 2, 2, -2, -2]], dtype=np.float64),
        cov0=0.001 * np.identity(3),
        num_particles=10000,
        num_gaussians=3,
        num_gmm_samples=10000,
    )


def plan_from_pickles(trajectory_path, odometry_path, as_sent=True):
    """Returns dict(trajectory[3,T], odometry[3,T-1]) + default parameters.  as_sent=True applies the
    12-digit rounding of the driver's string marshalling, so the arrays are what the plugin saw."""
    with open(trajectory_path, "rb") as f:
        tr = np.array(pickle.load(f, encoding="latin1"), dtype=np.float64).T
    with open(odometry_path, "rb") as f:
        od = np.array(pickle.load(f, encoding="latin1"), dtype=np.float64).T
    if as_sent:
        tr = np.array([[float(py2_str(v)) for v in row] for row in tr])
        od = np.array([[float(py2_str(v)) for v in row] for row in od])
    plan = default_params()
    plan.update(trajectory=np.ascontiguousarray(tr), odometry=np.ascontiguousarray(od))
    return plan


def load_plan_json(path):
    with open(path) as f:
        d = json.load(f)
    plan = {}
    for k, v in d.items():
        plan[k] = np.array(v, dtype=np.float64) if isinstance(v, list) else v
    return plan


def save_plan_json(plan, path):
    d = {k: (np.asarray(v).tolist() if isinstance(v, np.ndarray) else v) for k, v in plan.items()}
    with open(path, "w") as f:
        json.dump(d, f, indent=0)


def synthetic_em_problem(seed, n, K, centre=(3.0, 1.0, 1.5)):
    """BASELINE config 5 workload generator: n samples of a seeded 3-component mixture
    around `centre` (a trajectory pose; covariances of 1e-3 scale, like the propagated beliefs), and
    `keep_existing` initial parameters for a K-component fit (seeded perturbations of sample points).
    Returns (x float32 [3, n], hefts [K], means [K, 3], fcovs [K, 3, 3])."""
    rng = np.random.Generator(np.random.PCG64(seed))
    c = np.asarray(centre, dtype=np.float64)
    mus = c + rng.normal(0.0, 0.08, size=(3, 3))
    covs = []
    for _ in range(3):
        a = rng.normal(0.0, 1.0, size=(3, 3))
        covs.append(1e-3 * (a @ a.T / 3.0 + 0.5 * np.eye(3)))
    w = np.array([0.5, 0.3, 0.2])
    comp = rng.choice(3, size=n, p=w)
    x = np.empty((n, 3))
    for g in range(3):
        idx = np.nonzero(comp == g)[0]
        x[idx] = rng.multivariate_normal(mus[g], covs[g], size=idx.size)
    x32 = np.ascontiguousarray(x.T, dtype=np.float32)
    pick = rng.choice(n, size=K, replace=False)
    means = x32.T[pick].astype(np.float64) + rng.normal(0.0, 0.01, size=(K, 3))
    fcovs = np.tile(np.cov(x32.astype(np.float64)) + 1e-6 * np.eye(3), (K, 1, 1))
    hefts = np.full(K, 1.0 / K)
    return x32, hefts, means, fcovs


# ---- BASELINE config 4 workload: synthetic obstacle field + A*-style lattice plans ------
def synthetic_obstacle_field(seed=4096, n_boxes=256, room=40.0):
    """n_boxes z-prisms [cx, cy, cos, sin, hx, hy, zlo, zhi]: centre uniform in the room, half extents
    U(0.1, 0.5) m, yaw U(0, pi), z in [0, 0.2]."""
    rng = np.random.Generator(np.random.PCG64(seed))
    c = rng.uniform(-room / 2, room / 2, size=(n_boxes, 2))
    h = rng.uniform(0.1, 0.5, size=(n_boxes, 2))
    yaw = rng.uniform(0.0, np.pi, size=n_boxes)
    return np.column_stack([c, np.cos(yaw), np.sin(yaw), h, np.zeros(n_boxes), np.full(n_boxes, 0.2)])


def _wrap_2pi(a):
    """angleWrap (mcsimplugin/MCSimulator.h:56-65) for arrays: [0, 2 pi]."""
    a = np.where(a < 0, a + 2 * np.pi, a)
    a = np.where(a < 0, a + 2 * np.pi, a)
    a = np.where(a > 2 * np.pi, a - 2 * np.pi, a)
    return np.where(a > 2 * np.pi, a - 2 * np.pi, a)


def synthetic_plans(seed, n_plans, T, boxes, reach=0.4724, clear=0.05, step=0.15, room=40.0, persistence=0.9):
    """8-connected lattice walks (step 0.15 m, heading in {0, pi/2, pi, 3pi/2}: the discretisation of
    hw2_astar.py:122-123) whose nominal poses keep the robot's bounding circle `clear` metres away from
    every box's bounding circle; odometry by inverseOdometry (MCSimulator.h:434-449).
    Returns (trajectories [n_plans, 3, T], odometry [n_plans, 3, T-1])."""
    rng = np.random.Generator(np.random.PCG64(seed))
    bc, brad = boxes[:, :2], np.hypot(boxes[:, 4], boxes[:, 5])
    lim = (reach + brad + clear) ** 2

    def ok(p):
        d2 = ((p[:, None, :] - bc[None, :, :]) ** 2).sum(-1)
        inside = (np.abs(p) < room / 2 - 1.0).all(-1)
        return (d2 > lim[None, :]).all(-1) & inside

    dirs = np.array([[1, 0], [1, 1], [0, 1], [-1, 1], [-1, 0], [-1, -1], [0, -1], [1, -1]], dtype=np.float64) * step
    pos = np.zeros((n_plans, 2))
    todo = np.ones(n_plans, dtype=bool)
    while todo.any():
        cand = rng.uniform(-room / 2 + 2, room / 2 - 2, size=(int(todo.sum()), 2))
        good = ok(cand)
        idx = np.nonzero(todo)[0]
        pos[idx[good]] = cand[good]
        todo[idx[good]] = False
    d = rng.integers(0, 8, size=n_plans)
    traj = np.zeros((n_plans, 3, T))
    heading = (np.round(d / 2.0).astype(int) % 4) * (np.pi / 2)
    traj[:, 0, 0], traj[:, 1, 0], traj[:, 2, 0] = pos[:, 0], pos[:, 1], heading
    for t in range(1, T):
        change = rng.random(n_plans) > persistence
        d = np.where(change, rng.integers(0, 8, size=n_plans), d)
        moved = np.zeros(n_plans, dtype=bool)
        newpos = pos.copy()
        for attempt in range(8):
            trial = (d + np.where(moved, 0, attempt * rng.choice([1, -1], size=n_plans))) % 8
            cand = pos + dirs[trial]
            good = ok(cand) & ~moved
            newpos[good] = cand[good]
            d = np.where(good, trial, d)
            moved |= good
        pos = newpos
        heading = (np.round(d / 2.0).astype(int) % 4) * (np.pi / 2)
        traj[:, 0, t], traj[:, 1, t], traj[:, 2, t] = pos[:, 0], pos[:, 1], heading
    dx, dy = np.diff(traj[:, 0, :], axis=1), np.diff(traj[:, 1, :], axis=1)
    rot1 = _wrap_2pi(np.arctan2(dy, dx) - traj[:, 2, :-1])
    trans = np.hypot(dx, dy)
    rot2 = _wrap_2pi(traj[:, 2, 1:] - traj[:, 2, :-1] - rot1)
    return traj, np.stack([rot1, trans, rot2], axis=1)
