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
