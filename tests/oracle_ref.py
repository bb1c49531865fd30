"""TEST INFRASTRUCTURE: ctypes view of oracle/_ref/libmcsim_ref.so (the reference's own estimator
headers compiled against the OpenRAVE stand-in, oracle/ref/ref_oracle.cpp).  Only tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs use this."""
import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_LIB = os.environ.get("MCSIM_REF_LIB", os.path.join(ROOT, "oracle", "_ref", "libmcsim_ref.so"))
_dp = C.POINTER(C.c_double)


class RefParams(C.Structure):
    _fields_ = [
        ("alphas", C.c_double * 4), ("Q", C.c_double), ("n_landmarks", C.c_int), ("landmarks", _dp),
        ("n_particles", C.c_int), ("cov0", C.c_double * 9), ("path_length", C.c_int),
        ("trajectory", _dp), ("odometry", _dp), ("n_gaussians", C.c_int), ("n_gmm_samples", C.c_int),
        ("boxes", _dp), ("n_boxes", C.c_int), ("hulls", _dp), ("n_hulls", C.c_int),
        ("poly_hulls", _dp), ("n_poly_hulls", C.c_int),
    ]


class RefRecord(C.Structure):
    _fields_ = [
        ("poses", _dp), ("clearance", _dp), ("flags", C.POINTER(C.c_ubyte)), ("cap_calls", C.c_size_t),
        ("mu_path", _dp), ("cov_path", _dp), ("gmm_pre", _dp), ("gmm_counts", C.POINTER(C.c_int)),
        ("gmm_post", _dp),
    ]


_lib = None


def available():
    return os.path.isfile(REF_LIB)


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(REF_LIB)
        L.ref_mc_run.argtypes = [C.POINTER(RefParams), _dp, C.c_size_t, C.c_uint64, C.POINTER(RefRecord), _dp,
                                 C.POINTER(C.c_uint), _dp]
        L.ref_gmm_run.argtypes = [C.POINTER(RefParams), _dp, C.c_size_t, C.c_uint64, C.c_uint, C.POINTER(RefRecord),
                                  _dp, _dp]
        L.ref_em_fit.argtypes = [_dp, C.c_size_t, C.c_int, _dp, _dp, _dp, C.c_int, C.c_double, _dp]
        L.ref_cov.argtypes = [_dp, C.c_int, C.c_int, _dp]
        L.ref_cov.restype = None
        L.ref_clearance.argtypes = [C.POINTER(RefParams), _dp]
        L.ref_clearance.restype = C.c_double
        _lib = L
    return _lib


def _d(a):
    return a.ctypes.data_as(_dp)


def make_params(plan, boxes, hulls, n_particles=1000, n_gaussians=3, n_gmm_samples=1000, poly_hulls=None):
    keep = dict(
        tr=np.ascontiguousarray(plan["trajectory"], dtype=np.float64),
        od=np.ascontiguousarray(plan["odometry"], dtype=np.float64),
        lm=np.ascontiguousarray(plan["landmarks"], dtype=np.float64),
        bx=np.ascontiguousarray(boxes, dtype=np.float64).reshape(-1, 8),
        hu=np.ascontiguousarray(hulls, dtype=np.float64).reshape(-1, 8),
    )
    p = RefParams()
    p.alphas = (C.c_double * 4)(*[float(v) for v in plan["alphas"]])
    p.Q = float(plan["Q"])
    p.n_landmarks = keep["lm"].shape[1]
    p.landmarks = _d(keep["lm"])
    p.n_particles = int(n_particles)
    p.cov0 = (C.c_double * 9)(*[float(v) for v in np.asarray(plan["cov0"]).reshape(9)])
    p.path_length = keep["tr"].shape[1]
    p.trajectory, p.odometry = _d(keep["tr"]), _d(keep["od"])
    p.n_gaussians, p.n_gmm_samples = int(n_gaussians), int(n_gmm_samples)
    p.boxes, p.n_boxes = _d(keep["bx"]), keep["bx"].shape[0]
    p.hulls, p.n_hulls = _d(keep["hu"]), keep["hu"].shape[0]
    if poly_hulls is not None:
        keep["po"] = np.ascontiguousarray(poly_hulls, dtype=np.float64).reshape(-1, 19)
        p.poly_hulls, p.n_poly_hulls = _d(keep["po"]), keep["po"].shape[0]
    return p, keep


def mc_run(plan, boxes, hulls, n_particles, draws=None, mt_seed=0, record=True, poly_hulls=None):
    """Runs the reference's runSimulation().  Returns dict(P, counts[N], final[N,3], and when
    record: poses[T,N,3], clearance[T,N], flags[T,N], mu_path[T,3], cov_path[T,9])."""
    p, keep = make_params(plan, boxes, hulls, n_particles=n_particles, poly_hulls=poly_hulls)
    T, N = p.path_length, int(n_particles)
    rec = RefRecord()
    out = {}
    if record:
        out["poses"] = np.zeros((T, N, 3))
        out["clearance"] = np.zeros((T, N))
        out["flags"] = np.zeros((T, N), dtype=np.uint8)
        out["mu_path"] = np.zeros((T, 3))
        out["cov_path"] = np.zeros((T, 9))
        rec.poses, rec.clearance = _d(out["poses"]), _d(out["clearance"])
        rec.flags = out["flags"].ctypes.data_as(C.POINTER(C.c_ubyte))
        rec.cap_calls = T * N
        rec.mu_path, rec.cov_path = _d(out["mu_path"]), _d(out["cov_path"])
    P = C.c_double()
    counts = np.zeros(N, dtype=np.uint32)
    final = np.zeros((N, 3))
    dr = None if draws is None else np.ascontiguousarray(draws, dtype=np.float64)
    rc = lib().ref_mc_run(C.byref(p), _d(dr) if dr is not None else _dp(), 0 if dr is None else dr.size,
                          int(mt_seed), C.byref(rec), C.byref(P), counts.ctypes.data_as(C.POINTER(C.c_uint)), _d(final))
    if rc != 0:
        raise RuntimeError("ref_mc_run failed: %d" % rc)
    out.update(P=P.value, counts=counts, final=final)
    return out


def gmm_run(plan, boxes, hulls, K, n_samples, draws=None, mt_seed=0, multinomial_seed=1, record=True):
    """Runs the reference's runGMMEstimation()."""
    p, keep = make_params(plan, boxes, hulls, n_gaussians=K, n_gmm_samples=n_samples)
    T, N = p.path_length, int(n_samples)
    rec = RefRecord()
    out = {}
    if record:
        out["poses"] = np.zeros((T, N, 3))
        out["clearance"] = np.zeros((T, N))
        out["flags"] = np.zeros((T, N), dtype=np.uint8)
        out["gmm_pre"] = np.zeros((T, K, 13))
        out["gmm_counts"] = np.zeros((T, K), dtype=np.int32)
        out["gmm_post"] = np.zeros((K, 13))
        rec.poses, rec.clearance = _d(out["poses"]), _d(out["clearance"])
        rec.flags = out["flags"].ctypes.data_as(C.POINTER(C.c_ubyte))
        rec.cap_calls = T * N
        rec.gmm_pre, rec.gmm_post = _d(out["gmm_pre"]), _d(out["gmm_post"])
        rec.gmm_counts = out["gmm_counts"].ctypes.data_as(C.POINTER(C.c_int))
    P = C.c_double()
    prob = np.zeros(T)
    dr = None if draws is None else np.ascontiguousarray(draws, dtype=np.float64)
    rc = lib().ref_gmm_run(C.byref(p), _d(dr) if dr is not None else _dp(), 0 if dr is None else dr.size,
                           int(mt_seed), int(multinomial_seed), C.byref(rec), C.byref(P), _d(prob))
    if rc != 0:
        raise RuntimeError("ref_gmm_run failed: %d" % rc)
    out.update(P=P.value, probabilities=prob)
    return out


def em_fit(X, means, covs, hefts, em_iter, var_floor=1e-10):
    """Vendored arma::gmm_full::learn(keep_existing, km_iter=0).  X: [N,3].  Returns (means, covs, hefts, avg_log_p)."""
    Xc = np.ascontiguousarray(X, dtype=np.float64)  # N x 3 row-major == 3 x N column-major
    m = np.ascontiguousarray(means, dtype=np.float64).copy()
    c = np.ascontiguousarray(covs, dtype=np.float64).copy()
    h = np.ascontiguousarray(hefts, dtype=np.float64).copy()
    ll = C.c_double()
    rc = lib().ref_em_fit(_d(Xc), Xc.shape[0], m.shape[0], _d(m), _d(c), _d(h), int(em_iter), float(var_floor), C.byref(ll))
    if rc not in (0, 3):
        raise RuntimeError("ref_em_fit failed: %d" % rc)
    return m, c, h, ll.value
