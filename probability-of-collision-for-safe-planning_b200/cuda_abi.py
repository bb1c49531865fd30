"""ctypes view of include/mcsim_cuda.h (libmcsim_cuda.so) -- the thin C ABI over the sm_100a kernels.

This is the binding a Python maintainer of the reference would add; the structs
mirror the header field by field.  Loading fails loudly when the library has not been built, and
`Context()` fails loudly when there is no CUDA device: there is no CPU fallback.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "csrc", "libmcsim_cuda.so")

MAX_GAUSSIANS = 16
MAX_WAYPOINTS = 1024
MODE_REPLAY, MODE_PHILOX = 0, 1
GMM_KERNEL_AUTO, GMM_KERNEL_PER_WAYPOINT, GMM_KERNEL_CLUSTER = 0, 1, 2
GMM_REFIT_MOMENTS, GMM_REFIT_EM = 0, 1
EST_REFERENCE, EST_CLOSEDLOOP = 0, 1

_dp = C.POINTER(C.c_double)
_fp = C.POINTER(C.c_float)


class McParams(C.Structure):
    _fields_ = [
        ("T", C.c_int), ("trajectory", _dp), ("odometry", _dp),
        ("alphas", C.c_double * 4), ("Q", C.c_double),
        ("n_landmarks", C.c_int), ("landmarks", _dp), ("cov0", C.c_double * 9),
        ("n_rollouts", C.c_uint64), ("rollout_offset", C.c_uint64),
        ("mode", C.c_int), ("estimator", C.c_int), ("early_exit", C.c_int), ("seed", C.c_uint64),
        ("applied", _dp), ("z0", _dp), ("near_eps", C.c_double), ("precise", C.c_int), ("exhaustive", C.c_int),
        ("no_compaction", C.c_int), ("gmm_filter", C.c_int),
    ]


class McResult(C.Structure):
    _fields_ = [
        ("hits", C.c_uint64), ("near_contact", C.c_uint64), ("executed_steps", C.c_uint64),
        ("pair_tests", C.c_uint64),
        ("first_hit_hist", C.POINTER(C.c_uint64)), ("counts", C.POINTER(C.c_uint32)),
        ("poses", _fp), ("clearance", _fp), ("kernel_ms", C.c_float),
    ]


class EmParams(C.Structure):
    _fields_ = [
        ("K", C.c_int), ("n_fits", C.c_int), ("n", C.c_uint64), ("n_total", C.c_uint64),
        ("x", _fp), ("hefts", _dp), ("means", _dp), ("fcovs", _dp), ("iters", C.c_int), ("var_floor", C.c_double),
        ("x_row_stride", C.c_uint64),
    ]


class EmStaticParams(C.Structure):
    """mcsim_em_static_params (include/mcsim_cuda.h): every pointer is a DEVICE address owned by the caller."""
    _fields_ = [
        ("device", C.c_int), ("K", C.c_int), ("n_fits", C.c_int), ("iters", C.c_int), ("n", C.c_uint64),
        ("var_floor", C.c_double), ("x", C.c_void_p), ("hefts", C.c_void_p), ("means", C.c_void_p), ("fcovs", C.c_void_p),
        ("avg_log_p", C.c_void_p), ("iters_run", C.c_void_p), ("stream", C.c_void_p),
    ]


EM_STATIC_MAX_FITS, EM_STATIC_MAX_ITERS, EM_STATIC_SCRATCH_BYTES = 8, 64, 32768
EM_EXCHANGE_AUTO, EM_EXCHANGE_NCCL, EM_EXCHANGE_PEER = 0, 1, 2
FILTER_STATE, FILTER_DEVIATION = 0, 1
EM_MOMENTS_AUTO, EM_MOMENTS_CENTRED, EM_MOMENTS_SHIFTED = 0, 1, 2


class EmResult(C.Structure):
    _fields_ = [
        ("hefts", _dp), ("means", _dp), ("fcovs", _dp), ("avg_log_p", _dp), ("iters_run", C.POINTER(C.c_int)),
        ("kernel_ms", C.c_float),
    ]


class GmmBatchParams(C.Structure):
    _fields_ = [("K", C.c_int), ("n_samples", C.c_uint64), ("n_traj", C.c_int), ("trajectories", _dp), ("odometry", _dp),
                ("stream0", C.c_uint64), ("kernel", C.c_int)]


class GmmBatchResult(C.Structure):
    _fields_ = [("P", _dp), ("probabilities", _dp), ("degenerate", C.c_int), ("kernel_ms", C.c_float), ("pair_tests", C.c_uint64)]


class ScoreParams(C.Structure):
    _fields_ = [
        ("T", C.c_int), ("n_plans", C.c_int), ("trajectories", _dp), ("odometry", _dp),
        ("alphas", C.c_double * 4), ("Q", C.c_double), ("n_landmarks", C.c_int), ("landmarks", _dp),
        ("cov0", C.c_double * 9), ("rollouts", C.c_uint64), ("plan_offset", C.c_uint64),
        ("estimator", C.c_int), ("early_exit", C.c_int), ("seed", C.c_uint64), ("margin", C.c_double),
    ]


class ScoreResult(C.Structure):
    _fields_ = [
        ("hits", C.POINTER(C.c_uint64)), ("P", _dp), ("executed_steps", C.c_uint64), ("pair_tests", C.c_uint64),
        ("kernel_ms", C.c_float),
    ]


class GmmTruncateParams(C.Structure):
    _fields_ = [
        ("K", C.c_int), ("weights", _dp), ("means", _dp), ("covs", _dp),
        ("n_samples", C.c_uint64), ("mode", C.c_int), ("counts", C.POINTER(C.c_int)),
        ("normals", _dp), ("seed", C.c_uint64), ("stream_id", C.c_uint64), ("refit", C.c_int), ("em_iters", C.c_int),
    ]


class GmmTruncateResult(C.Structure):
    _fields_ = [
        ("weights", C.c_double * MAX_GAUSSIANS), ("means", C.c_double * (MAX_GAUSSIANS * 3)),
        ("covs", C.c_double * (MAX_GAUSSIANS * 9)),
        ("n_coll", C.c_uint64 * MAX_GAUSSIANS), ("n_free", C.c_uint64 * MAX_GAUSSIANS),
        ("p_collision", C.c_double), ("degenerate", C.c_int),
    ]


class GmmRunParams(C.Structure):
    _fields_ = [
        ("K", C.c_int), ("n_samples", C.c_uint64), ("normals", _dp), ("n_normals", C.c_size_t),
        ("counts", C.POINTER(C.c_int)), ("kernel", C.c_int), ("refit", C.c_int), ("em_iters", C.c_int),
    ]


class LinearParams(C.Structure):
    _fields_ = [("K", C.c_int), ("n_constraints", C.c_int), ("A", _dp), ("b", _dp), ("normals", _dp), ("n_normals", C.c_size_t)]


class GmmRunResult(C.Structure):
    _fields_ = [
        ("P", C.c_double), ("probabilities", _dp), ("gmm_pre", _dp), ("gmm_post", _dp),
        ("degenerate", C.c_int), ("kernel_ms", C.c_float), ("pair_tests", C.c_uint64),
    ]


_lib = None


def lib():
    """The loaded libmcsim_cuda.so; raises if it is missing (no silent fallback)."""
    global _lib
    if _lib is None:
        if not os.path.isfile(LIB_PATH):
            raise RuntimeError(
                "%s is missing: build it with __graft_entry__.build() or `make -C %s`; "
                "this package has no CPU fallback" % (LIB_PATH, os.path.dirname(LIB_PATH)))
        L = C.CDLL(LIB_PATH)
        vp = C.c_void_p
        L.mcsim_cuda_create.argtypes = [C.c_int, C.POINTER(vp)]
        L.mcsim_cuda_destroy.argtypes = [vp]
        L.mcsim_cuda_destroy.restype = None
        L.mcsim_cuda_last_error.restype = C.c_char_p
        L.mcsim_cuda_stream.argtypes = [vp]
        L.mcsim_cuda_stream.restype = vp
        L.mcsim_cuda_launch_count.argtypes = [vp]
        L.mcsim_cuda_launch_count.restype = C.c_uint64
        L.mcsim_cuda_measure_peaks.argtypes = [vp, C.c_double, _dp, _dp]
        L.mcsim_cuda_measure_peak.argtypes = [vp, C.c_int, C.c_double, _dp]
        L.mcsim_cuda_set_scene.argtypes = [vp, _dp, C.c_int, _dp, C.c_int]
        L.mcsim_cuda_set_scene_ex.argtypes = [vp, _dp, C.c_int, _dp, C.c_int, _dp, C.c_int]
        L.mcsim_cuda_group_set_scene_ex.argtypes = [vp, _dp, C.c_int, _dp, C.c_int, _dp, C.c_int]
        L.mcsim_cuda_mc_run.argtypes = [vp, C.POINTER(McParams), C.POINTER(McResult)]
        L.mcsim_cuda_mc_prepare.argtypes = [vp, C.POINTER(McParams)]
        L.mcsim_cuda_mc_launch.argtypes = [vp, C.c_uint64, C.c_uint64]
        L.mcsim_cuda_mc_collect.argtypes = [vp, C.POINTER(McResult)]
        L.mcsim_cuda_mc_counters.argtypes = [vp]
        L.mcsim_cuda_mc_counters.restype = vp
        L.mcsim_host_thread.argtypes = [C.POINTER(McParams), _dp, C.c_size_t, _dp, _dp, _dp, _dp, _dp, _dp]
        L.mcsim_host_normals.argtypes = [C.c_uint64, C.c_uint32, C.c_int, C.c_int, _dp]
        L.mcsim_step_normals.argtypes = [C.c_uint64, C.c_uint64, C.c_int, _dp]
        L.mcsim_step_normals.restype = C.c_int
        L.mcsim_cuda_step_normals.argtypes = [vp, C.c_uint64, C.c_uint64, C.c_uint64, C.c_int, _fp]
        L.mcsim_cuda_gmm_truncate.argtypes = [vp, C.POINTER(GmmTruncateParams), C.POINTER(GmmTruncateResult)]
        L.mcsim_cuda_gmm_run.argtypes = [vp, C.POINTER(McParams), C.POINTER(GmmRunParams), C.POINTER(GmmRunResult)]
        L.mcsim_cuda_gmm_run_batch.argtypes = [vp, C.POINTER(McParams), C.POINTER(GmmBatchParams), C.POINTER(GmmBatchResult)]
        L.mcsim_cuda_linear_run.argtypes = [vp, C.POINTER(McParams), C.POINTER(LinearParams), C.POINTER(GmmRunResult)]
        L.mcsim_truncate_linear.argtypes = [_dp, _dp, C.c_int, _dp, _dp, _dp, _dp, _dp]
        L.mcsim_cuda_score_plans.argtypes = [vp, C.POINTER(ScoreParams), C.POINTER(ScoreResult)]
        L.mcsim_cuda_score_prepare.argtypes = [vp, C.POINTER(ScoreParams)]
        L.mcsim_cuda_score_launch.argtypes = [vp]
        L.mcsim_cuda_score_collect.argtypes = [vp, C.POINTER(ScoreResult)]
        L.mcsim_cuda_em_fit.argtypes = [vp, C.POINTER(EmParams), C.POINTER(EmResult)]
        L.mcsim_cuda_em_prepare.argtypes = [vp, C.POINTER(EmParams)]
        L.mcsim_cuda_em_estep.argtypes = [vp]
        L.mcsim_cuda_em_mstep.argtypes = [vp]
        L.mcsim_cuda_em_collect.argtypes = [vp, C.POINTER(EmResult)]
        L.mcsim_cuda_em_stats.argtypes = [vp, C.POINTER(C.c_uint64)]
        L.mcsim_cuda_em_stats.restype = vp
        L.mcsim_cuda_em_set_moments.argtypes = [vp, C.c_int]
        L.mcsim_cuda_em_moments_in_use.argtypes = [vp, C.c_int, C.POINTER(C.c_int)]
        L.mcsim_cuda_em_fit_static.argtypes = [C.POINTER(EmStaticParams)]
        L.mcsim_cuda_em_fit_static.restype = C.c_int
        L.mcsim_cuda_io_bytes.argtypes = [vp, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
        L.mcsim_cuda_io_bytes.restype = None
        L.mcsim_cuda_device_count.restype = C.c_int
        L.mcsim_cuda_ctx_device.argtypes = [vp]
        L.mcsim_cuda_group_create.argtypes = [C.POINTER(C.c_int), C.c_int, C.POINTER(vp)]
        L.mcsim_cuda_group_destroy.argtypes = [vp]
        L.mcsim_cuda_group_destroy.restype = None
        L.mcsim_cuda_group_size.argtypes = [vp]
        L.mcsim_cuda_group_ctx.argtypes = [vp, C.c_int]
        L.mcsim_cuda_group_ctx.restype = vp
        L.mcsim_cuda_group_set_scene.argtypes = [vp, _dp, C.c_int, _dp, C.c_int]
        L.mcsim_cuda_group_mc_run.argtypes = [vp, C.POINTER(McParams), C.POINTER(McResult)]
        L.mcsim_cuda_group_em_fit.argtypes = [vp, C.POINTER(EmParams), C.POINTER(EmResult)]
        L.mcsim_cuda_group_set_em_exchange.argtypes = [vp, C.c_int]
        L.mcsim_cuda_group_em_exchange.argtypes = [vp]
        L.mcsim_cuda_group_score_plans.argtypes = [vp, C.POINTER(ScoreParams), C.POINTER(ScoreResult)]
        L.mcsim_cuda_group_gmm_run_batch.argtypes = [vp, C.POINTER(McParams), C.POINTER(GmmBatchParams), C.POINTER(GmmBatchResult)]
        L.mcsim_cuda_group_stats.argtypes = [vp, _fp, C.POINTER(C.c_uint64)]
        L.mcsim_cuda_group_stats.restype = None
        _lib = L
    return _lib


class McsimCudaError(RuntimeError):
    pass


def check(rc):
    if rc != 0:
        raise McsimCudaError("mcsim_cuda error %d: %s" % (rc, lib().mcsim_cuda_last_error().decode()))


def dptr(a):
    """double* of a C-contiguous float64 array (or NULL)."""
    if a is None:
        return _dp()
    assert a.dtype == np.float64 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(_dp)


def make_mc_params(plan, n_rollouts=0, mode=MODE_PHILOX, estimator=EST_REFERENCE, early_exit=True,
                   seed=0x5AFEC0DE, rollout_offset=0, applied=None, z0=None, near_eps=1e-4, precise=False, exhaustive=False, no_compaction=False,
                   gmm_filter=0):
    """plan: dict(trajectory[3,T], odometry[3,T-1], alphas[4], Q, landmarks[2,L], cov0[3,3]).
    Returns (McParams, keepalive list of the arrays its pointers refer to)."""
    tr = np.ascontiguousarray(plan["trajectory"], dtype=np.float64)
    od = np.ascontiguousarray(plan["odometry"], dtype=np.float64)
    lm = np.ascontiguousarray(plan["landmarks"], dtype=np.float64)
    keep = [tr, od, lm]
    p = McParams()
    p.T = tr.shape[1]
    p.trajectory, p.odometry, p.landmarks = dptr(tr), dptr(od), dptr(lm)
    p.alphas = (C.c_double * 4)(*[float(v) for v in plan["alphas"]])
    p.Q = float(plan["Q"])
    p.n_landmarks = lm.shape[1]
    p.cov0 = (C.c_double * 9)(*[float(v) for v in np.asarray(plan["cov0"], dtype=np.float64).reshape(9)])
    p.n_rollouts, p.rollout_offset = int(n_rollouts), int(rollout_offset)
    p.mode, p.estimator, p.early_exit, p.seed = mode, estimator, int(bool(early_exit)), int(seed)
    if applied is not None:
        ap = np.ascontiguousarray(applied, dtype=np.float64)
        keep.append(ap)
        p.applied = dptr(ap)
    if z0 is not None:
        zz = np.ascontiguousarray(z0, dtype=np.float64)
        keep.append(zz)
        p.z0 = dptr(zz)
    p.near_eps = float(near_eps)
    p.precise = int(bool(precise))
    p.exhaustive = int(bool(exhaustive))
    p.no_compaction = int(bool(no_compaction))
    p.gmm_filter = int(gmm_filter)
    return p, keep


def truncate_linear(mean, cov, A, b):
    """Analytic truncation of N(mean, cov) against the half-spaces A[k] . x <= b[k].
    Returns (mean', cov', p_violation[k])."""
    m = np.ascontiguousarray(mean, dtype=np.float64).reshape(3)
    cv = np.ascontiguousarray(cov, dtype=np.float64).reshape(9)
    Aa = np.ascontiguousarray(A, dtype=np.float64).reshape(-1, 3)
    bb = np.ascontiguousarray(b, dtype=np.float64).reshape(-1)
    mo, co, pv = np.zeros(3), np.zeros(9), np.zeros(Aa.shape[0])
    check(lib().mcsim_truncate_linear(dptr(m), dptr(cv), Aa.shape[0], dptr(Aa), dptr(bb), dptr(mo), dptr(co), dptr(pv)))
    return mo, co.reshape(3, 3), pv


def host_thread(plan, normals):
    """The reference's shared real-robot thread (FP64, host): returns dict of arrays."""
    p, keep = make_mc_params(plan)
    T, L = p.T, p.n_landmarks
    nz = np.ascontiguousarray(normals, dtype=np.float64)
    out = dict(applied_noisy=np.zeros((3, T - 1)), applied=np.zeros((3, T - 1)), real_path=np.zeros((3, T)),
               mu_path=np.zeros((3, T)), cov_path=np.zeros((T, 9)), obs=np.zeros((T - 1, L)))
    rc = lib().mcsim_host_thread(C.byref(p), dptr(nz), nz.size, dptr(out["applied_noisy"]), dptr(out["applied"]),
                                 dptr(out["real_path"]), dptr(out["mu_path"]), dptr(out["cov_path"]), dptr(out["obs"]))
    if rc != 0:
        raise McsimCudaError("mcsim_host_thread error %d" % rc)
    return out


def em_fit_static(device, x_ptr, n, K, n_fits, iters, hefts_ptr, means_ptr, fcovs_ptr, avg_log_p_ptr=None, iters_run_ptr=None,
                  stream=None, var_floor=1e-10):
    """mcsim_cuda_em_fit_static: the EM fit on the caller's DEVICE buffers (integer addresses), no allocation, asynchronous on
    `stream`.  hefts / means / fcovs are updated in place."""
    p = EmStaticParams()
    p.device, p.K, p.n_fits, p.iters, p.n, p.var_floor = int(device), int(K), int(n_fits), int(iters), int(n), float(var_floor)
    p.x, p.hefts, p.means, p.fcovs = x_ptr, hefts_ptr, means_ptr, fcovs_ptr
    p.avg_log_p, p.iters_run, p.stream = avg_log_p_ptr, iters_run_ptr, stream
    check(lib().mcsim_cuda_em_fit_static(C.byref(p)))


class Context:
    """One CUDA context handle (device + stream) of libmcsim_cuda.so."""

    def __init__(self, device=0):
        self._h = C.c_void_p()
        check(lib().mcsim_cuda_create(int(device), C.byref(self._h)))
        self.device = device

    def close(self):
        if self._h:
            lib().mcsim_cuda_destroy(self._h)
            self._h = C.c_void_p()

    # set_scene / mc_run / score_plans / em_fit / gmm_run_batch exist per context and per device group (class Group)
    def _entry(self, name):
        return getattr(lib(), "mcsim_cuda_" + name)

    @property
    def _target(self):
        return self._h

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def stream(self):
        return lib().mcsim_cuda_stream(self._h)

    @property
    def torch_stream(self):
        """The context's CUDA stream as a torch stream (collectives are issued on it, behind the kernels)."""
        import torch
        return torch.cuda.ExternalStream(self.stream, device=self.device)

    @property
    def launch_count(self):
        return int(lib().mcsim_cuda_launch_count(self._h))

    def measure_peaks(self, ms=200.0):
        """(FP32 TFLOP/s, SFU 1e12 op/s) measured with register-resident chains."""
        a, b = C.c_double(), C.c_double()
        check(lib().mcsim_cuda_measure_peaks(self._h, float(ms), C.byref(a), C.byref(b)))
        return a.value, b.value

    def measure_peak(self, which, ms=200.0):
        """which: 0 FFMA (TFLOP/s), 1 MUFU (1e12 op/s), 2 packed FFMA2 (TFLOP/s)."""
        a = C.c_double()
        check(lib().mcsim_cuda_measure_peak(self._h, int(which), float(ms), C.byref(a)))
        return a.value

    def set_scene(self, boxes, hulls, poly_hulls=None):
        """boxes [n, 8] (world), hulls [m, 8] (robot frame boxes, may be empty with polygons), poly_hulls [k, 19]:
        convex-polygon robot hulls n, zlo, zhi, x0, y0, ... (counter-clockwise)."""
        b = np.ascontiguousarray(boxes, dtype=np.float64).reshape(-1, 8)
        h = np.ascontiguousarray(hulls, dtype=np.float64).reshape(-1, 8)
        po = np.zeros((0, 19)) if poly_hulls is None else np.ascontiguousarray(poly_hulls, dtype=np.float64).reshape(-1, 19)
        check(self._entry("set_scene_ex")(self._target, dptr(b), b.shape[0], dptr(h), h.shape[0], dptr(po), po.shape[0]))
        # live (box hull, box) pairs: z-intervals overlap (the same rule as mcsim_cuda_set_scene)
        self.n_pairs = int(sum(((hh[6] < b[:, 7]) & (b[:, 6] < hh[7])).sum() for hh in h))

    def mc_run(self, params, want_counts=False, want_hist=False, audit=False):
        """Runs params.n_rollouts rollouts; returns a dict."""
        n, T = int(params.n_rollouts), int(params.T)
        r = McResult()
        out = {}
        if want_hist:
            out["first_hit_hist"] = np.zeros(T, dtype=np.uint64)
            r.first_hit_hist = out["first_hit_hist"].ctypes.data_as(C.POINTER(C.c_uint64))
        if want_counts:
            out["counts"] = np.zeros(n, dtype=np.uint32)
            r.counts = out["counts"].ctypes.data_as(C.POINTER(C.c_uint32))
        if audit:
            out["poses"] = np.zeros((T, 3, n), dtype=np.float32)
            out["clearance"] = np.zeros((T, n), dtype=np.float32)
            r.poses = out["poses"].ctypes.data_as(_fp)
            r.clearance = out["clearance"].ctypes.data_as(_fp)
        check(self._entry("mc_run")(self._target, C.byref(params), C.byref(r)))
        out.update(hits=int(r.hits), near_contact=int(r.near_contact), executed_steps=int(r.executed_steps),
                   pair_tests=int(r.pair_tests), kernel_ms=float(r.kernel_ms),
                   P=(r.hits / n if n else 0.0))
        return out

    def step_normals(self, n, t=0, id0=0, seed=0x5AFEC0DE):
        """[n, 12] float32: the step draws of rollouts id0 .. id0 + n - 1 at step t, computed on the device by the
        packed kernel's own device function."""
        out = np.zeros((int(n), 12), dtype=np.float32)
        check(lib().mcsim_cuda_step_normals(self._h, int(seed), int(id0), int(n), int(t), out.ctypes.data_as(_fp)))
        return out

    def mc_prepare(self, params):
        check(lib().mcsim_cuda_mc_prepare(self._h, C.byref(params)))

    def mc_launch(self, n, offset=0):
        check(lib().mcsim_cuda_mc_launch(self._h, int(n), int(offset)))

    def mc_counters_tensor(self, T):
        """torch int64 view (no copy) of the device counter block uint64[4 + T], for the NCCL all-reduce
        of a multi-GPU run; the block is allocated at its maximum size when the context is created."""
        import torch

        class _Raw:
            pass
        raw = _Raw()
        raw.__cuda_array_interface__ = dict(shape=(4 + int(T),), typestr="<i8", version=2,
                                            data=(int(lib().mcsim_cuda_mc_counters(self._h)), False))
        return torch.as_tensor(raw, device="cuda:%d" % self.device)

    def mc_collect(self, T=0, want_hist=False):
        r = McResult()
        out = {}
        if want_hist:
            out["first_hit_hist"] = np.zeros(T, dtype=np.uint64)
            r.first_hit_hist = out["first_hit_hist"].ctypes.data_as(C.POINTER(C.c_uint64))
        check(lib().mcsim_cuda_mc_collect(self._h, C.byref(r)))
        out.update(hits=int(r.hits), near_contact=int(r.near_contact), executed_steps=int(r.executed_steps),
                   pair_tests=int(r.pair_tests))
        return out

    def io_bytes(self):
        """(host->device, device->host) bytes copied by the last *_run call."""
        a, b = C.c_uint64(), C.c_uint64()
        lib().mcsim_cuda_io_bytes(self._h, C.byref(a), C.byref(b))
        return int(a.value), int(b.value)

    def gmm_run(self, plan, K, n_samples, mode=MODE_PHILOX, seed=0x5AFEC0DE, stream_id=0, draws=None, counts=None,
                record=False, exhaustive=False, kernel=GMM_KERNEL_AUTO, refit=GMM_REFIT_MOMENTS, em_iters=0, gmm_filter=0):
        """runGMMEstimation of one plan (MCSimulator.h:354-358 -> the GMM branch of EKF_GaussProp).
        Replay mode: draws = the reference's normal stream of the run, counts[T, K] = multinomial counts."""
        p, keep = make_mc_params(plan, mode=mode, seed=seed, rollout_offset=stream_id, exhaustive=exhaustive, gmm_filter=gmm_filter)
        T = int(p.T)
        g = GmmRunParams()
        g.K, g.n_samples, g.kernel, g.refit, g.em_iters = int(K), int(n_samples), int(kernel), int(refit), int(em_iters)
        if draws is not None:
            dr = np.ascontiguousarray(draws, dtype=np.float64)
            keep.append(dr)
            g.normals, g.n_normals = dptr(dr), dr.size
        if counts is not None:
            cn = np.ascontiguousarray(counts, dtype=np.int32)
            keep.append(cn)
            g.counts = cn.ctypes.data_as(C.POINTER(C.c_int))
        r = GmmRunResult()
        out = dict(probabilities=np.zeros(T))
        r.probabilities = dptr(out["probabilities"])
        if record:
            out["gmm_pre"] = np.zeros((T, int(K), 13))
            out["gmm_post"] = np.zeros((int(K), 13))
            r.gmm_pre, r.gmm_post = dptr(out["gmm_pre"]), dptr(out["gmm_post"])
        check(lib().mcsim_cuda_gmm_run(self._h, C.byref(p), C.byref(g), C.byref(r)))
        out.update(P=float(r.P), degenerate=int(r.degenerate), kernel_ms=float(r.kernel_ms), pair_tests=int(r.pair_tests))
        return out

    # ---- batched plan scoring (BASELINE config 4)
    def _score_params(self, model, trajectories, odometry, rollouts, plan_offset, estimator, early_exit, seed, margin):
        tr = np.ascontiguousarray(trajectories, dtype=np.float64)
        od = np.ascontiguousarray(odometry, dtype=np.float64)
        lm = np.ascontiguousarray(model["landmarks"], dtype=np.float64)
        assert tr.ndim == 3 and tr.shape[1] == 3 and od.shape == (tr.shape[0], 3, tr.shape[2] - 1)
        p = ScoreParams()
        p.T, p.n_plans = tr.shape[2], tr.shape[0]
        p.trajectories, p.odometry, p.landmarks = dptr(tr), dptr(od), dptr(lm)
        p.alphas = (C.c_double * 4)(*[float(v) for v in model["alphas"]])
        p.Q, p.n_landmarks = float(model["Q"]), lm.shape[1]
        p.cov0 = (C.c_double * 9)(*[float(v) for v in np.asarray(model["cov0"], dtype=np.float64).reshape(9)])
        p.rollouts, p.plan_offset = int(rollouts), int(plan_offset)
        p.estimator, p.early_exit, p.seed, p.margin = int(estimator), int(bool(early_exit)), int(seed), float(margin)
        return p, [tr, od, lm]

    def _score_result(self, n_plans):
        out = dict(hits=np.zeros(n_plans, dtype=np.uint64), P=np.zeros(n_plans))
        r = ScoreResult()
        r.hits, r.P = out["hits"].ctypes.data_as(C.POINTER(C.c_uint64)), dptr(out["P"])
        return r, out

    def score_plans(self, model, trajectories, odometry, rollouts, plan_offset=0, estimator=EST_REFERENCE,
                    early_exit=False, seed=0x5AFEC0DE, margin=3.0):
        """model: dict(alphas, Q, landmarks[2, 8], cov0); trajectories [n, 3, T], odometry [n, 3, T-1]."""
        p, keep = self._score_params(model, trajectories, odometry, rollouts, plan_offset, estimator, early_exit, seed, margin)
        r, out = self._score_result(int(p.n_plans))
        check(self._entry("score_plans")(self._target, C.byref(p), C.byref(r)))
        out.update(executed_steps=int(r.executed_steps), pair_tests=int(r.pair_tests), kernel_ms=float(r.kernel_ms))
        return out

    def score_prepare(self, model, trajectories, odometry, rollouts, **kw):
        p, keep = self._score_params(model, trajectories, odometry, rollouts, kw.get("plan_offset", 0),
                                     kw.get("estimator", EST_REFERENCE), kw.get("early_exit", False),
                                     kw.get("seed", 0x5AFEC0DE), kw.get("margin", 3.0))
        self._score_n = int(p.n_plans)
        check(lib().mcsim_cuda_score_prepare(self._h, C.byref(p)))

    def score_launch(self):
        check(lib().mcsim_cuda_score_launch(self._h))

    def score_collect(self):
        r, out = self._score_result(self._score_n)
        check(lib().mcsim_cuda_score_collect(self._h, C.byref(r)))
        out.update(executed_steps=int(r.executed_steps), pair_tests=int(r.pair_tests))
        return out

    # ---- EM fit (BASELINE config 5; replaces arma::gmm_full::learn(keep_existing), gmm_full_meat.hpp:764-890)
    def _em_params(self, x, hefts, means, fcovs, iters, var_floor, n_total):
        h = np.ascontiguousarray(hefts, dtype=np.float64)
        F, K = (1, h.shape[0]) if h.ndim == 1 else h.shape
        m = np.ascontiguousarray(means, dtype=np.float64).reshape(F, K, 3)
        cv = np.ascontiguousarray(fcovs, dtype=np.float64).reshape(F, K, 9)
        p = EmParams()
        keep = [h, m, cv]
        if x is not None:
            xx = np.ascontiguousarray(x, dtype=np.float32).reshape(F, 3, -1)
            keep.append(xx)
            p.x, p.n = xx.ctypes.data_as(_fp), xx.shape[2]
        else:
            p.n = int(self._em_n)
        self._em_n = int(p.n)
        p.K, p.n_fits, p.n_total = K, F, int(n_total)
        p.hefts, p.means, p.fcovs = dptr(h), dptr(m), dptr(cv)
        p.iters, p.var_floor = int(iters), float(var_floor)
        return p, keep, (F, K)

    def _em_result(self, F, K, iters):
        out = dict(hefts=np.zeros((F, K)), means=np.zeros((F, K, 3)), fcovs=np.zeros((F, K, 3, 3)),
                   avg_log_p=np.full((F, max(iters, 1)), np.nan), iters_run=np.zeros(F, dtype=np.int32))
        r = EmResult()
        r.hefts, r.means, r.fcovs, r.avg_log_p = dptr(out["hefts"]), dptr(out["means"]), dptr(out["fcovs"]), dptr(out["avg_log_p"])
        r.iters_run = out["iters_run"].ctypes.data_as(C.POINTER(C.c_int))
        return r, out

    def em_fit(self, x, hefts, means, fcovs, iters, var_floor=1e-10):
        """x: [F, 3, n] (or [3, n]) FP32 samples, component-major per fit, or None to reuse the resident ones;
        hefts [F, K], means [F, K, 3], fcovs [F, K, 3, 3]: initial parameters.  Returns a dict."""
        p, keep, (F, K) = self._em_params(x, hefts, means, fcovs, iters, var_floor, 0)
        r, out = self._em_result(F, K, iters)
        check(self._entry("em_fit")(self._target, C.byref(p), C.byref(r)))
        out["kernel_ms"] = float(r.kernel_ms)
        return out

    def em_prepare(self, x, hefts, means, fcovs, iters, var_floor=1e-10, n_total=0):
        p, keep, self._em_shape = self._em_params(x, hefts, means, fcovs, iters, var_floor, n_total)
        self._em_iters = int(iters)
        check(lib().mcsim_cuda_em_prepare(self._h, C.byref(p)))

    def em_estep(self):
        check(lib().mcsim_cuda_em_estep(self._h))

    def em_mstep(self):
        check(lib().mcsim_cuda_em_mstep(self._h))

    def em_set_moments(self, form):
        """Reference point of the EM statistics for the fits prepared from now on: EM_MOMENTS_AUTO (default) / _CENTRED / _SHIFTED."""
        check(lib().mcsim_cuda_em_set_moments(self._h, int(form)))

    def em_moments_in_use(self, fit=0):
        """EM_MOMENTS_CENTRED or EM_MOMENTS_SHIFTED: what the next E-step of `fit` uses (decided by the last M-step)."""
        form = C.c_int()
        check(lib().mcsim_cuda_em_moments_in_use(self._h, int(fit), C.byref(form)))
        return int(form.value)

    def em_collect(self):
        F, K = self._em_shape
        r, out = self._em_result(F, K, self._em_iters)
        check(lib().mcsim_cuda_em_collect(self._h, C.byref(r)))
        return out

    def em_stats_tensor(self):
        """torch float64 view (no copy) of the device statistics block [F * (K * 10 + 1)] for the all-reduce."""
        import torch
        n = C.c_uint64()
        ptr = lib().mcsim_cuda_em_stats(self._h, C.byref(n))

        class _Raw:
            pass
        raw = _Raw()
        raw.__cuda_array_interface__ = dict(shape=(int(n.value),), typestr="<f8", version=2, data=(int(ptr), False))
        return torch.as_tensor(raw, device="cuda:%d" % self.device)

    def gmm_run_batch(self, plan, K, n_samples, n_traj, trajectories=None, odometry=None, seed=0x5AFEC0DE, stream0=0,
                      exhaustive=False, kernel=GMM_KERNEL_AUTO, gmm_filter=0):
        """n_traj independent runGMMEstimation runs in the same launches (Philox mode); member b ==
        gmm_run(..., stream_id=stream0 + b) on plan b (trajectories [n_traj, 3, T]) or on `plan`."""
        p, keep = make_mc_params(plan, mode=MODE_PHILOX, seed=seed, exhaustive=exhaustive, gmm_filter=gmm_filter)
        T = int(p.T)
        g = GmmBatchParams()
        g.K, g.n_samples, g.n_traj, g.stream0, g.kernel = int(K), int(n_samples), int(n_traj), int(stream0), int(kernel)
        if trajectories is not None:
            tr = np.ascontiguousarray(trajectories, dtype=np.float64)
            od = np.ascontiguousarray(odometry, dtype=np.float64)
            assert tr.shape == (n_traj, 3, T) and od.shape == (n_traj, 3, T - 1)
            keep += [tr, od]
            g.trajectories, g.odometry = dptr(tr), dptr(od)
        out = dict(P=np.zeros(n_traj), probabilities=np.zeros((n_traj, T)))
        r = GmmBatchResult()
        r.P, r.probabilities = dptr(out["P"]), dptr(out["probabilities"])
        check(self._entry("gmm_run_batch")(self._target, C.byref(p), C.byref(g), C.byref(r)))
        out.update(degenerate=int(r.degenerate), kernel_ms=float(r.kernel_ms), pair_tests=int(r.pair_tests))
        return out

    def linear_run(self, plan, K, A, b, seed=0x5AFEC0DE, stream_id=0, normals=None, record=False, gmm_filter=0):
        """The GMM belief propagation with the analytic linear-constraint truncation (mcsim_cuda_linear_run):
        free region = all x with A[k] . x <= b[k]."""
        p, keep = make_mc_params(plan, mode=MODE_PHILOX, seed=seed, rollout_offset=stream_id, gmm_filter=gmm_filter)
        T = int(p.T)
        Aa = np.ascontiguousarray(A, dtype=np.float64).reshape(-1, 3)
        bb = np.ascontiguousarray(b, dtype=np.float64).reshape(-1)
        lp = LinearParams()
        lp.K, lp.n_constraints, lp.A, lp.b = int(K), Aa.shape[0], dptr(Aa), dptr(bb)
        if normals is not None:
            nz = np.ascontiguousarray(normals, dtype=np.float64)
            keep.append(nz)
            lp.normals, lp.n_normals = dptr(nz), nz.size
        r = GmmRunResult()
        out = dict(probabilities=np.zeros(T), gmm_post=np.zeros((int(K), 13)))
        r.probabilities, r.gmm_post = dptr(out["probabilities"]), dptr(out["gmm_post"])
        if record:
            out["gmm_pre"] = np.zeros((T, int(K), 13))
            r.gmm_pre = dptr(out["gmm_pre"])
        check(lib().mcsim_cuda_linear_run(self._h, C.byref(p), C.byref(lp), C.byref(r)))
        out.update(P=float(r.P), degenerate=int(r.degenerate), kernel_ms=float(r.kernel_ms))
        return out

    def gmm_truncate(self, weights, means, covs, n_samples, mode=MODE_PHILOX, counts=None, normals=None,
                     seed=0x5AFEC0DE, stream_id=0, refit=GMM_REFIT_MOMENTS, em_iters=0):
        """One truncateGMM (MCSimulator.h:570-642) of the mixture (weights[K], means[K,3], covs[K,3,3])."""
        w = np.ascontiguousarray(weights, dtype=np.float64)
        m = np.ascontiguousarray(means, dtype=np.float64).reshape(-1, 3)
        cv = np.ascontiguousarray(covs, dtype=np.float64).reshape(-1, 9)
        K = w.shape[0]
        p = GmmTruncateParams()
        p.K, p.weights, p.means, p.covs = K, dptr(w), dptr(m), dptr(cv)
        p.n_samples, p.mode, p.seed, p.stream_id = int(n_samples), int(mode), int(seed), int(stream_id)
        p.refit, p.em_iters = int(refit), int(em_iters)
        keep = [w, m, cv]
        if counts is not None:
            cn = np.ascontiguousarray(counts, dtype=np.int32)
            keep.append(cn)
            p.counts = cn.ctypes.data_as(C.POINTER(C.c_int))
        if normals is not None:
            nz = np.ascontiguousarray(normals, dtype=np.float64)
            keep.append(nz)
            p.normals = dptr(nz)
        r = GmmTruncateResult()
        check(lib().mcsim_cuda_gmm_truncate(self._h, C.byref(p), C.byref(r)))
        return dict(weights=np.array(r.weights[:K]), means=np.array(r.means[:3 * K]).reshape(K, 3),
                    covs=np.array(r.covs[:9 * K]).reshape(K, 3, 3), n_coll=np.array(r.n_coll[:K], dtype=np.uint64),
                    n_free=np.array(r.n_free[:K], dtype=np.uint64), p_collision=float(r.p_collision),
                    degenerate=int(r.degenerate))


class Group(Context):
    """Several devices behind one handle (mcsim_cuda_group_*): one context per device, the NCCL communicators owned by the
    group.  set_scene / mc_run / score_plans / em_fit / gmm_run_batch run on all devices (rollout ids / plans / samples /
    batch members split, counters and EM statistics all-reduced over NCCL); every other method addresses device 0's
    context.  No torch, no torch.distributed: this is what a C++ host gets."""

    def __init__(self, devices):
        devices = [int(d) for d in devices]
        arr = (C.c_int * len(devices))(*devices)
        self._g = C.c_void_p()
        self._h = C.c_void_p()
        check(lib().mcsim_cuda_group_create(arr, len(devices), C.byref(self._g)))
        self._h = C.c_void_p(lib().mcsim_cuda_group_ctx(self._g, 0))
        self.devices = devices
        self.device = devices[0]

    def close(self):
        if getattr(self, "_g", None):
            lib().mcsim_cuda_group_destroy(self._g)
            self._g = C.c_void_p()
            self._h = C.c_void_p()

    def _entry(self, name):
        return getattr(lib(), "mcsim_cuda_group_" + name)

    def set_em_exchange(self, mode):
        """EM_EXCHANGE_AUTO | EM_EXCHANGE_NCCL | EM_EXCHANGE_PEER: how em_fit moves the statistics every iteration."""
        check(lib().mcsim_cuda_group_set_em_exchange(self._g, int(mode)))

    @property
    def em_exchange(self):
        return int(lib().mcsim_cuda_group_em_exchange(self._g))

    @property
    def _target(self):
        return self._g

    def collective_stats(self):
        """(device time of the last call's collective in ms, collectives issued so far)"""
        ms, n = C.c_float(), C.c_uint64()
        lib().mcsim_cuda_group_stats(self._g, C.byref(ms), C.byref(n))
        return float(ms.value), int(n.value)
