"""Host-side mirror of the reference's plugin interface: `MCModule.SendCommand("<name> <args>")`.

In the reference, Python reaches the C++ plugin through OpenRAVE (`RaveCreateModule(env, 'MCModule')`,
MCSimulation.py:133-135) and every parameter crosses as a command string (MCSimulation.py:156-207).  Here the same
strings go to libmcsim.so (include/mcsim.h; C++ command dispatcher over the CUDA C ABI) through ctypes.  The
geometry the reference gets from `env.Load('data/pr2test2.env.xml')` (MCSimulation.py:123-124) is extracted from the
env XML by the C++ side when the module is created.
"""
import ctypes as C
import os

import numpy as np

from .inputs import py2_str

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "csrc", "libmcsim.so")
FIXTURE_DIR = os.path.join(os.path.dirname(_HERE), "data", "fixtures")
_dp = C.POINTER(C.c_double)
_lib = None


class McsimError(RuntimeError):
    pass


def lib():
    """The loaded libmcsim.so; raises if it has not been built (no fallback of any kind)."""
    global _lib
    if _lib is None:
        if not os.path.isfile(LIB_PATH):
            raise RuntimeError("%s is missing: build it with __graft_entry__.build() or `make -C %s`"
                               % (LIB_PATH, os.path.dirname(LIB_PATH)))
        L = C.CDLL(LIB_PATH)
        vp = C.c_void_p
        L.mcsim_create.argtypes = [C.c_char_p, C.c_char_p, C.c_int]
        L.mcsim_create.restype = vp
        L.mcsim_create_from_arrays.argtypes = [_dp, C.c_int, _dp, C.c_int, C.c_int]
        L.mcsim_create_from_arrays.restype = vp
        L.mcsim_send_command.argtypes = [vp, C.c_char_p, C.c_char_p, C.c_size_t, C.POINTER(C.c_size_t)]
        L.mcsim_last_output.argtypes = [vp, C.c_char_p, C.c_size_t]
        L.mcsim_last_output.restype = C.c_size_t
        L.mcsim_last_error.argtypes = [vp]
        L.mcsim_last_error.restype = C.c_char_p
        L.mcsim_get_boxes.argtypes = [vp, _dp, C.c_int]
        L.mcsim_get_hulls.argtypes = [vp, _dp, C.c_int]
        L.mcsim_get_poly_hulls.argtypes = [vp, _dp, C.c_int]
        L.mcsim_get_robot_pose.argtypes = [vp, _dp]
        L.mcsim_io_bytes.argtypes = [vp, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
        L.mcsim_io_bytes.restype = None
        L.mcsim_destroy.argtypes = [vp]
        L.mcsim_destroy.restype = None
        _lib = L
    return _lib


class MCModule:
    """One plugin module (mcsimplugin.cpp:7-46).  scene: path of an OpenRAVE env XML, or dict(boxes[n,8], hulls[m,8]),
    or None (send `setScene` / `setGeometry` later).  device = -1 gives a parse-only module (no GPU needed)."""

    def __init__(self, scene=None, device=0, fixture_dirs=(FIXTURE_DIR,)):
        L = lib()
        if isinstance(scene, (str, bytes, os.PathLike)):
            h = L.mcsim_create(os.fsencode(scene), ":".join(fixture_dirs).encode(), int(device))
        else:
            b = np.zeros((0, 8)) if scene is None else np.ascontiguousarray(scene["boxes"], dtype=np.float64).reshape(-1, 8)
            hu = np.zeros((0, 8)) if scene is None else np.ascontiguousarray(scene["hulls"], dtype=np.float64).reshape(-1, 8)
            h = L.mcsim_create_from_arrays(b.ctypes.data_as(_dp), b.shape[0], hu.ctypes.data_as(_dp), hu.shape[0], int(device))
        if not h:
            raise McsimError("mcsim_create failed: " + L.mcsim_last_error(None).decode())
        self._h = C.c_void_p(h)
        self._buf = C.create_string_buffer(1 << 16)

    def SendCommand(self, cmd):
        """Returns the handler's output text (OpenRAVE returns str on success).  A failing command raises."""
        L = lib()
        n = C.c_size_t()
        data = cmd.encode()
        rc = L.mcsim_send_command(self._h, data, self._buf, len(self._buf), C.byref(n))
        if rc == 0 and n.value >= len(self._buf):  # reply longer than the buffer: fetch it again, do NOT re-send
            self._buf = C.create_string_buffer(n.value + 1)  # (run* commands are not idempotent)
            L.mcsim_last_output(self._h, self._buf, len(self._buf))
        if rc != 0:
            raise McsimError("%s: %s" % (cmd.split(" ", 1)[0], L.mcsim_last_error(self._h).decode()))
        return self._buf.value.decode()

    def scene(self):
        L = lib()
        nb, nh = L.mcsim_get_boxes(self._h, None, 0), L.mcsim_get_hulls(self._h, None, 0)
        b, h, pose = np.zeros((nb, 8)), np.zeros((nh, 8)), np.zeros(3)
        L.mcsim_get_boxes(self._h, b.ctypes.data_as(_dp), nb)
        L.mcsim_get_hulls(self._h, h.ctypes.data_as(_dp), nh)
        has = L.mcsim_get_robot_pose(self._h, pose.ctypes.data_as(_dp))
        npo = L.mcsim_get_poly_hulls(self._h, None, 0)
        po = np.zeros((npo, 19))
        L.mcsim_get_poly_hulls(self._h, po.ctypes.data_as(_dp), npo)
        return dict(boxes=b, hulls=h, poly_hulls=po, robot_pose=pose if has else None)

    def io_bytes(self):
        a, b = C.c_uint64(), C.c_uint64()
        lib().mcsim_io_bytes(self._h, C.byref(a), C.byref(b))
        return int(a.value), int(b.value)

    def close(self):
        if self._h:
            lib().mcsim_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def list2string(values):
    """MCSimulation.py:81-85: Python-2 str() of every value followed by a space."""
    return "".join(py2_str(v) + " " for v in values)


def push_parameters(mod, plan, num_particles=None, num_gaussians=None, num_gmm_samples=None):
    """The parameter command sequence of MCSimulation.py:156-207, in its order."""
    lm = np.asarray(plan["landmarks"], dtype=np.float64)
    tr = np.asarray(plan["trajectory"], dtype=np.float64)
    od = np.asarray(plan["odometry"], dtype=np.float64)
    cov = np.asarray(plan["cov0"], dtype=np.float64).reshape(3, 3)
    mod.SendCommand("setAlphas " + list2string(plan["alphas"]))
    mod.SendCommand("setQ " + py2_str(plan["Q"]))
    mod.SendCommand("setNumLandmarks %d" % lm.shape[1])
    mod.SendCommand("setLandmarks " + list2string(lm[0]) + list2string(lm[1]))
    mod.SendCommand("setNumParticles %d" % int(plan.get("num_particles", 10000) if num_particles is None else num_particles))
    mod.SendCommand("setInitialCovariance " + list2string(cov[0]) + list2string(cov[1]) + list2string(cov[2]))
    mod.SendCommand("setPathLength %d" % tr.shape[1])
    mod.SendCommand("setTrajectory " + list2string(tr[0]) + list2string(tr[1]) + list2string(tr[2]))
    mod.SendCommand("setOdometry " + list2string(od[0]) + list2string(od[1]) + list2string(od[2]))
    if num_gaussians is not None:
        mod.SendCommand("setNumGaussians %d" % int(num_gaussians))
        mod.SendCommand("setNumGMMSamples %d" % int(num_particles if num_gmm_samples is None else num_gmm_samples))
