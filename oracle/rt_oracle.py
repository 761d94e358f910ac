"""TEST INFRASTRUCTURE ONLY -- ctypes wrapper around oracle/librt_oracle.so (rt_oracle.c).

Used by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
as the checker / timed CPU baseline.  Never imported by raytracing_b200/.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "librt_oracle.so")
_lib = None


class Spline(C.Structure):
    _fields_ = [("tu", C.c_void_p), ("tv", C.c_void_p), ("c", C.c_void_p),
                ("nu", C.c_int), ("nv", C.c_int), ("ku", C.c_int), ("kv", C.c_int)]


class Params(C.Structure):
    _fields_ = [("method", C.c_int), ("is_interface", C.c_int), ("max_size", C.c_int64),
                ("gamma", C.c_double), ("step", C.c_double),
                ("limx_i", C.c_double), ("limx_s", C.c_double), ("limy_i", C.c_double), ("limy_s", C.c_double)]


def build(force=False):
    src = os.path.join(_HERE, "rt_oracle.c")
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(["make", "-s", "-C", _HERE, "-B", "librt_oracle.so"])
    return _LIB_PATH


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_LIB_PATH)
        _lib.rto_bispev.restype = C.c_double
        _lib.rto_bispev.argtypes = [C.POINTER(Spline), C.c_double, C.c_double]
        _lib.rto_bispev_many.restype = None
        _lib.rto_bispev_many.argtypes = [C.POINTER(Spline), C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]
        _lib.rto_trace.restype = C.c_int64
        _lib.rto_trace.argtypes = [C.POINTER(Params)] + [C.POINTER(Spline)] * 3 + [C.c_void_p] * 3 + \
            [C.c_int64] + [C.c_void_p] * 4 + [C.c_int]
        _lib.rto_max_threads.restype = C.c_int
        _lib.rto_set_trig_noise.restype = None
        _lib.rto_set_trig_noise.argtypes = [C.c_int]
        _lib.rto_set_curv_tol_scale.restype = None
        _lib.rto_set_curv_tol_scale.argtypes = [C.c_double]
    return _lib


class HostSpline:
    """Holds (tu, tv, c, ku, kv) as contiguous float64 arrays + the C struct view."""

    def __init__(self, tu, tv, c, ku, kv):
        self.tu = np.ascontiguousarray(tu, dtype=np.float64)
        self.tv = np.ascontiguousarray(tv, dtype=np.float64)
        self.c = np.ascontiguousarray(c, dtype=np.float64).ravel()
        self.ku, self.kv = int(ku), int(kv)
        assert self.c.size == (self.tu.size - ku - 1) * (self.tv.size - kv - 1)
        self.struct = Spline(self.tu.ctypes.data, self.tv.ctypes.data, self.c.ctypes.data,
                             self.tu.size, self.tv.size, self.ku, self.kv)

    @classmethod
    def from_scipy(cls, spl):
        tu, tv, c = spl.tck
        ku, kv = spl.degrees
        return cls(tu, tv, c, ku, kv)

    def __call__(self, u, v):
        u = np.ascontiguousarray(np.atleast_1d(u), dtype=np.float64)
        v = np.ascontiguousarray(np.atleast_1d(v), dtype=np.float64)
        out = np.empty_like(u)
        lib().rto_bispev_many(C.byref(self.struct), u.ctypes.data, v.ctypes.data, u.size, out.ctypes.data)
        return out


def _as_host(s):
    return s if isinstance(s, HostSpline) else HostSpline.from_scipy(s)


def n_gradient(pos, grd, z):
    """RT_bench.py:141-156 on arrays of positions (N,2) -> n (N), grad (N,2)."""
    pos = np.atleast_2d(np.asarray(pos, dtype=np.float64))
    z, g0, g1 = _as_host(z), _as_host(grd[0]), _as_host(grd[1])
    x, y = pos[:, 0], pos[:, 1]
    return z(y, x), np.stack([g1(y, x), g0(y, x)], axis=1)


def trace(method, z, grd, step, max_size, gamma, box, is_interface, theta0, x0, y0,
          full=True, nthreads=1):
    """trazar (RT_bench.py:766-948) with an explicit fan.

    Returns dict(s_ray (M,6,R) or None, d_ray (3,R), errors (R), final (6,R), ray_steps int).
    """
    theta0 = np.ascontiguousarray(theta0, dtype=np.float64)
    R = theta0.size
    x0 = np.ascontiguousarray(np.broadcast_to(np.asarray(x0, dtype=np.float64), (R,)))
    y0 = np.ascontiguousarray(np.broadcast_to(np.asarray(y0, dtype=np.float64), (R,)))
    z, g0, g1 = _as_host(z), _as_host(grd[0]), _as_host(grd[1])
    p = Params(int(method), int(bool(is_interface)), int(max_size), float(gamma), float(step),
               float(box[0]), float(box[1]), float(box[2]), float(box[3]))
    s_ray = np.zeros((int(max_size), 6, R)) if full else None
    d_ray = np.zeros((3, R))
    errors = np.zeros(R)
    final = np.zeros((6, R))
    total = lib().rto_trace(C.byref(p), C.byref(z.struct), C.byref(g0.struct), C.byref(g1.struct),
                            theta0.ctypes.data, x0.ctypes.data, y0.ctypes.data, R,
                            s_ray.ctypes.data if full else None, d_ray.ctypes.data, errors.ctypes.data,
                            final.ctypes.data, int(nthreads))
    return dict(s_ray=s_ray, d_ray=d_ray, errors=errors, final=final, ray_steps=int(total))


def max_threads():
    return int(lib().rto_max_threads())


def set_trig_noise(seed):
    """0 = off (faithful oracle); >0 = every sin/cos result moved by +-1 ulp with probability 1/2
    (sensitivity probe of the reference algorithm, see rt_oracle.c); >= 1000 = the arguments of sin/cos as well."""
    lib().rto_set_trig_noise(int(seed))


def set_curv_tol_scale(scale):
    """1.0 = faithful oracle; 1 +- 1e-14 = probe of curvature_t's GOLD_TOL switch (RT:349): shows the rays whose
    curvature sits on the threshold, where the last bits of the curvature pick the branch (see rt_oracle.c)."""
    lib().rto_set_curv_tol_scale(float(scale))
