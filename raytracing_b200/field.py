"""Host pre-processing kept exactly as the reference does it (numpy + scipy), so the spline
coefficients handed to the GPU are the same bytes the reference evaluates.

genZ RT_bench.py:412-433, interpolacion RT_bench.py:435-464 (only z and grd are built: the
Hessian splines are dead code in the reference, RT_bench.py:462).
"""
import numpy as np
from scipy import interpolate

from . import constants as K


def genZ(xi, xs, yi, ys, f, delta=None):
    """RT_bench.py:412-433 (the module-global `f` is an explicit argument here)."""
    delta = K.DELTA if delta is None else delta
    qx = int((xs - xi + 6) / delta + 1)
    qy = int((ys - yi + 6) / delta + 1)
    x, y = np.linspace(xi - 3, xs + 3, qx), np.linspace(yi - 3, ys + 3, qy)
    X, Y = np.meshgrid(x, y)
    return x, y, X, Y, f(X, Y)


def interpolacion(x, y, Z, delta=None):
    """RT_bench.py:435-464: returns (z, grd).  Note np.gradient is fed DELTA, not the true grid
    spacing (a reference quirk that is part of the contract)."""
    delta = K.DELTA if delta is None else delta
    GradX, GradY = np.gradient(Z, delta, edge_order=2)
    z = interpolate.RectBivariateSpline(y, x, Z, kx=1, ky=1)
    funcgradX = interpolate.RectBivariateSpline(y, x, GradX, kx=3, ky=3)
    funcgradY = interpolate.RectBivariateSpline(y, x, GradY, kx=3, ky=3)
    return z, (funcgradX, funcgradY)


def spline_arrays(spl):
    """(tu, tv, c, ku, kv) of a scipy RectBivariateSpline (or pass-through of such a tuple)."""
    if isinstance(spl, (tuple, list)) and len(spl) == 5:
        tu, tv, c, ku, kv = spl
    else:
        tu, tv, c = spl.tck
        ku, kv = spl.degrees
    return (np.ascontiguousarray(tu, dtype=np.float64), np.ascontiguousarray(tv, dtype=np.float64),
            np.ascontiguousarray(c, dtype=np.float64).ravel(), int(ku), int(kv))


_FIELD_CACHE = {}


def build_field(user_choice):
    """Grid + splines of a scenario the way `__main__` builds them (RT_bench.py:1584-1588).

    Returns (z, grd) scipy objects; cached per scenario function (3 and 4 share one)."""
    uc = K.scenario_id(user_choice)
    f = K.SCENARIO_FUNCS[uc]
    lim = K.constants(uc)[5:9]
    key = (f.__name__, lim)
    if key not in _FIELD_CACHE:
        x, y, _X, _Y, Z = genZ(lim[0], lim[1], lim[2], lim[3], f)
        _FIELD_CACHE[key] = interpolacion(x, y, Z)
    return _FIELD_CACHE[key]
