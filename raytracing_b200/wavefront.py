"""Wavefront extraction: the arithmetic part of the reference's wavefront plot (graficar, RT_bench.py:987-1024)
as data, without any plotting.

  wavefront_points : per ray and travel time, the PCHIP-interpolated position and group angle (GPU, rtb_wavefront)
  wavefront        : per travel time, the points sorted by y, the PCHIP x(y) through them, its tangent / normal
                     angles and |ray angle - normal angle| (RT_bench.py:1009-1024; at most ray_count points per time,
                     so this cross-ray step runs on the host in numpy)
"""
import ctypes as C

import numpy as np

from . import _cabi
from .engine import _is_torch

DEFAULT_TIMES = np.arange(0.05, 0.6, 0.05)  # RT_bench.py:987


def wavefront_points(s_ray, d_ray, times=DEFAULT_TIMES, engine=None):
    """(x, y, angle, valid) arrays of shape (nt, R): PchipInterpolator(t_ray, .)(tau) of RT_bench.py:996-1001 for the
    rays with max(t_ray) >= tau.  s_ray is the full (max_size, 6, R) trajectory (numpy or torch CUDA)."""
    from .api import get_engine
    eng = engine or get_engine()
    dev = _is_torch(s_ray)
    rows, six, R = s_ray.shape
    assert six == 6
    if dev:
        import torch
        eng.use_torch_stream()
        times_a = torch.as_tensor(np.ascontiguousarray(times, dtype=np.float64)).to(s_ray.device)
        out = torch.empty(4, times_a.numel(), R, dtype=torch.float64, device=s_ray.device)
        ptr = lambda a: C.c_void_p(a.data_ptr())
        assert s_ray.is_contiguous() and d_ray.is_contiguous()
        nt = times_a.numel()
    else:
        s_ray = np.ascontiguousarray(s_ray, dtype=np.float64)
        d_ray = np.ascontiguousarray(d_ray, dtype=np.float64)
        times_a = np.ascontiguousarray(times, dtype=np.float64)
        out = np.empty((4, times_a.size, R))
        ptr = lambda a: C.c_void_p(a.ctypes.data)
        nt = times_a.size
    eng._check(eng._lib.rtb_wavefront(eng._ctx, _cabi.RTB_PTR_DEVICE if dev else _cabi.RTB_PTR_HOST, ptr(s_ray), ptr(d_ray),
                                      rows, R, ptr(times_a), nt, ptr(out)))
    return out[0], out[1], out[2], out[3] > 0.5


def pchip_knot_derivatives(x, y):
    """Derivatives of scipy's PchipInterpolator at its knots (PchipInterpolator._find_derivatives): what
    `pchip_interpolator.derivative()(knots)` returns in RT_bench.py:1018-1019."""
    x, y = np.asarray(x, dtype=np.float64), np.asarray(y, dtype=np.float64)
    hk, mk = np.diff(x), np.diff(y) / np.diff(x)
    if y.size == 2:
        return np.array([mk[0], mk[0]])
    smk = np.sign(mk)
    cond = (smk[1:] != smk[:-1]) | (mk[1:] == 0) | (mk[:-1] == 0)
    w1, w2 = 2 * hk[1:] + hk[:-1], hk[1:] + 2 * hk[:-1]
    with np.errstate(divide="ignore", invalid="ignore"):
        whmean = (w1 / mk[:-1] + w2 / mk[1:]) / (w1 + w2)
    dk = np.zeros_like(y)
    dk[1:-1] = np.where(cond, 0.0, 1.0 / np.where(cond, 1.0, whmean))

    def edge(h0, h1, m0, m1):
        d = ((2 * h0 + h1) * m0 - h0 * m1) / (h0 + h1)
        if np.sign(d) != np.sign(m0):
            return 0.0
        if np.sign(m0) != np.sign(m1) and abs(d) > 3 * abs(m0):
            return 3 * m0
        return d
    dk[0] = edge(hk[0], hk[1], mk[0], mk[1])
    dk[-1] = edge(hk[-1], hk[-2], mk[-1], mk[-2])
    return dk


def wavefront(s_ray, d_ray, times=DEFAULT_TIMES, engine=None):
    """Per travel time: dict(time, rays (indices of the valid rays, sorted by y), x, y, angle, tangent_angles,
    normal_angles, angle_diff) following RT_bench.py:1009-1024; times with fewer than two valid rays are skipped
    (RT_bench.py:1009)."""
    x, y, ang, valid = wavefront_points(s_ray, d_ray, times, engine)
    if _is_torch(x):
        x, y, ang, valid = (a.cpu().numpy() for a in (x, y, ang, valid))
    out = []
    for j, tau in enumerate(np.asarray(times)):
        idx = np.nonzero(valid[j])[0]
        if idx.size <= 1:
            continue
        order = np.argsort(y[j, idx])                   # RT:1013 sort by y
        k = idx[order]
        dy_dx = pchip_knot_derivatives(y[j, k], x[j, k])  # d x / d y of the wavefront through the points, RT:1017-1019
        tangent = np.pi / 2 - np.arctan(dy_dx)           # RT:1022
        normal = tangent - np.pi / 2                     # RT:1023
        # the reference pairs angle_vector (ray order) with normal_angles (sorted order), RT:1029: kept as is
        out.append(dict(time=float(tau), rays=k, x=x[j, k], y=y[j, k], angle=ang[j, idx], tangent_angles=tangent,
                        normal_angles=normal, angle_diff=np.abs(ang[j, idx] - normal)))
    return out
